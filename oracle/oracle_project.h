// oracle_project.h — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Row-at-a-time restatement of the filter + projection step in front of the hash aggregation / join operators:
//   MAIN/operator/project/PageProcessor.java:139-330   filter -> selected positions -> projections over them
//   MAIN/type/DecimalOperators.java:84-330             decimal +, -, * (result precision / scale, rescale, "Decimal overflow")
//   SPI/type/UnscaledDecimal128Arithmetic.java:206-250 (rescale), :333-420 (add / subtract, sign of the result),
//                                                :452-560 (multiply: sign = leftNegative != rightNegative, also for a zero product)
//   MAIN/type/BigintOperators.java:47-83, IntegerOperators.java  Math.addExact / subtractExact / multiplyExact
//   MAIN/type/DoubleOperators.java                      IEEE arithmetic and comparisons
//   MAIN/sql/gen/AndCodeGenerator.java, OrCodeGenerator.java   three-valued AND / OR
//   MAIN/operator/scalar/CombineHashFunction.java:25-30  31 * previous + hash
// Magnitudes are carried in unsigned __int128 here (the product path uses 64-bit limbs): two different routes to one answer.
#pragma once
#include "oracle_groupby.h"

namespace orc {

struct PVal {
    bool null = false;
    int64_t v = 0;        // BIGINT / INTEGER / DATE / SHORT_DECIMAL / BOOLEAN value, DOUBLE bits
    Dec128 d{0, 0};       // LONG_DECIMAL
    const uint8_t* p = nullptr; int64_t len = 0;   // VARCHAR
};
struct PType { int32_t type = 0, precision = 0, scale = 0; };

static inline bool is_decimal(int32_t t) { return t == PG_SHORT_DECIMAL || t == PG_LONG_DECIMAL; }
static inline bool is_integral(int32_t t) { return t == PG_BIGINT || t == PG_INTEGER || t == PG_DATE; }
static inline u128 pow10_u128(int n) { u128 r = 1; while (n-- > 0) r *= 10; return r; }
static const u128 TEN38 = ((u128)0x4B3B4CA85A86C47AULL << 64) | 0x098A224000000000ULL;

struct Mag { u128 m; bool neg; bool overflow; };
static inline Mag mag_of(const PVal& x, int32_t type) {
    Mag r; r.overflow = false;
    if (type == PG_LONG_DECIMAL) { r.m = ((u128)dec_hi_mag(x.d) << 64) | x.d.lo; r.neg = dec_is_negative(x.d); }
    else { r.neg = x.v < 0; r.m = r.neg ? (u128)((uint64_t)0 - (uint64_t)x.v) : (u128)(uint64_t)x.v; }
    return r;
}
// 128 x 128 -> overflow when the product needs more than 127 bits (UnscaledDecimal128Arithmetic.multiply :480-560)
static inline Mag mag_mul(const Mag& a, const Mag& b) {
    Mag r; r.neg = a.neg != b.neg; r.overflow = a.overflow || b.overflow;
    uint64_t a0 = (uint64_t)a.m, a1 = (uint64_t)(a.m >> 64), b0 = (uint64_t)b.m, b1 = (uint64_t)(b.m >> 64);
    if (a1 != 0 && b1 != 0) r.overflow = true;
    u128 lo = (u128)a0 * b0;
    u128 mid1 = (u128)a0 * b1, mid2 = (u128)a1 * b0;
    if ((mid1 >> 64) != 0 || (mid2 >> 64) != 0) r.overflow = true;
    u128 mid = (u128)(uint64_t)mid1 + (u128)(uint64_t)mid2;
    if ((mid >> 64) != 0) r.overflow = true;
    u128 sum = lo + (mid << 64);
    if (sum < lo) r.overflow = true;
    r.m = sum;
    if ((r.m >> 127) != 0) r.overflow = true;
    return r;
}
// signed addition of sign-magnitude values, sign rules of addWithOverflow (:354-379): equal magnitudes of opposite sign give +0
static inline Mag mag_add(const Mag& a, const Mag& b) {
    Mag r; r.overflow = a.overflow || b.overflow;
    if (a.neg == b.neg) { r.m = a.m + b.m; r.neg = a.neg; if (r.m < a.m || (r.m >> 127) != 0) r.overflow = true; }
    else if (a.m > b.m) { r.m = a.m - b.m; r.neg = a.neg; }
    else if (a.m < b.m) { r.m = b.m - a.m; r.neg = b.neg; }
    else { r.m = 0; r.neg = false; }
    return r;
}
static inline PVal mag_to_val(const Mag& m, int32_t type, bool* overflow) {
    PVal r;
    if (m.overflow || m.m >= TEN38) *overflow = true;
    if (type == PG_LONG_DECIMAL) r.d = dec_pack((uint64_t)m.m, (uint64_t)(m.m >> 64) & ~SIGN_MASK, m.neg);
    else r.v = m.neg ? (int64_t)((uint64_t)0 - (uint64_t)m.m) : (int64_t)(uint64_t)m.m;      // short results wrap like Java's long arithmetic
    return r;
}

struct Project {
    std::vector<pg_expr> nodes;
    std::vector<std::vector<uint8_t>> bytes;
    std::vector<PType> types;
    int32_t filter = -1;
    std::vector<int32_t> proj;
    std::vector<char> in_filter;     // node is part of the filter expression: evaluated for every row, before the projections
    std::vector<Mag> up0, up1;       // ADD / SUB / CAST: the power of ten each operand is multiplied by to reach the result scale
    bool overflow = false;

    static PType decimal_type(int p, int s) { PType t; t.precision = p > 38 ? 38 : p; t.scale = s; t.type = t.precision <= 18 ? PG_SHORT_DECIMAL : PG_LONG_DECIMAL; return t; }

    int init(const pg_expr* n, int32_t count, int32_t f, const int32_t* p, int32_t np) {
        nodes.assign(n, n + count); bytes.resize(count); types.resize(count); filter = f; proj.assign(p, p + np);
        for (int i = 0; i < count; i++) {
            pg_expr& e = nodes[i];
            if (e.op == PG_EXPR_CONST && e.bytes) { bytes[i].assign(e.bytes, e.bytes + e.nbytes); }
            PType t;
            auto ta = [&](int a) { return types[a]; };
            switch (e.op) {
                case PG_EXPR_INPUT: case PG_EXPR_CONST: case PG_EXPR_CAST: t.type = e.type; t.precision = e.precision; t.scale = e.scale; break;
                case PG_EXPR_ADD: case PG_EXPR_SUB: {
                    PType a = ta(e.arg0), b = ta(e.arg1);
                    if (is_decimal(a.type) && is_decimal(b.type)) {          // DecimalOperators.java:84-85, 184-185
                        int s = a.scale > b.scale ? a.scale : b.scale;
                        int ip = (a.precision - a.scale) > (b.precision - b.scale) ? (a.precision - a.scale) : (b.precision - b.scale);
                        t = decimal_type(ip + s + 1, s);
                    } else t = a;
                    break;
                }
                case PG_EXPR_MUL: {
                    PType a = ta(e.arg0), b = ta(e.arg1);
                    if (is_decimal(a.type) && is_decimal(b.type)) t = decimal_type(a.precision + b.precision, a.scale + b.scale);   // :261-262
                    else t = a;
                    break;
                }
                case PG_EXPR_HASH: t.type = PG_BIGINT; break;
                default: t.type = PG_BOOLEAN;
            }
            types[i] = t;
        }
        up0.assign(count, Mag{1, false, false}); up1.assign(count, Mag{1, false, false});
        for (int i = 0; i < count; i++) {
            const pg_expr& e = nodes[i];
            if ((e.op == PG_EXPR_ADD || e.op == PG_EXPR_SUB) && is_decimal(types[i].type)) {
                up0[i].m = pow10_u128(types[i].scale - types[e.arg0].scale); up1[i].m = pow10_u128(types[i].scale - types[e.arg1].scale);
            } else if (e.op == PG_EXPR_CAST && is_decimal(types[i].type)) {
                up0[i].m = pow10_u128(types[i].scale - (is_decimal(types[e.arg0].type) ? types[e.arg0].scale : 0));
                up1[i].m = pow10_u128(types[i].precision);
            }
        }
        in_filter.assign(count, 0);
        if (filter >= 0) {
            in_filter[filter] = 1;
            for (int i = count - 1; i >= 0; i--)
                if (in_filter[i]) { if (nodes[i].arg0 >= 0 && nodes[i].op != PG_EXPR_INPUT && nodes[i].op != PG_EXPR_CONST) in_filter[nodes[i].arg0] = 1;
                                    if (nodes[i].arg1 >= 0 && nodes[i].op != PG_EXPR_INPUT && nodes[i].op != PG_EXPR_CONST) in_filter[nodes[i].arg1] = 1; }
        }
        return PG_OK;
    }

    PVal input(const pg_col& c, int64_t r) const {
        PVal x; x.null = col_is_null(c, r);
        if (x.null) return x;
        if (c.type == PG_LONG_DECIMAL) x.d = col_dec(c, r);
        else if (c.type == PG_VARCHAR || c.type == PG_VARBINARY) { x.p = (const uint8_t*)c.values + c.offsets[r]; x.len = c.offsets[r + 1] - c.offsets[r]; }
        else x.v = col_long(c, r);
        return x;
    }

    int64_t hash_of(const PVal& x, const PType& t) const {   // per-type HASH_CODE, null -> 0 (see type_hash in oracle_groupby.h)
        if (x.null) return 0;
        switch (t.type) {
            case PG_BIGINT: return long_type_hash(x.v);
            case PG_INTEGER: case PG_DATE: return long_type_hash((int32_t)x.v);
            case PG_DOUBLE: { double d = bits_to_double(x.v); if (d == 0) d = 0; return long_type_hash(double_to_long_bits(d)); }
            case PG_SHORT_DECIMAL: return x.v;
            case PG_LONG_DECIMAL: return xxh64_long((int64_t)x.d.lo) ^ xxh64_long((int64_t)x.d.hi);
            case PG_VARCHAR: case PG_VARBINARY: return xxh64_bytes(x.p, x.len);
            case PG_BOOLEAN: return x.v ? 1231 : 1237;
        }
        return 0;
    }

    PVal arith(int node, int op, const PVal& a, const PType& ta, const PVal& b, const PType& tb, const PType& tr) {
        PVal r;
        if (a.null || b.null) { r.null = true; return r; }
        if (is_decimal(ta.type)) {
            if (tr.type == PG_SHORT_DECIMAL) {      // addShortShortShort / subtractShortShortShort / multiplyShortShortShort: plain long arithmetic
                if (op == PG_EXPR_MUL) r.v = (int64_t)((uint64_t)a.v * (uint64_t)b.v);
                else {
                    const uint64_t x = (uint64_t)a.v * (uint64_t)up0[node].m, y = (uint64_t)b.v * (uint64_t)up1[node].m;
                    r.v = (int64_t)(op == PG_EXPR_ADD ? x + y : x - y);
                }
                return r;
            }
            Mag x = mag_of(a, ta.type), y = mag_of(b, tb.type), z;
            if (op == PG_EXPR_MUL) z = mag_mul(x, y);
            else {   // rescale both operands to the result scale (calculateShortRescaleParameters / calculateLongRescaleParameters)
                Mag sx = up0[node].m == 1 ? x : mag_mul(x, up0[node]);
                Mag sy = up1[node].m == 1 ? y : mag_mul(y, up1[node]);
                if (op == PG_EXPR_SUB) sy.neg = !sy.neg;
                // subtract(left, right) of equal values gives +0; of different signs it is an addition that keeps left's sign
                z = mag_add(sx, sy);
            }
            bool ov = false;
            r = mag_to_val(z, tr.type, &ov);
            if (ov && tr.type == PG_LONG_DECIMAL) overflow = true;     // short results cannot overflow by construction of the precision
            return r;
        }
        if (ta.type == PG_DOUBLE) {
            double x = bits_to_double(a.v), y = bits_to_double(b.v), z = op == PG_EXPR_ADD ? x + y : op == PG_EXPR_SUB ? x - y : x * y;
            memcpy(&r.v, &z, 8);
            return r;
        }
        // bigint / integer: Math.*Exact (BigintOperators.java:47-83)
        bool ov;
        if (ta.type == PG_BIGINT) {
            int64_t z;
            ov = op == PG_EXPR_ADD ? __builtin_add_overflow(a.v, b.v, &z) : op == PG_EXPR_SUB ? __builtin_sub_overflow(a.v, b.v, &z) : __builtin_mul_overflow(a.v, b.v, &z);
            r.v = z;
        } else {
            int32_t z, x = (int32_t)a.v, y = (int32_t)b.v;
            ov = op == PG_EXPR_ADD ? __builtin_add_overflow(x, y, &z) : op == PG_EXPR_SUB ? __builtin_sub_overflow(x, y, &z) : __builtin_mul_overflow(x, y, &z);
            r.v = z;
        }
        if (ov) overflow = true;
        return r;
    }

    // -1 / 0 / +1 ; decimals of equal scale compare by value, doubles by IEEE (unordered -> 2)
    int compare(const PVal& a, const PType& ta, const PVal& b, const PType& tb) const {
        if (ta.type == PG_DOUBLE) {
            double x = bits_to_double(a.v), y = bits_to_double(b.v);
            if (x != x || y != y) return 2;
            return x < y ? -1 : x > y ? 1 : 0;
        }
        if (ta.type == PG_VARCHAR || ta.type == PG_VARBINARY) {
            int64_t m = a.len < b.len ? a.len : b.len;
            int c = m ? memcmp(a.p, b.p, (size_t)m) : 0;
            if (c != 0) return c < 0 ? -1 : 1;
            return a.len < b.len ? -1 : a.len > b.len ? 1 : 0;
        }
        if (is_decimal(ta.type) && (ta.type == PG_LONG_DECIMAL || tb.type == PG_LONG_DECIMAL)) {
            Mag x = mag_of(a, ta.type), y = mag_of(b, tb.type);
            if (x.m == 0 && y.m == 0) return 0;
            if (x.neg != y.neg) return x.neg ? -1 : 1;
            int c = x.m < y.m ? -1 : x.m > y.m ? 1 : 0;
            return x.neg ? -c : c;
        }
        return a.v < b.v ? -1 : a.v > b.v ? 1 : 0;
    }

    PVal eval_node(int i, const std::vector<PVal>& vals, const pg_col* cols, int64_t r) {
        const pg_expr& e = nodes[i];
        PVal x;
        switch (e.op) {
            case PG_EXPR_INPUT: return input(cols[e.channel], r);
            case PG_EXPR_CONST:
                x.null = e.is_null != 0;
                if (e.type == PG_LONG_DECIMAL) { x.d.lo = (uint64_t)e.lo; x.d.hi = (uint64_t)e.hi; }
                else if (e.type == PG_VARCHAR || e.type == PG_VARBINARY) { x.p = bytes[i].data(); x.len = (int64_t)bytes[i].size(); }
                else x.v = e.lo;
                return x;
            case PG_EXPR_ADD: case PG_EXPR_SUB: case PG_EXPR_MUL:
                return arith(i, e.op, vals[e.arg0], types[e.arg0], vals[e.arg1], types[e.arg1], types[i]);
            case PG_EXPR_LT: case PG_EXPR_LE: case PG_EXPR_GT: case PG_EXPR_GE: case PG_EXPR_EQ: case PG_EXPR_NE: {
                const PVal &a = vals[e.arg0], &b = vals[e.arg1];
                if (a.null || b.null) { x.null = true; return x; }
                int c = compare(a, types[e.arg0], b, types[e.arg1]);
                bool res;
                if (c == 2) res = e.op == PG_EXPR_NE;
                else res = e.op == PG_EXPR_LT ? c < 0 : e.op == PG_EXPR_LE ? c <= 0 : e.op == PG_EXPR_GT ? c > 0 : e.op == PG_EXPR_GE ? c >= 0 : e.op == PG_EXPR_EQ ? c == 0 : c != 0;
                x.v = res ? 1 : 0;
                return x;
            }
            case PG_EXPR_AND: {      // FALSE wins over NULL
                const PVal &a = vals[e.arg0], &b = vals[e.arg1];
                bool af = !a.null && !a.v, bf = !b.null && !b.v;
                if (af || bf) { x.v = 0; return x; }
                if (a.null || b.null) { x.null = true; return x; }
                x.v = 1; return x;
            }
            case PG_EXPR_OR: {       // TRUE wins over NULL
                const PVal &a = vals[e.arg0], &b = vals[e.arg1];
                bool at = !a.null && a.v, bt = !b.null && b.v;
                if (at || bt) { x.v = 1; return x; }
                if (a.null || b.null) { x.null = true; return x; }
                x.v = 0; return x;
            }
            case PG_EXPR_NOT: x = vals[e.arg0]; if (!x.null) x.v = x.v ? 0 : 1; return x;
            case PG_EXPR_IS_NULL: x.v = vals[e.arg0].null ? 1 : 0; return x;
            case PG_EXPR_CAST: {
                const PVal& a = vals[e.arg0]; const PType& ta = types[e.arg0]; const PType& tr = types[i];
                if (a.null) { x.null = true; return x; }
                if (tr.type == PG_DOUBLE) { double d = (double)a.v; memcpy(&x.v, &d, 8); return x; }
                if (tr.type == PG_BIGINT) { x.v = a.v; return x; }
                // -> decimal: value * 10^(scale difference), checked against the target precision (DecimalCasts / rescale)
                Mag m = mag_of(a, ta.type);
                Mag z = mag_mul(m, up0[i]);
                if (z.overflow || z.m >= up1[i].m) overflow = true;
                bool ov = false;
                return mag_to_val(z, tr.type, &ov);
            }
            case PG_EXPR_HASH: {
                uint64_t prev = e.arg0 >= 0 ? (uint64_t)vals[e.arg0].v : 0;
                x.v = (int64_t)(31 * prev + (uint64_t)hash_of(vals[e.arg1], types[e.arg1]));
                return x;
            }
        }
        return x;
    }

    static void store(const PVal& v, const PType& t, pg_col_out* o, int64_t i) {
        if (o->nulls) o->nulls[i] = v.null ? 1 : 0;
        switch (t.type) {
            case PG_INTEGER: case PG_DATE: ((int32_t*)o->values)[i] = v.null ? 0 : (int32_t)v.v; break;
            case PG_BOOLEAN: ((uint8_t*)o->values)[i] = v.null ? 0 : (uint8_t)v.v; break;
            case PG_LONG_DECIMAL: ((uint64_t*)o->values)[2 * i] = v.null ? 0 : v.d.lo; ((uint64_t*)o->values)[2 * i + 1] = v.null ? 0 : v.d.hi; break;
            case PG_VARCHAR: case PG_VARBINARY: {
                if (i == 0) o->offsets[0] = 0;
                const int32_t off = o->offsets[i]; const int64_t len = v.null ? 0 : v.len;
                if (len) memcpy((uint8_t*)o->values + off, v.p, (size_t)len);
                o->offsets[i + 1] = off + (int32_t)len; break;
            }
            default: ((int64_t*)o->values)[i] = v.null ? 0 : v.v;
        }
    }

    // PageProcessor.process: the projections of the rows that pass the filter, written to out[] in page order; returns their number.
    // Like the reference, the filter produces the selected positions first (PageProcessor.java:139-180); input references are then
    // copied column by column over those positions (the reference wraps them in a DictionaryBlock view, InputPageProjection), and
    // the computed projections are evaluated row by row over the selected positions only.
    int64_t process(const pg_col* cols, int64_t n, pg_col_out* out) {
        std::vector<PVal> vals(nodes.size());
        std::vector<int> first, second;
        std::vector<char> needed(nodes.size(), 0);
        for (size_t k = 0; k < proj.size(); k++) if (nodes[proj[k]].op != PG_EXPR_INPUT) needed[proj[k]] = 1;
        for (int i = (int)nodes.size() - 1; i >= 0; i--)
            if (needed[i] && nodes[i].op != PG_EXPR_INPUT && nodes[i].op != PG_EXPR_CONST) {
                if (nodes[i].arg0 >= 0) needed[nodes[i].arg0] = 1;
                if (nodes[i].arg1 >= 0) needed[nodes[i].arg1] = 1;
            }
        for (size_t i = 0; i < nodes.size(); i++) {
            if (nodes[i].op == PG_EXPR_CONST) vals[i] = eval_node((int)i, vals, cols, 0);
            else if (in_filter[i]) first.push_back((int)i);
            else if (needed[i]) second.push_back((int)i);
        }
        std::vector<int64_t> sel;
        sel.reserve((size_t)n);
        for (int64_t r = 0; r < n; r++) {
            for (int i : first) vals[i] = eval_node(i, vals, cols, r);
            if (filter >= 0 && (vals[filter].null || !vals[filter].v)) continue;
            sel.push_back(r);
        }
        const int64_t m = (int64_t)sel.size();
        for (size_t k = 0; k < proj.size(); k++) {
            const pg_expr& e = nodes[proj[k]];
            if (e.op != PG_EXPR_INPUT) continue;
            const pg_col& c = cols[e.channel]; pg_col_out* o = &out[k];
            if (o->nulls) for (int64_t i = 0; i < m; i++) o->nulls[i] = c.nulls ? c.nulls[sel[i]] : 0;
            switch (c.type) {
                case PG_INTEGER: case PG_DATE: for (int64_t i = 0; i < m; i++) ((int32_t*)o->values)[i] = ((const int32_t*)c.values)[sel[i]]; break;
                case PG_BOOLEAN: for (int64_t i = 0; i < m; i++) ((uint8_t*)o->values)[i] = ((const uint8_t*)c.values)[sel[i]]; break;
                case PG_LONG_DECIMAL:
                    for (int64_t i = 0; i < m; i++) { ((uint64_t*)o->values)[2 * i] = ((const uint64_t*)c.values)[2 * sel[i]]; ((uint64_t*)o->values)[2 * i + 1] = ((const uint64_t*)c.values)[2 * sel[i] + 1]; }
                    break;
                case PG_VARCHAR: case PG_VARBINARY: {
                    int32_t off = 0; o->offsets[0] = 0;
                    for (int64_t i = 0; i < m; i++) {
                        const int64_t r = sel[i];
                        const int32_t len = (c.nulls && c.nulls[r]) ? 0 : c.offsets[r + 1] - c.offsets[r];
                        const uint8_t* src = (const uint8_t*)c.values + c.offsets[r];
                        uint8_t* dst = (uint8_t*)o->values + off;
                        if (len == 1) *dst = *src; else if (len) memcpy(dst, src, (size_t)len);
                        off += len; o->offsets[i + 1] = off;
                    }
                    break;
                }
                default: for (int64_t i = 0; i < m; i++) ((int64_t*)o->values)[i] = ((const int64_t*)c.values)[sel[i]];
            }
            // nulls of a pass-through keep zeros in the value slots, like the row-at-a-time store
            if (c.nulls && c.type != PG_VARCHAR && c.type != PG_VARBINARY)
                for (int64_t i = 0; i < m; i++) if (c.nulls[sel[i]]) {
                    if (c.type == PG_LONG_DECIMAL) { ((uint64_t*)o->values)[2 * i] = 0; ((uint64_t*)o->values)[2 * i + 1] = 0; }
                    else if (c.type == PG_INTEGER || c.type == PG_DATE) ((int32_t*)o->values)[i] = 0;
                    else if (c.type == PG_BOOLEAN) ((uint8_t*)o->values)[i] = 0;
                    else ((int64_t*)o->values)[i] = 0;
                }
        }
        if (!second.empty() || true) {
            bool any_computed = false;
            for (size_t k = 0; k < proj.size(); k++) any_computed = any_computed || nodes[proj[k]].op != PG_EXPR_INPUT;
            if (any_computed)
                for (int64_t i = 0; i < m; i++) {
                    const int64_t r = sel[i];
                    for (int j : first) if (needed[j]) vals[j] = eval_node(j, vals, cols, r);      // a projection may reuse a node of the filter
                    for (int j : second) vals[j] = eval_node(j, vals, cols, r);
                    for (size_t k = 0; k < proj.size(); k++) if (nodes[proj[k]].op != PG_EXPR_INPUT) store(vals[proj[k]], types[proj[k]], &out[k], i);
                }
        }
        return m;
    }
};

}  // namespace orc
