// oracle_accum.h — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Restates the grouped accumulators for count/sum/avg/min/max:
//   MAIN/operator/aggregation/{CountAggregation,CountColumn,LongSumAggregation,DoubleSumAggregation,
//   AverageAggregations,DecimalSumAggregation,DecimalAverageAggregation,AbstractMinMaxAggregationFunction}.java
// and the generated per-row loop of AccumulatorCompiler.java:487-566
//   for pos: if testMask(mask,pos) && !arg.isNull(pos) { state.setGroupId(ids[pos]); input(state, arg[pos]) }
#pragma once
#include "oracle_groupby.h"

namespace orc {

// T = overflow * 2^127 + sign * magnitude, as a 192-bit two's complement (w0 low).
struct I192 { uint64_t w0, w1, w2; };
static inline I192 i192_from_state(const Dec128& d, int64_t overflow) {
    // magnitude (127 bit) -> signed 192
    uint64_t m0 = d.lo, m1 = dec_hi_mag(d), m2 = 0;
    if (dec_is_negative(d)) {  // negate
        m0 = ~m0; m1 = ~m1; m2 = ~m2;
        if (++m0 == 0) if (++m1 == 0) ++m2;
    }
    // add overflow << 127
    uint64_t o1 = (uint64_t)overflow << 63;            // bit 127 lives in w1 bit 63
    uint64_t o2 = (uint64_t)(overflow >> 1);           // arithmetic shift keeps the sign
    uint64_t r1 = m1 + o1; uint64_t c = r1 < m1 ? 1 : 0;
    I192 r; r.w0 = m0; r.w1 = r1; r.w2 = m2 + o2 + c; return r;
}
// divide |T| by count with ROUND_HALF_UP (BigDecimal.divide(count, scale, ROUND_HALF_UP),
// DecimalAverageAggregation.java:209-220).  Returns sign-magnitude 128; *fits=false when > 127 bit.
static inline Dec128 i192_div_half_up(I192 t, uint64_t count, bool* fits) {
    bool neg = (t.w2 >> 63) != 0;
    if (neg) { t.w0 = ~t.w0; t.w1 = ~t.w1; t.w2 = ~t.w2; if (++t.w0 == 0) if (++t.w1 == 0) ++t.w2; }
    u128 rem = 0; uint64_t q[3]; uint64_t w[3] = {t.w2, t.w1, t.w0};
    for (int k = 0; k < 3; k++) { u128 cur = (rem << 64) | w[k]; q[k] = (uint64_t)(cur / count); rem = cur % count; }
    // half-up on the magnitude: 2*rem >= count
    if (rem * 2 >= (u128)count) { if (++q[2] == 0) if (++q[1] == 0) ++q[0]; }
    *fits = (q[0] == 0) && ((q[1] >> 63) == 0);
    bool zero = (q[1] == 0 && q[2] == 0);
    return dec_pack(q[2], q[1] & ~SIGN_MASK, neg && !zero);
}

enum { AGG_COUNT = 1, AGG_COUNT_COL = 2, AGG_SUM = 3, AGG_AVG = 4, AGG_MIN = 5, AGG_MAX = 6 };

struct Accum {
    int kind = 0; int in_type = 0;
    std::vector<int64_t> cnt;      // count / first
    std::vector<int64_t> lsum;     // bigint sum / min-max long value
    std::vector<double> dsum;      // double sum / min-max double
    std::vector<Dec128> dec;       // decimal sum
    std::vector<int64_t> ovf;      // decimal overflow counter
    std::vector<uint8_t> set;      // decimal "longDecimal != null" / min-max "!isNull"
    bool overflow_error = false;   // BigintOperators.add overflow (NUMERIC_VALUE_OUT_OF_RANGE)

    bool is_dec() const { return in_type == PG_SHORT_DECIMAL || in_type == PG_LONG_DECIMAL; }
    void ensure(int64_t groups) {
        if ((int64_t)cnt.size() >= groups) return;
        cnt.resize(groups, 0); lsum.resize(groups, 0); dsum.resize(groups, 0.0);
        Dec128 z; z.lo = 0; z.hi = 0; dec.resize(groups, z); ovf.resize(groups, 0); set.resize(groups, 0);
    }
    void dec_accumulate(int64_t g, const Dec128& v) {  // DecimalSumAggregation.java:151-157
        set[g] = 1;
        Dec128 out; ovf[g] += dec_add_with_overflow(dec[g], v, out); dec[g] = out;
    }
    void input_row(int64_t g, const pg_col& c, int64_t i) {
        switch (kind) {
            case AGG_COUNT: case AGG_COUNT_COL: cnt[g] += 1; break;
            case AGG_SUM:
                if (in_type == PG_BIGINT || in_type == PG_INTEGER) {         // LongSumAggregation.java:34-39
                    cnt[g] += 1; int64_t r;
                    if (__builtin_add_overflow(lsum[g], col_long(c, i), &r)) overflow_error = true; else lsum[g] = r;
                } else if (in_type == PG_DOUBLE) {                           // DoubleSumAggregation.java:33-38
                    cnt[g] += 1; dsum[g] += bits_to_double(col_long(c, i));
                } else if (in_type == PG_SHORT_DECIMAL) dec_accumulate(g, dec_from_long(col_long(c, i)));
                else dec_accumulate(g, col_dec(c, i));
                break;
            case AGG_AVG:
                if (in_type == PG_BIGINT || in_type == PG_INTEGER) { cnt[g] += 1; dsum[g] += (double)col_long(c, i); }  // AverageAggregations.java:34-39
                else if (in_type == PG_DOUBLE) { cnt[g] += 1; dsum[g] += bits_to_double(col_long(c, i)); }
                else { dec_accumulate(g, in_type == PG_SHORT_DECIMAL ? dec_from_long(col_long(c, i)) : col_dec(c, i)); cnt[g] += 1; }
                break;
            case AGG_MIN: case AGG_MAX: minmax(g, c, i); break;
        }
    }
    // AbstractMinMaxAggregationFunction.java:274-306 + MAIN/util/MinMaxCompare.java
    void minmax(int64_t g, const pg_col& c, int64_t i) {
        bool mn = kind == AGG_MIN;
        if (in_type == PG_DOUBLE) {
            double v = bits_to_double(col_long(c, i));
            if (!set[g]) { set[g] = 1; dsum[g] = v; return; }
            bool take = mn ? (double_compare(v, dsum[g]) < 0) : ((v > dsum[g]) || (dsum[g] != dsum[g]));
            if (take) dsum[g] = v;
        } else if (in_type == PG_LONG_DECIMAL) {
            Dec128 v = col_dec(c, i);
            if (!set[g]) { set[g] = 1; dec[g] = v; return; }
            int cmp = dec_compare_signed(v, dec[g]);
            if (mn ? cmp < 0 : cmp > 0) dec[g] = v;
        } else {
            int64_t v = col_long(c, i);
            if (!set[g]) { set[g] = 1; lsum[g] = v; return; }
            if (mn ? v < lsum[g] : v > lsum[g]) lsum[g] = v;
        }
    }
    // UnscaledDecimal128Arithmetic.compare :880-902 (negative zero == zero)
    static int dec_compare_signed(const Dec128& a, const Dec128& b) {
        bool az = a.lo == 0 && dec_hi_mag(a) == 0, bz = b.lo == 0 && dec_hi_mag(b) == 0;
        bool an = dec_is_negative(a) && !az, bn = dec_is_negative(b) && !bz;
        if (an != bn) return an ? -1 : 1;
        int c = dec_compare_abs(a, b);
        return an ? -c : c;
    }

    void add_input(const int32_t* ids, int64_t groups, const pg_col* arg, const pg_col* mask, int64_t n) {
        ensure(groups);
        for (int64_t i = 0; i < n; i++) {
            if (mask && (col_is_null(*mask, i) || !((const uint8_t*)mask->values)[i])) continue;
            if (kind != AGG_COUNT && arg && col_is_null(*arg, i)) continue;
            input_row(ids[i], *arg, i);
        }
    }

    // number of intermediate columns (see include/prestogpu.h for the layout table)
    int n_inter() const {
        if (kind == AGG_COUNT || kind == AGG_COUNT_COL || kind == AGG_MIN || kind == AGG_MAX) return 1;
        if (is_dec()) return 1;
        return 2;
    }
    // addIntermediate == combine() per row (AccumulatorCompiler.java:671-783)
    void add_intermediate(const int32_t* ids, int64_t groups, const pg_col* st, int64_t n) {
        ensure(groups);
        for (int64_t i = 0; i < n; i++) {
            int64_t g = ids[i];
            switch (kind) {
                case AGG_COUNT: case AGG_COUNT_COL: if (!col_is_null(st[0], i)) cnt[g] += col_long(st[0], i); break;
                case AGG_SUM: case AGG_AVG:
                    if (is_dec()) {
                        if (col_is_null(st[0], i)) break;
                        const uint8_t* p = (const uint8_t*)st[0].values + st[0].offsets[i];
                        int64_t o; Dec128 d;
                        if (kind == AGG_AVG) { int64_t c; memcpy(&c, p, 8); p += 8; cnt[g] += c; }
                        memcpy(&o, p, 8); memcpy(&d.lo, p + 8, 8); memcpy(&d.hi, p + 16, 8);
                        ovf[g] += o;                                     // DecimalSumAggregation.combine :165-175
                        if (!set[g]) { set[g] = 1; dec[g] = d; } else dec_accumulate(g, d);
                    } else if (kind == AGG_SUM) {                        // [count BIGINT, sum BIGINT|DOUBLE]
                        int64_t c = col_long(st[0], i); cnt[g] += c;
                        if (in_type == PG_DOUBLE) dsum[g] += bits_to_double(col_long(st[1], i));
                        else { int64_t r; if (__builtin_add_overflow(lsum[g], col_long(st[1], i), &r)) overflow_error = true; else lsum[g] = r; }
                    } else {                                             // avg: [sum DOUBLE, count BIGINT]
                        dsum[g] += bits_to_double(col_long(st[0], i)); cnt[g] += col_long(st[1], i);
                    }
                    break;
                case AGG_MIN: case AGG_MAX: if (!col_is_null(st[0], i)) minmax(g, st[0], i); break;
            }
        }
    }
};

}  // namespace orc
