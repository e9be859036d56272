// oracle_groupby.h — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Restates MAIN/operator/{BigintGroupByHash,MultiChannelGroupByHash,InterpretedHashGenerator}.java
#pragma once
#include "oracle_common.h"

namespace orc {

// ---- Block accessors over pg_col (host memory only) ----
static inline bool col_is_null(const pg_col& c, int64_t i) { return c.nulls && c.nulls[i]; }
static inline int64_t col_long(const pg_col& c, int64_t i) {
    switch (c.type) {
        case PG_INTEGER: case PG_DATE: return ((const int32_t*)c.values)[i];
        case PG_BOOLEAN: return ((const int8_t*)c.values)[i];
        default: return ((const int64_t*)c.values)[i];
    }
}
static inline Dec128 col_dec(const pg_col& c, int64_t i) {
    const uint64_t* p = (const uint64_t*)c.values;
    Dec128 d; d.lo = p[2 * i]; d.hi = p[2 * i + 1]; return d;
}

// per-type HASH_CODE (falls back to XX_HASH_64 when the type declares none,
// SPI/type/TypeOperators.java:222-236); null -> NULL_HASH_CODE = 0 (MAIN/type/TypeUtils.java:42)
static inline int64_t type_hash(const pg_col& c, int64_t i) {
    if (col_is_null(c, i)) return 0;
    switch (c.type) {
        case PG_BIGINT: return long_type_hash(col_long(c, i));                 // AbstractLongType.java:138-142
        case PG_INTEGER: case PG_DATE: return long_type_hash((int32_t)col_long(c, i)); // AbstractIntType.java:142-145
        case PG_DOUBLE: {                                                          // DoubleType.java:164-170
            double d = bits_to_double(col_long(c, i));
            if (d == 0) d = 0;
            return long_type_hash(double_to_long_bits(d));
        }
        case PG_SHORT_DECIMAL: return col_long(c, i);                            // ShortDecimalType.java:128-131 (identity)
        case PG_LONG_DECIMAL: { Dec128 d = col_dec(c, i); return xxh64_long((int64_t)d.lo) ^ xxh64_long((int64_t)d.hi); } // LongDecimalType.java:165-179
        case PG_VARCHAR: case PG_VARBINARY: {                                      // VarcharType.java:246-256
            const uint8_t* b = (const uint8_t*)c.values;
            return xxh64_bytes(b + c.offsets[i], c.offsets[i + 1] - c.offsets[i]);
        }
        case PG_BOOLEAN: return col_long(c, i) ? 1231 : 1237;                    // BooleanType hashCode
    }
    return 0;
}
// MAIN/operator/InterpretedHashGenerator.java:56-70 + scalar/CombineHashFunction.java:25-30
static inline int64_t row_hash(const pg_col* cols, int n, int64_t i) {
    uint64_t h = 0;  // INITIAL_HASH_VALUE, HashGenerationOptimizer.java:98
    for (int k = 0; k < n; k++) h = 31 * h + (uint64_t)type_hash(cols[k], i);
    return (int64_t)h;
}

// A growable typed column (the reference appends new keys to PageBuilders; we keep one
// vector per channel).  Also used for build-side storage in the join oracle.
struct Store {
    int32_t type = 0;
    std::vector<int64_t> v;     // fixed width (2 per row for LONG_DECIMAL)
    std::vector<uint8_t> nulls;
    std::vector<uint8_t> bytes;
    std::vector<int32_t> offs{0};
    int64_t size() const { return (int64_t)nulls.size(); }
    void append(const pg_col& c, int64_t i) {
        bool isn = col_is_null(c, i);
        nulls.push_back(isn ? 1 : 0);
        if (type == PG_VARCHAR || type == PG_VARBINARY) {
            if (!isn) {
                const uint8_t* b = (const uint8_t*)c.values;
                bytes.insert(bytes.end(), b + c.offsets[i], b + c.offsets[i + 1]);
            }
            offs.push_back((int32_t)bytes.size());
        } else if (type == PG_LONG_DECIMAL) {
            if (isn) { v.push_back(0); v.push_back(0); }
            else { Dec128 d = col_dec(c, i); v.push_back((int64_t)d.lo); v.push_back((int64_t)d.hi); }
        } else {
            v.push_back(isn ? 0 : col_long(c, i));
        }
    }
    // IS_DISTINCT_FROM == false (null == null; NaN == NaN for doubles, DoubleType.java:180-190)
    bool not_distinct(int64_t row, const pg_col& c, int64_t i) const {
        bool a = nulls[row], b = col_is_null(c, i);
        if (a || b) return a == b;
        return equal_nonnull(row, c, i);
    }
    // nan_eq=true: IS_DISTINCT_FROM semantics (group-by); false: EQUAL operator (join, NaN != NaN)
    bool equal_nonnull(int64_t row, const pg_col& c, int64_t i, bool nan_eq = true) const {
        if (type == PG_VARCHAR || type == PG_VARBINARY) {
            int32_t la = offs[row + 1] - offs[row], lb = c.offsets[i + 1] - c.offsets[i];
            if (la != lb) return false;
            return memcmp(bytes.data() + offs[row], (const uint8_t*)c.values + c.offsets[i], la) == 0;
        }
        if (type == PG_LONG_DECIMAL) { Dec128 d = col_dec(c, i); return (uint64_t)v[2 * row] == d.lo && (uint64_t)v[2 * row + 1] == d.hi; }
        if (type == PG_DOUBLE) {
            double x = bits_to_double(v[row]), y = bits_to_double(col_long(c, i));
            if (x != x && y != y) return nan_eq;
            return x == y;
        }
        return v[row] == col_long(c, i);
    }
};

// ---------------------------------------------------------------------------------
// GroupByHash: one object restating both implementations; `bigint_mode` selects
// BigintGroupByHash (GroupByHash.createGroupByHash, MAIN/operator/GroupByHash.java:45-59).
// ---------------------------------------------------------------------------------
struct GroupBy {
    bool bigint_mode = false;
    bool has_hash_channel = false;  // precomputed $hashvalue channel present
    int nkeys = 0;
    std::vector<int32_t> types;
    int64_t capacity = 0, mask = 0, max_fill = 0;
    int64_t next_group = 0;
    int64_t hash_collisions = 0;
    int64_t rehash_count = 0;
    // bigint mode
    std::vector<int64_t> values;      // by slot
    std::vector<int32_t> group_ids;   // by slot, -1 empty
    std::vector<int64_t> values_by_group;
    int64_t null_group = -1;
    // multi-channel mode
    std::vector<int64_t> group_by_slot;  // groupId per slot or -1 (stands for groupAddressByHash)
    std::vector<uint8_t> raw_hash_byte;  // rawHashByHashPosition
    std::vector<Store> keys;             // channelBuilders
    std::vector<int64_t> hash_by_group;  // precomputed hash channel copy (or recomputed)

    void init(const int32_t* t, int n, bool hash_channel, int64_t expected) {
        nkeys = n; types.assign(t, t + n); has_hash_channel = hash_channel;
        bigint_mode = (n == 1 && t[0] == PG_BIGINT);
        capacity = array_size(expected, 0.75f);
        max_fill = calculate_max_fill(capacity);
        mask = capacity - 1;
        if (bigint_mode) { values.assign(capacity, 0); group_ids.assign(capacity, -1); }
        else {
            group_by_slot.assign(capacity, -1); raw_hash_byte.assign(capacity, 0);
            keys.resize(n); for (int k = 0; k < n; k++) keys[k].type = t[k];
        }
    }
    bool need_rehash() const { return next_group >= max_fill; }

    // BigintGroupByHash.tryRehash :262-313 / MultiChannelGroupByHash.tryRehash :364-424
    int rehash() {
        int64_t nc = capacity * 2;
        if (nc > 0x7fffffffLL) return PG_E_INSUFFICIENT_RESOURCES;
        int64_t nm = nc - 1;
        if (bigint_mode) {
            std::vector<int64_t> nv(nc, 0); std::vector<int32_t> ng(nc, -1);
            for (int64_t g = 0; g < next_group; g++) {
                if (g == null_group) continue;
                int64_t val = values_by_group[g];
                int64_t p = (int64_t)(fmix64((uint64_t)val) & (uint64_t)nm);
                while (ng[p] != -1) { p = (p + 1) & nm; hash_collisions++; }
                nv[p] = val; ng[p] = (int32_t)g;
            }
            values.swap(nv); group_ids.swap(ng);
        } else {
            std::vector<int64_t> ns(nc, -1); std::vector<uint8_t> nb(nc, 0);
            // the reference walks old slots in slot order (:384-404)
            for (int64_t s = 0; s < capacity; s++) {
                int64_t g = group_by_slot[s];
                if (g == -1) continue;
                int64_t h = hash_by_group[g];
                int64_t p = (int64_t)(fmix64((uint64_t)h) & (uint64_t)nm);
                while (ns[p] != -1) { p = (p + 1) & nm; hash_collisions++; }
                ns[p] = g; nb[p] = (uint8_t)h;
            }
            group_by_slot.swap(ns); raw_hash_byte.swap(nb);
        }
        capacity = nc; mask = nm; max_fill = calculate_max_fill(nc); rehash_count++;
        return PG_OK;
    }

    // BigintGroupByHash.putIfAbsent :213-260
    int64_t put_bigint(const pg_col& c, int64_t i) {
        if (col_is_null(c, i)) {
            // note: no rehash check here in the reference (:215-222); the Work loop does it
            if (null_group < 0) { null_group = next_group++; values_by_group.push_back(0); }
            return null_group;
        }
        int64_t val = col_long(c, i);
        int64_t p = (int64_t)(fmix64((uint64_t)val) & (uint64_t)mask);
        while (true) {
            int32_t g = group_ids[p];
            if (g == -1) break;
            if (values[p] == val) return g;
            p = (p + 1) & mask; hash_collisions++;
        }
        int64_t g = next_group++;
        values[p] = val; values_by_group.push_back(val); group_ids[p] = (int32_t)g;
        if (need_rehash()) rehash();
        return g;
    }
    // MultiChannelGroupByHash.putIfAbsent :281-342, positionNotDistinctFromCurrentRow :441-447
    int64_t put_multi(const pg_col* cols, const pg_col* hash_col, int64_t i) {
        int64_t raw = hash_col ? col_long(*hash_col, i) : row_hash(cols, nkeys, i);
        int64_t p = (int64_t)(fmix64((uint64_t)raw) & (uint64_t)mask);
        while (group_by_slot[p] != -1) {
            if (raw_hash_byte[p] == (uint8_t)raw) {
                int64_t g = group_by_slot[p]; bool eq = true;
                for (int k = 0; k < nkeys && eq; k++) eq = keys[k].not_distinct(g, cols[k], i);
                if (eq) return g;
            }
            p = (p + 1) & mask; hash_collisions++;
        }
        for (int k = 0; k < nkeys; k++) keys[k].append(cols[k], i);
        hash_by_group.push_back(raw);
        int64_t g = next_group++;
        group_by_slot[p] = g; raw_hash_byte[p] = (uint8_t)raw;
        if (need_rehash()) rehash();
        return g;
    }
    int64_t put(const pg_col* cols, const pg_col* hash_col, int64_t i) {
        return bigint_mode ? put_bigint(cols[0], i) : put_multi(cols, hash_col, i);
    }
    // contains(): BigintGroupByHash.java:178-203, MultiChannelGroupByHash.java:253-266
    bool contains(const pg_col* cols, const pg_col* hash_col, int64_t i) const {
        if (bigint_mode) {
            const pg_col& c = cols[0];
            if (col_is_null(c, i)) return null_group >= 0;
            int64_t val = col_long(c, i);
            int64_t p = (int64_t)(fmix64((uint64_t)val) & (uint64_t)mask);
            while (true) {
                if (group_ids[p] == -1) return false;
                if (values[p] == val) return true;
                p = (p + 1) & mask;
            }
        }
        int64_t raw = hash_col ? col_long(*hash_col, i) : row_hash(cols, nkeys, i);
        int64_t p = (int64_t)(fmix64((uint64_t)raw) & (uint64_t)mask);
        while (group_by_slot[p] != -1) {
            if (raw_hash_byte[p] == (uint8_t)raw) {
                int64_t g = group_by_slot[p]; bool eq = true;
                for (int k = 0; k < nkeys && eq; k++) eq = keys[k].not_distinct(g, cols[k], i);
                if (eq) return true;
            }
            p = (p + 1) & mask;
        }
        return false;
    }
    // getRawHash(groupId): BigintGroupByHash.java:205-209, MultiChannelGroupByHash.java:159-166
    int64_t raw_hash(int64_t g) const {
        if (bigint_mode) return g == null_group ? 0 : long_type_hash(values_by_group[g]);
        return hash_by_group[g];
    }
};

}  // namespace orc
