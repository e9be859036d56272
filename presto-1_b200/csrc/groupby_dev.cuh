// groupby_dev.cuh — device helpers of the GroupByHash table shared by groupby.cu and the fused aggregation kernel.
#pragma once
#include "groupby.h"

namespace pg {

// ------------------------------------------------------------------ key helpers (multi-channel mode)
PG_D int64_t keys_row_hash(const KeyCols& kc, int nkeys, int64_t i) {
    uint64_t h = 0;  // INITIAL_HASH_VALUE (HashGenerationOptimizer.java:98)
    for (int k = 0; k < nkeys; k++) h = 31 * h + (uint64_t)dcol_type_hash(kc.c[k], i);
    return (int64_t)h;
}

PG_D bool double_not_distinct(int64_t a, int64_t b) {  // SPI/type/DoubleType.java:181-192
    if (bits_is_nan(a) && bits_is_nan(b)) return true;
    return __longlong_as_double(a) == __longlong_as_double(b);
}

// IS NOT DISTINCT FROM between input rows i and j (same page)
PG_D bool keys_equal_rows(const KeyCols& kc, const KeyLayout& L, int64_t i, int64_t j) {
    for (int k = 0; k < L.nkeys; k++) {
        const DCol& c = kc.c[k];
        bool ni = dcol_is_null(c, i), nj = dcol_is_null(c, j);
        if (ni || nj) { if (ni != nj) return false; continue; }
        switch (c.type) {
            case PG_VARCHAR: case PG_VARBINARY: {
                int32_t oi = c.offsets[i], li = c.offsets[i + 1] - oi, oj = c.offsets[j], lj = c.offsets[j + 1] - oj;
                if (li != lj) return false;
                const uint8_t* b = (const uint8_t*)c.values;
                for (int32_t t = 0; t < li; t++) if (b[oi + t] != b[oj + t]) return false;
                break;
            }
            case PG_LONG_DECIMAL: {
                const int64_t* p = (const int64_t*)c.values;
                if (p[2 * i] != p[2 * j] || p[2 * i + 1] != p[2 * j + 1]) return false;
                break;
            }
            case PG_DOUBLE:
                if (!double_not_distinct(dcol_long(c, i), dcol_long(c, j))) return false;
                break;
            default:
                if (dcol_long(c, i) != dcol_long(c, j)) return false;
        }
    }
    return true;
}

// IS NOT DISTINCT FROM between input row i and the stored key row of group g
PG_D bool keys_equal_store(const KeyCols& kc, const KeyLayout& L, int64_t i, const uint64_t* key_rows, const uint8_t* arena,
                           int64_t g) {
    const uint64_t* row = key_rows + g * L.row_words;
    uint64_t nullbits = row[0];
    for (int k = 0; k < L.nkeys; k++) {
        const DCol& c = kc.c[k];
        bool ni = dcol_is_null(c, i), ng = (nullbits >> k) & 1;
        if (ni || ng) { if (ni != ng) return false; continue; }
        const uint64_t* w = row + L.word_off[k];
        switch (c.type) {
            case PG_VARCHAR: case PG_VARBINARY: {
                int32_t oi = c.offsets[i], li = c.offsets[i + 1] - oi;
                int64_t og = (int64_t)(w[0] >> 24); int32_t lg = (int32_t)(w[0] & 0xFFFFFF);
                if (li != lg) return false;
                const uint8_t* b = (const uint8_t*)c.values;
                for (int32_t t = 0; t < li; t++) if (b[oi + t] != arena[og + t]) return false;
                break;
            }
            case PG_LONG_DECIMAL: {
                const uint64_t* p = (const uint64_t*)c.values;
                if (p[2 * i] != w[0] || p[2 * i + 1] != w[1]) return false;
                break;
            }
            case PG_DOUBLE:
                if (!double_not_distinct(dcol_long(c, i), (int64_t)w[0])) return false;
                break;
            default:
                if (dcol_long(c, i) != (int64_t)w[0]) return false;
        }
    }
    return true;
}


// read-only view of a GroupByHash for look-ups from other kernels
struct GbView {
    const Slot* slots;
    int64_t cap;
    int64_t mask;
    const uint64_t* key_rows;
    const uint8_t* arena;
    int32_t bigint_mode;
    int32_t has_hash;
    DCol hash_col;
    KeyLayout L;
    KeyCols kc;
};

// group id of input row `row`, or -1 when its key has no numbered group yet (slots are read through L1: see gb_insert_kernel)
PG_D int32_t gb_find_gid(const GbView& v, int64_t row) {
    if (v.bigint_mode) {
        const DCol& c = v.kc.c[0];
        if (dcol_is_null(c, row)) return v.slots[v.cap + 1].gid;
        unsigned long long key = (unsigned long long)((const int64_t*)c.values)[row];
        if (key == EMPTY_KEY) return v.slots[v.cap].gid;
        int64_t p = (int64_t)(fmix64(key) & (uint64_t)v.mask);
        while (true) {
            unsigned long long k = v.slots[p].word;
            if (k == key) return v.slots[p].gid;
            if (k == EMPTY_KEY) return -1;
            p = (p + 1) & v.mask;
        }
    }
    int64_t raw = v.has_hash ? ((const int64_t*)v.hash_col.values)[row] : keys_row_hash(v.kc, v.L.nkeys, row);
    uint64_t f = fmix64((uint64_t)raw);
    uint32_t tag = (uint32_t)(f >> 32);
    int64_t p = (int64_t)(f & (uint64_t)v.mask);
    while (true) {
        unsigned long long w = v.slots[p].word;
        if (w == EMPTY_WORD) return -1;
        uint32_t ref = (uint32_t)w;
        if ((uint32_t)(w >> 32) == tag && !(ref & REF_PENDING) && keys_equal_store(v.kc, v.L, row, v.key_rows, v.arena, (int64_t)ref))
            return (int32_t)ref;
        p = (p + 1) & v.mask;
    }
}

}  // namespace pg
