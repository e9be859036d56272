// common.cuh — device/host helpers shared by every kernel of the hash-aggregation / hash-join path.
//
// Arithmetic that must agree bit-for-bit with the reference (paths relative to /root/reference,
// MAIN = core/trino-main/src/main/java/io/trino, SPI = core/trino-spi/src/main/java/io/trino/spi):
//   long_type_hash   SPI/type/AbstractLongType.java:126-130
//   fmix64           fastutil HashCommon.murmurHash3(long), restated at MAIN/operator/PagesHash.java:224-240
//   xxh64_*          io.airlift.slice.XxHash64 (standard XXH64, seed 0)
//   row hash fold    MAIN/operator/InterpretedHashGenerator.java:56-70, scalar/CombineHashFunction.java:25-30
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/prestogpu.h"

#define PG_HD __host__ __device__ __forceinline__
#define PG_D __device__ __forceinline__

namespace pg {

// ------------------------------------------------------------------ hashing
static const uint64_t XP1 = 0x9E3779B185EBCA87ULL;
static const uint64_t XP2 = 0xC2B2AE3D27D4EB4FULL;
static const uint64_t XP3 = 0x165667B19E3779F9ULL;
static const uint64_t XP4 = 0x85EBCA77C2B2AE63ULL;
static const uint64_t XP5 = 0x27D4EB2F165667C5ULL;

PG_HD uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
PG_HD int64_t long_type_hash(int64_t v) { return (int64_t)(rotl64((uint64_t)v * XP2, 31) * XP1); }
PG_HD uint64_t fmix64(uint64_t h) {
    h ^= h >> 33; h *= 0xff51afd7ed558ccdULL;
    h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 33;
    return h;
}
PG_HD uint64_t xxh_round(uint64_t acc, uint64_t v) { return rotl64(acc + v * XP2, 31) * XP1; }
PG_HD uint64_t xxh_merge(uint64_t h, uint64_t v) { return (h ^ xxh_round(0, v)) * XP1 + XP4; }
PG_HD uint64_t xxh_final(uint64_t h) {
    h ^= h >> 33; h *= XP2; h ^= h >> 29; h *= XP3; h ^= h >> 32; return h;
}
PG_HD int64_t xxh64_long(int64_t v) {
    uint64_t h = XP5 + 8;
    h ^= xxh_round(0, (uint64_t)v);
    h = rotl64(h, 27) * XP1 + XP4;
    return (int64_t)xxh_final(h);
}
PG_HD uint64_t rd64u(const uint8_t* p) {
    uint64_t v = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) v |= (uint64_t)p[i] << (8 * i);
    return v;
}
PG_HD uint32_t rd32u(const uint8_t* p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
PG_HD int64_t xxh64_bytes(const uint8_t* p, int64_t len) {
    const uint8_t* end = p + len;
    uint64_t h;
    if (len >= 32) {
        uint64_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0 - XP1;
        const uint8_t* lim = end - 32;
        do {
            v1 = xxh_round(v1, rd64u(p)); v2 = xxh_round(v2, rd64u(p + 8));
            v3 = xxh_round(v3, rd64u(p + 16)); v4 = xxh_round(v4, rd64u(p + 24));
            p += 32;
        } while (p <= lim);
        h = rotl64(v1, 1) + rotl64(v2, 7) + rotl64(v3, 12) + rotl64(v4, 18);
        h = xxh_merge(h, v1); h = xxh_merge(h, v2); h = xxh_merge(h, v3); h = xxh_merge(h, v4);
    } else {
        h = XP5;
    }
    h += (uint64_t)len;
    while (p + 8 <= end) { h ^= xxh_round(0, rd64u(p)); h = rotl64(h, 27) * XP1 + XP4; p += 8; }
    if (p + 4 <= end) { h ^= (uint64_t)rd32u(p) * XP1; h = rotl64(h, 23) * XP2 + XP3; p += 4; }
    while (p < end) { h ^= (uint64_t)(*p) * XP5; h = rotl64(h, 11) * XP1; p++; }
    return (int64_t)xxh_final(h);
}
// LocalPartitionGenerator.getPartition (MAIN/operator/exchange/LocalPartitionGenerator.java:50-63)
PG_HD uint64_t bit_reverse64(uint64_t x) {
    x = ((x >> 1) & 0x5555555555555555ULL) | ((x & 0x5555555555555555ULL) << 1);
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
    x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
    return (x >> 32) | (x << 32);
}

// fastutil HashCommon.arraySize(expected, 0.75f) and BigintGroupByHash.calculateMaxFill (:325-334)
static inline int64_t array_size_075(int64_t expected) {
    float q = (float)expected / 0.75f;
    int64_t need = (int64_t)ceil((double)q);
    int64_t s = 1;
    while (s < need) s <<= 1;
    return s < 2 ? 2 : s;
}
static inline int64_t max_fill_075(int64_t cap) {
    int64_t mf = (int64_t)ceil((double)((float)cap * 0.75f));
    if (mf == cap) mf--;
    return mf;
}

// ------------------------------------------------------------------ doubles
PG_HD int64_t canon_double_bits(int64_t b) {  // Double.doubleToLongBits: canonical NaN
    uint64_t u = (uint64_t)b;
    if ((u & 0x7ff0000000000000ULL) == 0x7ff0000000000000ULL && (u & 0x000fffffffffffffULL)) return 0x7ff8000000000000LL;
    return b;
}
PG_HD bool bits_is_nan(int64_t b) {
    uint64_t u = (uint64_t)b;
    return (u & 0x7ff0000000000000ULL) == 0x7ff0000000000000ULL && (u & 0x000fffffffffffffULL);
}
// Total order of Double.compare as a signed int64 key (-0 < +0, canonical NaN largest).
PG_HD int64_t double_order_key(int64_t b) {
    b = canon_double_bits(b);
    return b ^ ((b >> 63) & 0x7fffffffffffffffLL);
}
PG_HD int64_t double_order_key_inv(int64_t k) { return k ^ ((k >> 63) & 0x7fffffffffffffffLL); }

// ------------------------------------------------------------------ device column view (one Trino Block)
struct DCol {
    int32_t type;
    int32_t pad;
    const void* values;
    const uint8_t* nulls;
    const int32_t* offsets;
};

PG_D bool dcol_is_null(const DCol& c, int64_t i) { return c.nulls != nullptr && c.nulls[i] != 0; }
PG_D int64_t dcol_long(const DCol& c, int64_t i) {
    switch (c.type) {
        case PG_INTEGER: case PG_DATE: return ((const int32_t*)c.values)[i];
        case PG_BOOLEAN: return ((const int8_t*)c.values)[i];
        default: return ((const int64_t*)c.values)[i];
    }
}
// per-type HASH_CODE, null -> 0 (MAIN/type/TypeUtils.java:42); see oracle/oracle_groupby.h type_hash
PG_D int64_t dcol_type_hash(const DCol& c, int64_t i) {
    if (dcol_is_null(c, i)) return 0;
    switch (c.type) {
        case PG_BIGINT: return long_type_hash(((const int64_t*)c.values)[i]);
        case PG_INTEGER: case PG_DATE: return long_type_hash((int64_t)((const int32_t*)c.values)[i]);
        case PG_DOUBLE: {
            int64_t b = ((const int64_t*)c.values)[i];
            if ((b & 0x7fffffffffffffffLL) == 0) b = 0;  // -0 -> +0 (SPI/type/DoubleType.java:164-170)
            return long_type_hash(canon_double_bits(b));
        }
        case PG_SHORT_DECIMAL: return ((const int64_t*)c.values)[i];  // identity (ShortDecimalType.java:128-131)
        case PG_LONG_DECIMAL: {
            const int64_t* p = (const int64_t*)c.values + 2 * i;
            return xxh64_long(p[0]) ^ xxh64_long(p[1]);
        }
        case PG_VARCHAR: case PG_VARBINARY: {
            const uint8_t* b = (const uint8_t*)c.values;
            int32_t o = c.offsets[i];
            return xxh64_bytes(b + o, c.offsets[i + 1] - o);
        }
        case PG_BOOLEAN: return ((const int8_t*)c.values)[i] ? 1231 : 1237;
    }
    return 0;
}

static inline int type_width(int32_t t) {
    switch (t) {
        case PG_INTEGER: case PG_DATE: return 4;
        case PG_BOOLEAN: return 1;
        case PG_LONG_DECIMAL: return 16;
        case PG_VARCHAR: case PG_VARBINARY: return 0;
        default: return 8;
    }
}
static inline bool type_is_var(int32_t t) { return t == PG_VARCHAR || t == PG_VARBINARY; }
static inline bool type_valid(int32_t t) { return t >= PG_BIGINT && t <= PG_BOOLEAN; }

}  // namespace pg
