// oracle_common.h — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the reference's (phd3/presto-1 = Trino 352-SNAPSHOT) hash
// aggregation / hash join arithmetic.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may call into oracle/.  The
// product path (presto-1_b200/) never links or imports it.
//
// Parity pinning: the reference is Java and cannot be compiled in this image
// (no JDK), so this restatement is pinned against the known-answer vectors in
// the reference's own tests (see tests/test_oracle_golden.py, tests/golden/).
// Hash constants of third-party code (airlift slice XxHash64, fastutil
// HashCommon) are restated from their published algorithms (XXH64 seed 0,
// MurmurHash3 fmix64) and pinned against the public XXH64 test vectors.
//
// Paths cited are relative to /root/reference/.  MAIN = core/trino-main/src/main/java/io/trino,
// SPI = core/trino-spi/src/main/java/io/trino/spi.
#pragma once
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <string>

#include "../include/prestogpu.h"

namespace orc {

static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

static const uint64_t P1 = 0x9E3779B185EBCA87ULL;
static const uint64_t P2 = 0xC2B2AE3D27D4EB4FULL;
static const uint64_t P3 = 0x165667B19E3779F9ULL;
static const uint64_t P4 = 0x85EBCA77C2B2AE63ULL;
static const uint64_t P5 = 0x27D4EB2F165667C5ULL;

// SPI/type/AbstractLongType.java:126-130  ("xxHash64 mix")
static inline int64_t long_type_hash(int64_t v) {
    return (int64_t)(rotl64((uint64_t)v * P2, 31) * P1);
}

// fastutil HashCommon.murmurHash3(long) == MurmurHash3 fmix64; restated in-tree at
// MAIN/operator/PagesHash.java:224-240
static inline uint64_t fmix64(uint64_t h) {
    h ^= h >> 33; h *= 0xff51afd7ed558ccdULL;
    h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 33;
    return h;
}

static inline uint64_t xxh_round(uint64_t acc, uint64_t v) { return rotl64(acc + v * P2, 31) * P1; }
static inline uint64_t xxh_merge(uint64_t h, uint64_t v) { return (h ^ xxh_round(0, v)) * P1 + P4; }
static inline uint64_t xxh_final(uint64_t h) {
    h ^= h >> 33; h *= P2; h ^= h >> 29; h *= P3; h ^= h >> 32; return h;
}
// io.airlift.slice.XxHash64.hash(long): XXH64 (seed 0) of the 8 little-endian bytes.
static inline int64_t xxh64_long(int64_t v) {
    uint64_t h = P5 + 8;
    h ^= xxh_round(0, (uint64_t)v);
    h = rotl64(h, 27) * P1 + P4;
    return (int64_t)xxh_final(h);
}
static inline uint64_t rd64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; }
static inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
// io.airlift.slice.XxHash64.hash(Slice, off, len): standard XXH64, seed 0.
static inline int64_t xxh64_bytes(const uint8_t* p, int64_t len) {
    const uint8_t* end = p + len;
    uint64_t h;
    if (len >= 32) {
        uint64_t v1 = P1 + P2, v2 = P2, v3 = 0, v4 = 0 - P1;
        const uint8_t* lim = end - 32;
        do {
            v1 = xxh_round(v1, rd64(p)); v2 = xxh_round(v2, rd64(p + 8));
            v3 = xxh_round(v3, rd64(p + 16)); v4 = xxh_round(v4, rd64(p + 24));
            p += 32;
        } while (p <= lim);
        h = rotl64(v1, 1) + rotl64(v2, 7) + rotl64(v3, 12) + rotl64(v4, 18);
        h = xxh_merge(h, v1); h = xxh_merge(h, v2); h = xxh_merge(h, v3); h = xxh_merge(h, v4);
    } else {
        h = P5;
    }
    h += (uint64_t)len;
    while (p + 8 <= end) { h ^= xxh_round(0, rd64(p)); h = rotl64(h, 27) * P1 + P4; p += 8; }
    if (p + 4 <= end) { h ^= (uint64_t)rd32(p) * P1; h = rotl64(h, 23) * P2 + P3; p += 4; }
    while (p < end) { h ^= (uint64_t)(*p) * P5; h = rotl64(h, 11) * P1; p++; }
    return (int64_t)xxh_final(h);
}

// fastutil 8.3.0 HashCommon.arraySize(int expected, float f):
//   max(2, nextPowerOfTwo((long) Math.ceil(expected / f)))   (float division)
static inline int64_t array_size(int64_t expected, float f) {
    float q = (float)expected / f;
    int64_t need = (int64_t)std::ceil((double)q);
    int64_t s = 1;
    while (s < need) s <<= 1;
    if (s < 2) s = 2;
    return s;
}
// MAIN/operator/BigintGroupByHash.java:325-334
static inline int64_t calculate_max_fill(int64_t hash_size) {
    int64_t mf = (int64_t)std::ceil((double)((float)hash_size * 0.75f));
    if (mf == hash_size) mf--;
    return mf;
}

// java.lang.Double.doubleToLongBits: canonical NaN
static inline int64_t double_to_long_bits(double d) {
    if (d != d) return 0x7ff8000000000000LL;
    int64_t b; memcpy(&b, &d, 8); return b;
}
static inline double bits_to_double(int64_t b) { double d; memcpy(&d, &b, 8); return d; }
// java.lang.Double.compare
static inline int double_compare(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    int64_t x = double_to_long_bits(a), y = double_to_long_bits(b);
    return x == y ? 0 : (x < y ? -1 : 1);
}

// ---------------------------------------------------------------------------------
// Sign-magnitude 128-bit decimal (SPI/type/UnscaledDecimal128Arithmetic.java)
// lo = low 64 bits of magnitude, hi = high 63 bits | sign in bit 63.
// ---------------------------------------------------------------------------------
struct Dec128 { uint64_t lo; uint64_t hi; };
static const uint64_t SIGN_MASK = 0x8000000000000000ULL;

// :165-176 unscaledDecimal(long)
static inline Dec128 dec_from_long(int64_t v) {
    Dec128 d;
    if (v < 0) { d.lo = (uint64_t)0 - (uint64_t)v; d.hi = SIGN_MASK; }
    else { d.lo = (uint64_t)v; d.hi = 0; }
    return d;
}
static inline bool dec_is_negative(const Dec128& d) { return (d.hi & SIGN_MASK) != 0; }
static inline uint64_t dec_hi_mag(const Dec128& d) { return d.hi & ~SIGN_MASK; }
static inline Dec128 dec_pack(uint64_t lo, uint64_t hi, bool neg) {
    Dec128 d; d.lo = lo; d.hi = hi | (neg ? SIGN_MASK : 0); return d;
}
// :904-920 compareAbsolute (on magnitudes)
static inline int dec_compare_abs(const Dec128& a, const Dec128& b) {
    uint64_t ah = dec_hi_mag(a), bh = dec_hi_mag(b);
    if (ah != bh) return ah < bh ? -1 : 1;
    if (a.lo != b.lo) return a.lo < b.lo ? -1 : 1;
    return 0;
}
// :354-379 addWithOverflow + :420-450 helpers.  Returns overflow in {-1,0,1}.
static inline int64_t dec_add_with_overflow(const Dec128& l, const Dec128& r, Dec128& out) {
    bool ln = dec_is_negative(l), rn = dec_is_negative(r);
    int64_t overflow = 0;
    if (ln == rn) {
        uint64_t l1 = dec_hi_mag(l), r1 = dec_hi_mag(r);
        uint64_t z0 = l.lo + r.lo;
        uint64_t carry = z0 < l.lo ? 1 : 0;
        uint64_t inter = l1 + r1 + carry;
        uint64_t z1 = inter & ~SIGN_MASK;
        out = dec_pack(z0, z1, ln);
        overflow = (int64_t)(inter >> 63);
        if (ln) overflow = -overflow;
    } else {
        int c = dec_compare_abs(l, r);
        if (c == 0) { out.lo = 0; out.hi = 0; }
        else {
            const Dec128& a = c > 0 ? l : r;
            const Dec128& b = c > 0 ? r : l;
            bool neg = c > 0 ? ln : !ln;
            uint64_t z0 = a.lo - b.lo;
            uint64_t borrow = a.lo < z0 ? 1 : 0;
            uint64_t z1 = dec_hi_mag(a) - dec_hi_mag(b) - borrow;
            out = dec_pack(z0, z1, neg);
        }
    }
    return overflow;
}
// :1002-1022 exceedsOrEqualTenToThirtyEight:  10^38 = 0x4B3B4CA85A86C47A_098A224000000000
static inline bool dec_exceeds_1e38(const Dec128& d) {
    uint64_t hi = dec_hi_mag(d);
    const uint64_t H = 0x4B3B4CA85A86C47AULL, L = 0x098A224000000000ULL;
    if (hi != H) return hi > H;
    return d.lo >= L;
}

typedef __int128 i128;
typedef unsigned __int128 u128;

}  // namespace orc
