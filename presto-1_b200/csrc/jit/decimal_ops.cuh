// decimal_ops.cuh — sign-magnitude 128-bit decimal arithmetic for filter / projection expressions, shared by the interpreting
// kernels (project.cu) and, pasted in front of the NVRTC source, by the run-time specialised kernels.  Builtin types only.
// Semantics follow SPI/type/UnscaledDecimal128Arithmetic.java: multiply (:452-560; the product's sign is leftNegative !=
// rightNegative even when it is zero, more than 127 bits overflow), add / subtract (:333-420; equal magnitudes of opposite
// sign give +0), throwIfOverflows (>= 10^38, :1002-1022).
#ifndef PGX_DECIMAL_OPS_CUH
#define PGX_DECIMAL_OPS_CUH

typedef unsigned long long pgx_u64;
typedef long long pgx_i64;

struct PgxMag { pgx_u64 lo, hi; int neg, ovf; };

__device__ __forceinline__ PgxMag pgx_mag_short(pgx_i64 v) {
    PgxMag m; m.neg = v < 0; m.lo = m.neg ? 0ULL - (pgx_u64)v : (pgx_u64)v; m.hi = 0; m.ovf = 0; return m;
}
__device__ __forceinline__ PgxMag pgx_mag_long(pgx_u64 lo, pgx_u64 hi) {
    PgxMag m; m.neg = (int)(hi >> 63); m.lo = lo; m.hi = hi & 0x7fffffffffffffffULL; m.ovf = 0; return m;
}
__device__ __forceinline__ PgxMag pgx_mag_const(pgx_u64 lo, pgx_u64 hi) { PgxMag m; m.neg = 0; m.lo = lo; m.hi = hi; m.ovf = 0; return m; }
// magnitude x 64-bit magnitude (the common case: one operand is a short decimal or a power of ten below 2^64)
__device__ __forceinline__ PgxMag pgx_mul64(PgxMag a, pgx_u64 b, int bneg) {
    PgxMag r; r.neg = a.neg != bneg; r.ovf = a.ovf;
    r.lo = a.lo * b;
    const pgx_u64 c = __umul64hi(a.lo, b);
    const pgx_u64 h = a.hi * b;
    r.hi = h + c;
    if (__umul64hi(a.hi, b) != 0 || r.hi < h || (r.hi >> 63)) r.ovf = 1;
    return r;
}
__device__ __forceinline__ PgxMag pgx_mul(PgxMag a, PgxMag b) {
    if (b.hi == 0) { PgxMag r = pgx_mul64(a, b.lo, b.neg); r.ovf |= b.ovf; return r; }
    if (a.hi == 0) { PgxMag r = pgx_mul64(b, a.lo, a.neg); r.ovf |= a.ovf; return r; }
    PgxMag r; r.neg = a.neg != b.neg; r.ovf = 1; r.lo = a.lo * b.lo; r.hi = 0; return r;   // both above 2^64: more than 127 bits
}
// a + b on signed magnitudes (b already negated by the caller for a subtraction)
__device__ __forceinline__ PgxMag pgx_add(PgxMag a, PgxMag b) {
    PgxMag r; r.ovf = a.ovf | b.ovf;
    if (a.neg == b.neg) {
        r.neg = a.neg; r.lo = a.lo + b.lo;
        const pgx_u64 c = r.lo < a.lo ? 1ULL : 0ULL;
        r.hi = a.hi + b.hi + c;
        if (r.hi >> 63) r.ovf = 1;
        return r;
    }
    const bool a_bigger = a.hi != b.hi ? a.hi > b.hi : a.lo > b.lo;
    if (a.hi == b.hi && a.lo == b.lo) { r.neg = 0; r.lo = 0; r.hi = 0; return r; }
    const PgxMag& x = a_bigger ? a : b;
    const PgxMag& y = a_bigger ? b : a;
    r.neg = x.neg; r.lo = x.lo - y.lo;
    r.hi = x.hi - y.hi - (x.lo < y.lo ? 1ULL : 0ULL);
    return r;
}
__device__ __forceinline__ int pgx_ge_1e38(PgxMag m) {   // 10^38 = 0x4B3B4CA85A86C47A_098A224000000000
    if (m.hi != 0x4B3B4CA85A86C47AULL) return m.hi > 0x4B3B4CA85A86C47AULL;
    return m.lo >= 0x098A224000000000ULL;
}
__device__ __forceinline__ int pgx_overflows(PgxMag m) { return m.ovf | pgx_ge_1e38(m); }
__device__ __forceinline__ pgx_u64 pgx_long_hi(PgxMag m) { return m.hi | ((pgx_u64)m.neg << 63); }
__device__ __forceinline__ pgx_i64 pgx_short_value(PgxMag m) { return m.neg ? (pgx_i64)(0ULL - m.lo) : (pgx_i64)m.lo; }
// -1 / 0 / +1 by value (+0 == -0)
__device__ __forceinline__ int pgx_compare(PgxMag a, PgxMag b) {
    if ((a.lo | a.hi) == 0 && (b.lo | b.hi) == 0) return 0;
    if (a.neg != b.neg) return a.neg ? -1 : 1;
    int c = a.hi != b.hi ? (a.hi < b.hi ? -1 : 1) : (a.lo != b.lo ? (a.lo < b.lo ? -1 : 1) : 0);
    return a.neg ? -c : c;
}


// ---- two's complement form, used inside the specialised aggregation kernels where a derived long decimal only ever feeds a sum
// or a comparison (the sign of a zero is not observable there).  A value is (w0, w1) = signed 128 bits; "err" collects overflow.
struct Pgx2c { pgx_u64 w0; pgx_i64 w1; };

__device__ __forceinline__ Pgx2c pgx2c_from_short(pgx_i64 v) { Pgx2c r; r.w0 = (pgx_u64)v; r.w1 = v >> 63; return r; }
__device__ __forceinline__ Pgx2c pgx2c_from_long(pgx_u64 lo, pgx_u64 hi) {   // sign-magnitude (Int128ArrayBlock) -> two's complement
    Pgx2c r; r.w0 = lo; r.w1 = (pgx_i64)(hi & 0x7fffffffffffffffULL);
    if ((pgx_i64)hi < 0) { r.w0 = 0ULL - lo; r.w1 = (pgx_i64)(~(pgx_u64)r.w1 + (lo == 0 ? 1ULL : 0ULL)); }
    return r;
}
// short x short: cannot overflow (|a|, |b| < 2^63 gives |product| < 2^126 < 10^38).  Branch-free: one signed 64 x 64 -> 128 product.
__device__ __forceinline__ Pgx2c pgx2c_mul_ss(pgx_i64 a, pgx_i64 b) {
    Pgx2c r;
    r.w0 = (pgx_u64)a * (pgx_u64)b; r.w1 = __mul64hi(a, b);
    return r;
}
// |value| >= 10^38 (or beyond 127 bits, flagged by the caller): the reference's throwIfOverflows
__device__ __forceinline__ int pgx2c_ge_1e38(Pgx2c a) {
    const pgx_u64 h = (pgx_u64)(a.w1 ^ (a.w1 >> 63));                  // |hi| for non-negatives, |hi| - 1 .. |hi| for negatives
    if (h < 0x4B3B4CA85A86C479ULL) return 0;                           // the common case: one compare
    pgx_u64 m0 = a.w0, m1 = (pgx_u64)a.w1;
    if (a.w1 < 0) { m0 = 0ULL - m0; m1 = ~m1 + (m0 == 0 ? 1ULL : 0ULL); }
    if (m1 != 0x4B3B4CA85A86C47AULL) return m1 > 0x4B3B4CA85A86C47AULL;
    return m0 >= 0x098A224000000000ULL;
}
// long x short
__device__ __forceinline__ Pgx2c pgx2c_mul_ls(Pgx2c a, pgx_i64 b, int& err) {
    if (a.w1 == ((pgx_i64)a.w0 >> 63)) {                               // the long operand fits 64 bits: a short x short product
        Pgx2c r = pgx2c_mul_ss((pgx_i64)a.w0, b);
        err |= pgx2c_ge_1e38(r);
        return r;
    }
    // general case on magnitudes
    PgxMag m; m.neg = a.w1 < 0; m.ovf = 0; m.lo = a.w0; m.hi = (pgx_u64)a.w1;
    if (m.neg) { m.lo = 0ULL - m.lo; m.hi = ~m.hi + (m.lo == 0 ? 1ULL : 0ULL); }
    const PgxMag bm = pgx_mag_short(b);
    const PgxMag z = pgx_mul64(m, bm.lo, bm.neg);
    err |= pgx_overflows(z);
    Pgx2c r; r.w0 = z.lo; r.w1 = (pgx_i64)z.hi;
    if (z.neg) { r.w0 = 0ULL - z.lo; r.w1 = (pgx_i64)(~z.hi + (z.lo == 0 ? 1ULL : 0ULL)); }
    return r;
}
__device__ __forceinline__ Pgx2c pgx2c_add(Pgx2c a, Pgx2c b, int& err) {
    Pgx2c r; r.w0 = a.w0 + b.w0;
    r.w1 = (pgx_i64)((pgx_u64)a.w1 + (pgx_u64)b.w1 + (r.w0 < a.w0 ? 1ULL : 0ULL));
    if (((a.w1 ^ r.w1) & (b.w1 ^ r.w1)) < 0) err |= 1;                // signed 128-bit wrap
    err |= pgx2c_ge_1e38(r);
    return r;
}
__device__ __forceinline__ Pgx2c pgx2c_neg(Pgx2c a) { Pgx2c r; r.w0 = 0ULL - a.w0; r.w1 = (pgx_i64)(~(pgx_u64)a.w1 + (a.w0 == 0 ? 1ULL : 0ULL)); return r; }
__device__ __forceinline__ int pgx2c_compare(Pgx2c a, Pgx2c b) {
    if (a.w1 != b.w1) return a.w1 < b.w1 ? -1 : 1;
    return a.w0 != b.w0 ? (a.w0 < b.w0 ? -1 : 1) : 0;
}

#endif
