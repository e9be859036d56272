// accum_dev.cuh — per-row value extraction and state update for the grouped accumulators.
// Reference arithmetic: LongSumAggregation.java:34-39, DoubleSumAggregation.java:33-38, AverageAggregations.java:34-39,
// DecimalSumAggregation.java:141-157 (-> UnscaledDecimal128Arithmetic.addWithOverflow :354-379),
// DecimalAverageAggregation.java:152-170, AbstractMinMaxAggregationFunction.java:274-306, util/MinMaxCompare.java.
#pragma once
#include "accum.h"

namespace pg {

static const unsigned long long SIGN64 = 0x8000000000000000ULL;

// One row's contribution, normalised so that every op class sees the same shape.
struct Contribution {
    unsigned long long cnt;         // count increment
    unsigned long long a0, a1, a2;  // integer addend, 192-bit two's complement (sums) / keys (min, max)
    double d;                       // double addend
};

PG_D void negate128(unsigned long long& lo, unsigned long long& hi) {
    lo = ~lo; hi = ~hi;
    lo += 1; if (lo == 0) hi += 1;
}

// sign-magnitude (lo, hi|sign) -> 192-bit two's complement
PG_D void signmag_to_twos(unsigned long long lo, unsigned long long hi, unsigned long long& a0, unsigned long long& a1,
                          unsigned long long& a2) {
    bool neg = (hi & SIGN64) != 0;
    hi &= ~SIGN64;
    a0 = lo; a1 = hi; a2 = 0;
    if (neg && (lo | hi)) { negate128(a0, a1); a2 = ~0ULL; }
}

PG_D long long max_double_key(long long bits) {   // maxDouble: NaN loses against every number (MinMaxCompare.java:100-106)
    return bits_is_nan(bits) ? (long long)0x8000000000000000LL : double_order_key(bits);
}

// raw input row (addInput).  Returns false when the row does not contribute (masked out / null argument).
PG_D bool extract_input(const AggOp& op, int64_t row, Contribution& c) {
    if (op.mask.values != nullptr) {
        if (dcol_is_null(op.mask, row) || ((const uint8_t*)op.mask.values)[row] == 0) return false;
    }
    if (op.skip_null && dcol_is_null(op.arg, row)) return false;
    c.cnt = 1;
    switch (op.op) {
        case OP_COUNT: break;
        case OP_SUM_I64: {
            long long v = dcol_long(op.arg, row);
            c.a0 = (unsigned long long)v; c.a1 = (unsigned long long)(v >> 63);
            break;
        }
        case OP_SUM_F64: {
            long long v = dcol_long(op.arg, row);
            c.d = op.convert_i2d ? (double)v : __longlong_as_double(v);
            break;
        }
        case OP_SUM_DEC:
            if (op.in_type == PG_LONG_DECIMAL) {
                const unsigned long long* p = (const unsigned long long*)op.arg.values + 2 * row;
                signmag_to_twos(p[0], p[1], c.a0, c.a1, c.a2);
            } else {
                long long v = dcol_long(op.arg, row);
                c.a0 = (unsigned long long)v; c.a1 = c.a2 = (unsigned long long)(v >> 63);
            }
            break;
        case OP_MIN64: {
            long long v = dcol_long(op.arg, row);
            c.a0 = (unsigned long long)(op.in_type == PG_DOUBLE ? double_order_key(v) : v);
            break;
        }
        case OP_MAX64: {
            long long v = dcol_long(op.arg, row);
            c.a0 = (unsigned long long)(op.in_type == PG_DOUBLE ? max_double_key(v) : v);
            break;
        }
        case OP_MIN128: case OP_MAX128: {
            const unsigned long long* p = (const unsigned long long*)op.arg.values + 2 * row;
            unsigned long long lo, hi, ext;
            signmag_to_twos(p[0], p[1], lo, hi, ext);
            c.a1 = hi;            // compared as signed
            c.a0 = lo ^ SIGN64;   // compared as signed after flipping the top bit
            break;
        }
    }
    return true;
}

// ---- global state update (atomics)
PG_D void atomic_add_chain2(unsigned long long* w, unsigned long long a0, unsigned long long a1) {
    unsigned long long old0 = atomicAdd(&w[0], a0);
    unsigned long long add1 = a1 + ((old0 + a0) < a0 ? 1ULL : 0ULL);
    if (add1) atomicAdd(&w[1], add1);
}
PG_D void atomic_add_chain3(unsigned long long* w, unsigned long long a0, unsigned long long a1, unsigned long long a2) {
    unsigned long long old0 = atomicAdd(&w[0], a0);
    unsigned long long c0 = (old0 + a0) < a0 ? 1ULL : 0ULL;
    unsigned long long add1 = a1 + c0;
    unsigned long long c1 = add1 < c0 ? 1ULL : 0ULL;       // a1 == ~0 and c0 == 1
    if (add1) {
        unsigned long long old1 = atomicAdd(&w[1], add1);
        c1 += (old1 + add1) < add1 ? 1ULL : 0ULL;
    }
    unsigned long long add2 = a2 + c1;
    if (add2) atomicAdd(&w[2], add2);
}

PG_D void update_global(const AggOp& op, int64_t g, const Contribution& c) {
    unsigned long long* st = op.state + g * op.words;
    switch (op.op) {
        case OP_COUNT: atomicAdd(&st[0], c.cnt); break;
        case OP_SUM_I64: atomicAdd(&st[0], c.cnt); atomic_add_chain2(st + 1, c.a0, c.a1); break;
        case OP_SUM_F64: atomicAdd(&st[0], c.cnt); atomicAdd((double*)&st[1], c.d); break;
        case OP_SUM_DEC: atomicAdd(&st[0], c.cnt); atomic_add_chain3(st + 1, c.a0, c.a1, c.a2); break;
        case OP_MIN64: st[0] = 1; atomicMin((long long*)&st[1], (long long)c.a0); break;
        case OP_MAX64: st[0] = 1; atomicMax((long long*)&st[1], (long long)c.a0); break;
        case OP_MIN128: st[0] = 1; atomicMin((long long*)&st[1], (long long)c.a1); break;   // pass 1: high word only
        case OP_MAX128: st[0] = 1; atomicMax((long long*)&st[1], (long long)c.a1); break;
    }
}

// ---- block-private state update (plain arithmetic on a thread's own words; stride = distance between words)
PG_D void update_private(const AggOp& op, unsigned long long* t, int stride, const Contribution& c) {
    switch (op.op) {
        case OP_COUNT: t[0] += c.cnt; break;
        case OP_SUM_I64: {
            t[0] += c.cnt;
            unsigned long long s0 = t[stride] + c.a0;
            t[stride] = s0;
            t[2 * stride] += c.a1 + (s0 < c.a0 ? 1ULL : 0ULL);
            break;
        }
        case OP_SUM_F64: {
            t[0] += c.cnt;
            t[stride] = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)t[stride]) + c.d);
            break;
        }
        case OP_SUM_DEC: {
            t[0] += c.cnt;
            unsigned long long s0 = t[stride] + c.a0;
            t[stride] = s0;
            unsigned long long c0 = s0 < c.a0 ? 1ULL : 0ULL;
            unsigned long long add1 = c.a1 + c0;
            unsigned long long c1 = add1 < c0 ? 1ULL : 0ULL;
            unsigned long long s1 = t[2 * stride] + add1;
            t[2 * stride] = s1;
            c1 += s1 < add1 ? 1ULL : 0ULL;
            t[3 * stride] += c.a2 + c1;
            break;
        }
        case OP_MIN64: t[0] = 1; if ((long long)c.a0 < (long long)t[stride]) t[stride] = c.a0; break;
        case OP_MAX64: t[0] = 1; if ((long long)c.a0 > (long long)t[stride]) t[stride] = c.a0; break;
        default: break;   // 128-bit min/max never run in private mode
    }
}

PG_D unsigned long long identity_word(int op, int w) {
    if (w == 0) return 0;
    switch (op) {
        case OP_MIN64: return 0x7fffffffffffffffULL;
        case OP_MAX64: return 0x8000000000000000ULL;
        case OP_MIN128: return 0x7fffffffffffffffULL;
        case OP_MAX128: return 0x8000000000000000ULL;
        default: return 0;
    }
}

}  // namespace pg
