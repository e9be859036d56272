// accum_eval.cu — addIntermediate (combine) and evaluateFinal / evaluateIntermediate kernels.
// Intermediate layouts follow the reference serializers (see include/prestogpu.h):
//   state/LongDecimalWithOverflowStateSerializer.java:38-66       [overflow][lo][hi]           24 B VARBINARY
//   state/LongDecimalWithOverflowAndLongStateSerializer.java:38-68 [count][overflow][lo][hi]    32 B VARBINARY
//   StateCompiler.java:207-217 (fields sorted by name) for LongLongState / LongAndDoubleState / LongDoubleState
#include <vector>

#include "accum_dev.cuh"
#include "scan.cuh"

namespace pg {

typedef unsigned __int128 u128;

// ------------------------------------------------------------------ addIntermediate
PG_D unsigned long long load_u64_unaligned(const uint8_t* p) { return rd64u(p); }

PG_D void intermediate_rows(const AggOp& op, const DCol& s0, const DCol& s1, int kind, const int32_t* __restrict__ gids, int64_t n) {
    int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; row < n; row += stride) {
        int64_t g = gids[row];
        Contribution c;
        c.cnt = 0; c.a0 = c.a1 = c.a2 = 0; c.d = 0;
        switch (op.op) {
            case OP_COUNT:
                if (dcol_is_null(s0, row)) continue;
                c.cnt = (unsigned long long)dcol_long(s0, row);
                break;
            case OP_SUM_I64: {   // [count BIGINT, sum BIGINT]
                c.cnt = (unsigned long long)dcol_long(s0, row);
                long long v = dcol_long(s1, row);
                c.a0 = (unsigned long long)v; c.a1 = (unsigned long long)(v >> 63);
                break;
            }
            case OP_SUM_F64:
                if (kind == PG_AGG_SUM) {   // [count BIGINT, sum DOUBLE]
                    c.cnt = (unsigned long long)dcol_long(s0, row); c.d = __longlong_as_double(dcol_long(s1, row));
                } else {                    // avg: [sum DOUBLE, count BIGINT]
                    c.d = __longlong_as_double(dcol_long(s0, row)); c.cnt = (unsigned long long)dcol_long(s1, row);
                }
                break;
            case OP_SUM_DEC: {
                if (dcol_is_null(s0, row)) continue;
                const uint8_t* p = (const uint8_t*)s0.values + s0.offsets[row];
                if (kind == PG_AGG_AVG) { c.cnt = load_u64_unaligned(p); p += 8; } else c.cnt = 1;
                long long ovf = (long long)load_u64_unaligned(p);
                unsigned long long lo = load_u64_unaligned(p + 8), hi = load_u64_unaligned(p + 16);
                signmag_to_twos(lo, hi, c.a0, c.a1, c.a2);
                // + overflow * 2^127
                unsigned long long o1 = (unsigned long long)ovf << 63, o2 = (unsigned long long)(ovf >> 1);
                unsigned long long n1 = c.a1 + o1;
                c.a2 += o2 + (n1 < o1 ? 1ULL : 0ULL);
                c.a1 = n1;
                break;
            }
            default: {   // min / max: nullable value of the input type
                AggOp tmp = op; tmp.arg = s0; tmp.mask.values = nullptr; tmp.skip_null = 1;
                if (!extract_input(tmp, row, c)) continue;
            }
        }
        update_global(op, g, c);
    }
}
// every aggregate of the operator in one launch: blockIdx.y selects the aggregate
struct IntermBatch {
    AggOp op[MAX_AGGS];
    DCol s0[MAX_AGGS], s1[MAX_AGGS];
    int32_t kind[MAX_AGGS];
};
__global__ void acc_intermediate_kernel(const __grid_constant__ IntermBatch b, const int32_t* __restrict__ gids, int64_t n) {
    const int a = blockIdx.y;
    intermediate_rows(b.op[a], b.s0[a], b.s1[a], b.kind[a], gids, n);
}

// ------------------------------------------------------------------ evaluate
struct DecState {
    long long overflow;
    unsigned long long lo, hi;   // sign-magnitude
};
// 192-bit two's complement -> the reference's (overflow, sign-magnitude) representation of the same value
PG_D DecState twos_to_state(unsigned long long w0, unsigned long long w1, unsigned long long w2) {
    bool neg = (w2 >> 63) != 0;
    if (neg) {
        w0 = ~w0; w1 = ~w1; w2 = ~w2;
        if (++w0 == 0) if (++w1 == 0) ++w2;
    }
    unsigned long long ovf_mag = (w2 << 1) | (w1 >> 63);
    DecState s;
    s.overflow = neg ? -(long long)ovf_mag : (long long)ovf_mag;
    s.lo = w0;
    s.hi = (w1 & ~SIGN64) | (neg ? SIGN64 : 0);
    return s;
}
PG_D bool exceeds_1e38(unsigned long long lo, unsigned long long hi) {   // UnscaledDecimal128Arithmetic.java:1002-1022
    hi &= ~SIGN64;
    const unsigned long long H = 0x4B3B4CA85A86C47AULL, L = 0x098A224000000000ULL;
    if (hi != H) return hi > H;
    return lo >= L;
}
// |T| / count, ROUND_HALF_UP (DecimalAverageAggregation.java:209-220).  fits=false when the quotient needs > 127 bits.
PG_D void avg_192(unsigned long long w0, unsigned long long w1, unsigned long long w2, unsigned long long count, unsigned long long& qlo,
                  unsigned long long& qhi, bool& fits) {
    bool neg = (w2 >> 63) != 0;
    if (neg) {
        w0 = ~w0; w1 = ~w1; w2 = ~w2;
        if (++w0 == 0) if (++w1 == 0) ++w2;
    }
    unsigned long long w[3] = {w2, w1, w0}, q[3];
    u128 rem = 0;
    for (int k = 0; k < 3; k++) {
        u128 cur = (rem << 64) | w[k];
        q[k] = (unsigned long long)(cur / count);
        rem = cur % count;
    }
    if (rem * 2 >= (u128)count) { if (++q[2] == 0) if (++q[1] == 0) ++q[0]; }
    fits = q[0] == 0 && (q[1] >> 63) == 0;
    bool zero = q[1] == 0 && q[2] == 0;
    qlo = q[2];
    qhi = (q[1] & ~SIGN64) | ((neg && !zero) ? SIGN64 : 0);
}

struct EvalArgs {
    const unsigned long long* state;
    int32_t op, kind, in_type, words, final_step;
    int64_t first, n;
    void* out0; uint8_t* nulls0;
    void* out1; uint8_t* nulls1;
    long long* error_group;   // device, min over failing groups
};

PG_D void put64(void* out, uint8_t* nulls, int64_t i, long long v, bool isnull) {
    ((long long*)out)[i] = isnull ? 0 : v;
    if (nulls) nulls[i] = isnull ? 1 : 0;
}

PG_D void evaluate_rows(const EvalArgs& a) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < a.n; i += stride) {
        int64_t g = a.first + i;
        const unsigned long long* st = a.state + g * a.words;
        unsigned long long cnt = st[0];
        bool err = false;
        switch (a.op) {
            case OP_COUNT: put64(a.out0, a.nulls0, i, (long long)cnt, false); break;
            case OP_SUM_I64: {
                long long lo = (long long)st[1], hi = (long long)st[2];
                if (hi != (lo >> 63)) err = true;   // BigintOperators.add -> NUMERIC_VALUE_OUT_OF_RANGE
                if (a.final_step) put64(a.out0, a.nulls0, i, lo, cnt == 0);
                else { put64(a.out0, a.nulls0, i, (long long)cnt, false); put64(a.out1, a.nulls1, i, lo, false); }
                break;
            }
            case OP_SUM_F64: {
                double s = __longlong_as_double((long long)st[1]);
                long long sbits = canon_double_bits((long long)st[1]);
                if (a.kind == PG_AGG_SUM) {
                    if (a.final_step) put64(a.out0, a.nulls0, i, sbits, cnt == 0);
                    else { put64(a.out0, a.nulls0, i, (long long)cnt, false); put64(a.out1, a.nulls1, i, sbits, false); }
                } else {
                    if (a.final_step) put64(a.out0, a.nulls0, i, canon_double_bits(__double_as_longlong(s / (double)(long long)cnt)), cnt == 0);
                    else { put64(a.out0, a.nulls0, i, sbits, false); put64(a.out1, a.nulls1, i, (long long)cnt, false); }
                }
                break;
            }
            case OP_SUM_DEC: {
                // intermediate output is written by acc_evaluate_dec_inter_kernel
                if (a.kind == PG_AGG_SUM) {   // DecimalSumAggregation.outputLongDecimal :178-190
                    DecState s = twos_to_state(st[1], st[2], st[3]);
                    bool isnull = cnt == 0;
                    if (!isnull && (s.overflow != 0 || exceeds_1e38(s.lo, s.hi))) err = true;
                    ((unsigned long long*)a.out0)[2 * i] = isnull ? 0 : s.lo;
                    ((unsigned long long*)a.out0)[2 * i + 1] = isnull ? 0 : s.hi;
                    if (a.nulls0) a.nulls0[i] = isnull ? 1 : 0;
                } else {                      // DecimalAverageAggregation.output* :188-220
                    bool isnull = cnt == 0, fits = true;
                    unsigned long long qlo = 0, qhi = 0;
                    if (!isnull) avg_192(st[1], st[2], st[3], cnt, qlo, qhi, fits);
                    if (a.in_type == PG_SHORT_DECIMAL) {
                        bool neg = (qhi & SIGN64) != 0;
                        bool ok = fits && (qhi & ~SIGN64) == 0 && (qlo <= 0x7fffffffffffffffULL || (neg && qlo == SIGN64));
                        if (!ok) err = true;
                        long long v = neg ? (long long)(0 - qlo) : (long long)qlo;
                        put64(a.out0, a.nulls0, i, v, isnull);
                    } else {
                        if (!fits || exceeds_1e38(qlo, qhi)) err = true;
                        ((unsigned long long*)a.out0)[2 * i] = qlo; ((unsigned long long*)a.out0)[2 * i + 1] = qhi;
                        if (a.nulls0) a.nulls0[i] = isnull ? 1 : 0;
                    }
                }
                break;
            }
            case OP_MIN64: case OP_MAX64: {
                bool isnull = cnt == 0;
                long long k = (long long)st[1];
                if (a.in_type == PG_DOUBLE) {
                    long long bits = (a.op == OP_MAX64 && k == (long long)0x8000000000000000LL) ? 0x7ff8000000000000LL : double_order_key_inv(k);
                    put64(a.out0, a.nulls0, i, bits, isnull);
                } else if (a.in_type == PG_INTEGER || a.in_type == PG_DATE) {
                    ((int32_t*)a.out0)[i] = isnull ? 0 : (int32_t)k;
                    if (a.nulls0) a.nulls0[i] = isnull ? 1 : 0;
                } else if (a.in_type == PG_BOOLEAN) {
                    ((uint8_t*)a.out0)[i] = isnull ? 0 : (uint8_t)k;
                    if (a.nulls0) a.nulls0[i] = isnull ? 1 : 0;
                } else put64(a.out0, a.nulls0, i, k, isnull);
                break;
            }
            default: {   // 128-bit min/max: [cnt, hi, lo^sign] two's complement -> sign-magnitude
                bool isnull = cnt == 0;
                unsigned long long hi = st[1], lo = st[2] ^ SIGN64;
                bool neg = (hi >> 63) != 0;
                if (neg) negate128(lo, hi);
                ((unsigned long long*)a.out0)[2 * i] = isnull ? 0 : lo;
                ((unsigned long long*)a.out0)[2 * i + 1] = isnull ? 0 : (hi | (neg ? SIGN64 : 0));
                if (a.nulls0) a.nulls0[i] = isnull ? 1 : 0;
            }
        }
        if (err && cnt != 0) atomicMin(a.error_group, (long long)g);
    }
}
struct EvalBatch { EvalArgs a[MAX_AGGS]; };
__global__ void acc_evaluate_kernel(const __grid_constant__ EvalBatch b) { evaluate_rows(b.a[blockIdx.y]); }

// decimal intermediate state -> VARBINARY rows (null and zero-length when the state was never touched)
__global__ void __launch_bounds__(SCAN_THREADS) acc_dec_inter_count_kernel(const unsigned long long* state, int64_t first, int64_t n,
                                                                           int entry, Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i < n && state[(first + i) * 4] != 0) acc.a += entry;
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
PG_D void dec_inter_write_tile(const unsigned long long* state, int64_t first, int64_t n, int entry, const Pair64* tile_prefix,
                               uint8_t* out_bytes, int32_t* out_offsets, uint8_t* out_nulls, Pair64* smem) {
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run{0, 0};
    if (tile_prefix) run = tile_prefix[blockIdx.x];   // null: a single tile
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (tile_base + (int64_t)j * SCAN_THREADS >= n) break;
        Pair64 v{0, 0};
        const unsigned long long* st = nullptr;
        if (i < n) { st = state + (first + i) * 4; if (st[0] != 0) v.a = entry; }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (i < n) {
            int64_t dst = run.a + ex.a;
            out_offsets[i] = (int32_t)dst;
            if (i == n - 1) out_offsets[n] = (int32_t)(dst + v.a);
            if (out_nulls) out_nulls[i] = v.a ? 0 : 1;
            if (v.a) {
                DecState s = twos_to_state(st[1], st[2], st[3]);
                unsigned long long words[4]; int nw = 0;
                if (entry == 32) words[nw++] = st[0];
                words[nw++] = (unsigned long long)s.overflow; words[nw++] = s.lo; words[nw++] = s.hi;
                for (int w = 0; w < nw; w++)
                    for (int b = 0; b < 8; b++) out_bytes[dst + w * 8 + b] = (uint8_t)(words[w] >> (8 * b));
            }
        }
        run = run + total;
    }
}
__global__ void __launch_bounds__(SCAN_THREADS) acc_dec_inter_write_kernel(const unsigned long long* state, int64_t first, int64_t n, int entry,
                                                                           const Pair64* tile_prefix, uint8_t* out_bytes, int32_t* out_offsets,
                                                                           uint8_t* out_nulls) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    dec_inter_write_tile(state, first, n, entry, tile_prefix, out_bytes, out_offsets, out_nulls, smem);
}
// few groups: every decimal aggregate's VARBINARY column in one launch of single-tile blocks (blockIdx.y = aggregate)
struct DecInterArgs {
    const unsigned long long* state;
    uint8_t* out_bytes; int32_t* out_offsets; uint8_t* out_nulls;
    int32_t entry, pad;
};
struct DecInterBatch { DecInterArgs a[MAX_AGGS]; };
__global__ void __launch_bounds__(SCAN_THREADS) acc_dec_inter_small_kernel(const __grid_constant__ DecInterBatch b, int64_t first, int64_t n) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const DecInterArgs& a = b.a[blockIdx.y];
    dec_inter_write_tile(a.state, first, n, a.entry, nullptr, a.out_bytes, a.out_offsets, a.out_nulls, smem);
}


// ------------------------------------------------------------------ host side
static int op_for(int32_t kind, int32_t in_type) {
    bool dec = in_type == PG_SHORT_DECIMAL || in_type == PG_LONG_DECIMAL;
    bool integral = in_type == PG_BIGINT || in_type == PG_INTEGER;
    switch (kind) {
        case PG_AGG_COUNT: case PG_AGG_COUNT_COL: return OP_COUNT;
        case PG_AGG_SUM:
            if (integral) return OP_SUM_I64;
            if (in_type == PG_DOUBLE) return OP_SUM_F64;
            if (dec) return OP_SUM_DEC;
            break;
        case PG_AGG_AVG:
            if (integral || in_type == PG_DOUBLE) return OP_SUM_F64;
            if (dec) return OP_SUM_DEC;
            break;
        case PG_AGG_MIN: case PG_AGG_MAX:
            if (in_type == PG_LONG_DECIMAL) return kind == PG_AGG_MIN ? OP_MIN128 : OP_MAX128;
            if (in_type == PG_BIGINT || in_type == PG_INTEGER || in_type == PG_DATE || in_type == PG_DOUBLE ||
                in_type == PG_SHORT_DECIMAL || in_type == PG_BOOLEAN)
                return kind == PG_AGG_MIN ? OP_MIN64 : OP_MAX64;
            break;
    }
    fail(PG_E_INVALID_ARGUMENT, "unsupported aggregate kind %d over type %d", kind, in_type);
}

void agg_create(pg_ctx* ctx, int32_t kind, int32_t in_type, pg_agg** out) {
    PG_CHECK_ARG(out != nullptr, "null argument");
    PG_CHECK_ARG(kind == PG_AGG_COUNT || type_valid(in_type), "unknown input type %d", in_type);
    pg_agg* a = new pg_agg();
    try {
        a->ctx = ctx; a->kind = kind; a->in_type = kind == PG_AGG_COUNT ? 0 : in_type;
        a->op = op_for(kind, in_type);
        a->words = state_words(a->op);
    } catch (...) { delete a; throw; }
    *out = a;
}

void agg_ensure_capacity(pg_agg* agg, int64_t group_count) {   // GroupedAccumulator state.ensureCapacity(groupCount)
    if (group_count <= agg->capacity) return;
    int64_t cap = agg->capacity ? agg->capacity : 1024;
    while (cap < group_count) cap *= 2;
    agg->state.grow(agg->ctx, cap * agg->words * 8, agg->capacity * agg->words * 8);
    init_state(agg->ctx, agg->op, agg->state.as<unsigned long long>(), agg->capacity, cap - agg->capacity);
    agg->capacity = cap;
}

AggOp agg_make_op(pg_agg* agg, const DCol* arg, const DCol* mask) {
    AggOp op;
    memset(&op, 0, sizeof(op));
    op.op = agg->op; op.in_type = agg->in_type; op.words = agg->words;
    op.skip_null = agg->kind != PG_AGG_COUNT;
    op.convert_i2d = (agg->kind == PG_AGG_AVG && (agg->in_type == PG_BIGINT || agg->in_type == PG_INTEGER)) ? 1 : 0;
    if (arg) op.arg = *arg;
    if (mask) op.mask = *mask;
    op.state = agg->state.as<unsigned long long>();
    return op;
}

// addIntermediate for every aggregate of an operator: s0[i]/s1[i] are aggregate i's intermediate columns (see the layout table
// in prestogpu.h).  One launch for all of them; 128-bit min/max keep their two-pass protocol.
void accumulate_intermediate_batch(pg_ctx* ctx, pg_agg* const* aggs, const DCol* s0, const DCol* s1, int na, const int32_t* d_ids, int64_t n,
                                   int64_t group_count) {
    if (n <= 0 || na <= 0) return;
    IntermBatch b;
    memset(&b, 0, sizeof(b));
    int nb = 0;
    for (int i = 0; i < na; i++) {
        pg_agg* agg = aggs[i];
        if (agg->op == OP_MIN128 || agg->op == OP_MAX128) {   // same two-pass protocol as addInput
            AggProgram prog; memset(&prog, 0, sizeof(prog)); prog.naggs = 1;
            prog.ops[0] = agg_make_op(agg, &s0[i], nullptr);
            accumulate(ctx, prog, d_ids, n, group_count);
            continue;
        }
        b.op[nb] = agg_make_op(agg, nullptr, nullptr);
        b.s0[nb] = s0[i]; b.s1[nb] = s1[i]; b.kind[nb] = agg->kind;
        nb++;
    }
    if (nb == 0) return;
    dim3 grid((unsigned)grid_for(ctx, n, 256), (unsigned)nb);
    PG_LAUNCH(ctx, acc_intermediate_kernel, grid, 256, 0, b, d_ids, n);
}
void accumulate_intermediate(pg_agg* agg, const DCol& s0, const DCol& s1, const int32_t* d_ids, int64_t n, int64_t group_count) {
    accumulate_intermediate_batch(agg->ctx, &agg, &s0, &s1, 1, d_ids, n, group_count);
}

int agg_n_intermediate(const pg_agg* agg) {
    if (agg->op == OP_SUM_I64 || agg->op == OP_SUM_F64) return 2;
    return 1;
}

static void out_types(const pg_agg* agg, int final_step, int32_t* t) {
    t[0] = t[1] = 0;
    bool dec = agg->op == OP_SUM_DEC;
    switch (agg->kind) {
        case PG_AGG_COUNT: case PG_AGG_COUNT_COL: t[0] = PG_BIGINT; break;
        case PG_AGG_SUM:
            if (dec) t[0] = final_step ? PG_LONG_DECIMAL : PG_VARBINARY;
            else if (agg->op == OP_SUM_I64) { t[0] = PG_BIGINT; if (!final_step) t[1] = PG_BIGINT; }
            else { if (final_step) t[0] = PG_DOUBLE; else { t[0] = PG_BIGINT; t[1] = PG_DOUBLE; } }
            break;
        case PG_AGG_AVG:
            if (dec) t[0] = final_step ? agg->in_type : PG_VARBINARY;
            else { t[0] = PG_DOUBLE; if (!final_step) t[1] = PG_BIGINT; }
            break;
        default: t[0] = agg->in_type;
    }
}

void agg_evaluate(pg_agg* agg, int64_t first, int64_t n, int32_t final_step, pg_col_out* out, int64_t* error_group, bool do_sync) {
    pg_ctx* ctx = agg->ctx;
    PG_CHECK_ARG(first >= 0 && n >= 0 && out != nullptr, "bad arguments");
    if (n == 0) return;
    DevBuf err(ctx, 8);
    memset_async(ctx, err.p, 0x7f, 8);
    agg_evaluate_async(agg, first, n, final_step, out, err.as<long long>());
    if (do_sync || error_group) {
        int64_t eg;
        read_scalars(ctx, err.p, 1, &eg);
        if (eg != 0x7f7f7f7f7f7f7f7fLL) {
            if (error_group) *error_group = eg;
            fail(PG_E_NUMERIC_OVERFLOW, agg->op == OP_SUM_I64 ? "bigint addition overflow" : "Decimal overflow");
        }
    }
}

// Launches the evaluation of `na` aggregates over groups [first, first+n) and the copies into host outputs without waiting.
// out: the aggregates' output columns back to back (one per aggregate for the final step, agg_n_intermediate otherwise);
// d_errs[i] (device, preset to 0x7f..7f) receives the smallest overflowing group id of aggregate i.
// Fixed-width outputs of all aggregates go out in one launch; so do the decimal VARBINARY states when the groups fit one tile.
void agg_evaluate_batch(pg_ctx* ctx, pg_agg* const* aggs, int na, int64_t first, int64_t n, int32_t final_step, pg_col_out* out,
                        long long* d_errs) {
    PG_CHECK_ARG(first >= 0 && n >= 0 && out != nullptr && na >= 0 && na <= MAX_AGGS, "bad arguments");
    if (n == 0 || na == 0) return;
    std::vector<OutStage> st(2 * (size_t)na);
    std::vector<int64_t> vbytes(2 * (size_t)na, 0);
    std::vector<int> cols_of(na, 1);
    EvalBatch eb; DecInterBatch db;
    int ne = 0, nd = 0, col = 0;
    const int64_t ntiles = num_tiles(n);
    // Two passes: columns that need a read-back of their byte total are produced and handed over first, because a
    // synchronisation flushes and recycles the output staging arena the others are about to share.
    std::vector<char> done(na, 0);
    for (int pass = 0; pass < 2; pass++) {
        col = 0;
        for (int i = 0; i < na; col += cols_of[i], i++) {
            pg_agg* agg = aggs[i];
            int32_t t[2];
            out_types(agg, final_step, t);
            const int ncols = t[1] ? 2 : 1;
            cols_of[i] = ncols;
            pg_col_out* o = out + col;
            OutStage* s = &st[2 * (size_t)i];
            int64_t* vb = &vbytes[2 * (size_t)i];
            const int entry = agg->kind == PG_AGG_AVG ? 32 : 24;
            const bool needs_total = t[0] == PG_VARBINARY && !(ntiles == 1 && o[0].values_capacity_bytes >= (int64_t)entry * n);
            if (pass == 0) {
                agg_ensure_capacity(agg, first + n);
                for (int k = 0; k < ncols; k++)
                    PG_CHECK_ARG(o[k].type == t[k], "output column %d must have type %d, got %d", col + k, t[k], o[k].type);
            }
            if (needs_total != (pass == 0)) continue;
            const unsigned long long* state = agg->state.as<unsigned long long>();
            if (needs_total) {
                DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
                Pair64* d_total = tiles.as<Pair64>() + ntiles;
                PG_LAUNCH(ctx, acc_dec_inter_count_kernel, (int)ntiles, SCAN_THREADS, 0, state, first, n, entry, tiles.as<Pair64>());
                scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, d_total);
                int64_t tot[2];
                read_scalars(ctx, d_total, 2, tot);
                vb[0] = tot[0];
                s[0].begin(ctx, &o[0], n, vb[0]);
                PG_LAUNCH(ctx, acc_dec_inter_write_kernel, (int)ntiles, SCAN_THREADS, 0, state, first, n, entry, tiles.as<Pair64>(),
                          (uint8_t*)s[0].d_values, s[0].d_offsets, s[0].d_nulls);
                s[0].finish(vb[0]);
                done[i] = 1;
            } else if (t[0] == PG_VARBINARY) {
                vb[0] = (int64_t)entry * n;   // upper bound: no read-back of the exact total
                s[0].begin(ctx, &o[0], n, vb[0]);
                DecInterArgs& d = db.a[nd++];
                d.state = state; d.entry = entry; d.pad = 0;
                d.out_bytes = (uint8_t*)s[0].d_values; d.out_offsets = s[0].d_offsets; d.out_nulls = s[0].d_nulls;
            } else {
                for (int k = 0; k < ncols; k++) {
                    vb[k] = n * type_width(t[k]);
                    s[k].begin(ctx, &o[k], n, vb[k]);
                }
                EvalArgs& a = eb.a[ne++];
                a.state = state; a.op = agg->op; a.kind = agg->kind; a.in_type = agg->in_type; a.words = agg->words; a.final_step = final_step;
                a.first = first; a.n = n;
                a.out0 = s[0].d_values; a.nulls0 = s[0].d_nulls;
                a.out1 = ncols > 1 ? s[1].d_values : nullptr; a.nulls1 = ncols > 1 ? s[1].d_nulls : nullptr;
                a.error_group = d_errs + i;
            }
        }
    }
    if (ne > 0) {
        dim3 grid((unsigned)grid_for(ctx, n, 256), (unsigned)ne);
        PG_LAUNCH(ctx, acc_evaluate_kernel, grid, 256, 0, eb);
    }
    if (nd > 0) {
        dim3 grid(1, (unsigned)nd);
        PG_LAUNCH(ctx, acc_dec_inter_small_kernel, grid, SCAN_THREADS, 0, db, first, n);
    }
    for (int i = 0; i < na; i++) {
        if (done[i]) continue;
        for (int k = 0; k < cols_of[i]; k++) st[2 * (size_t)i + k].finish(vbytes[2 * (size_t)i + k]);
    }
}
void agg_evaluate_async(pg_agg* agg, int64_t first, int64_t n, int32_t final_step, pg_col_out* out, long long* d_err) {
    agg_evaluate_batch(agg->ctx, &agg, 1, first, n, final_step, out, d_err);
}

}  // namespace pg

using namespace pg;

extern "C" {

int32_t pg_agg_create(pg_ctx* ctx, int32_t kind, int32_t in_type, pg_agg** out) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] { agg_create(ctx, kind, in_type, out); });
}
int32_t pg_agg_destroy(pg_agg* agg) {
    if (!agg) return PG_OK;
    return guard(agg->ctx, [&] { delete agg; });
}
int32_t pg_agg_add_input(pg_agg* agg, const int32_t* ids, int32_t ids_mem, int64_t group_count, const pg_col* arg, const pg_col* mask,
                         int64_t n) {
    if (!agg) return PG_E_INVALID_ARGUMENT;
    return guard(agg->ctx, [&] {
        pg_ctx* ctx = agg->ctx;
        PG_CHECK_ARG(n >= 0 && group_count >= 0 && (ids != nullptr || n == 0), "bad arguments");
        PG_CHECK_ARG(agg->kind == PG_AGG_COUNT || arg != nullptr, "aggregate needs an argument column");
        agg_ensure_capacity(agg, group_count);
        if (n == 0) return;
        DevBuf idbuf;
        const int32_t* d_ids = ids;
        if (ids_mem != PG_MEM_DEVICE) { idbuf.alloc(ctx, n * 4); copy_async(ctx, idbuf.p, ids, n * 4, cudaMemcpyHostToDevice); d_ids = idbuf.as<int32_t>(); }
        ColImport ca, cm;
        if (arg && agg->kind != PG_AGG_COUNT) {
            PG_CHECK_ARG(arg->type == agg->in_type, "argument type %d does not match the aggregate's input type %d", arg->type, agg->in_type);
            import_col(ctx, *arg, n, ca);
        }
        if (mask) { PG_CHECK_ARG(mask->type == PG_BOOLEAN, "mask must be BOOLEAN"); import_col(ctx, *mask, n, cm); }
        AggProgram prog;
        memset(&prog, 0, sizeof(prog));
        prog.naggs = 1;
        prog.ops[0] = agg_make_op(agg, (arg && agg->kind != PG_AGG_COUNT) ? &ca.d : nullptr, mask ? &cm.d : nullptr);
        accumulate(ctx, prog, d_ids, n, group_count);
        sync(ctx);
    });
}
int32_t pg_agg_add_intermediate(pg_agg* agg, const int32_t* ids, int32_t ids_mem, int64_t group_count, const pg_col* state, int64_t n) {
    if (!agg) return PG_E_INVALID_ARGUMENT;
    return guard(agg->ctx, [&] {
        pg_ctx* ctx = agg->ctx;
        PG_CHECK_ARG(n >= 0 && group_count >= 0 && state != nullptr && (ids != nullptr || n == 0), "bad arguments");
        agg_ensure_capacity(agg, group_count);
        if (n == 0) return;
        DevBuf idbuf;
        const int32_t* d_ids = ids;
        if (ids_mem != PG_MEM_DEVICE) { idbuf.alloc(ctx, n * 4); copy_async(ctx, idbuf.p, ids, n * 4, cudaMemcpyHostToDevice); d_ids = idbuf.as<int32_t>(); }
        int nc = agg_n_intermediate(agg);
        ColImport c0, c1;
        import_col(ctx, state[0], n, c0);
        DCol d1; memset(&d1, 0, sizeof(d1));
        if (nc > 1) { import_col(ctx, state[1], n, c1); d1 = c1.d; }
        accumulate_intermediate(agg, c0.d, d1, d_ids, n, group_count);
        sync(ctx);
    });
}
int32_t pg_agg_n_intermediate(pg_agg* agg, int32_t* out) {
    if (!agg || !out) return PG_E_INVALID_ARGUMENT;
    *out = agg_n_intermediate(agg);
    return PG_OK;
}
int32_t pg_agg_evaluate(pg_agg* agg, int64_t first, int64_t n, int32_t final_step, pg_col_out* out, int64_t* error_group) {
    if (!agg) return PG_E_INVALID_ARGUMENT;
    return guard(agg->ctx, [&] {
        if (error_group) *error_group = -1;
        try { agg_evaluate(agg, first, n, final_step, out, error_group, true); } catch (...) { cudaStreamSynchronize(agg->ctx->stream); throw; }
        sync(agg->ctx);
    });
}
int32_t pg_agg_device_bytes(pg_agg* agg, int64_t* out) {
    if (!agg || !out) return PG_E_INVALID_ARGUMENT;
    *out = agg->state.bytes;
    return PG_OK;
}

}  // extern "C"
