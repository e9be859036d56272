// accum_private.cu — low-cardinality grouped accumulation (the "shared-memory pre-aggregation" path of north_star).
//
// Every thread owns a private copy of the (few) groups' partial states in shared memory, laid out
// table[(word * G + group) * T + thread] so that a warp's accesses to one word hit 32 different banks: updates are
// plain conflict-free LDS/ADD/STS, no atomics in the row loop.  Identical pieces of state are shared between
// aggregates (sum(x) and avg(x) share one sum, all aggregates under the same mask share one count), and partial sums use
// the narrowest exact width (128 bit for 64-bit inputs, 192 bit for 128-bit inputs).  At the end each warp tree-reduces
// its 32 columns with shuffles (sum + carry count per word) and lane 0 folds the result into the global states with the
// carry-chain atomics of accum_dev.cuh.
#include "accum_dev.cuh"

namespace pg {

enum { PS_CNT = 0, PS_ADD2 = 1, PS_ADD3 = 2, PS_F64 = 3, PS_MIN = 4, PS_MAX = 5 };
enum { CV_INT = 0, CV_I2D = 1, CV_F64 = 2, CV_DEC128 = 3, CV_MINKEY = 4, CV_MAXKEY = 5 };
static const int MAX_PCOLS = 20;
static const int MAX_PCONDS = 8;
static const int MAX_PSLOTS = 12;

struct PrivSlot {
    int32_t kind, conv, word_off, nwords, src, cond;
};
struct PrivCond {
    int32_t mask_col, null_col;   // indices into cols[] or -1
};
struct PrivAgg {
    int32_t op, words, cnt_slot, val_slot;
    unsigned long long* state;
};
struct PrivPlan {
    int32_t ncols, nconds, nslots, words, naggs, pad;
    DCol cols[MAX_PCOLS];
    PrivCond conds[MAX_PCONDS];
    PrivSlot slots[MAX_PSLOTS];
    PrivAgg aggs[MAX_AGGS];
};

PG_D void wreduce_sum_carry(unsigned long long& s, unsigned long long& carry) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long os = __shfl_down_sync(0xffffffffu, s, d);
        unsigned long long oc = __shfl_down_sync(0xffffffffu, carry, d);
        unsigned long long ns = s + os;
        carry += oc + (ns < s ? 1ULL : 0ULL);
        s = ns;
    }
}

template <int T>
__global__ void __launch_bounds__(T) acc_private2_kernel(const PrivPlan plan, const int32_t* __restrict__ gids, int64_t n, int G) {
    extern __shared__ unsigned long long table[];
    const int tid = threadIdx.x;
    const int GT = G * T;
    // ---- init own columns
    for (int s = 0; s < plan.nslots; s++) {
        const PrivSlot& ps = plan.slots[s];
        unsigned long long idv = ps.kind == PS_MIN ? 0x7fffffffffffffffULL : (ps.kind == PS_MAX ? 0x8000000000000000ULL : 0ULL);
        for (int w = 0; w < ps.nwords; w++)
            for (int g = 0; g < G; g++) table[(ps.word_off + w) * GT + g * T + tid] = idv;
    }
    // ---- row loop
    const int64_t stride = (int64_t)gridDim.x * T;
    for (int64_t row = (int64_t)blockIdx.x * T + tid; row < n; row += stride) {
        const int g = gids[row];
        unsigned condbits = 0;
#pragma unroll
        for (int c = 0; c < MAX_PCONDS; c++) {
            if (c < plan.nconds) {
                const PrivCond& pc = plan.conds[c];
                bool ok = true;
                if (pc.mask_col >= 0) {
                    const DCol& m = plan.cols[pc.mask_col];
                    ok = !dcol_is_null(m, row) && ((const uint8_t*)m.values)[row] != 0;
                }
                if (ok && pc.null_col >= 0) ok = plan.cols[pc.null_col].nulls[row] == 0;
                condbits |= (ok ? 1u : 0u) << c;
            }
        }
        // load phase: every distinct argument is read once; the loads are independent of the shared-memory updates below
        unsigned long long v0[MAX_PSLOTS], v1[MAX_PSLOTS];
#pragma unroll
        for (int s = 0; s < MAX_PSLOTS; s++) {
            v0[s] = 0; v1[s] = 0;
            if (s < plan.nslots) {
                const PrivSlot& ps = plan.slots[s];
                if (ps.kind != PS_CNT && ((condbits >> ps.cond) & 1)) {
                    const DCol& c = plan.cols[ps.src];
                    if (ps.conv == CV_DEC128) {
                        const ulonglong2 p = ((const ulonglong2*)c.values)[row];
                        v0[s] = p.x; v1[s] = p.y;
                    } else {
                        v0[s] = (unsigned long long)dcol_long(c, row);
                    }
                }
            }
        }
        // update phase
        unsigned long long* base = table + g * T + tid;
#pragma unroll
        for (int s = 0; s < MAX_PSLOTS; s++) {
            if (s < plan.nslots) {
                const PrivSlot& ps = plan.slots[s];
                if ((condbits >> ps.cond) & 1) {
                    unsigned long long* t = base + ps.word_off * GT;
                    switch (ps.kind) {
                        case PS_CNT: t[0] += 1; break;
                        case PS_ADD2: {
                            unsigned long long a0 = v0[s], a1 = (unsigned long long)((long long)a0 >> 63);
                            unsigned long long s0 = t[0] + a0;
                            t[0] = s0;
                            t[GT] += a1 + (s0 < a0 ? 1ULL : 0ULL);
                            break;
                        }
                        case PS_ADD3: {
                            unsigned long long a0, a1, a2;
                            signmag_to_twos(v0[s], v1[s], a0, a1, a2);
                            unsigned long long s0 = t[0] + a0;
                            t[0] = s0;
                            unsigned long long c0 = s0 < a0 ? 1ULL : 0ULL;
                            unsigned long long add1 = a1 + c0;
                            unsigned long long c1 = add1 < c0 ? 1ULL : 0ULL;
                            unsigned long long s1 = t[GT] + add1;
                            t[GT] = s1;
                            c1 += s1 < add1 ? 1ULL : 0ULL;
                            t[2 * GT] += a2 + c1;
                            break;
                        }
                        case PS_F64: {
                            double d = ps.conv == CV_I2D ? (double)(long long)v0[s] : __longlong_as_double((long long)v0[s]);
                            t[0] = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)t[0]) + d);
                            break;
                        }
                        case PS_MIN: {
                            long long k = ps.conv == CV_MINKEY ? double_order_key((long long)v0[s]) : (long long)v0[s];
                            if (k < (long long)t[0]) t[0] = (unsigned long long)k;
                            break;
                        }
                        case PS_MAX: {
                            long long k = ps.conv == CV_MAXKEY ? max_double_key((long long)v0[s]) : (long long)v0[s];
                            if (k > (long long)t[0]) t[0] = (unsigned long long)k;
                            break;
                        }
                    }
                }
            }
        }
    }
    // ---- flush: per aggregate and group, reduce the warp's 32 columns and fold into the global state
    const int lane = tid & 31;
    for (int a = 0; a < plan.naggs; a++) {
        const PrivAgg& pa = plan.aggs[a];
        const PrivSlot& cs = plan.slots[pa.cnt_slot];
        for (int g = 0; g < G; g++) {
            unsigned long long cnt = table[cs.word_off * GT + g * T + tid], zero = 0;
            wreduce_sum_carry(cnt, zero);
            cnt = __shfl_sync(0xffffffffu, cnt, 0);
            if (cnt == 0) continue;   // warp-uniform
            unsigned long long* st = pa.state + (int64_t)g * pa.words;
            if (pa.op == OP_COUNT) { if (lane == 0) atomicAdd(&st[0], cnt); continue; }
            const PrivSlot& vs = plan.slots[pa.val_slot];
            const unsigned long long* col = table + vs.word_off * GT + g * T + tid;
            switch (vs.kind) {
                case PS_ADD2: {
                    unsigned long long s0 = col[0], c0 = 0, s1 = col[GT], c1 = 0;
                    wreduce_sum_carry(s0, c0);
                    wreduce_sum_carry(s1, c1);
                    if (lane == 0) {
                        unsigned long long hi = s1 + c0;   // 128-bit signed partial (never overflows: < 2^31 rows of 64-bit values per thread)
                        atomicAdd(&st[0], cnt);
                        if (pa.op == OP_SUM_I64) atomic_add_chain2(st + 1, s0, hi);
                        else atomic_add_chain3(st + 1, s0, hi, (unsigned long long)((long long)hi >> 63));
                    }
                    break;
                }
                case PS_ADD3: {
                    unsigned long long s0 = col[0], c0 = 0, s1 = col[GT], c1 = 0, s2 = col[2 * GT], c2 = 0;
                    wreduce_sum_carry(s0, c0);
                    wreduce_sum_carry(s1, c1);
                    wreduce_sum_carry(s2, c2);
                    if (lane == 0) {
                        unsigned long long a1 = s1 + c0;
                        unsigned long long k1 = c1 + (a1 < s1 ? 1ULL : 0ULL);
                        atomicAdd(&st[0], cnt);
                        atomic_add_chain3(st + 1, s0, a1, s2 + k1);
                    }
                    break;
                }
                case PS_F64: {
                    double d = __longlong_as_double((long long)col[0]);
#pragma unroll
                    for (int k = 16; k > 0; k >>= 1) d += __shfl_down_sync(0xffffffffu, d, k);
                    if (lane == 0) { atomicAdd(&st[0], cnt); atomicAdd((double*)&st[1], d); }
                    break;
                }
                case PS_MIN: case PS_MAX: {
                    long long v = (long long)col[0];
#pragma unroll
                    for (int k = 16; k > 0; k >>= 1) {
                        long long o = __shfl_down_sync(0xffffffffu, v, k);
                        v = vs.kind == PS_MIN ? (o < v ? o : v) : (o > v ? o : v);
                    }
                    if (lane == 0) {
                        st[0] = 1;
                        if (vs.kind == PS_MIN) atomicMin((long long*)&st[1], v); else atomicMax((long long*)&st[1], v);
                    }
                    break;
                }
                default: break;
            }
        }
    }
}

// ---- host: build the plan (dedupe columns, conditions and state), pick the block size, launch
static int find_col(PrivPlan& p, const DCol& c) {
    for (int i = 0; i < p.ncols; i++)
        if (p.cols[i].values == c.values && p.cols[i].nulls == c.nulls && p.cols[i].type == c.type) return i;
    if (p.ncols >= MAX_PCOLS) return -1;
    p.cols[p.ncols] = c;
    return p.ncols++;
}

bool accumulate_private(pg_ctx* ctx, const AggProgram& prog, const int32_t* gids, int64_t n, int64_t group_count) {
    if (group_count < 1 || group_count > 64) return false;
    PrivPlan p;
    memset(&p, 0, sizeof(p));
    p.naggs = prog.naggs;
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        if (op.op == OP_MIN128 || op.op == OP_MAX128) return false;
        // condition = (mask column, null mask of the argument)
        int mask_col = -1, null_col = -1;
        if (op.mask.values) { mask_col = find_col(p, op.mask); if (mask_col < 0) return false; }
        int src = -1;
        if (op.op != OP_COUNT || (op.skip_null && op.arg.nulls)) { src = find_col(p, op.arg); if (src < 0) return false; }
        if (op.skip_null && op.arg.nulls) null_col = src;
        int cond = -1;
        for (int c = 0; c < p.nconds; c++) if (p.conds[c].mask_col == mask_col && p.conds[c].null_col == null_col) cond = c;
        if (cond < 0) {
            if (p.nconds >= MAX_PCONDS) return false;
            cond = p.nconds++;
            p.conds[cond].mask_col = mask_col; p.conds[cond].null_col = null_col;
        }
        auto find_slot = [&](int kind, int conv, int src_col, int nwords) -> int {
            for (int s = 0; s < p.nslots; s++) {
                const PrivSlot& ps = p.slots[s];
                if (ps.kind == kind && ps.conv == conv && ps.src == src_col && ps.cond == cond) return s;
            }
            if (p.nslots >= MAX_PSLOTS) return -1;
            PrivSlot& ps = p.slots[p.nslots];
            ps.kind = kind; ps.conv = conv; ps.src = src_col; ps.cond = cond; ps.nwords = nwords; ps.word_off = p.words;
            p.words += nwords;
            return p.nslots++;
        };
        PrivAgg& pa = p.aggs[a];
        pa.op = op.op; pa.words = op.words; pa.state = op.state;
        pa.cnt_slot = find_slot(PS_CNT, 0, -1, 1);
        if (pa.cnt_slot < 0) return false;
        pa.val_slot = -1;
        switch (op.op) {
            case OP_COUNT: break;
            case OP_SUM_I64: pa.val_slot = find_slot(PS_ADD2, CV_INT, src, 2); break;
            case OP_SUM_F64: pa.val_slot = find_slot(PS_F64, op.convert_i2d ? CV_I2D : CV_F64, src, 1); break;
            case OP_SUM_DEC:
                pa.val_slot = op.in_type == PG_LONG_DECIMAL ? find_slot(PS_ADD3, CV_DEC128, src, 3) : find_slot(PS_ADD2, CV_INT, src, 2);
                break;
            case OP_MIN64: pa.val_slot = find_slot(PS_MIN, op.in_type == PG_DOUBLE ? CV_MINKEY : CV_INT, src, 1); break;
            case OP_MAX64: pa.val_slot = find_slot(PS_MAX, op.in_type == PG_DOUBLE ? CV_MAXKEY : CV_INT, src, 1); break;
            default: return false;
        }
        if (op.op != OP_COUNT && pa.val_slot < 0) return false;
    }
    const int G = (int)group_count;
    const int64_t per_thread = (int64_t)p.words * G * 8;
    const int64_t budget = 220 * 1024;
    int T = 256;
    if (per_thread * 256 > budget / 2) T = 128;           // keep at least two blocks (or one 256-thread block) per SM
    if (per_thread * T > budget) return false;
    int64_t smem = per_thread * T;
    int blocks_per_sm = (int)(budget / smem);
    if (blocks_per_sm > 2048 / T) blocks_per_sm = 2048 / T;
    int64_t want = (n + T - 1) / T;
    int64_t grid = (int64_t)ctx->sm_count * blocks_per_sm;
    if (grid > want) grid = want;
    if (T == 256) {
        PG_CUDA(cudaFuncSetAttribute(acc_private2_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        PG_LAUNCH(ctx, acc_private2_kernel<256>, (int)grid, 256, (size_t)smem, p, gids, n, G);
    } else {
        PG_CUDA(cudaFuncSetAttribute(acc_private2_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        PG_LAUNCH(ctx, acc_private2_kernel<128>, (int)grid, 128, (size_t)smem, p, gids, n, G);
    }
    return true;
}

}  // namespace pg
