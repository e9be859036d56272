// accum.cu — grouped accumulator kernels: addInput (global-atomic and block-private variants), state init.
#include "accum_dev.cuh"

namespace pg {

static const int ACC_THREADS = 256;
static const int PRIV_THREADS = 128;
static const int PRIV_SMEM_BUDGET = 200 * 1024;

// ------------------------------------------------------------------ GLOBAL: one row per thread, atomics on HBM/L2 state
__global__ void __launch_bounds__(ACC_THREADS) acc_global_kernel(AggProgram prog, const int32_t* __restrict__ gids, int64_t n) {
    int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; row < n; row += stride) {
        int64_t g = gids[row];
        for (int a = 0; a < prog.naggs; a++) {
            const AggOp& op = prog.ops[a];
            Contribution c;
            if (extract_input(op, row, c)) update_global(op, g, c);
        }
    }
}

// second pass of 128-bit min/max: rows whose high word equals the group's extreme compete on the low word
__global__ void acc_minmax128_low_kernel(AggOp op, const int32_t* __restrict__ gids, int64_t n) {
    int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; row < n; row += stride) {
        Contribution c;
        if (!extract_input(op, row, c)) continue;
        unsigned long long* st = op.state + (int64_t)gids[row] * op.words;
        if (st[1] != c.a1) continue;
        if (op.op == OP_MIN128) atomicMin((long long*)&st[2], (long long)c.a0);
        else atomicMax((long long*)&st[2], (long long)c.a0);
    }
}
// between the passes: snapshot the high words before pass 1, reset the low word where the high word moved
__global__ void acc_minmax128_snapshot_kernel(const unsigned long long* state, int64_t groups, unsigned long long* snap) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; g < groups; g += stride) snap[g] = state[g * 3 + 1];
}
__global__ void acc_minmax128_reset_kernel(unsigned long long* state, int64_t groups, const unsigned long long* snap, int op) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; g < groups; g += stride)
        if (state[g * 3 + 1] != snap[g]) state[g * 3 + 2] = identity_word(op, 2);
}

// ------------------------------------------------------------------ PRIVATE: per-thread tables in shared memory
// Layout: table[(word * G + group) * PRIV_THREADS + thread] — every thread owns one 8-byte column, so updates are
// conflict-free plain loads/stores (no atomics).  At the end each warp tree-reduces its 32 columns with shuffles
// (sum + carry count per word) and lane 0 applies the result to the global state with the carry-chain atomics.
PG_D void warp_reduce_sum_carry(unsigned long long& s, unsigned long long& carry) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long os = __shfl_down_sync(0xffffffffu, s, d);
        unsigned long long oc = __shfl_down_sync(0xffffffffu, carry, d);
        unsigned long long ns = s + os;
        carry += oc + (ns < s ? 1ULL : 0ULL);
        s = ns;
    }
}

__global__ void __launch_bounds__(PRIV_THREADS) acc_private_kernel(AggProgram prog, const int32_t* __restrict__ gids, int64_t n, int G) {
    extern __shared__ unsigned long long table[];
    const int T = PRIV_THREADS;
    const int tid = threadIdx.x;
    // init own columns
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        for (int w = 0; w < op.words; w++)
            for (int g = 0; g < G; g++) table[((op.priv_off + w) * G + g) * T + tid] = identity_word(op.op, w);
    }
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t row = (int64_t)blockIdx.x * blockDim.x + tid; row < n; row += stride) {
        int g = gids[row];
        for (int a = 0; a < prog.naggs; a++) {
            const AggOp& op = prog.ops[a];
            Contribution c;
            if (extract_input(op, row, c)) update_private(op, &table[(op.priv_off * G + g) * T + tid], G * T, c);
        }
    }
    // flush: warp-level reduction, then lane 0 -> global state
    const int lane = tid & 31;
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        for (int g = 0; g < G; g++) {
            unsigned long long* col = &table[(op.priv_off * G + g) * T + tid];
            const int S = G * T;
            unsigned long long cnt = col[0], zero = 0;
            warp_reduce_sum_carry(cnt, zero);
            unsigned long long* st = op.state + (int64_t)g * op.words;
            switch (op.op) {
                case OP_COUNT:
                    if (lane == 0 && cnt) atomicAdd(&st[0], cnt);
                    break;
                case OP_SUM_I64: {
                    unsigned long long s0 = col[S], c0 = 0, s1 = col[2 * S], c1 = 0;
                    warp_reduce_sum_carry(s0, c0);
                    warp_reduce_sum_carry(s1, c1);
                    if (lane == 0 && cnt) { atomicAdd(&st[0], cnt); atomic_add_chain2(st + 1, s0, s1 + c0); }
                    break;
                }
                case OP_SUM_F64: {
                    double d = __longlong_as_double((long long)col[S]);
#pragma unroll
                    for (int k = 16; k > 0; k >>= 1) d += __shfl_down_sync(0xffffffffu, d, k);
                    if (lane == 0 && cnt) { atomicAdd(&st[0], cnt); atomicAdd((double*)&st[1], d); }
                    break;
                }
                case OP_SUM_DEC: {
                    unsigned long long s0 = col[S], c0 = 0, s1 = col[2 * S], c1 = 0, s2 = col[3 * S], c2 = 0;
                    warp_reduce_sum_carry(s0, c0);
                    warp_reduce_sum_carry(s1, c1);
                    warp_reduce_sum_carry(s2, c2);
                    if (lane == 0 && cnt) {
                        // fold the carry counts into the next word (192-bit arithmetic)
                        unsigned long long a1 = s1 + c0;
                        unsigned long long k1 = c1 + (a1 < s1 ? 1ULL : 0ULL);
                        unsigned long long a2 = s2 + k1;
                        atomicAdd(&st[0], cnt);
                        atomic_add_chain3(st + 1, s0, a1, a2);
                    }
                    break;
                }
                case OP_MIN64: case OP_MAX64: {
                    long long v = (long long)col[S];
#pragma unroll
                    for (int k = 16; k > 0; k >>= 1) {
                        long long o = __shfl_down_sync(0xffffffffu, v, k);
                        v = op.op == OP_MIN64 ? (o < v ? o : v) : (o > v ? o : v);
                    }
                    if (lane == 0 && cnt) {
                        st[0] = 1;
                        if (op.op == OP_MIN64) atomicMin((long long*)&st[1], v); else atomicMax((long long*)&st[1], v);
                    }
                    break;
                }
                default: break;
            }
        }
    }
}

// ------------------------------------------------------------------ state init
__global__ void acc_init_state_kernel(unsigned long long* state, int op, int words, int64_t first, int64_t count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t total = count * words;
    for (; i < total; i += stride) state[first * words + i] = identity_word(op, (int)(i % words));
}

void init_state(pg_ctx* ctx, int op, unsigned long long* state, int64_t first, int64_t count) {
    if (count <= 0) return;
    int words = state_words(op);
    if (op == OP_COUNT || op == OP_SUM_I64 || op == OP_SUM_F64 || op == OP_SUM_DEC) {
        memset_async(ctx, state + first * words, 0, count * words * 8);
        return;
    }
    PG_LAUNCH(ctx, acc_init_state_kernel, grid_for(ctx, count * words, 256, 4), 256, 0, state, op, words, first, count);
}

void accumulate(pg_ctx* ctx, const AggProgram& prog_in, const int32_t* gids, int64_t n, int64_t group_count) {
    if (n <= 0 || prog_in.naggs == 0) return;
    AggProgram prog = prog_in;
    bool has128 = false;
    int priv_words = 0;
    for (int a = 0; a < prog.naggs; a++) {
        AggOp& op = prog.ops[a];
        if (op.op == OP_MIN128 || op.op == OP_MAX128) has128 = true;
        op.priv_off = priv_words;
        priv_words += op.words;
    }
    prog.priv_words = priv_words;
    // strategy: tiny group counts -> block-private tables (no atomics in the row loop); otherwise global atomics
    int64_t smem = (int64_t)priv_words * group_count * PRIV_THREADS * 8;
    bool use_private = !has128 && group_count >= 1 && smem <= PRIV_SMEM_BUDGET && n >= 64 * 1024;
    std::vector<DevBuf> snaps;
    if (has128) {
        for (int a = 0; a < prog.naggs; a++) {
            AggOp& op = prog.ops[a];
            if (op.op != OP_MIN128 && op.op != OP_MAX128) continue;
            snaps.emplace_back(ctx, group_count * 8);
            PG_LAUNCH(ctx, acc_minmax128_snapshot_kernel, grid_for(ctx, group_count, 256), 256, 0, op.state, group_count,
                      snaps.back().as<unsigned long long>());
        }
    }
    if (use_private) {
        static bool attr_set = false;
        if (!attr_set) {
            PG_CUDA(cudaFuncSetAttribute(acc_private_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PRIV_SMEM_BUDGET));
            attr_set = true;
        }
        int blocks_per_sm = (int)(PRIV_SMEM_BUDGET / (smem > 0 ? smem : 1));
        if (blocks_per_sm < 1) blocks_per_sm = 1;
        if (blocks_per_sm > 8) blocks_per_sm = 8;
        int64_t want = (n + PRIV_THREADS - 1) / PRIV_THREADS;
        int64_t grid = (int64_t)ctx->sm_count * blocks_per_sm;
        if (grid > want) grid = want;
        PG_LAUNCH(ctx, acc_private_kernel, (int)grid, PRIV_THREADS, (size_t)smem, prog, gids, n, (int)group_count);
    } else {
        PG_LAUNCH(ctx, acc_global_kernel, grid_for(ctx, n, ACC_THREADS), ACC_THREADS, 0, prog, gids, n);
    }
    if (has128) {
        int si = 0;
        for (int a = 0; a < prog.naggs; a++) {
            AggOp& op = prog.ops[a];
            if (op.op != OP_MIN128 && op.op != OP_MAX128) continue;
            PG_LAUNCH(ctx, acc_minmax128_reset_kernel, grid_for(ctx, group_count, 256), 256, 0, op.state, group_count,
                      snaps[si++].as<unsigned long long>(), op.op);
            PG_LAUNCH(ctx, acc_minmax128_low_kernel, grid_for(ctx, n, 256), 256, 0, op, gids, n);
        }
    }
}

}  // namespace pg
