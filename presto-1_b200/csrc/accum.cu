// accum.cu — grouped accumulator kernels: addInput (global-atomic and block-private variants), state init.
#include "accum_dev.cuh"

namespace pg {

static const int ACC_THREADS = 256;

// ------------------------------------------------------------------ GLOBAL: one row per thread, atomics on HBM/L2 state
__global__ void __launch_bounds__(ACC_THREADS) acc_global_kernel(AggProgram prog, const int32_t* __restrict__ gids, int64_t n, const uint8_t* __restrict__ only_rows) {
    int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; row < n; row += stride) {
        int64_t g = gids[row];
        if (only_rows != nullptr && only_rows[row] == 0) continue;
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

void accumulate(pg_ctx* ctx, const AggProgram& prog_in, const int32_t* gids, int64_t n, int64_t group_count, const uint8_t* only_rows) {
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
    // strategy: few groups -> thread-private tables in shared memory (accum_private.cu); otherwise global atomics
    bool use_private = !has128 && n >= 4 * 1024;
    PG_CHECK_STATE(!(has128 && only_rows), "128-bit min/max cannot be combined with a row filter");
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
    if (!(use_private && accumulate_private(ctx, prog, gids, n, group_count, only_rows, nullptr)))
        PG_LAUNCH(ctx, acc_global_kernel, grid_for(ctx, n, ACC_THREADS), ACC_THREADS, 0, prog, gids, n, only_rows);
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
