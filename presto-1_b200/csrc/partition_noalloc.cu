// partition_noalloc.cu — pg_partition_positions_noalloc: the hash-partition positions pass (partition id + stable destination row
// of every row, rows per partition) WITHOUT any device allocation.  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is a variant of pg_partition_create's three steps (partition.cu: per-tile histogram -> scan -> stable rank) whose scratch —
// the (partition-major, tile-minor) histogram, which pg_partition_create takes from the stream-ordered pool — lives in ONE
// file-scope `static __device__` array of fixed size.  Consequences, all part of the contract (include/prestogpu.h):
//   * at most PG_NOALLOC_MAX_ROWS rows per call (the array holds 64 partitions x 1024 tiles of 4096 rows + the grand total:
//     (65536 + 1) x 16 B, about 1 MiB, held by every GPU that loads the library until the process exits);
//   * every call on a device shares the array: calls that overlap on the GPU (different streams / contexts) race.  Calls on one
//     stream are ordered and safe;
//   * the array keeps the last call's contents.  Nothing has to start at zero: the histogram kernel writes every word
//     [0, nparts * ntiles) before the scan reads it and the scan writes the total behind them, so stale words are never read.
// In HOST code the array's name is only a stand-in for the symbol, not a device address; the address is resolved once per device
// with cudaGetSymbolAddress (scratch_for_device) and that pointer is what the kernels receive.
#include <mutex>

#include "scan.cuh"

namespace pg {

namespace {

const int NA_THREADS = 256;
const int NA_ITEMS = 16;
const int NA_TILE = NA_THREADS * NA_ITEMS;           // 4096 rows per block
const int NA_MAX_PARTS = 64;
const int64_t NA_MAX_TILES = 1024;
const int64_t NA_MAX_ROWS = NA_MAX_TILES * NA_TILE;   // 4 194 304 = PG_NOALLOC_MAX_ROWS
const int64_t NA_SCRATCH_WORDS = NA_MAX_PARTS * NA_MAX_TILES + 1;

static __device__ __align__(256) Pair64 g_noalloc_scratch[NA_SCRATCH_WORDS];

// (rawHash & Long.MAX_VALUE) % n  /  xxhash64(reverse(rawHash)) & (n - 1): HashGenerator.java:24-35, LocalPartitionGenerator.java:50-63
PG_D int na_partition_of(int64_t raw, int nparts, int mode) {
    if (mode == PG_PARTITION_REMOTE) return (int)(((unsigned long long)raw & 0x7fffffffffffffffULL) % (unsigned long long)nparts);
    return (int)xxh64_long((int64_t)bit_reverse64((uint64_t)raw)) & (nparts - 1);
}
// page order inside a tile = (warp, item, lane): a warp owns 512 consecutive rows and visits them 32 at a time
PG_D int64_t na_row_of(int64_t tile_base, int warp, int j, int lane) { return tile_base + ((int64_t)warp * NA_ITEMS + j) * 32 + lane; }

__global__ void __launch_bounds__(NA_THREADS) na_hist_kernel(const int64_t* __restrict__ raw_hash, int64_t n, int nparts, int mode,
                                                             uint8_t* __restrict__ part_of_row, Pair64* __restrict__ counts, int64_t ntiles) {
    __shared__ int hist[NA_MAX_PARTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < NA_MAX_PARTS) hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * NA_TILE;
    for (int j = 0; j < NA_ITEMS; j++) {
        const int64_t row = na_row_of(base, warp, j, lane);
        int p = -1;
        if (row < n) { p = na_partition_of(raw_hash[row], nparts, mode); part_of_row[row] = (uint8_t)p; }
        const unsigned peers = __match_any_sync(0xffffffffu, p);
        if (p >= 0 && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&hist[p], __popc(peers));
    }
    __syncthreads();
    if (threadIdx.x < nparts) counts[(int64_t)threadIdx.x * ntiles + blockIdx.x] = Pair64{hist[threadIdx.x], 0};
}

// counts[] holds the exclusive scan: counts[p * ntiles + t] = first destination of tile t's rows of partition p; counts[nparts * ntiles] = n
__global__ void __launch_bounds__(NA_THREADS) na_rank_kernel(const uint8_t* __restrict__ part_of_row, int64_t n, int nparts,
                                                             const Pair64* __restrict__ offsets, int64_t ntiles, int32_t* __restrict__ dest,
                                                             int64_t* __restrict__ out_counts) {
    __shared__ int warp_cnt[NA_THREADS / 32][NA_MAX_PARTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t base = (int64_t)blockIdx.x * NA_TILE;
    if (blockIdx.x == 0 && threadIdx.x < nparts) {      // rows per partition = distance between the partitions' first destinations
        const int p = threadIdx.x;
        const int64_t start = offsets[(int64_t)p * ntiles].a;
        const int64_t end = p + 1 < nparts ? offsets[(int64_t)(p + 1) * ntiles].a : n;
        out_counts[p] = end - start;
    }
    for (int q = lane; q < nparts; q += 32) warp_cnt[warp][q] = 0;
    __syncwarp();
    int mine[NA_ITEMS];
#pragma unroll
    for (int j = 0; j < NA_ITEMS; j++) {
        const int64_t row = na_row_of(base, warp, j, lane);
        const int p = row < n ? (int)part_of_row[row] : -1;
        mine[j] = p;
        const unsigned peers = __match_any_sync(0xffffffffu, p);
        if (p >= 0 && (peers & ((1u << lane) - 1u)) == 0) warp_cnt[warp][p] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    for (int q = threadIdx.x; q < nparts; q += NA_THREADS) {
        int run = (int)offsets[(int64_t)q * ntiles + blockIdx.x].a;
        for (int w = 0; w < NA_THREADS / 32; w++) { const int c = warp_cnt[w][q]; warp_cnt[w][q] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NA_ITEMS; j++) {
        const int p = mine[j];
        const unsigned peers = __match_any_sync(0xffffffffu, p);
        const int rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (p >= 0) dest[na_row_of(base, warp, j, lane)] = warp_cnt[warp][p] + rank_in_warp;
        __syncwarp();
        if (p >= 0 && rank_in_warp == 0) warp_cnt[warp][p] += __popc(peers);
        __syncwarp();
    }
}

__global__ void na_zero_counts_kernel(int64_t* out_counts, int nparts) {
    if ((int)threadIdx.x < nparts) out_counts[threadIdx.x] = 0;
}

// Device address of the scratch array on `device` (which must be current).  Resolved once per device; the array's name itself
// must never reach a kernel from host code.
Pair64* scratch_for_device(int device) {
    static std::mutex mu;
    static Pair64* ptr[64] = {nullptr};
    PG_CHECK_ARG(device >= 0 && device < 64, "device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            fail(PG_E_CUDA, "cudaGetSymbolAddress(no-alloc scratch) failed on device %d: %s", device, cudaGetErrorString(e));
        }
        ptr[device] = (Pair64*)p;
    }
    return ptr[device];
}

}  // namespace

}  // namespace pg

using namespace pg;

extern "C" int32_t pg_partition_positions_noalloc(pg_ctx* ctx, const int64_t* raw_hash, int64_t n, int32_t nparts, int32_t mode,
                                                  uint8_t* out_part_of_row, int32_t* out_dest, int64_t* out_counts) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        PG_CHECK_ARG(n >= 0 && out_counts && (n == 0 || (raw_hash && out_part_of_row && out_dest)), "bad arguments (all pointers are device pointers)");
        PG_CHECK_ARG(nparts >= 1 && nparts <= NA_MAX_PARTS, "between 1 and %d partitions are supported", NA_MAX_PARTS);
        PG_CHECK_ARG(mode == PG_PARTITION_REMOTE || (mode == PG_PARTITION_LOCAL && (nparts & (nparts - 1)) == 0),
                     "mode must be PG_PARTITION_REMOTE, or PG_PARTITION_LOCAL with a power-of-two partition count");
        PG_CHECK_ARG(n <= NA_MAX_ROWS, "pg_partition_positions_noalloc handles at most %lld rows per call (fixed scratch), got %lld; use pg_partition_create",
                     (long long)NA_MAX_ROWS, (long long)n);
        if (n == 0) { PG_LAUNCH(ctx, na_zero_counts_kernel, 1, NA_MAX_PARTS, 0, out_counts, nparts); return; }
        Pair64* scratch = scratch_for_device(ctx->device);      // guard() made ctx->device current
        const int64_t ntiles = (n + NA_TILE - 1) / NA_TILE;
        PG_LAUNCH(ctx, na_hist_kernel, (int)ntiles, NA_THREADS, 0, raw_hash, n, nparts, mode, out_part_of_row, scratch, ntiles);
        // (the single-block scan kernel directly: scan_tile_totals() takes scratch memory from the pool for long inputs)
        PG_LAUNCH(ctx, scan_tile_totals_kernel, 1, SCAN_THREADS, 0, scratch, (int64_t)nparts * ntiles, scratch + (int64_t)nparts * ntiles);
        PG_LAUNCH(ctx, na_rank_kernel, (int)ntiles, NA_THREADS, 0, out_part_of_row, n, nparts, scratch, ntiles, out_dest, out_counts);
        // asynchronous: the results are ready when the context's stream reaches this point (pg_ctx_synchronize / stream order)
    });
}
