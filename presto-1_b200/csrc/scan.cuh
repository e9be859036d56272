// scan.cuh — block-level and device-level exclusive scans used for first-seen group-id assignment,
// match-position compaction and partition offsets.  Two accumulators (count, bytes) are scanned together.
#pragma once
#include "context.h"

namespace pg {

struct Pair64 {
    int64_t a, b;
};
PG_HD Pair64 operator+(const Pair64& x, const Pair64& y) { return Pair64{x.a + y.a, x.b + y.b}; }

static const int SCAN_THREADS = 256;
static const int SCAN_ITEMS = 8;                          // rows per thread per tile (striped)
static const int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // rows per block tile

PG_D Pair64 shfl_up_pair(Pair64 v, int d) {
    Pair64 r;
    r.a = __shfl_up_sync(0xffffffffu, v.a, d);
    r.b = __shfl_up_sync(0xffffffffu, v.b, d);
    return r;
}

// Exclusive scan of one value per thread across a SCAN_THREADS block.  Returns the exclusive prefix of this
// thread; *total receives the block total.  smem: Pair64[SCAN_THREADS/32 + 1].  Ends with a __syncthreads().
PG_D Pair64 block_exclusive_scan(Pair64 v, Pair64* smem, Pair64* total) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Pair64 inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Pair64 o = shfl_up_pair(inc, d);
        if (lane >= d) inc = inc + o;
    }
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const int nw = SCAN_THREADS / 32;
        Pair64 w = lane < nw ? smem[lane] : Pair64{0, 0};
        Pair64 winc = w;
#pragma unroll
        for (int d = 1; d < nw; d <<= 1) {
            Pair64 o = shfl_up_pair(winc, d);
            if (lane >= d) winc = winc + o;
        }
        if (lane < nw) smem[lane] = Pair64{winc.a - w.a, winc.b - w.b};   // exclusive warp offsets
        if (lane == nw - 1) smem[nw] = winc;                              // block total
    }
    __syncthreads();
    Pair64 base = smem[warp];
    *total = smem[SCAN_THREADS / 32];
    Pair64 ex = Pair64{base.a + inc.a - v.a, base.b + inc.b - v.b};
    __syncthreads();
    return ex;
}

// Single-block exclusive scan over per-tile totals (in place).  meta_total (optional) receives the grand total.
__global__ void scan_tile_totals_kernel(Pair64* tiles, int64_t ntiles, Pair64* grand_total);

// host helper: exclusive-scan `tiles[ntiles]` in place, writing the grand total to *d_total (device)
void scan_tile_totals(pg_ctx* ctx, Pair64* tiles, int64_t ntiles, Pair64* d_total);

static inline int64_t num_tiles(int64_t n) { return (n + SCAN_TILE - 1) / SCAN_TILE; }

}  // namespace pg
