// scan.cu — single-block scan over per-tile totals (second level of the two-level device scan).
#include "scan.cuh"

namespace pg {

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_totals_kernel(Pair64* tiles, int64_t ntiles, Pair64* grand_total) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    Pair64 carry{0, 0};
    const int PER = 8;
    for (int64_t base = 0; base < ntiles; base += (int64_t)SCAN_THREADS * PER) {
        int64_t i0 = base + (int64_t)threadIdx.x * PER;
        Pair64 v[PER];
        Pair64 s{0, 0};
#pragma unroll
        for (int j = 0; j < PER; j++) {
            v[j] = (i0 + j < ntiles) ? tiles[i0 + j] : Pair64{0, 0};
            s = s + v[j];
        }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(s, smem, &total);
        Pair64 run = carry + ex;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            if (i0 + j < ntiles) tiles[i0 + j] = run;
            run = run + v[j];
        }
        carry = carry + total;
    }
    if (threadIdx.x == 0 && grand_total) *grand_total = carry;
}

// ---- many entries (one block walking them serially took 1.5 ms for 586 K): chunks of SCAN_TILE entries are scanned by one block
// each, the chunk totals by the single-block kernel above, and a third pass adds each chunk's offset.
__global__ void __launch_bounds__(SCAN_THREADS) scan_chunks_kernel(Pair64* tiles, int64_t ntiles, Pair64* chunk_totals) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t i0 = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    Pair64 v[SCAN_ITEMS];
    Pair64 s{0, 0};
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        v[j] = (i0 + j < ntiles) ? tiles[i0 + j] : Pair64{0, 0};
        s = s + v[j];
    }
    Pair64 total;
    Pair64 run = block_exclusive_scan(s, smem, &total);
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        if (i0 + j < ntiles) tiles[i0 + j] = run;
        run = run + v[j];
    }
    if (threadIdx.x == 0) chunk_totals[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan_add_offsets_kernel(Pair64* tiles, int64_t ntiles, const Pair64* chunk_offsets) {
    const Pair64 off = chunk_offsets[blockIdx.x];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    for (int j = threadIdx.x; j < SCAN_TILE; j += SCAN_THREADS) {
        const int64_t i = base + j;
        if (i < ntiles) tiles[i] = tiles[i] + off;
    }
}

void scan_tile_totals(pg_ctx* ctx, Pair64* tiles, int64_t ntiles, Pair64* d_total) {
    if (ntiles <= 8 * SCAN_TILE) {
        PG_LAUNCH(ctx, scan_tile_totals_kernel, 1, SCAN_THREADS, 0, tiles, ntiles, d_total);
        return;
    }
    const int64_t nchunks = (ntiles + SCAN_TILE - 1) / SCAN_TILE;
    DevBuf chunk_totals(ctx, (nchunks + 1) * (int64_t)sizeof(Pair64));
    PG_LAUNCH(ctx, scan_chunks_kernel, (int)nchunks, SCAN_THREADS, 0, tiles, ntiles, chunk_totals.as<Pair64>());
    PG_LAUNCH(ctx, scan_tile_totals_kernel, 1, SCAN_THREADS, 0, chunk_totals.as<Pair64>(), nchunks, d_total ? d_total : chunk_totals.as<Pair64>() + nchunks);
    PG_LAUNCH(ctx, scan_add_offsets_kernel, (int)nchunks, SCAN_THREADS, 0, tiles, ntiles, chunk_totals.as<Pair64>());
}

}  // namespace pg
