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

void scan_tile_totals(pg_ctx* ctx, Pair64* tiles, int64_t ntiles, Pair64* d_total) {
    PG_LAUNCH(ctx, scan_tile_totals_kernel, 1, SCAN_THREADS, 0, tiles, ntiles, d_total);
}

}  // namespace pg
