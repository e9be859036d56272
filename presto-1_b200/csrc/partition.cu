// partition.cu — hash partitioning of a page for the multi-GPU exchange (replaces the row routing of
// MAIN/operator/exchange/LocalPartitionGenerator.java:50-63, PartitioningExchanger.java:79-95 and
// MAIN/operator/PartitionedOutputOperator / HashGenerator.getPartition, MAIN/operator/HashGenerator.java:24-35).
//
// Three kernels: (1) per-tile histogram of partition ids, (2) scan of the (partition-major, tile-minor) counts,
// (3) stable ranking: dest[row] = offset[partition][tile] + rank of the row among the tile's rows of that partition.
// Columns are then scattered with out[dest[row]] = in[row]; rows keep their relative order inside a partition.
// The scatter can target local memory (NCCL all-to-all follows) or peer GPUs' receive buffers directly (P2P over NVLink).
#include <algorithm>
#include <vector>

#include "groupby_dev.cuh"
#include "scan.cuh"

namespace pg {

static const int PART_THREADS = 256;
static const int PART_ITEMS = 16;
static const int PART_TILE = PART_THREADS * PART_ITEMS;
static const int MAX_PARTS = 64;

// mode 0: HashGenerator.getPartition = (rawHash & Long.MAX_VALUE) % partitionCount
// mode 1: LocalPartitionGenerator = (int) XxHash64.hash(Long.reverse(rawHash)) & (partitionCount - 1)
// (mode 0 with `magic` = floor(2^64 / nparts), 0 for a power of two: x % n without a 64-bit division per row.  mulhi(x, magic) is
//  the quotient or one less, so one conditional subtraction finishes it.)
PG_D int partition_of(int64_t raw, int nparts, int mode, unsigned long long magic = 0) {
    if (mode == 0) {
        const unsigned long long x = (unsigned long long)raw & 0x7fffffffffffffffULL;
        if ((nparts & (nparts - 1)) == 0) return (int)(x & (unsigned long long)(nparts - 1));
        if (magic == 0) return (int)(x % (unsigned long long)nparts);
        unsigned long long r = x - __umul64hi(x, magic) * (unsigned long long)nparts;
        if (r >= (unsigned long long)nparts) r -= (unsigned long long)nparts;
        return (int)r;
    }
    return (int)xxh64_long((int64_t)bit_reverse64((uint64_t)raw)) & (nparts - 1);
}
// Lanes of the warp whose row goes to the same partition as this lane's.  Up to 8 partitions: one ballot per partition (cheaper than
// __match_any_sync); more: __match_any_sync.  p = -1 (row beyond the page) never matches a real partition.
PG_D unsigned part_peers(int p, int nparts) {
    if (nparts <= 8) {
        unsigned mine = 0;
        for (int q = 0; q < nparts; q++) { const unsigned b = __ballot_sync(0xffffffffu, p == q); if (p == q) mine = b; }
        return mine;
    }
    return __match_any_sync(0xffffffffu, p);
}
PG_D int64_t part_row_hash(const KeyCols& kc, int nkeys, const DCol& hash_col, int has_hash, int64_t row) {
    return has_hash ? ((const int64_t*)hash_col.values)[row] : keys_row_hash(kc, nkeys, row);
}

// A tile is PART_TILE consecutive rows; warp w of the block owns the PART_ITEMS * 32 consecutive rows starting at w * PART_ITEMS * 32
// and visits them 32 at a time, so "rows in page order" = (tile, warp, item, lane) order and ranking needs no block-wide step per item.
PG_D int64_t part_row_of(int64_t tile_base, int warp, int j, int lane) { return tile_base + ((int64_t)warp * PART_ITEMS + j) * 32 + lane; }

__global__ void __launch_bounds__(PART_THREADS) part_hist_kernel(KeyCols kc, int nkeys, DCol hash_col, int has_hash, int64_t n, int nparts,
                                                                 int mode, uint8_t* part_of_row, Pair64* counts, int64_t ntiles) {
    __shared__ int hist[MAX_PARTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < MAX_PARTS) hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * PART_TILE;
    const unsigned long long magic = 0xFFFFFFFFFFFFFFFFULL / (unsigned long long)nparts;   // = floor(2^64 / nparts) unless nparts is a power of two
#pragma unroll 4
    for (int j = 0; j < PART_ITEMS; j++) {
        const int64_t row = part_row_of(base, warp, j, lane);
        int p = -1;
        if (row < n) {
            p = partition_of(part_row_hash(kc, nkeys, hash_col, has_hash, row), nparts, mode, magic);
            part_of_row[row] = (uint8_t)p;
        }
        // one shared-memory atomic per distinct partition in the warp, not one per row
        const unsigned peers = part_peers(p, nparts);
        if (p >= 0 && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&hist[p], __popc(peers));
    }
    __syncthreads();
    if (threadIdx.x < nparts) counts[(int64_t)threadIdx.x * ntiles + blockIdx.x] = Pair64{hist[threadIdx.x], 0};
}

// The common case — one BIGINT key without nulls, no precomputed hash channel: all 16 key loads of a thread are issued before the
// first hash is computed, and no per-row type dispatch.  Same tile / warp / item order and the same outputs as part_hist_kernel.
__global__ void __launch_bounds__(PART_THREADS) part_hist_bigint_kernel(const int64_t* __restrict__ keys, int64_t n, int nparts, int mode,
                                                                        uint8_t* __restrict__ part_of_row, Pair64* __restrict__ counts, int64_t ntiles) {
    __shared__ int hist[MAX_PARTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < MAX_PARTS) hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * PART_TILE;
    const unsigned long long magic = 0xFFFFFFFFFFFFFFFFULL / (unsigned long long)nparts;
    int64_t k[PART_ITEMS];
#pragma unroll
    for (int j = 0; j < PART_ITEMS; j++) {
        const int64_t row = part_row_of(base, warp, j, lane);
        k[j] = row < n ? __ldg(keys + row) : 0;
    }
#pragma unroll
    for (int j = 0; j < PART_ITEMS; j++) {
        const int64_t row = part_row_of(base, warp, j, lane);
        int p = -1;
        if (row < n) {
            p = partition_of((int64_t)(31ULL * 0ULL + (uint64_t)long_type_hash(k[j])), nparts, mode, magic);     // keys_row_hash with one channel
            part_of_row[row] = (uint8_t)p;
        }
        const unsigned peers = part_peers(p, nparts);
        if (p >= 0 && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&hist[p], __popc(peers));
    }
    __syncthreads();
    if (threadIdx.x < nparts) counts[(int64_t)threadIdx.x * ntiles + blockIdx.x] = Pair64{hist[threadIdx.x], 0};
}

// after the scan counts[p * ntiles + t] = first destination of tile t inside the partition-major output order.
// Two block-wide steps per tile: the warps count their rows per partition, an exclusive sum over the warps gives every warp its
// first destination per partition, then each warp ranks its own rows (32 at a time, __match_any_sync) against its running counters.
__global__ void __launch_bounds__(PART_THREADS) part_rank_kernel(const uint8_t* part_of_row, int64_t n, int nparts, const Pair64* offsets,
                                                                 int64_t ntiles, int32_t* dest) {
    __shared__ int warp_cnt[PART_THREADS / 32][MAX_PARTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t base = (int64_t)blockIdx.x * PART_TILE;
    for (int q = lane; q < nparts; q += 32) warp_cnt[warp][q] = 0;
    __syncwarp();
    int mine[PART_ITEMS];
#pragma unroll
    for (int j = 0; j < PART_ITEMS; j++) {
        const int64_t row = part_row_of(base, warp, j, lane);
        const int p = row < n ? (int)part_of_row[row] : -1;
        mine[j] = p;
        const unsigned peers = part_peers(p, nparts);
        if (p >= 0 && (peers & ((1u << lane) - 1u)) == 0) warp_cnt[warp][p] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // warp_cnt[w][p] -> first destination of warp w's rows of partition p
    for (int q = threadIdx.x; q < nparts; q += PART_THREADS) {
        int run = (int)offsets[(int64_t)q * ntiles + blockIdx.x].a;
        for (int w = 0; w < PART_THREADS / 32; w++) { const int c = warp_cnt[w][q]; warp_cnt[w][q] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < PART_ITEMS; j++) {
        const int p = mine[j];
        const unsigned peers = part_peers(p, nparts);
        const int rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (p >= 0) {
            dest[part_row_of(base, warp, j, lane)] = warp_cnt[warp][p] + rank_in_warp;
        }
        __syncwarp();
        if (p >= 0 && rank_in_warp == 0) warp_cnt[warp][p] += __popc(peers);
        __syncwarp();
    }
}

// out[dest[i]] = in[i] for one fixed width column; peer_base != null: the destination row d of partition p lives at
// peer_base[p] + (d - part_start[p] + peer_row_offset[p]) * width (direct P2P store into the receiver's buffer)
struct ScatterPeers {
    void* base[MAX_PARTS];
    uint8_t* nulls[MAX_PARTS];
    int64_t part_start[MAX_PARTS];
    int64_t peer_row_offset[MAX_PARTS];
};
template <typename T>
__global__ void part_scatter_kernel(const T* __restrict__ in, const uint8_t* __restrict__ in_nulls, const int32_t* __restrict__ dest, int64_t n,
                                    T* out, uint8_t* out_nulls) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int32_t d = dest[i];
        out[d] = in[i];
        if (out_nulls) out_nulls[d] = in_nulls ? in_nulls[i] : 0;
    }
}
template <typename T>
__global__ void part_scatter_peer_kernel(const T* __restrict__ in, const uint8_t* __restrict__ in_nulls, const int32_t* __restrict__ dest,
                                         const uint8_t* __restrict__ part_of_row, int64_t n, ScatterPeers peers) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int p = part_of_row[i];
        int64_t d = (int64_t)dest[i] - peers.part_start[p] + peers.peer_row_offset[p];
        ((T*)peers.base[p])[d] = in[i];
        if (peers.nulls[p]) peers.nulls[p][d] = in_nulls ? in_nulls[i] : 0;
    }
}

// Peer scatter through shared memory.  The rows of one tile that go to partition p occupy CONSECUTIVE destination rows (the rank is
// stable), so a block first lays its tile out partition-major in shared memory and then stores it with consecutive threads writing
// consecutive addresses: runs of tile_rows / nparts elements per destination instead of the 1..4-element fragments a row-order
// scatter produces (over NVLink at 8 GPUs: 250 GB/s with fragments).  offsets = the scanned (partition, tile) counts of the handle.
template <typename T>
__global__ void __launch_bounds__(PART_THREADS) part_scatter_staged_kernel(const T* __restrict__ in, const uint8_t* __restrict__ in_nulls,
                                                                           const int32_t* __restrict__ dest, const uint8_t* __restrict__ part_of_row,
                                                                           int64_t n, int nparts, const Pair64* __restrict__ offsets, int64_t ntiles,
                                                                           ScatterPeers peers) {
    extern __shared__ __align__(16) unsigned char stage[];
    T* vals = (T*)stage;
    uint8_t* nl = (uint8_t*)(vals + PART_TILE);
    __shared__ long long s_first[MAX_PARTS];      // first destination (partition-major order) of this tile's rows of partition p
    __shared__ int s_start[MAX_PARTS + 1];        // where partition p starts inside the staged tile
    const int64_t tile = blockIdx.x, base = tile * PART_TILE;
    const int rows = (int)((n - base) < PART_TILE ? (n - base) : PART_TILE);
    if (threadIdx.x < nparts) {
        const int q = threadIdx.x;
        const long long f = offsets[(int64_t)q * ntiles + tile].a;
        const long long nxt = tile + 1 < ntiles ? offsets[(int64_t)q * ntiles + tile + 1].a : (q + 1 < nparts ? offsets[(int64_t)(q + 1) * ntiles].a : n);
        s_first[q] = f;
        s_start[q + 1] = (int)(nxt - f);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        s_start[0] = 0;
        for (int q = 0; q < nparts; q++) s_start[q + 1] += s_start[q];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < rows; i += PART_THREADS) {
        const int64_t r = base + i;
        const int q = part_of_row[r];
        const int l = (int)((long long)dest[r] - s_first[q]) + s_start[q];
        vals[l] = in[r];
        nl[l] = in_nulls ? in_nulls[r] : 0;
    }
    __syncthreads();
    for (int l = threadIdx.x; l < rows; l += PART_THREADS) {
        int lo = 0, hi = nparts - 1;                 // the partition whose staged range holds l
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_start[mid] <= l) lo = mid; else hi = mid - 1; }
        const int q = lo;
        const int64_t d = s_first[q] + (l - s_start[q]) - peers.part_start[q] + peers.peer_row_offset[q];
        ((T*)peers.base[q])[d] = vals[l];
        if (peers.nulls[q]) peers.nulls[q][d] = nl[l];
    }
}

// ---- every column of the page in ONE launch, without the dest[] array.
// A block owns one tile.  It ranks the tile's rows from part_of_row (the same stable order part_rank_kernel produces), so that row i of
// the tile gets a place lpos inside a partition-major staging area; then, column after column, the values are staged in shared
// memory and stored with consecutive threads writing consecutive destination rows of one receiver.  Per row the kernel reads
// 1 + sum(widths) bytes and writes sum(widths): no dest[] round trip (4 B written + 4 B per column read in the per-column form).
static const int SC_MAX_COLS = 16;
struct ScatterTable {                       // lives in device memory (too large for kernel parameters at 64 partitions)
    int32_t ncols, nparts;
    int64_t part_start[MAX_PARTS];          // first partition-major row of partition p on this rank
    int64_t peer_row_offset[MAX_PARTS];     // first row this rank writes in receiver p's buffers
    const void* in[SC_MAX_COLS];
    const uint8_t* in_nulls[SC_MAX_COLS];
    int32_t width[SC_MAX_COLS];
    void* base[SC_MAX_COLS][MAX_PARTS];
    uint8_t* nulls[SC_MAX_COLS][MAX_PARTS];
};

template <typename T, int THREADS>
PG_D void scatter_all_column(const ScatterTable* tab, int c, int64_t base, int rows, int warp, int lane, const unsigned short* s_l, unsigned char* stage,
                             const uint8_t* s_q, const long long* s_delta) {
    constexpr int ITEMS = PART_TILE / THREADS;
    T* vals = (T*)stage;
    uint8_t* nl = (uint8_t*)(vals + PART_TILE);
    const T* in = (const T*)tab->in[c];
    const uint8_t* in_nulls = tab->in_nulls[c];
    bool out_nulls = false;
    for (int q = 0; q < tab->nparts; q++) out_nulls = out_nulls || tab->nulls[c][q] != nullptr;
    T v[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {            // all loads of the thread are issued before the first store to shared memory
        const int i = (warp * ITEMS + j) * 32 + lane;
        if (i < rows) v[j] = in[base + i];
    }
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        const int i = (warp * ITEMS + j) * 32 + lane;
        if (i < rows) {
            const int l = s_l[i];
            vals[l] = v[j];
            if (out_nulls) nl[l] = in_nulls ? in_nulls[base + i] : 0;
        }
    }
    __syncthreads();
    for (int l = threadIdx.x; l < rows; l += THREADS) {
        const int q = s_q[l];
        const long long d = (long long)l + s_delta[q];
        ((T*)tab->base[c][q])[d] = vals[l];       // (the table's pointers stay in L1; copying them to shared memory per column was slower)
        if (out_nulls) tab->nulls[c][q][d] = nl[l];
    }
    __syncthreads();
}

// THREADS threads share the 4096 rows of a tile: warp w owns the PART_TILE / (THREADS / 32) consecutive rows starting at
// w * that many and visits them 32 at a time, so ranks follow row order whatever THREADS is (the histogram's tiles are the same
// 4096 rows; only their division among warps differs, which a stable rank does not see).
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) part_scatter_all_kernel(const uint8_t* __restrict__ part_of_row, int64_t n, int nparts,
                                                                         const Pair64* __restrict__ offsets, int64_t ntiles,
                                                                         const ScatterTable* __restrict__ tab) {
    constexpr int ITEMS = PART_TILE / THREADS;
    extern __shared__ __align__(16) unsigned char stage[];
    __shared__ long long s_delta[MAX_PARTS];      // destination row in the receiver's buffer = staged place + s_delta[partition]
    __shared__ int s_start[MAX_PARTS + 1];        // where partition p starts inside the staged tile
    __shared__ int warp_cnt[THREADS / 32][MAX_PARTS];
    __shared__ uint8_t s_q[PART_TILE];            // partition of every staged place
    __shared__ unsigned short s_l[PART_TILE];     // staged place of every row of the tile (kept here, not in registers: every column pass reads it)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t tile = blockIdx.x, base = tile * PART_TILE;
    const int rows = (int)((n - base) < PART_TILE ? (n - base) : PART_TILE);
    if (threadIdx.x < nparts) {
        const int q = threadIdx.x;
        const long long f = offsets[(int64_t)q * ntiles + tile].a;
        const long long nxt = tile + 1 < ntiles ? offsets[(int64_t)q * ntiles + tile + 1].a : (q + 1 < nparts ? offsets[(int64_t)(q + 1) * ntiles].a : n);
        s_delta[q] = f;
        s_start[q + 1] = (int)(nxt - f);
    }
    for (int q = lane; q < nparts; q += 32) warp_cnt[warp][q] = 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        s_start[0] = 0;
        for (int q = 0; q < nparts; q++) s_start[q + 1] += s_start[q];
    }
    int mine[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        const int i = (warp * ITEMS + j) * 32 + lane;
        mine[j] = i < rows ? (int)part_of_row[base + i] : -1;
    }
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        const int p = mine[j];
        const unsigned peers = part_peers(p, nparts);
        if (p >= 0 && (peers & ((1u << lane) - 1u)) == 0) warp_cnt[warp][p] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    for (int q = threadIdx.x; q < nparts; q += THREADS) {
        int run = s_start[q];
        for (int w = 0; w < THREADS / 32; w++) { const int cnt = warp_cnt[w][q]; warp_cnt[w][q] = run; run += cnt; }
        s_delta[q] = s_delta[q] - s_start[q] - tab->part_start[q] + tab->peer_row_offset[q];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        const int p = mine[j];
        const unsigned peers = part_peers(p, nparts);
        const int rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (p >= 0) {
            const int l = warp_cnt[warp][p] + rank_in_warp;
            s_l[(warp * ITEMS + j) * 32 + lane] = (unsigned short)l;
            s_q[l] = (uint8_t)p;
        }
        __syncwarp();
        if (p >= 0 && rank_in_warp == 0) warp_cnt[warp][p] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    const int ncols = tab->ncols;
    for (int c = 0; c < ncols; c++) {
        switch (tab->width[c]) {
            case 1: scatter_all_column<uint8_t, THREADS>(tab, c, base, rows, warp, lane, s_l, stage, s_q, s_delta); break;
            case 4: scatter_all_column<int32_t, THREADS>(tab, c, base, rows, warp, lane, s_l, stage, s_q, s_delta); break;
            case 8: scatter_all_column<int64_t, THREADS>(tab, c, base, rows, warp, lane, s_l, stage, s_q, s_delta); break;
            default: scatter_all_column<int4, THREADS>(tab, c, base, rows, warp, lane, s_l, stage, s_q, s_delta);
        }
    }
}

}  // namespace pg

using namespace pg;

struct pg_partitioner {
    pg_ctx* ctx = nullptr;
    int32_t nparts = 0, mode = 0;
    int64_t n = 0;
    DevBuf part_of_row, dest, counts;
    bool has_dest = false;      // dest[] is only needed by the per-column scatters: computed on first use (ensure_dest)
    int64_t part_counts[MAX_PARTS];
    int64_t part_start[MAX_PARTS];
};

// dest[row] = place of the row in partition-major order (stable); one more pass over part_of_row
static void ensure_dest(pg_partitioner* pt) {
    if (pt->has_dest || pt->n == 0) return;
    pg_ctx* ctx = pt->ctx;
    const int64_t ntiles = (pt->n + PART_TILE - 1) / PART_TILE;
    pt->dest.alloc(ctx, pt->n * 4);
    PG_LAUNCH(ctx, part_rank_kernel, (int)ntiles, PART_THREADS, 0, pt->part_of_row.as<uint8_t>(), pt->n, pt->nparts, pt->counts.as<Pair64>(), ntiles,
              pt->dest.as<int32_t>());
    pt->has_dest = true;
}

extern "C" {

/* Computes the partition of every row and its destination position in partition-major order.
 * out_counts[nparts] (host) = rows per partition.  The handle keeps dest[] / part_of_row[] on the device for the scatters. */
int32_t pg_partition_create(pg_ctx* ctx, const pg_col* keys, int32_t nkeys, const pg_col* hash_col, int64_t n, int32_t nparts, int32_t mode,
                            int64_t* out_counts, pg_partitioner** out) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        PG_CHECK_ARG(out && out_counts && (keys || hash_col) && n >= 0, "bad arguments");
        PG_CHECK_ARG(nparts >= 1 && nparts <= MAX_PARTS, "between 1 and %d partitions are supported", MAX_PARTS);
        PG_CHECK_ARG(mode == 0 || (nparts & (nparts - 1)) == 0, "the local partition function needs a power-of-two partition count");
        PG_CHECK_ARG(n <= 0x7fffffffLL, "page too large");
        PG_CHECK_ARG(nkeys >= 0 && nkeys <= MAX_KEYS, "too many key channels");
        pg_partitioner* pt = new pg_partitioner();
        try {
            pt->ctx = ctx; pt->nparts = nparts; pt->mode = mode; pt->n = n;
            for (int p = 0; p < nparts; p++) { pt->part_counts[p] = 0; pt->part_start[p] = 0; }
            if (n > 0) {
                KeyCols kc; memset(&kc, 0, sizeof(kc));
                std::vector<ColImport> imp(nkeys);
                ColImport himp;
                DCol hc; memset(&hc, 0, sizeof(hc));
                int has_hash = 0;
                if (hash_col) { import_col(ctx, *hash_col, n, himp); hc = himp.d; has_hash = 1; }
                else for (int k = 0; k < nkeys; k++) { import_col(ctx, keys[k], n, imp[k]); kc.c[k] = imp[k].d; }
                int64_t ntiles = (n + PART_TILE - 1) / PART_TILE;
                pt->part_of_row.alloc(ctx, n);
                pt->counts.alloc(ctx, ((int64_t)nparts * ntiles + 1) * (int64_t)sizeof(Pair64));
                Pair64* counts = pt->counts.as<Pair64>();
                if (!has_hash && nkeys == 1 && kc.c[0].type == PG_BIGINT && !kc.c[0].nulls)
                    PG_LAUNCH(ctx, part_hist_bigint_kernel, (int)ntiles, PART_THREADS, 0, (const int64_t*)kc.c[0].values, n, nparts, mode,
                              pt->part_of_row.as<uint8_t>(), counts, ntiles);
                else
                    PG_LAUNCH(ctx, part_hist_kernel, (int)ntiles, PART_THREADS, 0, kc, nkeys, hc, has_hash, n, nparts, mode,
                              pt->part_of_row.as<uint8_t>(), counts, ntiles);
                scan_tile_totals(ctx, counts, (int64_t)nparts * ntiles, counts + (int64_t)nparts * ntiles);
                // partition starts = scanned count of each partition's first tile
                std::vector<int64_t> starts(nparts);
                for (int p = 0; p < nparts; p++) copy_async(ctx, &ctx->h_scalar[p], &counts[(int64_t)p * ntiles].a, 8, cudaMemcpyDeviceToHost);
                sync(ctx);
                for (int p = 0; p < nparts; p++) pt->part_start[p] = ctx->h_scalar[p];
                for (int p = 0; p < nparts; p++) pt->part_counts[p] = (p + 1 < nparts ? pt->part_start[p + 1] : n) - pt->part_start[p];
            }
            for (int p = 0; p < nparts; p++) out_counts[p] = pt->part_counts[p];
        } catch (...) { delete pt; throw; }
        *out = pt;
    });
}

int32_t pg_partition_destroy(pg_partitioner* pt) {
    if (!pt) return PG_OK;
    return guard(pt->ctx, [&] { delete pt; });
}

/* out[dest[i]] = in[i]: the column in partition-major order (fixed width types; device output) */
int32_t pg_partition_scatter(pg_partitioner* pt, const pg_col* in, pg_col_out* out) {
    if (!pt) return PG_E_INVALID_ARGUMENT;
    return guard(pt->ctx, [&] {
        pg_ctx* ctx = pt->ctx;
        PG_CHECK_ARG(in && out && in->n == pt->n && out->type == in->type, "bad arguments");
        PG_CHECK_ARG(!type_is_var(in->type), "variable width columns cannot be scattered");
        PG_CHECK_ARG(out->mem == PG_MEM_DEVICE && out->capacity >= pt->n, "output must be a device column with capacity");
        if (pt->n == 0) return;
        ColImport ci;
        import_col(ctx, *in, pt->n, ci);
        ensure_dest(pt);
        int w = type_width(in->type);
        int grid = grid_for(ctx, pt->n, 256, 2);
        const int32_t* dest = pt->dest.as<int32_t>();
        uint8_t* on = out->nulls;
        switch (w) {
            case 1: PG_LAUNCH(ctx, part_scatter_kernel<uint8_t>, grid, 256, 0, (const uint8_t*)ci.d.values, ci.d.nulls, dest, pt->n, (uint8_t*)out->values, on); break;
            case 4: PG_LAUNCH(ctx, part_scatter_kernel<int32_t>, grid, 256, 0, (const int32_t*)ci.d.values, ci.d.nulls, dest, pt->n, (int32_t*)out->values, on); break;
            case 8: PG_LAUNCH(ctx, part_scatter_kernel<int64_t>, grid, 256, 0, (const int64_t*)ci.d.values, ci.d.nulls, dest, pt->n, (int64_t*)out->values, on); break;
            default: PG_LAUNCH(ctx, part_scatter_kernel<int4>, grid, 256, 0, (const int4*)ci.d.values, ci.d.nulls, dest, pt->n, (int4*)out->values, on);
        }
        sync(ctx);
    });
}

/* Fused exchange: scatter the column straight into the receivers' buffers over NVLink peer mappings.
 * peer_values[p] / peer_nulls[p]: device pointers valid on THIS device that alias rank p's receive column (P2P / symmetric
 * memory); peer_row_offset[p]: first row this rank may write in rank p's buffer (= rows sent to p by lower ranks). */
int32_t pg_partition_scatter_peers(pg_partitioner* pt, const pg_col* in, void* const* peer_values, uint8_t* const* peer_nulls,
                                   const int64_t* peer_row_offset) {
    if (!pt) return PG_E_INVALID_ARGUMENT;
    return guard(pt->ctx, [&] {
        pg_ctx* ctx = pt->ctx;
        PG_CHECK_ARG(in && peer_values && peer_row_offset && in->n == pt->n, "bad arguments");
        PG_CHECK_ARG(!type_is_var(in->type), "variable width columns cannot be scattered");
        if (pt->n == 0) return;
        ColImport ci;
        import_col(ctx, *in, pt->n, ci);
        ensure_dest(pt);
        ScatterPeers sp;
        memset(&sp, 0, sizeof(sp));
        for (int p = 0; p < pt->nparts; p++) {
            sp.base[p] = peer_values[p]; sp.nulls[p] = peer_nulls ? peer_nulls[p] : nullptr;
            sp.part_start[p] = pt->part_start[p]; sp.peer_row_offset[p] = peer_row_offset[p];
        }
        int w = type_width(in->type);
        const int32_t* dest = pt->dest.as<int32_t>();
        const uint8_t* por = pt->part_of_row.as<uint8_t>();
        const int64_t ntiles = (pt->n + PART_TILE - 1) / PART_TILE;
        const Pair64* offs = pt->counts.as<Pair64>();
        const size_t smem = (size_t)PART_TILE * (w + 1);
        if (getenv("PRESTOGPU_PEER_SCATTER_DIRECT")) {
            int grid = grid_for(ctx, pt->n, 256, 2);
            switch (w) {
                case 1: PG_LAUNCH(ctx, part_scatter_peer_kernel<uint8_t>, grid, 256, 0, (const uint8_t*)ci.d.values, ci.d.nulls, dest, por, pt->n, sp); break;
                case 4: PG_LAUNCH(ctx, part_scatter_peer_kernel<int32_t>, grid, 256, 0, (const int32_t*)ci.d.values, ci.d.nulls, dest, por, pt->n, sp); break;
                case 8: PG_LAUNCH(ctx, part_scatter_peer_kernel<int64_t>, grid, 256, 0, (const int64_t*)ci.d.values, ci.d.nulls, dest, por, pt->n, sp); break;
                default: PG_LAUNCH(ctx, part_scatter_peer_kernel<int4>, grid, 256, 0, (const int4*)ci.d.values, ci.d.nulls, dest, por, pt->n, sp);
            }
        } else {
            switch (w) {
                case 1:
                    PG_LAUNCH(ctx, part_scatter_staged_kernel<uint8_t>, (int)ntiles, PART_THREADS, smem, (const uint8_t*)ci.d.values, ci.d.nulls, dest, por,
                              pt->n, pt->nparts, offs, ntiles, sp);
                    break;
                case 4:
                    PG_LAUNCH(ctx, part_scatter_staged_kernel<int32_t>, (int)ntiles, PART_THREADS, smem, (const int32_t*)ci.d.values, ci.d.nulls, dest, por,
                              pt->n, pt->nparts, offs, ntiles, sp);
                    break;
                case 8:
                    PG_LAUNCH(ctx, part_scatter_staged_kernel<int64_t>, (int)ntiles, PART_THREADS, smem, (const int64_t*)ci.d.values, ci.d.nulls, dest, por,
                              pt->n, pt->nparts, offs, ntiles, sp);
                    break;
                default: {
                    static bool attr_set = false;
                    if (!attr_set) {
                        PG_CUDA(cudaFuncSetAttribute((const void*)part_scatter_staged_kernel<int4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                        attr_set = true;
                    }
                    PG_LAUNCH(ctx, part_scatter_staged_kernel<int4>, (int)ntiles, PART_THREADS, smem, (const int4*)ci.d.values, ci.d.nulls, dest, por, pt->n,
                              pt->nparts, offs, ntiles, sp);
                }
            }
        }
        sync(ctx);
    });
}

/* Fused exchange of a whole page: every (fixed width) column is stored into the receivers' buffers by ONE kernel
 * (part_scatter_all_kernel).  peer_values / peer_nulls: ncols x nparts pointers, column-major ([c * nparts + p]); a column whose
 * peer_nulls entries are null travels without a mask.  Stream-ordered: returns after the launch (the caller's barrier on the same
 * stream orders the stores before the receivers' reads). */
int32_t pg_partition_scatter_peers_all(pg_partitioner* pt, const pg_col* cols, int32_t ncols, void* const* peer_values, uint8_t* const* peer_nulls,
                                       const int64_t* peer_row_offset) {
    if (!pt) return PG_E_INVALID_ARGUMENT;
    return guard(pt->ctx, [&] {
        pg_ctx* ctx = pt->ctx;
        PG_CHECK_ARG(cols && peer_values && peer_row_offset && ncols >= 1 && ncols <= SC_MAX_COLS, "between 1 and %d columns per call", SC_MAX_COLS);
        if (pt->n == 0) return;
        std::vector<ColImport> imp(ncols);
        ScatterTable tab;
        memset(&tab, 0, sizeof(tab));
        tab.ncols = ncols; tab.nparts = pt->nparts;
        int maxw = 1;
        for (int c = 0; c < ncols; c++) {
            PG_CHECK_ARG(cols[c].n == pt->n, "column %d has %lld rows, the partitioner %lld", c, (long long)cols[c].n, (long long)pt->n);
            PG_CHECK_ARG(!type_is_var(cols[c].type), "variable width columns cannot be scattered");
            import_col(ctx, cols[c], pt->n, imp[c]);
            tab.in[c] = imp[c].d.values; tab.in_nulls[c] = imp[c].d.nulls; tab.width[c] = type_width(cols[c].type);
            maxw = std::max(maxw, tab.width[c]);
            bool with_nulls = false;        // a column travels with a mask when any receiver gave a mask buffer (empty partitions may pass NULL)
            for (int p = 0; peer_nulls && p < pt->nparts; p++) with_nulls = with_nulls || peer_nulls[(size_t)c * pt->nparts + p] != nullptr;
            for (int p = 0; p < pt->nparts; p++) {
                tab.base[c][p] = peer_values[(size_t)c * pt->nparts + p];
                tab.nulls[c][p] = with_nulls ? peer_nulls[(size_t)c * pt->nparts + p] : nullptr;
                PG_CHECK_ARG(pt->part_counts[p] == 0 || (tab.base[c][p] && (!with_nulls || tab.nulls[c][p])), "column %d: missing receive buffer of partition %d", c, p);
            }
        }
        for (int p = 0; p < pt->nparts; p++) { tab.part_start[p] = pt->part_start[p]; tab.peer_row_offset[p] = peer_row_offset[p]; }
        DevBuf own_tab;
        const ScatterTable* d_tab = (const ScatterTable*)in_arena_stage(ctx, &tab, sizeof(tab));     // reaches the device with the launch's flush
        if (!d_tab) {
            own_tab.alloc(ctx, sizeof(tab));
            copy_async(ctx, own_tab.p, &tab, sizeof(tab), cudaMemcpyHostToDevice);
            d_tab = (const ScatterTable*)own_tab.p;
        }
        const int64_t ntiles = (pt->n + PART_TILE - 1) / PART_TILE;
        const size_t smem = (size_t)PART_TILE * (maxw + 1);
        // thread layout / register budget, 300 M rows x 3 columns into 8 local buffers, steady state: 256 threads x 16 rows per thread with
        // the rows' places in registers 6.6 ms at 64 registers (spills), 11.3 ms at 128 (two blocks per SM); places in shared memory and
        // 512 threads x 8 rows: 5.1 ms at 64 registers, 4.2 ms at 40 (three blocks = 75 % of the warp slots), 4.5 ms at 32.  The kernel
        // waits between its block-wide phases: resident warps are what it needs.  PRESTOGPU_SCATTER_SHAPE: 0 256x4, 1 512x2, 2 1024x1,
        // 3 512x3 (default), 4 512x4, 5 1024x2, 6 256x6
        int shape = 3;
        if (const char* e = getenv("PRESTOGPU_SCATTER_SHAPE")) shape = atoi(e);
#define PG_SCATTER_ALL(T, M)                                                                                                          \
        do {                                                                                                                           \
            static bool prepared = false;                                                                                              \
            if (!prepared) {                                                                                                           \
                PG_CUDA(cudaFuncSetAttribute((const void*)part_scatter_all_kernel<T, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(PART_TILE * 17))); \
                prepared = true;                                                                                                       \
            }                                                                                                                          \
            auto part_scatter_all_kernel_ = part_scatter_all_kernel<T, M>;                                                             \
            PG_LAUNCH(ctx, part_scatter_all_kernel_, (int)ntiles, T, smem, pt->part_of_row.as<uint8_t>(), pt->n, pt->nparts, pt->counts.as<Pair64>(), ntiles, d_tab); \
        } while (0)
        if (shape == 0) PG_SCATTER_ALL(256, 4);
        else if (shape == 2) PG_SCATTER_ALL(1024, 1);
        else if (shape == 3) PG_SCATTER_ALL(512, 3);
        else if (shape == 4) PG_SCATTER_ALL(512, 4);
        else if (shape == 5) PG_SCATTER_ALL(1024, 2);
        else if (shape == 6) PG_SCATTER_ALL(256, 6);
        else PG_SCATTER_ALL(512, 2);
#undef PG_SCATTER_ALL
    });
}

/* partition id of every row (uint8) — parity checks against LocalPartitionGenerator / HashGenerator.getPartition */
int32_t pg_partition_ids(pg_partitioner* pt, uint8_t* out, int32_t out_mem) {
    if (!pt || !out) return PG_E_INVALID_ARGUMENT;
    return guard(pt->ctx, [&] {
        if (pt->n == 0) return;
        copy_async(pt->ctx, out, pt->part_of_row.p, pt->n, out_mem == PG_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost);
        sync(pt->ctx);
    });
}

}  // extern "C"
