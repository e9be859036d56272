// hashagg_wide.cu — host side of the many-groups GROUP BY path (see hashagg_wide.h, jit/wide_body.cuh).
//   wide_add_page        one specialised kernel per page: probe/claim + atomicMax(first row) + every aggregate's update
//   grow                 when a block ran out of its claim quota: bigger table, buckets re-inserted, left-out rows re-run
//   wide_prepare_output  first-seen numbering without a sort: one bit per input row marks "a group was first seen here",
//                        prefix popcounts turn the bit position into the group id, then keys and states are scattered
//                        into the dense per-group arrays the existing output code (groupby.cu export, accum_eval.cu) reads.
#include "hashagg_wide.h"

#include <stdio.h>
#include <algorithm>
#include <stdlib.h>

#include "jit.h"
#include "jit/wide_args.h"
#include "jit_embedded_gen.h"

namespace pg {

namespace {

enum { IN_NONE = 0, IN_I64 = 1, IN_I32 = 2, IN_F64 = 3, IN_I2F = 4, IN_DEC128 = 5 };
const char* in_name(int k) { static const char* n[] = {"NONE", "I64", "I32", "F64", "I2F", "DEC128"}; return n[k]; }
const char* op_name(int op) {
    switch (op) { case OP_COUNT: return "CNT"; case OP_SUM_I64: return "SUMI64"; case OP_SUM_F64: return "SUMF64"; default: return "SUMDEC"; }
}
const int WIDE_THREADS = 256;
const int WIDE_CHUNK = 32;     // buckets per chunk of the front region (jit/wide_body.cuh: PGW_CHUNK)

__device__ __forceinline__ unsigned long long w_fmix64(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return k;
}

// re-insert every occupied bucket of the old table into the (zeroed) new one; the two extra buckets are copied as they are
__global__ void wide_rebuild_kernel(const unsigned long long* old_table, int64_t old_cap, unsigned long long* new_table, int64_t new_cap, int bw) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < old_cap + 2; i += stride) {
        const unsigned long long* src = old_table + i * bw;
        unsigned long long* dst;
        if (i >= old_cap) {
            if (src[1] == 0) continue;
            dst = new_table + (new_cap + (i - old_cap)) * bw;
        } else {
            const unsigned long long key = src[0];
            if (key == 0) continue;
            uint32_t idx = (uint32_t)(((w_fmix64(key) >> 32) * (unsigned long long)new_cap) >> 32);
            while (atomicCAS(new_table + (int64_t)idx * bw, 0ULL, key) != 0ULL) idx = idx + 1 == (uint32_t)new_cap ? 0u : idx + 1;
            dst = new_table + (int64_t)idx * bw;
        }
        for (int w = 1; w < bw; w++) dst[w] = src[w];
    }
}

// ---- output numbering.  Buckets are addressed by one index: [0, F) the front region, [F, F + cap) the table, then the two extra buckets
struct WideView {
    const unsigned long long* front;
    const unsigned long long* fkeys;      // the front buckets' packed keys: 8 bytes say whether a 64..160-byte bucket is in use
    const unsigned long long* table;
    int64_t F, cap;
    int bw;
    __device__ __forceinline__ const unsigned long long* at(int64_t i) const { return i < F ? front + i * bw : table + (i - F) * bw; }
    __device__ __forceinline__ int64_t total() const { return F + cap + 2; }
};
__global__ void wide_mark_first_rows_kernel(WideView v, uint32_t* bitmap) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < v.total(); i += stride) {
        if (i < v.F && v.fkeys[i] == 0) continue;
        const unsigned long long* b = v.at(i);
        if (i >= v.F && (i < v.F + v.cap ? b[0] == 0 : b[1] == 0)) continue;
        const unsigned long long row = ~b[1];
        atomicOr(bitmap + (row >> 5), 1u << (row & 31));
    }
}
// tile totals of popcounts (SCAN_TILE words per tile), then an exclusive prefix per word
__global__ void __launch_bounds__(SCAN_THREADS) wide_popc_tiles_kernel(const uint32_t* bitmap, int64_t nwords, Pair64* tiles) {
    __shared__ long long red[SCAN_THREADS / 32];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    long long s = 0;
    for (int j = threadIdx.x; j < SCAN_TILE; j += SCAN_THREADS) {
        int64_t w = base + j;
        if (w < nwords) s += __popc(bitmap[w]);
    }
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long t = 0;
        for (int k = 0; k < SCAN_THREADS / 32; k++) t += red[k];
        tiles[blockIdx.x].a = t; tiles[blockIdx.x].b = 0;
    }
}
// (the prefix and the word's bits are stored side by side: numbering a bucket then costs one random 8-byte read instead of two 4-byte ones)
__global__ void __launch_bounds__(SCAN_THREADS) wide_word_prefix_kernel(const uint32_t* bitmap, int64_t nwords, const Pair64* tile_prefix,
                                                                        uint2* word_prefix) {
    // SCAN_ITEMS consecutive words per thread, block-wide exclusive scan of the per-thread sums
    __shared__ long long warp_sums[SCAN_THREADS / 32];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int cnt[SCAN_ITEMS];
    long long mine = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        int64_t w = base + k;
        cnt[k] = w < nwords ? __popc(bitmap[w]) : 0;
        mine += cnt[k];
    }
    long long incl = mine;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { long long o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    long long warp_off = 0;
    for (int k = 0; k < wid; k++) warp_off += warp_sums[k];
    long long run = tile_prefix[blockIdx.x].a + warp_off + incl - mine;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        int64_t w = base + k;
        if (w < nwords) word_prefix[w] = make_uint2((uint32_t)run, bitmap[w]);
        run += cnt[k];
    }
}

struct GatherPlan {
    int32_t naggs, bw, flag_word, pad;
    int32_t op[PGW_MAX_AGGS], in_kind[PGW_MAX_AGGS], cnt_mode[PGW_MAX_AGGS], cnt_word[PGW_MAX_AGGS], val_word[PGW_MAX_AGGS], flag_bit[PGW_MAX_AGGS],
        words[PGW_MAX_AGGS];
    unsigned long long* state[PGW_MAX_AGGS];
};
// 192-bit two's complement accumulator helpers for recombining limb sums
__device__ __forceinline__ void add192(unsigned long long* w, unsigned long long a0, unsigned long long a1, unsigned long long a2) {
    unsigned long long s0 = w[0] + a0;
    unsigned long long c0 = s0 < a0 ? 1ULL : 0ULL;
    unsigned long long t1 = a1 + c0;
    unsigned long long c1 = t1 < c0 ? 1ULL : 0ULL;
    unsigned long long s1 = w[1] + t1;
    c1 += s1 < t1 ? 1ULL : 0ULL;
    w[0] = s0; w[1] = s1; w[2] = w[2] + a2 + c1;
}
__device__ __forceinline__ void add_limb(unsigned long long* w, unsigned long long x, int shift, bool is_signed) {
    const unsigned long long ext = is_signed ? (unsigned long long)((long long)x >> 63) : 0ULL;
    const unsigned long long up = is_signed ? (unsigned long long)((long long)x >> 32) : (x >> 32);
    switch (shift) {
        case 0: add192(w, x, ext, ext); break;
        case 32: add192(w, x << 32, up, ext); break;
        case 64: add192(w, 0, x, ext); break;
        default: add192(w, 0, x << 32, up); break;
    }
}
// one group's key and recombined states into the dense per-group arrays (row g)
__device__ __forceinline__ void wide_emit_group(const unsigned long long* b, unsigned long long key, int64_t g, const GatherPlan& p, int64_t* values_by_group) {
    values_by_group[g] = (int64_t)key;
    const unsigned long long flags = p.flag_word >= 0 ? b[p.flag_word] : 0;
    for (int a = 0; a < p.naggs; a++) {
        unsigned long long* st = p.state[a] + g * p.words[a];
        st[0] = p.cnt_mode[a] == 0 ? b[p.cnt_word[a]] : ((flags >> p.flag_bit[a]) & 1ULL);
        const unsigned long long* v = b + p.val_word[a];
        if (p.op[a] == OP_SUM_F64) st[1] = v[0];
        else if (p.op[a] != OP_COUNT) {
            unsigned long long w[3] = {0, 0, 0};
            if (p.in_kind[a] == IN_DEC128) { add_limb(w, v[0], 0, false); add_limb(w, v[1], 32, false); add_limb(w, v[2], 64, false); add_limb(w, v[3], 96, true); }
            else { add_limb(w, v[0], 0, false); add_limb(w, v[1], 32, true); }
            st[1] = w[0]; st[2] = w[1];
            if (p.op[a] == OP_SUM_DEC) st[3] = w[2];
        }
    }
}
// group id of every occupied bucket.  Buckets of the front region lie in (nearly) first-seen order when the rows arrived sorted,
// so their keys and states go straight to row gid of the dense arrays (a sequential sweep on both sides); buckets of the hashed
// table only record bucket_of_group[gid] and are fetched in group order by wide_gather_kernel (random reads, coalesced writes).
__global__ void wide_number_buckets_kernel(WideView v, GatherPlan p, const uint2* __restrict__ word_prefix,
                                           uint32_t* bucket_of_group, Slot* special, int64_t* values_by_group) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t cap = v.F + v.cap;
    for (; i < v.total(); i += stride) {
        if (i < v.F && v.fkeys[i] == 0) continue;
        const unsigned long long* b = v.at(i);
        if (i >= v.F && (i < cap ? b[0] == 0 : b[1] == 0)) continue;
        const unsigned long long row = ~b[1];
        const uint2 pw = __ldg(word_prefix + (row >> 5));
        const int64_t gid = (int64_t)pw.x + __popc(pw.y & ((1u << (row & 31)) - 1u));
        if (i < v.F) wide_emit_group(b, v.fkeys[i], gid, p, values_by_group);
        else bucket_of_group[gid] = (uint32_t)i;
        if (i == cap) special[1].gid = (int32_t)gid;     // the null group (BigintGroupByHash.java:215-222)
    }
}
// in group-id order: read the group's bucket (one random 64..160 byte read), write keys and dense states (coalesced)
__global__ void wide_gather_kernel(WideView v, GatherPlan p, const uint32_t* bucket_of_group, int64_t groups,
                                   int64_t* values_by_group) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t cap = v.F + v.cap;
    for (; g < groups; g += stride) {
        const int64_t i = bucket_of_group[g];
        if (i == 0xFFFFFFFFLL) continue;                 // a front bucket: already written by wide_number_buckets_kernel
        const unsigned long long* b = v.at(i);
        wide_emit_group(b, i >= cap ? 0ULL : b[0], g, p, values_by_group);
    }
}

// ---- reordering a page for locality (rows in random key order against a table far larger than the L2).
// A row's bucket index is ((fmix64(key) >> 32) * cap) >> 32, monotone in the hash, so the top bits of fmix64(key) name a contiguous
// slice of the table.  The page's columns are rewritten partition-major by those bits (a counting pass and a scattering pass, order
// inside a partition left to the atomics): the fused kernel then works through one L2-sized slice of the table at a time instead of
// taking an HBM (and TLB) miss per row.  First-seen order survives because every row keeps its page index (row_ids).
const int RE_THREADS = 512, RE_ITEMS = 16, RE_TILE = RE_THREADS * RE_ITEMS, RE_MAX_COLS = 72, RE_MAX_P = 1024;
struct ReorderCols {
    int32_t ncols, pad;
    const void* in[RE_MAX_COLS];
    void* out[RE_MAX_COLS];
    int32_t width[RE_MAX_COLS];      // 1, 4 or 8 (a 16-byte column is listed as two 8-byte halves with stride 2)
    int32_t stride[RE_MAX_COLS];     // elements of `width` bytes per row
    int32_t offset[RE_MAX_COLS];     // element inside the row
};
__device__ __forceinline__ int re_part(const long long* keys, const uint8_t* key_nulls, int64_t row, int shift) {
    if (key_nulls && key_nulls[row]) return 0;
    return (int)(w_fmix64((unsigned long long)__ldg(keys + row)) >> shift);
}
__global__ void __launch_bounds__(RE_THREADS) wide_reorder_hist_kernel(const long long* keys, const uint8_t* key_nulls, int64_t n, int shift, int P,
                                                                       unsigned long long* hist) {
    __shared__ unsigned int h[RE_MAX_P];
    for (int t = threadIdx.x; t < P; t += RE_THREADS) h[t] = 0;
    __syncthreads();
    const int64_t ntiles = (n + RE_TILE - 1) / RE_TILE;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * RE_TILE;
#pragma unroll 8
        for (int j = 0; j < RE_ITEMS; j++) {
            const int64_t row = base + (int64_t)j * RE_THREADS + threadIdx.x;
            if (row < n) atomicAdd(&h[re_part(keys, key_nulls, row, shift)], 1u);
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < P; t += RE_THREADS) if (h[t]) atomicAdd(&hist[t], (unsigned long long)h[t]);
}
// exclusive prefix of the P partition sizes -> first destination row of every partition (one block of RE_MAX_P threads)
__global__ void __launch_bounds__(RE_MAX_P) wide_reorder_scan_kernel(const unsigned long long* hist, int P, unsigned long long* cursor) {
    __shared__ unsigned long long warp_tot[RE_MAX_P / 32];
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const unsigned long long v = t < P ? hist[t] : 0;
    unsigned long long incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    unsigned long long off = 0;
    for (int k = 0; k < wid; k++) off += warp_tot[k];
    if (t < P) cursor[t] = off + incl - v;
}
// One tile of RE_TILE rows per block.  The tile's rows are ranked per partition with shared-memory atomics, the tile reserves a
// run of destination rows in every partition with one global atomic each, and every column then passes through a shared-memory
// staging area laid out partition-major, so that consecutive threads store consecutive destination rows (runs of
// RE_TILE / P rows: whole sectors instead of one partial sector per row).
template <typename T>
__device__ __forceinline__ void re_move_column(const T* __restrict__ in, T* __restrict__ out, int stride, int offset, int64_t base, int rows,
                                               const unsigned short* lpos, T* stage, const uint32_t* s_dest) {
#pragma unroll
    for (int j = 0; j < RE_ITEMS; j++) {
        const int i = j * RE_THREADS + threadIdx.x;
        if (i < rows) stage[lpos[j]] = in[(base + i) * stride + offset];
    }
    __syncthreads();
    for (int l = threadIdx.x; l < rows; l += RE_THREADS) out[(int64_t)s_dest[l] * stride + offset] = stage[l];
    __syncthreads();
}
__global__ void __launch_bounds__(RE_THREADS, 2) wide_reorder_scatter_kernel(ReorderCols cols, const long long* keys, const uint8_t* key_nulls, int64_t n,
                                                                          int shift, int P, unsigned long long* cursor, uint32_t* row_ids) {
    extern __shared__ __align__(16) unsigned char re_smem[];
    unsigned long long* stage = (unsigned long long*)re_smem;                  // RE_TILE x 8 bytes
    uint32_t* s_dest = (uint32_t*)(re_smem + (size_t)RE_TILE * 8);             // destination row of every staged place
    unsigned int* cnt = s_dest + RE_TILE;                                      // P: rows of the tile per partition, then its first staged place
    unsigned int* gfirst = cnt + RE_MAX_P;                                     // P: first destination row reserved for the tile
    __shared__ unsigned int warp_tot[RE_THREADS / 32];
    for (int t = threadIdx.x; t < P; t += RE_THREADS) cnt[t] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RE_TILE;
    const int rows = (int)((n - base) < RE_TILE ? (n - base) : RE_TILE);
    uint32_t pr[RE_ITEMS];       // partition << 16 | rank inside the tile's rows of that partition
#pragma unroll
    for (int j = 0; j < RE_ITEMS; j++) {
        const int i = j * RE_THREADS + threadIdx.x;
        pr[j] = 0;
        if (i < rows) {
            const int p = re_part(keys, key_nulls, base + i, shift);
            pr[j] = ((uint32_t)p << 16) | atomicAdd(&cnt[p], 1u);
        }
    }
    __syncthreads();
    // exclusive scan of cnt[0..P) (two bins per thread), and the tile's reservation in every partition
    {
        const int t0 = threadIdx.x * 2;
        const unsigned int c0 = t0 < P ? cnt[t0] : 0u, c1 = t0 + 1 < P ? cnt[t0 + 1] : 0u;
        if (c0) gfirst[t0] = (unsigned int)atomicAdd(&cursor[t0], (unsigned long long)c0);
        if (c1) gfirst[t0 + 1] = (unsigned int)atomicAdd(&cursor[t0 + 1], (unsigned long long)c1);
        const unsigned int mine = c0 + c1;
        unsigned int incl = mine;
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        unsigned int off = 0;
        for (int k = 0; k < wid; k++) off += warp_tot[k];
        if (t0 < P) cnt[t0] = off + incl - mine;
        if (t0 + 1 < P) cnt[t0 + 1] = off + incl - mine + c0;
    }
    __syncthreads();
    unsigned short lpos[RE_ITEMS];
    uint32_t* stage32 = (uint32_t*)stage;
#pragma unroll
    for (int j = 0; j < RE_ITEMS; j++) {
        const int i = j * RE_THREADS + threadIdx.x;
        lpos[j] = 0;
        if (i < rows) {
            const uint32_t p = pr[j] >> 16, r = pr[j] & 0xffffu;
            const uint32_t l = cnt[p] + r;
            lpos[j] = (unsigned short)l;
            s_dest[l] = gfirst[p] + r;
            stage32[l] = (uint32_t)(base + i);
        }
    }
    __syncthreads();
    for (int l = threadIdx.x; l < rows; l += RE_THREADS) row_ids[s_dest[l]] = stage32[l];
    __syncthreads();
    for (int c = 0; c < cols.ncols; c++) {
        switch (cols.width[c]) {
            case 1: re_move_column<uint8_t>((const uint8_t*)cols.in[c], (uint8_t*)cols.out[c], cols.stride[c], cols.offset[c], base, rows, lpos, (uint8_t*)stage, s_dest); break;
            case 4: re_move_column<uint32_t>((const uint32_t*)cols.in[c], (uint32_t*)cols.out[c], cols.stride[c], cols.offset[c], base, rows, lpos, (uint32_t*)stage, s_dest); break;
            default: re_move_column<unsigned long long>((const unsigned long long*)cols.in[c], (unsigned long long*)cols.out[c], cols.stride[c], cols.offset[c], base, rows, lpos, stage, s_dest);
        }
    }
}
const size_t RE_SMEM = (size_t)RE_TILE * 8 + (size_t)RE_TILE * 4 + 2 * RE_MAX_P * 4;

struct Reordered {
    std::vector<DevBuf> bufs;
    DevBuf row_ids;
    DCol key;
    std::vector<DCol> args, masks;
};
int in_kind_width(int k) { return k == IN_NONE ? 0 : k == IN_I32 ? 4 : k == IN_DEC128 ? 16 : 8; }

// true when the page was rewritten (out holds the reordered columns and the row ids)
bool reorder_for_locality(WideTable* w, const DCol& key, const DCol* args_cols, const DCol* masks, int64_t n, Reordered& out) {
    pg_ctx* ctx = w->ctx;
    const char* env = getenv("PRESTOGPU_WIDE_REORDER");
    if (env && env[0] == '0') return false;
    const int64_t table_bytes = w->cap * (int64_t)w->bucket_words * 8;
    int64_t min_rows = 4 << 20, slice = 32 << 20;
    if (const char* e = getenv("PRESTOGPU_WIDE_REORDER_MIN_ROWS")) min_rows = atoll(e);
    if (const char* e = getenv("PRESTOGPU_WIDE_REORDER_SLICE_MB")) slice = atoll(e) << 20;
    if (w->range || w->split || n < min_rows || n >= (1LL << 32) || table_bytes < 8 * slice) return false;
    int logp = 4;
    while (logp < 10 && (table_bytes >> logp) > slice) logp++;
    const int P = 1 << logp, shift = 64 - logp;
    ReorderCols rc;
    memset(&rc, 0, sizeof(rc));
    auto add = [&](const void* in, int width) -> void* {
        if (!in) return nullptr;
        const int pieces = width == 16 ? 2 : 1;          // a 16-byte column moves as two 8-byte halves
        PG_CHECK_STATE(rc.ncols + pieces <= RE_MAX_COLS, "too many columns to reorder");
        out.bufs.emplace_back(ctx, (n + 4) * (int64_t)width);
        void* o = out.bufs.back().p;
        for (int h = 0; h < pieces; h++) {
            rc.in[rc.ncols] = in; rc.out[rc.ncols] = o; rc.width[rc.ncols] = width == 16 ? 8 : width; rc.stride[rc.ncols] = pieces; rc.offset[rc.ncols] = h;
            rc.ncols++;
        }
        return o;
    };
    out.bufs.reserve(RE_MAX_COLS);
    out.key = key;
    out.key.values = add(key.values, 8);
    out.key.nulls = (const uint8_t*)add(key.nulls, 1);
    out.args.assign(args_cols, args_cols + w->aggs.size());
    out.masks.assign(masks, masks + w->aggs.size());
    for (size_t i = 0; i < w->aggs.size(); i++) {
        const int width = in_kind_width(w->aggs[i].in_kind);
        if (width) out.args[i].values = add(args_cols[i].values, width);
        out.args[i].nulls = (const uint8_t*)add(args_cols[i].nulls, 1);       // count(column) reads the null mask without the values
        out.masks[i].values = add(masks[i].values, 1);
        out.masks[i].nulls = (const uint8_t*)add(masks[i].nulls, 1);
    }
    out.row_ids.alloc(ctx, (n + 4) * 4);
    DevBuf hist(ctx, 2 * RE_MAX_P * 8);
    memset_async(ctx, hist.p, 0, 2 * RE_MAX_P * 8);
    unsigned long long* d_hist = hist.as<unsigned long long>();
    unsigned long long* d_cursor = d_hist + RE_MAX_P;
    const int64_t ntiles = (n + RE_TILE - 1) / RE_TILE;
    const int hgrid = (int)std::min<int64_t>(ntiles, (int64_t)ctx->sm_count * 8);
    PG_LAUNCH(ctx, wide_reorder_hist_kernel, hgrid, RE_THREADS, 0, (const long long*)key.values, key.nulls, n, shift, P, d_hist);
    PG_LAUNCH(ctx, wide_reorder_scan_kernel, 1, RE_MAX_P, 0, d_hist, P, d_cursor);
    static bool smem_set = false;
    if (!smem_set) {
        PG_CUDA(cudaFuncSetAttribute((const void*)wide_reorder_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RE_SMEM));
        smem_set = true;
    }
    PG_LAUNCH(ctx, wide_reorder_scatter_kernel, (int)ntiles, RE_THREADS, RE_SMEM, rc, (const long long*)key.values, key.nulls, n, shift, P, d_cursor,
              out.row_ids.as<uint32_t>());
    return true;
}

// ---- split mode (the default with a front region; PRESTOGPU_WIDE_SPLIT=0/1 overrides), first kernel: where does every row's state live?
// wide_locate_kernel reads ONLY the keys, finds or claims the bucket of every run of equal keys (the same warp-uniform rounds as
// jit/wide_body.cuh) and writes one 32-bit bucket number per row: [0, F) front region, [F, F + cap) table, F + cap the null key,
// F + cap + 1 key 0, WIDE_SKIP for rows this launch does not handle, WIDE_LEFT for rows left out because the table may not grow
// inside a launch.  The specialised update kernel (PGW_BIDX) then streams the bucket numbers and the arguments and only issues reds.
// Differences to the fused kernel's front probing, found with the ncu source page (4.5 claim rounds per warp trip because keys
// that arrive together go for the same free slot): the sector is chosen by multiplicative hashing (consecutive keys spread evenly
// over a chunk's 8 sectors), the free slot by a per-key rotation, and a lost claim moves on to the next slot of the rotation
// without reading the sector again.
static const uint32_t WIDE_SKIP = 0xFFFFFFFFu, WIDE_LEFT = 0xFFFFFFFEu;
struct LocateArgs {
    const long long* keys;
    const uint8_t* key_nulls;
    const uint8_t* only_rows;
    int64_t n;
    uint32_t* bidx;
    uint8_t* overflow_flag;
    unsigned long long* table;
    int64_t cap;
    int bw;
    int range, quota, vec_ok;
    unsigned long long* front_keys;
    int64_t F;
    long long k0;
    unsigned long long nseg;
    int shift;
    unsigned long long room;
    unsigned long long* counters;      // [0] table claims, [1] rows left out, [2] front claims
};
__device__ __forceinline__ ulonglong2 loc_ld_ca2(const unsigned long long* p) {
    ulonglong2 v; asm volatile("ld.global.ca.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p)); return v;
}
__device__ __forceinline__ ulonglong2 loc_ld_cg2(const unsigned long long* p) {
    ulonglong2 v; asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p)); return v;
}
__device__ __forceinline__ uint32_t wide_split_sector(unsigned long long key) { return (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 61) << 2; }
__global__ void __launch_bounds__(256, 3) wide_locate_kernel(const LocateArgs a) {
    const int V = 4;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * V;
    unsigned long long left_out = 0, claims = 0, fclaims = 0;
    for (int64_t r0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V; r0 - (int64_t)lane * V < a.n; r0 += stride) {
        unsigned long long key[V];
        bool valid[V], isnull[V], live[V];
        uint32_t res[V];
        if (a.vec_ok && r0 + V <= a.n) {
            const ulonglong2 k01 = __ldg((const ulonglong2*)(a.keys + r0)), k23 = __ldg((const ulonglong2*)(a.keys + r0 + 2));
            key[0] = k01.x; key[1] = k01.y; key[2] = k23.x; key[3] = k23.y;
            uint32_t kn = 0, on = 0x01010101u;
            if (a.key_nulls) kn = __ldg((const uint32_t*)(a.key_nulls + r0));
            if (a.only_rows) on = __ldg((const uint32_t*)(a.only_rows + r0));
#pragma unroll
            for (int j = 0; j < V; j++) { isnull[j] = ((kn >> (8 * j)) & 0xffu) != 0; valid[j] = ((on >> (8 * j)) & 0xffu) != 0; }
        } else {
#pragma unroll
            for (int j = 0; j < V; j++) {
                const int64_t r = r0 + j;
                const bool in = r < a.n;
                key[j] = in ? (unsigned long long)a.keys[r] : 0ULL;
                isnull[j] = in && a.key_nulls && a.key_nulls[r] != 0;
                valid[j] = in && (!a.only_rows || a.only_rows[r] != 0);
            }
        }
        // runs: the last row of a run of equal keys does the work, the rows before it copy its answer
#pragma unroll
        for (int j = 0; j < V; j++) {
            live[j] = valid[j];
            res[j] = WIDE_SKIP;
            if (j + 1 < V && valid[j] && valid[j + 1] && isnull[j] == isnull[j + 1] && (isnull[j] || key[j] == key[j + 1])) live[j] = false;
        }
        uint32_t idx[V], st[V];          // st: 0 settled, 1 walking its front chunk, 2 at a free front slot, 3 probing the table
        unsigned long long old[V];
#pragma unroll
        for (int j = 0; j < V; j++) {
            idx[j] = 0; st[j] = 0; old[j] = 1;
            if (!live[j]) continue;
            if (isnull[j] || key[j] == 0) res[j] = (uint32_t)(a.F + a.cap + (isnull[j] ? 0 : 1));
            else st[j] = 3;
        }
        if (a.range) {
            uint32_t fleft[V], usecg = 0;
            int tries[V];
#pragma unroll
            for (int j = 0; j < V; j++) {
                fleft[j] = WIDE_CHUNK / 4; tries[j] = 0;
                if (st[j] != 3) continue;
                const unsigned long long seg = (key[j] - (unsigned long long)a.k0) >> a.shift;
                if (seg < a.nseg) { st[j] = 1; idx[j] = (uint32_t)(seg * WIDE_CHUNK) + wide_split_sector(key[j]); }
            }
            while (true) {
                while (true) {
                    ulonglong2 ka[V], kb[V];
#pragma unroll
                    for (int j = 0; j < V; j++) {
                        if (st[j] != 1) continue;
                        const unsigned long long* sp = a.front_keys + (idx[j] & ~3u);
                        if (!((usecg >> j) & 1u)) { ka[j] = loc_ld_ca2(sp); kb[j] = loc_ld_ca2(sp + 2); }
                        else { ka[j] = loc_ld_cg2(sp); kb[j] = loc_ld_cg2(sp + 2); }
                    }
                    bool walking = false;
#pragma unroll
                    for (int j = 0; j < V; j++) {
                        if (st[j] != 1) continue;
                        const unsigned long long k0 = ka[j].x, k1 = ka[j].y, k2 = kb[j].x, k3 = kb[j].y, kk = key[j];
                        const int hit = k0 == kk ? 0 : k1 == kk ? 1 : k2 == kk ? 2 : k3 == kk ? 3 : -1;
                        const int rot = (int)(((kk * 0x9E3779B97F4A7C15ULL) >> 59) & 3);
                        const unsigned zmask = (k0 == 0 ? 1u : 0u) | (k1 == 0 ? 2u : 0u) | (k2 == 0 ? 4u : 0u) | (k3 == 0 ? 8u : 0u);
                        const unsigned zrot = ((zmask >> rot) | (zmask << (4 - rot))) & 15u;
                        if (hit >= 0) { res[j] = (idx[j] & ~3u) | (uint32_t)hit; st[j] = 0; }
                        else if (zrot) {
                            const int first = __ffs(zrot) - 1;               // position in the rotation of the first free slot
                            idx[j] = (idx[j] & ~3u) | (uint32_t)((first + rot) & 3);
                            tries[j] = 4 - first;                             // slots from there to the end of the rotation
                            st[j] = 2;
                        } else {
                            idx[j] = (idx[j] & ~(uint32_t)(WIDE_CHUNK - 1)) | (((idx[j] & ~3u) + 4) & (WIDE_CHUNK - 4));
                            if (--fleft[j] == 0) st[j] = 3; else walking = true;
                        }
                    }
                    if (!__any_sync(0xffffffffu, walking)) break;
                }
#pragma unroll
                for (int j = 0; j < V; j++) if (st[j] == 2) old[j] = atomicCAS(a.front_keys + idx[j], 0ULL, key[j]);
                bool again = false;
#pragma unroll
                for (int j = 0; j < V; j++) {
                    if (st[j] != 2) continue;
                    if (old[j] == 0 || old[j] == key[j]) { res[j] = idx[j]; st[j] = 0; if (old[j] == 0) fclaims++; }
                    else if (--tries[j] > 0) { idx[j] = (idx[j] & ~3u) | ((idx[j] + 1) & 3u); again = true; }    // next slot of the rotation, no re-read
                    else {                                                                                   // sector full meanwhile: next one
                        idx[j] = (idx[j] & ~(uint32_t)(WIDE_CHUNK - 1)) | (((idx[j] & ~3u) + 4) & (WIDE_CHUNK - 4));
                        if (--fleft[j] == 0) st[j] = 3; else { st[j] = 1; usecg |= 1u << j; }
                        again = true;
                    }
                }
                if (!__any_sync(0xffffffffu, again)) break;
            }
        }
        {
            bool allowed = true;
            uint32_t mine = 0;
            if (a.quota) allowed = __ldcg(a.counters) < a.room;
#pragma unroll
            for (int j = 0; j < V; j++) if (st[j] == 3) idx[j] = (uint32_t)(((w_fmix64(key[j]) >> 32) * (unsigned long long)a.cap) >> 32);
            while (true) {
#pragma unroll
                for (int j = 0; j < V; j++) {
                    if (st[j] != 3) continue;
                    unsigned long long* bk = a.table + (int64_t)idx[j] * a.bw;
                    old[j] = allowed ? atomicCAS(bk, 0ULL, key[j]) : __ldcg(bk);
                }
                bool more = false;
#pragma unroll
                for (int j = 0; j < V; j++) {
                    if (st[j] != 3) continue;
                    if (allowed && old[j] == 0) mine++;
                    if (old[j] == key[j] || (allowed && old[j] == 0)) { res[j] = (uint32_t)(a.F + idx[j]); st[j] = 0; }
                    else if (old[j] == 0) { res[j] = WIDE_LEFT; st[j] = 0; }
                    else { idx[j] = idx[j] + 1 == (uint32_t)a.cap ? 0u : idx[j] + 1; more = true; }
                }
                if (!__any_sync(0xffffffffu, more)) break;
            }
            if (a.quota) {
                const uint32_t all = __reduce_add_sync(0xffffffffu, mine);
                if (all && lane == 0) atomicAdd(a.counters, (unsigned long long)all);
            } else claims += mine;
        }
#pragma unroll
        for (int j = V - 2; j >= 0; j--) if (valid[j] && !live[j]) res[j] = res[j + 1];
#pragma unroll
        for (int j = 0; j < V; j++)
            if (valid[j] && res[j] == WIDE_LEFT) { left_out++; a.overflow_flag[r0 + j] = 1; }
        if (r0 + V <= a.n) *(uint4*)(a.bidx + r0) = make_uint4(res[0], res[1], res[2], res[3]);
        else {
#pragma unroll
            for (int j = 0; j < V; j++) if (r0 + j < a.n) a.bidx[r0 + j] = res[j];
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        left_out += __shfl_down_sync(0xffffffffu, left_out, d); claims += __shfl_down_sync(0xffffffffu, claims, d);
        fclaims += __shfl_down_sync(0xffffffffu, fclaims, d);
    }
    if (lane == 0) {
        if (left_out) atomicAdd(a.counters + 1, left_out);
        if (claims) atomicAdd(a.counters, claims);
        if (fclaims) atomicAdd(a.counters + 2, fclaims);
    }
}

// distinct keys among the first rows: one block, open addressing in shared memory
__global__ void __launch_bounds__(1024) wide_sample_distinct_kernel(const long long* keys, const uint8_t* nulls, int n, unsigned long long* out) {
    __shared__ unsigned long long set[4096];   // sample <= 2048 keys: load factor <= 0.5
    __shared__ unsigned int distinct;
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) set[i] = 0;
    if (threadIdx.x == 0) distinct = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (nulls && nulls[i]) continue;
        unsigned long long k = (unsigned long long)keys[i];
        if (k == 0) continue;
        uint32_t h = (uint32_t)(w_fmix64(k) >> 52);
        while (true) {
            unsigned long long old = atomicCAS(&set[h], 0ULL, k);
            if (old == 0) { atomicAdd(&distinct, 1u); break; }
            if (old == k) break;
            h = (h + 1) & 4095;
        }
    }
    // rows that repeat the key of the row before them: input clustered by key (TPC-H lineitem by order key);
    // rows whose key is not below their predecessor's: input (nearly) sorted by key; the sample's key range gives the key density
    unsigned int rep = 0, asc = 0;
    long long lo = 0x7fffffffffffffffLL, hi = -0x7fffffffffffffffLL - 1;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (nulls && nulls[i]) continue;
        const long long k = keys[i];
        lo = k < lo ? k : lo; hi = k > hi ? k : hi;
        if (i > 0 && !(nulls && nulls[i - 1])) { rep += k == keys[i - 1] ? 1u : 0u; asc += k >= keys[i - 1] ? 1u : 0u; }
    }
    __shared__ unsigned int repeats, ascending;
    __shared__ long long s_lo, s_hi;
    if (threadIdx.x == 0) { repeats = 0; ascending = 0; s_lo = 0x7fffffffffffffffLL; s_hi = -0x7fffffffffffffffLL - 1; }
    __syncthreads();
    if (rep) atomicAdd(&repeats, rep);
    if (asc) atomicAdd(&ascending, asc);
    atomicMin(&s_lo, lo); atomicMax(&s_hi, hi);
    __syncthreads();
    if (threadIdx.x == 0) { out[0] = distinct; out[1] = repeats; out[2] = (unsigned long long)s_lo; out[3] = (unsigned long long)s_hi; out[4] = ascending; }
}

int64_t pick_cap(int64_t groups) {
    int64_t cap = (int64_t)((double)groups / 0.55) + 1024;
    if (cap < (1 << 16)) cap = 1 << 16;
    return cap;
}

void alloc_table(WideTable* w, int64_t cap) {
    PG_CHECK_STATE(cap < (1LL << 31), "hash table of %lld buckets exceeds the supported size", (long long)cap);   // GENERIC_INSUFFICIENT_RESOURCES analogue
    w->table.alloc(w->ctx, (cap + 2) * w->bucket_words * 8);
    memset_async(w->ctx, w->table.p, 0, w->table.bytes);
    w->cap = cap;
    w->max_fill = (int64_t)(cap * 0.75);
}

void grow(WideTable* w, int64_t min_groups) {
    pg_ctx* ctx = w->ctx;
    int64_t new_cap = pick_cap(min_groups > 2 * w->claimed ? min_groups : 2 * w->claimed);
    if (new_cap < 2 * w->cap) new_cap = 2 * w->cap;
    DevBuf old = std::move(w->table);
    int64_t old_cap = w->cap;
    alloc_table(w, new_cap);
    PG_LAUNCH(ctx, wide_rebuild_kernel, grid_for(ctx, old_cap + 2, 256), 256, 0, old.as<unsigned long long>(), old_cap,
              w->table.as<unsigned long long>(), new_cap, w->bucket_words);
}

}  // namespace

int64_t sample_distinct_keys(pg_ctx* ctx, const DCol& key, int64_t sample, int64_t* adjacent_repeats, KeySample* more) {
    if (sample > 2048) sample = 2048;
    DevBuf out(ctx, 40);
    PG_LAUNCH(ctx, wide_sample_distinct_kernel, 1, 1024, 0, (const long long*)key.values, key.nulls, (int)sample, out.as<unsigned long long>());
    int64_t d[5];
    read_scalars(ctx, out.p, 5, d);
    if (adjacent_repeats) *adjacent_repeats = d[1];
    if (more) { more->distinct = d[0]; more->repeats = d[1]; more->min_key = d[2]; more->max_key = d[3]; more->ascending = d[4]; more->rows = sample; }
    return d[0];
}

bool wide_supported(const std::vector<pg_agg_spec>& specs, const std::vector<pg_agg*>& aggs) {
    if (aggs.empty() || aggs.size() > PGW_MAX_AGGS) return false;
    for (size_t i = 0; i < aggs.size(); i++) {
        const pg_agg* a = aggs[i];
        const bool is32 = a->in_type == PG_INTEGER || a->in_type == PG_DATE;
        switch (a->op) {
            case OP_COUNT: break;
            case OP_SUM_I64: break;
            case OP_SUM_F64: if (is32) return false; break;
            case OP_SUM_DEC: break;
            default: return false;
        }
        (void)specs;
    }
    return true;
}

bool wide_range_applicable(const KeySample& s) {
    if (s.rows < 1024 || s.distinct < 128 || s.max_key <= s.min_key) return false;
    if (s.ascending * 100 < (s.rows - 1) * 98) return false;                                // not sorted by key
    const double density = (double)s.distinct / ((double)s.max_key - (double)s.min_key + 1.0);
    return density >= 1.0 / 4096.0;
}

WideTable* wide_create(pg_ctx* ctx, const std::vector<pg_agg_spec>& specs, const std::vector<pg_agg*>& aggs, int32_t step, int64_t expected_groups,
                       const KeySample* sorted_sample) {
    WideTable* w = new WideTable();
    try {
        w->ctx = ctx;
        int words = 2, flags = 0;
        // flag bits first (one shared word), then the 64-bit counts and the value words
        for (size_t i = 0; i < aggs.size(); i++) {
            const pg_agg* a = aggs[i];
            WideAgg wa;
            memset(&wa, 0, sizeof(wa));
            wa.op = a->op; wa.words = a->words; wa.skip_null = a->kind != PG_AGG_COUNT ? 1 : 0;
            const bool is32 = a->in_type == PG_INTEGER || a->in_type == PG_DATE;
            switch (a->op) {
                case OP_COUNT: wa.in_kind = IN_NONE; break;
                case OP_SUM_I64: wa.in_kind = is32 ? IN_I32 : IN_I64; break;
                case OP_SUM_F64: wa.in_kind = (a->kind == PG_AGG_AVG && a->in_type == PG_BIGINT) ? IN_I2F : IN_F64; break;
                default: wa.in_kind = a->in_type == PG_LONG_DECIMAL ? IN_DEC128 : IN_I64;
            }
            // a plain sum only needs to know whether any row arrived (NULL result otherwise); partial steps ship real counts
            wa.val_words = a->op == OP_COUNT ? 0 : a->op == OP_SUM_F64 ? 1 : wa.in_kind == IN_DEC128 ? 4 : 2;
            wa.cnt_mode = (step == PG_STEP_SINGLE && a->kind == PG_AGG_SUM) ? 1 : 0;
            if (wa.cnt_mode == 1) wa.flag_bit = flags++;
            w->aggs.push_back(wa);
        }
        if (flags > 0) w->flag_word = words++;
        for (WideAgg& wa : w->aggs) {
            if (wa.cnt_mode == 0) wa.cnt_word = words++;
            wa.val_word = words;
            words += wa.val_words;
        }
        w->bucket_words = (words + 3) & ~3;
        w->counters.alloc(ctx, 32);
        const char* renv = getenv("PRESTOGPU_WIDE_RANGE");
        if (sorted_sample && expected_groups > 0 && !(renv && renv[0] == '0') && wide_range_applicable(*sorted_sample)) {
            // between 11 and 22 keys per chunk of 32 buckets at the sample's key density (a chunk that fills up spills into the table);
            // the region spans the key range `expected_groups` keys need
            const double density = (double)sorted_sample->distinct / ((double)sorted_sample->max_key - (double)sorted_sample->min_key + 1.0);
            const double width = (WIDE_CHUNK * 0.6875) / density;
            int shift = 0;
            while (shift < 40 && (double)(2ULL << shift) <= width) shift++;      // largest power of two <= width
            const double span = (double)expected_groups / density * 1.02;
            const int64_t nseg = (int64_t)(span / (double)(1ULL << shift)) + 2;
            const int64_t fb = nseg * WIDE_CHUNK;
            if (fb < (1LL << 31) && (double)fb <= 2.5 * ((double)expected_groups / 0.55)) {
                w->range = true; w->k0 = sorted_sample->min_key; w->shift = shift; w->nseg = nseg; w->front_buckets = fb;
                w->front.alloc(ctx, fb * w->bucket_words * 8);
                memset_async(ctx, w->front.p, 0, w->front.bytes);
                w->front_keys.alloc(ctx, fb * 8);
                memset_async(ctx, w->front_keys.p, 0, w->front_keys.bytes);
            }
        }
        // two kernels (locate, then update) when the accesses are sequential - measured 1.9 ms against 2.35 ms fused on lineitem SF10 in key
        // order; one fused kernel when every row is a random access anyway (12.8 ms against 16.6 ms).  PRESTOGPU_WIDE_SPLIT=0/1 overrides.
        w->split = w->range;
        if (const char* e = getenv("PRESTOGPU_WIDE_SPLIT")) w->split = e[0] != '0';
        // with a front region the table only takes the keys the region cannot: it starts small and grows on demand
        alloc_table(w, pick_cap(expected_groups > 0 ? (w->range ? std::max<int64_t>(expected_groups / 32, 2 * 1024 * 1024) : expected_groups) : 0));
        (void)specs;
    } catch (...) { delete w; throw; }
    return w;
}

bool wide_add_page(WideTable* w, const DCol& key_in, const DCol* args_in, const DCol* masks_in, int64_t n, int64_t row_base) {
    pg_ctx* ctx = w->ctx;
    if (n <= 0) return true;
    PG_CHECK_STATE(!w->prepared, "the operator already produced output");
    if (row_base + n >= (1LL << 32))   // limb sums are exact for < 2^32 rows per group
        fail(PG_E_INSUFFICIENT_RESOURCES, "more than 2^32 rows in one aggregation operator instance");
    Reordered re;
    const bool reordered = reorder_for_locality(w, key_in, args_in, masks_in, n, re);
    const DCol& key = reordered ? re.key : key_in;
    const DCol* args_cols = reordered ? re.args.data() : args_in;
    const DCol* masks = reordered ? re.masks.data() : masks_in;
    // ---- source
    std::string pre;
    char buf[320];
    // rows per thread and trip: 4 when every column allows 16-byte (4-byte for byte columns) vector loads
    auto al = [](const void* p, uintptr_t a) { return p == nullptr || ((uintptr_t)p % a) == 0; };
    int vec = 4;
    if (!al(key.values, 16) || !al(key.nulls, 4)) vec = 1;
    for (size_t i = 0; i < w->aggs.size(); i++)
        if (!al(args_cols[i].values, 16) || !al(args_cols[i].nulls, 4) || !al(masks[i].values, 4) || !al(masks[i].nulls, 4)) vec = 1;
    if (const char* e = getenv("PRESTOGPU_WIDE_VEC")) { if (atoi(e) == 1) vec = 1; }
    snprintf(buf, sizeof(buf), "#define PGW_BUCKET_WORDS %d\n#define PGW_HASKEYNULL %d\n#define PGW_FLAGWORD %d\n#define PGW_VEC %d\n", w->bucket_words,
             key.nulls ? 1 : 0, w->flag_word, vec);
    pre += buf;
    pre += w->range ? "#define PGW_RANGE 1\n" : "#define PGW_RANGE 0\n";
    // reordered pages: first-seen order from the carried row ids, and tiles handed out in page order (an atomic counter) so that all
    // resident blocks work inside the same few table slices
    pre += reordered ? (getenv("PRESTOGPU_WIDE_NOPEEK") ? "#define PGW_ROWIDS 1\n#define PGW_DYNAMIC 1\n#define PGW_PEEK 0\n" : "#define PGW_ROWIDS 1\n#define PGW_DYNAMIC 1\n#define PGW_PEEK 1\n")
                     : "#define PGW_ROWIDS 0\n#define PGW_DYNAMIC 0\n#define PGW_PEEK 0\n";
    int minb = 3;
    if (const char* e = getenv("PRESTOGPU_WIDE_MINB")) { if (atoi(e) >= 1 && atoi(e) <= 8) minb = atoi(e); }
    snprintf(buf, sizeof(buf), "#define PGW_MINB %d\n#define PGW_PREFETCH %d\n#define PGW_LDCA %d\n", minb, getenv("PRESTOGPU_WIDE_NOPF") ? 0 : 1,
             getenv("PRESTOGPU_WIDE_LDCG") ? 0 : 1);
    pre += buf;
    std::string aggs_macro = "#define PGW_AGGS(X)";
    for (size_t i = 0; i < w->aggs.size(); i++) {
        const WideAgg& a = w->aggs[i];
        const DCol& m = masks[i];
        const DCol& c = args_cols[i];
        snprintf(buf, sizeof(buf), " X(%d, %s, %s, %d, %d, %d, %s, %d, %d, %d)", (int)i, op_name(a.op), in_name(a.in_kind), m.values ? 1 : 0,
                 (m.values && m.nulls) ? 1 : 0, (a.skip_null && c.nulls) ? 1 : 0, a.cnt_mode == 0 ? "C64" : "FLAG", a.cnt_word, a.val_word,
                 a.flag_bit);
        aggs_macro += buf;
    }
    pre += aggs_macro + "\n";
    bool flags_unconditional = !key.nulls;     // (rows with a null key share one bucket with every kind of row; keep the atomic there)
    for (size_t i = 0; i < w->aggs.size(); i++)
        if (w->aggs[i].cnt_mode == 1 && (masks[i].values || (w->aggs[i].skip_null && args_cols[i].nulls))) flags_unconditional = false;
    pre += flags_unconditional ? "#define PGW_FLAGS_UNCONDITIONAL 1\n" : "#define PGW_FLAGS_UNCONDITIONAL 0\n";
    // variants: [only] honours only_rows (re-run of left-out rows); [quota] bounds the claims of every block
    JitKernel* kern[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
    auto get_kernel = [&](int only, int quota) -> JitKernel* {
        if (!kern[only][quota]) {
            std::string p2 = pre + (only ? "#define PGW_HASONLY 1\n" : "#define PGW_HASONLY 0\n") + (quota ? "#define PGW_QUOTA 1\n" : "#define PGW_QUOTA 0\n");
            kern[only][quota] = jit_get(ctx, "wide:" + p2, p2 + WIDE_ARGS_SRC + WIDE_BODY_SRC, "pg_wide_agg");
        }
        return kern[only][quota];
    };
    // split mode (default with a front region): wide_locate_kernel + the update-only variant of the specialised kernel, else one fused kernel
    bool split = w->split;
    JitKernel* kern_bidx = nullptr;
    if (split) {
        std::string p2 = pre + "#define PGW_HASONLY 0\n#define PGW_QUOTA 0\n#define PGW_BIDX 1\n";
        kern_bidx = jit_get(ctx, "wide:" + p2, p2 + WIDE_ARGS_SRC + WIDE_BODY_SRC, "pg_wide_agg");
        if (!kern_bidx) return false;
    } else if (!get_kernel(0, 0)) return false;
    DevBuf bidx;

    PgwArgs a;
    memset(&a, 0, sizeof(a));
    a.keys = (const long long*)key.values; a.key_nulls = key.nulls;
    for (size_t i = 0; i < w->aggs.size(); i++) {
        a.arg[i] = args_cols[i].values; a.arg_nulls[i] = args_cols[i].nulls;
        a.mask[i] = (const unsigned char*)masks[i].values; a.mask_nulls[i] = masks[i].values ? masks[i].nulls : nullptr;
    }
    a.n = n; a.row_base = row_base;
    a.row_ids = reordered ? re.row_ids.as<uint32_t>() : nullptr;
    a.counters = w->counters.as<unsigned long long>();
    a.front = w->front.as<unsigned long long>(); a.front_keys = w->front_keys.as<unsigned long long>(); a.k0 = w->k0; a.nseg = (unsigned long long)w->nseg; a.shift = w->shift;
    DevBuf flags[2];               // left-out markers: written by launch i, read as the row filter by launch i + 1
    int cur = 0;
    const uint8_t* only = nullptr;
    int64_t pending = n;           // upper bound of the buckets the rows still to run can claim
    // more blocks than are resident: the kernel waits on memory, SMs do not all progress at the same speed, and equal static shares
    // would make every launch as slow as its slowest SM (measured with rows in random order: 9.8 ms against 7.9 ms)
    int per_sm = 32;
    if (const char* e = getenv("PRESTOGPU_WIDE_GRID")) { if (atoi(e) > 0) per_sm = atoi(e); }
    int grid = grid_for(ctx, n, WIDE_THREADS, vec, 4);
    if (grid > ctx->sm_count * per_sm) grid = ctx->sm_count * per_sm;
    // what the warps in flight can claim between two looks at the global claim counter (PGW_QUOTA): 8 resident warps per block slot,
    // `vec` rows per lane; 4 block slots per SM is more than the kernel's registers allow
    const int64_t in_flight = (int64_t)ctx->sm_total * 4 * (WIDE_THREADS / 32) * 32 * vec;
    while (true) {
        int64_t room = w->max_fill - w->claimed;
        // (with a front region `pending` overstates the table's share by far: grow geometrically, the left-out re-run does the rest)
        if (room < pending && room < 2 * in_flight) { grow(w, w->range ? 2 * w->claimed + 4 * in_flight : w->claimed + pending); continue; }
        // claims beyond max_fill are tolerated up to 0.9 of the buckets within one launch (the table is grown before the next page)
        const int64_t hard_room = (int64_t)(w->cap * 0.9) - w->claimed - in_flight;
        const int64_t quota = room >= pending ? (int64_t)1 << 62 : hard_room;
        if (!flags[cur].p) flags[cur].alloc(ctx, n);
        memset_async(ctx, flags[cur].p, 0, n);
        memset_async(ctx, w->counters.p, 0, 32);
        a.table = w->table.as<unsigned long long>(); a.cap = w->cap; a.room = (unsigned long long)quota;
        a.only_rows = only; a.overflow_flag = flags[cur].as<uint8_t>();
        if (split) {
            // first kernel: one bucket number per row (keys only); second kernel: the specialised update over bucket numbers
            if (!bidx.p) bidx.alloc(ctx, (n + 4) * 4);
            LocateArgs la;
            memset(&la, 0, sizeof(la));
            la.keys = (const long long*)key.values; la.key_nulls = key.nulls; la.only_rows = only; la.n = n;
            la.bidx = bidx.as<uint32_t>(); la.overflow_flag = flags[cur].as<uint8_t>();
            la.table = w->table.as<unsigned long long>(); la.cap = w->cap; la.bw = w->bucket_words;
            la.range = w->range ? 1 : 0; la.quota = room >= pending ? 0 : 1;
            la.vec_ok = (al(key.values, 16) && al(key.nulls, 4) && al(only, 4)) ? 1 : 0;
            la.front_keys = w->front_keys.as<unsigned long long>(); la.F = w->front_buckets; la.k0 = w->k0;
            la.nseg = (unsigned long long)w->nseg; la.shift = w->shift; la.room = (unsigned long long)quota;
            la.counters = w->counters.as<unsigned long long>();
            PG_LAUNCH(ctx, wide_locate_kernel, grid, WIDE_THREADS, 0, la);
            a.bidx = bidx.as<uint32_t>(); a.front_buckets = (unsigned long long)w->front_buckets;
            void* kargs[] = {&a};
            jit_launch(ctx, kern_bidx, "wide_update_jit", grid, WIDE_THREADS, 0, kargs);
        } else {
            JitKernel* k = get_kernel(only ? 1 : 0, room >= pending ? 0 : 1);
            if (!k) fail(PG_E_INTERNAL, "specialised kernel unavailable for the re-run: %s", ctx->jit ? ctx->jit->last_error.c_str() : "");
            void* kargs[] = {&a};
            jit_launch(ctx, k, "wide_agg_jit", grid, WIDE_THREADS, 0, kargs);
        }
        // one read-back: claims, left-out rows, and whether the two extra buckets are in use
        copy_async(ctx, ctx->h_scalar, w->counters.p, 24, cudaMemcpyDeviceToHost);
        copy_async(ctx, ctx->h_scalar + 3, w->table.as<unsigned long long>() + w->cap * w->bucket_words + 1, 8, cudaMemcpyDeviceToHost);
        copy_async(ctx, ctx->h_scalar + 4, w->table.as<unsigned long long>() + (w->cap + 1) * w->bucket_words + 1, 8, cudaMemcpyDeviceToHost);
        sync(ctx);
        w->claimed += ctx->h_scalar[0];
        w->front_claimed += ctx->h_scalar[2];
        const int64_t left_out = ctx->h_scalar[1];
        w->null_seen = ctx->h_scalar[3] != 0;
        w->zero_seen = ctx->h_scalar[4] != 0;
        if (getenv("PRESTOGPU_DEBUG"))
            fprintf(stderr, "[wide] n=%lld cap=%lld room=%lld pending=%lld quota=%lld grid=%d -> claimed=%lld front=%lld left_out=%lld\n", (long long)n,
                    (long long)w->cap, (long long)room, (long long)pending, (long long)quota, grid, (long long)w->claimed, (long long)w->front_claimed,
                    (long long)left_out);
        if (left_out == 0) break;
        // some block ran out of quota: grow, then re-run exactly the rows that were left out
        grow(w, w->claimed + left_out);
        pending = left_out;
        only = flags[cur].as<uint8_t>();
        cur ^= 1;
    }
    return true;
}

void wide_prepare_output(WideTable* w, pg_groupby* gb, const std::vector<pg_agg*>& aggs, int64_t total_rows) {
    pg_ctx* ctx = w->ctx;
    if (w->prepared) return;
    const int64_t groups = w->groups();
    const int64_t nwords = (total_rows + 31) / 32 + 1;
    DevBuf bitmap(ctx, nwords * 4), prefix(ctx, nwords * 8);
    memset_async(ctx, bitmap.p, 0, nwords * 4);
    WideView view;
    view.front = w->front.as<unsigned long long>(); view.fkeys = w->front_keys.as<unsigned long long>(); view.table = w->table.as<unsigned long long>();
    view.F = w->front_buckets; view.cap = w->cap; view.bw = w->bucket_words;
    PG_CHECK_STATE(w->front_buckets + w->cap + 2 < (1LL << 32), "hash table exceeds the supported size");
    const int grid = grid_for(ctx, w->front_buckets + w->cap + 2, 256);
    PG_LAUNCH(ctx, wide_mark_first_rows_kernel, grid, 256, 0, view, bitmap.as<uint32_t>());
    const int64_t ntiles = num_tiles(nwords);
    DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
    PG_LAUNCH(ctx, wide_popc_tiles_kernel, (int)ntiles, SCAN_THREADS, 0, bitmap.as<uint32_t>(), nwords, tiles.as<Pair64>());
    scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, tiles.as<Pair64>() + ntiles);
    PG_LAUNCH(ctx, wide_word_prefix_kernel, (int)ntiles, SCAN_THREADS, 0, bitmap.as<uint32_t>(), nwords, tiles.as<Pair64>(), prefix.as<uint2>());
    // dense destinations
    gb->values_by_group.grow(ctx, (groups + 2) * 8, 0);
    GatherPlan p;
    memset(&p, 0, sizeof(p));
    p.naggs = (int32_t)w->aggs.size(); p.bw = w->bucket_words; p.flag_word = w->flag_word;
    for (size_t i = 0; i < w->aggs.size(); i++) {
        agg_ensure_capacity(aggs[i], groups);
        p.op[i] = w->aggs[i].op; p.in_kind[i] = w->aggs[i].in_kind;
        p.cnt_mode[i] = w->aggs[i].cnt_mode; p.cnt_word[i] = w->aggs[i].cnt_word; p.val_word[i] = w->aggs[i].val_word;
        p.flag_bit[i] = w->aggs[i].flag_bit; p.words[i] = w->aggs[i].words;
        p.state[i] = aggs[i]->state.as<unsigned long long>();
    }
    DevBuf bucket_of_group(ctx, (groups + 1) * 4);
    memset_async(ctx, bucket_of_group.p, 0xff, (groups + 1) * 4);
    PG_LAUNCH(ctx, wide_number_buckets_kernel, grid, 256, 0, view, p, prefix.as<uint2>(), bucket_of_group.as<uint32_t>(),
              gb->slots.as<Slot>() + gb->cap, gb->values_by_group.as<int64_t>());
    // (with a front region the hashed table holds few groups, often none: the gather then has nothing to fetch)
    if (groups > 0 && w->claimed + (w->null_seen ? 1 : 0) + (w->zero_seen ? 1 : 0) > 0)
        PG_LAUNCH(ctx, wide_gather_kernel, grid_for(ctx, groups, 256), 256, 0, view, p,
                  bucket_of_group.as<uint32_t>(), groups, gb->values_by_group.as<int64_t>());
    gb->groups = groups;
    while (gb->groups >= max_fill_075(gb->ref_capacity) && gb->ref_capacity < (1LL << 30)) gb->ref_capacity *= 2;
    sync(ctx);
    w->prepared = true;
    w->table.release();
    w->front.release();
    w->front_keys.release();
}

}  // namespace pg
