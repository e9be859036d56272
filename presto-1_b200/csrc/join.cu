// join.cu — join build (PagesIndex + PagesHash + ArrayPositionLinks) and probe (JoinProbe + LookupJoinOperator emit).
#include <stdlib.h>

#include "join.h"
#include "project.h"

namespace pg {

static const int J_THREADS = 256;

// ------------------------------------------------------------------ build-side storage (PagesIndex.addPage)
__global__ void j_rebase_offsets_kernel(int32_t* dst, const int32_t* src, int64_t n, int32_t delta, int first_page) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // dst[0] of the whole column is written once (first page); later pages write dst[1..n] relative to their base
    for (; i <= n; i += stride) {
        if (i == 0 && !first_page) continue;
        dst[i] = src[i] - src[0] + delta;
    }
}

static void store_append(pg_ctx* ctx, ColStore& s, const ColImport& in, int64_t old_rows, int64_t n) {
    int64_t new_rows = old_rows + n;
    if (new_rows > s.rows_cap) {
        int64_t cap = s.rows_cap ? s.rows_cap : 1024;
        while (cap < new_rows) cap *= 2;
        int w = type_width(s.type);
        if (w) s.values.grow(ctx, cap * w, old_rows * w);
        if (s.has_nulls) s.nulls.grow(ctx, cap, old_rows);
        if (type_is_var(s.type)) s.offsets.grow(ctx, (cap + 1) * 4, (old_rows + 1) * 4);
        s.rows_cap = cap;
    }
    if (in.d.nulls && !s.has_nulls) {
        s.nulls.alloc(ctx, s.rows_cap);
        memset_async(ctx, s.nulls.p, 0, s.rows_cap);
        s.has_nulls = true;
    }
    if (s.has_nulls) {
        if (in.d.nulls) copy_async(ctx, s.nulls.as<uint8_t>() + old_rows, in.d.nulls, n, cudaMemcpyDeviceToDevice);
        else memset_async(ctx, s.nulls.as<uint8_t>() + old_rows, 0, n);
    }
    if (type_is_var(s.type)) {
        // bytes referenced by this page: [src_off[0], src_off[n])
        int32_t ends[2] = {0, 0};
        copy_async(ctx, ctx->h_scalar, in.d.offsets, 4, cudaMemcpyDeviceToHost);
        copy_async(ctx, (char*)ctx->h_scalar + 8, in.d.offsets + n, 4, cudaMemcpyDeviceToHost);
        sync(ctx);
        memcpy(&ends[0], ctx->h_scalar, 4); memcpy(&ends[1], (char*)ctx->h_scalar + 8, 4);
        int64_t nbytes = (int64_t)ends[1] - ends[0];
        if (s.bytes_used + nbytes > 0x7fffffffLL) fail(PG_E_INSUFFICIENT_RESOURCES, "variable width build column exceeds 2 GB");
        if (s.bytes_used + nbytes > s.values.bytes) s.values.grow(ctx, (s.bytes_used + nbytes) * 2 + 1024, s.bytes_used);
        copy_async(ctx, s.values.as<uint8_t>() + s.bytes_used, (const uint8_t*)in.d.values + ends[0], nbytes, cudaMemcpyDeviceToDevice);
        PG_LAUNCH(ctx, j_rebase_offsets_kernel, grid_for(ctx, n + 1, 256), 256, 0, s.offsets.as<int32_t>() + old_rows, in.d.offsets, n,
                  (int32_t)s.bytes_used, old_rows == 0 ? 1 : 0);
        s.bytes_used += nbytes;
    } else {
        int w = type_width(s.type);
        copy_async(ctx, s.values.as<char>() + old_rows * w, in.d.values, n * w, cudaMemcpyDeviceToDevice);
    }
}

// ------------------------------------------------------------------ key helpers
PG_D bool jrow_has_null(const JoinCols& jc, int64_t i) {   // SimplePagesHashStrategy.isPositionNull :268-276 / JoinProbe :109-116
    for (int k = 0; k < jc.n; k++) if (dcol_is_null(jc.c[k], i)) return true;
    return false;
}
PG_D int64_t jrow_hash(const JoinCols& jc, int64_t i) {
    uint64_t h = 0;
    for (int k = 0; k < jc.n; k++) h = 31 * h + (uint64_t)dcol_type_hash(jc.c[k], i);
    return (int64_t)h;
}
// EQUAL between row i of `a` and row j of `b` (no nulls on either side; NaN != NaN, DoubleType.java:158-162)
PG_D bool jrows_equal(const JoinCols& a, int64_t i, const JoinCols& b, int64_t j) {
    for (int k = 0; k < a.n; k++) {
        const DCol& x = a.c[k]; const DCol& y = b.c[k];
        switch (x.type) {
            case PG_VARCHAR: case PG_VARBINARY: {
                int32_t oi = x.offsets[i], li = x.offsets[i + 1] - oi, oj = y.offsets[j], lj = y.offsets[j + 1] - oj;
                if (li != lj) return false;
                const uint8_t* p = (const uint8_t*)x.values; const uint8_t* q = (const uint8_t*)y.values;
                for (int32_t t = 0; t < li; t++) if (p[oi + t] != q[oj + t]) return false;
                break;
            }
            case PG_LONG_DECIMAL: {
                const int64_t* p = (const int64_t*)x.values; const int64_t* q = (const int64_t*)y.values;
                if (p[2 * i] != q[2 * j] || p[2 * i + 1] != q[2 * j + 1]) return false;
                break;
            }
            case PG_DOUBLE:
                if (!(__longlong_as_double(dcol_long(x, i)) == __longlong_as_double(dcol_long(y, j)))) return false;
                break;
            default:
                if (dcol_long(x, i) != dcol_long(y, j)) return false;
        }
    }
    return true;
}

// find (probe side) or claim (build side) the slot of row i.  Returns -1 when absent (FIND only).
template <bool CLAIM>
PG_D int64_t j_locate(JSlot* slots, int64_t cap, int64_t mask, const JoinCols& build, const JoinCols& rowcols, int64_t i,
                      const DCol& hash_col, int has_hash) {
    if (build.bigint_fast) {
        unsigned long long key = (unsigned long long)((const int64_t*)rowcols.c[0].values)[i];
        if (key == EMPTY_KEY) return cap;
        int64_t p = (int64_t)(fmix64(key) & (uint64_t)mask);
        while (true) {
            unsigned long long k = CLAIM ? *((volatile unsigned long long*)&slots[p].word) : slots[p].word;
            if (k == key) return p;
            if (k == EMPTY_KEY) {
                if (!CLAIM) return -1;
                unsigned long long old = atomicCAS(&slots[p].word, EMPTY_KEY, key);
                if (old == EMPTY_KEY || old == key) return p;
            }
            p = (p + 1) & mask;
        }
    }
    int64_t raw = has_hash ? ((const int64_t*)hash_col.values)[i] : jrow_hash(rowcols, i);
    uint64_t f = fmix64((uint64_t)raw);
    uint32_t tag = (uint32_t)(f >> 32);
    unsigned long long mine = ((unsigned long long)tag << 32) | (unsigned long long)(uint32_t)i;
    int64_t p = (int64_t)(f & (uint64_t)mask);
    while (true) {
        unsigned long long w = CLAIM ? *((volatile unsigned long long*)&slots[p].word) : slots[p].word;
        if (w == EMPTY_WORD) {
            if (!CLAIM) return -1;
            unsigned long long old = atomicCAS(&slots[p].word, EMPTY_WORD, mine);
            if (old == EMPTY_WORD) return p;
            w = old;
        }
        if ((uint32_t)(w >> 32) == tag && jrows_equal(rowcols, i, build, (int64_t)(uint32_t)w)) return p;
        p = (p + 1) & mask;
    }
}

// build pass 1: every non-null-key row claims/finds its slot, count++, head = max(row)
__global__ void __launch_bounds__(J_THREADS) j_build_kernel(JSlot* slots, int64_t cap, JoinCols build, DCol hash_col, int has_hash, int64_t rows,
                                                            int32_t* row_slot, int* has_dups) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    bool dup = false;
    for (; i < rows; i += stride) {
        if (jrow_has_null(build, i)) { row_slot[i] = -1; continue; }   // PagesHash.java:94-96
        int64_t s = j_locate<true>(slots, cap, cap - 1, build, build, i, hash_col, has_hash);
        if (atomicAdd(&slots[s].count, 1) > 0) dup = true;
        atomicMax(&slots[s].head, (int32_t)i);
        row_slot[i] = (int32_t)s;
    }
    if (dup) *has_dups = 1;
}

// ------------------------------------------------------------------ single BIGINT join key: slot placement
// The table has two regions.  The range-chunked region places key k in chunk (k - kmin) >> shift (32 slots = four 128-byte lines):
// inside the chunk the key's hash picks ONE line and a start slot, and probing stays inside that line.  Build and probe rows that
// arrive in key order (orders by o_orderkey, lineitem clustered by l_orderkey) therefore walk the table front to back and every line
// comes from HBM once, instead of one random 32-byte sector per row.  The chunk width is chosen from the build keys' range so that a
// chunk holds 8..16 keys.  A key whose line is full, or outside [kmin, kmin + (nchunks << shift)), lives in the hashed region behind
// it (linear probing, power-of-two size); without a range region (range == 0) that is the whole table.
PG_D void jl_start(const JLayout& L, unsigned long long key, uint64_t h, int64_t& p, int& left) {
    if (L.range) {
        const unsigned long long seg = (key - L.kmin) >> L.shift;      // keys below kmin wrap around to a huge chunk number
        if (seg < (unsigned long long)L.nchunks) {
            // line and start slot inside the chunk: multiplicative (Fibonacci) hashing of the key itself.  A chunk holds consecutive
            // keys when the rows arrive sorted, and those spread almost evenly over the 32 slots - with the murmur bits two of the ~4
            // keys of a line started at the same slot every other line, and the loser of the claim had to read the slot again.
            const unsigned long long f = L.fib ? (key * 0x9E3779B97F4A7C15ULL) >> 59 : (((h & 3) << 3) | ((h >> 2) & 7));
            p = (int64_t)(seg << 5) + (int64_t)f; left = 8; return;
        }
    }
    p = L.R + (int64_t)((L.range ? h >> 8 : h) & (uint64_t)L.omask);      // (without a range region: the generic kernels' fmix64 & mask)
    left = 0;
}
// next slot of the probe sequence: the rest of the line (cyclic), then the hashed region
PG_D void jl_next(const JLayout& L, uint64_t h, int64_t& p, int& left) {
    if (left > 1) { left--; p = (p & ~(int64_t)7) | ((p + 1) & 7); }
    else if (left == 1) { left = 0; p = L.R + (int64_t)((h >> 8) & (uint64_t)L.omask); }      // (left > 0 only with a range region)
    else p = L.R + ((p - L.R + 1) & L.omask);
}
PG_D int4 ld_slot_cg(const JSlot* s) { return __ldcg((const int4*)s); }
PG_D unsigned long long slot_word(const int4& v) { return ((unsigned long long)(uint32_t)v.y << 32) | (uint32_t)v.x; }
// 128-bit compare-and-swap of a whole slot: an empty slot becomes {key, count 1, head row} in one atomic
PG_D void cas_slot(JSlot* s, unsigned long long clo, unsigned long long chi, unsigned long long nlo, unsigned long long nhi,
                   unsigned long long& olo, unsigned long long& ohi) {
    asm volatile("{\n\t.reg .b128 c, s, d;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 s, {%4, %5};\n\tatom.global.cas.b128 d, [%6], c, s;\n\tmov.b128 {%0, %1}, d;\n\t}"
                 : "=l"(olo), "=l"(ohi) : "l"(clo), "l"(chi), "l"(nlo), "l"(nhi), "l"(s) : "memory");
}

// min / max of the non-null build keys (meta[6], meta[7]; preset to +inf / -inf by the host)
__global__ void __launch_bounds__(256) j_key_range_kernel(const int64_t* __restrict__ keys, const uint8_t* __restrict__ nulls, int64_t rows,
                                                          long long* meta) {
    long long lo = 0x7fffffffffffffffLL, hi = -0x7fffffffffffffffLL - 1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += (int64_t)gridDim.x * blockDim.x) {
        if (nulls != nullptr && nulls[i] != 0) continue;
        const long long k = keys[i];
        lo = k < lo ? k : lo; hi = k > hi ? k : hi;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const long long l2 = __shfl_down_sync(0xffffffffu, lo, d), h2 = __shfl_down_sync(0xffffffffu, hi, d);
        lo = l2 < lo ? l2 : lo; hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(meta + 6, lo); atomicMax(meta + 7, hi); }
}

// Build.  Four rows per thread (a warp's lanes hold consecutive rows, so sorted build keys touch consecutive lines); every probe
// round is warp-uniform: all slots of the round are read first (a plain read: an atomic on a sector that is not in L2 yet is far
// slower), then every lane that saw a free slot claims it with ONE 128-bit compare-and-swap that also sets count = 1 and head = row.
// Only rows whose key is already there (duplicate build keys) need the separate count / head atomics.  Lanes leaving a
// data-dependent loop at different iterations are not brought back together by the hardware, hence the __any_sync rounds.
// meta: [0] duplicate flag (int), [4] slots claimed, [5] rows inserted, [8] set when the hashed region behind a range region is too full,
//       [9] slots claimed in that hashed region (the host enforces a load factor of 3/4 on it)
__global__ void __launch_bounds__(J_THREADS) j_build_fast_kernel(JSlot* slots, JLayout L, const int64_t* __restrict__ keys,
                                                                 const uint8_t* __restrict__ nulls, int64_t rows, int32_t* row_slot,
                                                                 unsigned long long* meta) {
    const int R = 4;
    const int64_t threads = (int64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    const unsigned long long EMPTY_HI = 0xFFFFFFFF00000000ULL;          // count 0, head -1
    unsigned long long claims = 0, inserted = 0, oclaims = 0;      // oclaims: claims inside the hashed region behind a range region
    bool special_dup = false, failed = false;
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base - lane < rows; base += threads * R) {
        unsigned long long key[R];
        uint64_t h[R];
        int64_t p[R];
        int left[R], st[R], steps[R];       // st: 0 settled, 1 probing, 2 at a free slot, 3 key found (duplicate)
#pragma unroll
        for (int k = 0; k < R; k++) {
            const int64_t i = base + k * threads;
            const bool valid = i < rows && !(nulls != nullptr && nulls[i] != 0);          // PagesHash.java:94-96
            key[k] = valid ? (unsigned long long)keys[i] : 0ULL;
            st[k] = 0; p[k] = -1; left[k] = 0; steps[k] = 0; h[k] = 0;
            if (!valid) continue;
            if (key[k] == EMPTY_KEY) {
                p[k] = L.total;
                if (atomicAdd(&slots[L.total].count, 1) > 0) special_dup = true;
                atomicMax(&slots[L.total].head, (int32_t)i);
                continue;
            }
            h[k] = fmix64(key[k]);
            jl_start(L, key[k], h[k], p[k], left[k]);
            st[k] = 1;
        }
        while (true) {
            int4 sv[R];
#pragma unroll
            for (int k = 0; k < R; k++) if (st[k] == 1) sv[k] = ld_slot_cg(slots + p[k]);
#pragma unroll
            for (int k = 0; k < R; k++) {
                if (st[k] != 1) continue;
                const unsigned long long w = slot_word(sv[k]);
                if (w == key[k]) st[k] = 3;
                else if (w == EMPTY_KEY) st[k] = 2;
                else {
                    jl_next(L, h[k], p[k], left[k]);
                    if (left[k] == 0 && L.range && ++steps[k] > 2048) { failed = true; st[k] = 0; p[k] = -1; }
                }
            }
            unsigned long long olo[R], ohi[R];
#pragma unroll
            for (int k = 0; k < R; k++)
                if (st[k] == 2) cas_slot(slots + p[k], EMPTY_KEY, EMPTY_HI, key[k], 1ULL | ((unsigned long long)(uint32_t)(base + k * threads) << 32), olo[k], ohi[k]);
            bool more = false;
#pragma unroll
            for (int k = 0; k < R; k++) {
                if (st[k] == 2) {
                    if (olo[k] == EMPTY_KEY) { claims++; inserted++; st[k] = 0; if (L.range && p[k] >= L.R) oclaims++; }
                    else if (olo[k] == key[k]) st[k] = 3;
                    else { jl_next(L, h[k], p[k], left[k]); st[k] = 1; }
                }
                more |= st[k] == 1;
            }
            if (!__any_sync(0xffffffffu, more)) break;
        }
#pragma unroll
        for (int k = 0; k < R; k++) {
            const int64_t i = base + k * threads;
            if (st[k] == 3) {
                inserted++;
                atomicAdd(&slots[p[k]].count, 1);
                atomicMax(&slots[p[k]].head, (int32_t)i);
            }
            if (i < rows) row_slot[i] = (int32_t)p[k];
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        claims += __shfl_down_sync(0xffffffffu, claims, d);
        inserted += __shfl_down_sync(0xffffffffu, inserted, d);
        oclaims += __shfl_down_sync(0xffffffffu, oclaims, d);
    }
    if (lane == 0) {
        if (claims) atomicAdd(meta + 4, claims);       // words 0..3 of meta belong to the duplicate flag and the segment sort
        if (inserted) atomicAdd(meta + 5, inserted);
        if (oclaims) atomicAdd(meta + 9, oclaims);
    }
    if (special_dup) *(int*)meta = 1;
    if (failed) meta[8] = 1;
}

// ------------------------------------------------------------------ duplicate keys: CSR segments in descending row order
__global__ void __launch_bounds__(SCAN_THREADS) j_seg_count_kernel(const JSlot* slots, int64_t nslots, Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i < nslots) { int c = slots[i].count; if (c > 1) acc.a += c; }
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_THREADS) j_seg_offsets_kernel(const JSlot* slots, int64_t nslots, const Pair64* tile_prefix,
                                                                     int32_t* seg_off) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = tile_prefix[blockIdx.x];
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        Pair64 v{0, 0};
        if (i < nslots) { int c = slots[i].count; if (c > 1) v.a = c; }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (i < nslots) seg_off[i] = (int32_t)(run.a + ex.a);
        run = run + total;
    }
}
__global__ void j_seg_fill_kernel(const JSlot* slots, const int32_t* row_slot, int64_t rows, const int32_t* seg_off, int32_t* seg_fill,
                                  int32_t* seg_rows) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < rows; i += stride) {
        int32_t s = row_slot[i];
        if (s < 0 || slots[s].count <= 1) continue;
        int pos = atomicAdd(&seg_fill[s], 1);
        seg_rows[seg_off[s] + pos] = (int32_t)i;
    }
}
static const int SMALL_SEG = 48;
// small segments: one thread sorts descending (insertion sort); large ones are queued for the block-wide sorter
__global__ void j_seg_sort_small_kernel(const JSlot* slots, int64_t nslots, const int32_t* seg_off, int32_t* seg_rows, int32_t* large_list,
                                        int* large_count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < nslots; i += stride) {
        int c = slots[i].count;
        if (c <= 1) continue;
        if (c > SMALL_SEG) { large_list[atomicAdd(large_count, 1)] = (int32_t)i; continue; }
        int32_t* seg = seg_rows + seg_off[i];
        for (int a = 1; a < c; a++) {
            int32_t v = seg[a];
            int b = a - 1;
            while (b >= 0 && seg[b] < v) { seg[b + 1] = seg[b]; b--; }
            seg[b + 1] = v;
        }
    }
}
// one block per large segment: bitonic sort (descending) directly in global memory
__global__ void __launch_bounds__(256) j_seg_sort_large_kernel(const JSlot* slots, const int32_t* seg_off, int32_t* seg_rows,
                                                               const int32_t* large_list) {
    int32_t slot = large_list[blockIdx.x];
    int64_t c = slots[slot].count;
    int32_t* seg = seg_rows + seg_off[slot];
    int64_t m = 1;
    while (m < c) m <<= 1;
    // bitonic network in its single-direction form (first step of every merge mirrors the partner), so that the
    // virtual padding beyond c behaves as -inf and comparisons with it can simply be skipped
    for (int64_t k = 2; k <= m; k <<= 1) {
        for (int64_t t = threadIdx.x; t < m; t += blockDim.x) {
            int64_t p = t ^ (k - 1);
            if (p > t && p < c) { int32_t a = seg[t], b = seg[p]; if (a < b) { seg[t] = b; seg[p] = a; } }
        }
        __syncthreads();
        for (int64_t j = k >> 2; j > 0; j >>= 1) {
            for (int64_t t = threadIdx.x; t < m; t += blockDim.x) {
                int64_t p = t ^ j;
                if (p > t && p < c) { int32_t a = seg[t], b = seg[p]; if (a < b) { seg[t] = b; seg[p] = a; } }
            }
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------ probe
struct ProbeArgs {
    const JSlot* slots;
    int64_t cap;
    const int32_t* seg_off;
    const int32_t* seg_rows;
    int32_t* probe_slot;
    int64_t n;
    int32_t join_type;
    int32_t build_empty;
    JLayout L;                 // single BIGINT key fast path
};

__global__ void __launch_bounds__(SCAN_THREADS) j_probe_count_kernel(ProbeArgs a, JoinCols build, JoinCols probe, DCol hash_col, int has_hash,
                                                                     Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
#pragma unroll 2
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t row = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (row < a.n) {
            int64_t s = -1;
            if (!a.build_empty && !jrow_has_null(probe, row))
                s = j_locate<false>((JSlot*)a.slots, a.cap, a.cap - 1, build, probe, row, hash_col, has_hash);
            int c = 0;
            if (s >= 0) { c = a.slots[s].count; if (c == 0) s = -1; }
            a.probe_slot[row] = (int32_t)s;
            if (c == 0 && a.join_type == 1) c = 1;   // PROBE_OUTER: appendNullForBuild (LookupJoinOperator.java:354-361)
            acc.a += c;
        }
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS) j_probe_emit_kernel(ProbeArgs a, const Pair64* tile_prefix, int32_t* out_probe,
                                                                    int64_t* out_build) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = tile_prefix[blockIdx.x];
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t row = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        Pair64 v{0, 0};
        int32_t s = -1; int c = 0;
        if (row < a.n) {
            s = a.probe_slot[row];
            if (s >= 0) c = a.slots[s].count;
            v.a = (c == 0 && a.join_type == 1) ? 1 : c;
        }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (v.a) {
            int64_t o = run.a + ex.a;
            if (c == 0) { out_probe[o] = (int32_t)row; out_build[o] = -1; }
            else if (c == 1) { out_probe[o] = (int32_t)row; out_build[o] = a.slots[s].head; }
            else {
                const int32_t* seg = a.seg_rows + a.seg_off[s];
                for (int k = 0; k < c; k++) { out_probe[o + k] = (int32_t)row; out_build[o + k] = seg[k]; }
            }
        }
        run = run + total;
    }
}

// ---- single BIGINT join key (BigintPagesHash-like fast path).  Probe pages usually arrive clustered by key (lineitem by
// order key): a run of adjacent lanes with the same key probes once (its first lane) and shares the answer by shuffle, so
// the random slot reads shrink by the run length.  The count pass stores per probe row either the build position itself
// (exactly one match: the emit pass then needs no table access at all), -1 (no match), or -(slot + 2) (several matches).
__global__ void __launch_bounds__(SCAN_THREADS) j_probe_count_fast_kernel(ProbeArgs a, const int64_t* __restrict__ keys, const uint8_t* __restrict__ nulls,
                                                                          Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    const int lane = threadIdx.x & 31;
    Pair64 acc{0, 0};
#pragma unroll 2
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t row = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        const bool in_range = row < a.n;
        const bool valid = in_range && !a.build_empty && !(nulls != nullptr && nulls[row] != 0);   // JoinProbe.java:109-116
        const unsigned long long key = valid ? (unsigned long long)keys[row] : 0ULL;
        const unsigned long long prev = __shfl_up_sync(0xffffffffu, key, 1);
        const int pv = __shfl_up_sync(0xffffffffu, valid ? 1 : 0, 1);
        const bool start = lane == 0 || pv != (valid ? 1 : 0) || prev != key;
        const unsigned starts = __ballot_sync(0xffffffffu, start);
        const int leader = 31 - __clz(starts & (0xffffffffu >> (31 - lane)));
        int code = -1, c = 0;
        // warp-uniform probe rounds: every leader reads one slot per round
        bool probing = valid && lane == leader;
        const uint64_t h = fmix64(key);
        int64_t p = a.L.total;
        int left = 0;
        if (probing && key != EMPTY_KEY) jl_start(a.L, key, h, p, left);
        while (true) {
            if (probing) {
                const int4 sv = __ldg((const int4*)(a.slots + p));        // {word lo, word hi, count, head}
                const unsigned long long w = slot_word(sv);
                if (w == key) {
                    c = sv.z;
                    code = c == 1 ? sv.w : (c > 1 ? -(int)(p + 2) : -1);
                    probing = false;
                } else if (w == EMPTY_KEY || p == a.L.total) probing = false;
                else jl_next(a.L, h, p, left);
            }
            if (!__any_sync(0xffffffffu, probing)) break;
        }
        code = __shfl_sync(0xffffffffu, code, leader);
        c = __shfl_sync(0xffffffffu, c, leader);
        if (in_range) {
            if (!valid) { code = -1; c = 0; }
            a.probe_slot[row] = code;
            if (c == 0 && a.join_type == 1) c = 1;   // PROBE_OUTER: appendNullForBuild (LookupJoinOperator.java:354-361)
            acc.a += c;
        }
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS) j_probe_emit_fast_kernel(ProbeArgs a, const Pair64* tile_prefix, int32_t* out_probe,
                                                                         int64_t* out_build) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = tile_prefix[blockIdx.x];
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t row = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        Pair64 v{0, 0};
        int code = -1, c = 0;
        int64_t s = -1;
        if (row < a.n) {
            code = a.probe_slot[row];
            if (code >= 0) c = 1;
            else if (code < -1) { s = -(int64_t)code - 2; c = a.slots[s].count; }
            v.a = (c == 0 && a.join_type == 1) ? 1 : c;
        }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (v.a) {
            const int64_t o = run.a + ex.a;
            if (c == 0) { out_probe[o] = (int32_t)row; out_build[o] = -1; }
            else if (c == 1) { out_probe[o] = (int32_t)row; out_build[o] = code; }
            else {
                const int32_t* seg = a.seg_rows + a.seg_off[s];
                for (int k = 0; k < c; k++) { out_probe[o + k] = (int32_t)row; out_build[o + k] = seg[k]; }
            }
        }
        run = run + total;
    }
}

// the same for unique build keys (every probe row emits at most one row): a warp owns 32 * SCAN_ITEMS consecutive rows, ballot + popc
// place the emitting rows, and the block needs one exchange of warp totals instead of a block-wide scan per item
__global__ void __launch_bounds__(SCAN_THREADS) j_probe_emit_unique_kernel(ProbeArgs a, const Pair64* tile_prefix, int32_t* out_probe,
                                                                           int64_t* out_build) {
    __shared__ int s_warp[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t wbase = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)warp * (32 * SCAN_ITEMS);
    int code[SCAN_ITEMS];
    unsigned b[SCAN_ITEMS];
    int mine = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t row = wbase + j * 32 + lane;
        code[j] = row < a.n ? a.probe_slot[row] : -1;
        b[j] = __ballot_sync(0xffffffffu, row < a.n && (code[j] >= 0 || a.join_type == 1));   // PROBE_OUTER: a miss emits (row, -1)
        mine += __popc(b[j]);
    }
    if (lane == 0) s_warp[warp] = mine;
    __syncthreads();
    int64_t d = tile_prefix[blockIdx.x].a;
    for (int w = 0; w < warp; w++) d += s_warp[w];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        if ((b[j] >> lane) & 1u) {
            const int64_t o = d + __popc(b[j] & ((1u << lane) - 1u));
            out_probe[o] = (int32_t)(wbase + j * 32 + lane);
            out_build[o] = code[j] >= 0 ? code[j] : -1;
        }
        d += __popc(b[j]);
    }
}

// ---- single BIGINT key, build keys unique (no position links), caller's buffers hold at least n rows.  Every probe row yields at
// most one output row, and in the common foreign-key join (lineitem -> orders) and in every PROBE_OUTER join each row yields exactly
// one: the output is then the identity on the probe side.  This kernel resolves the rows (four items of a thread together: keys
// loaded first, then warp-uniform probe rounds that read all four slots), stores the per-row code for the compaction passes, and
// ALSO writes (row, build position) to output position `row` on the assumption that nothing will be filtered out.  It counts the
// emitting rows per tile and in total (meta[0]); when the total equals n the output is complete after this one pass and no scan or
// second pass runs.  (A single-pass version with a decoupled look-back for the tile offsets was slower: 0.82 ms against 0.45 ms,
// the prefix could not travel through 29 K tiles fast enough.)
// GATHER: the build channels of the matched row are copied to the output row as well (LookupSource.appendTo fused into the probe:
// the 64-bit build positions are never written or read back); out_build may then be null.
static const int JG_MAX_COLS = 8;
struct GatherTab {
    int32_t n, pad;
    const void* src[JG_MAX_COLS];
    const uint8_t* src_nulls[JG_MAX_COLS];
    void* out[JG_MAX_COLS];
    uint8_t* out_nulls[JG_MAX_COLS];
    int32_t width[JG_MAX_COLS];
};
// LookupSource.appendTo for G rows of one lane, channel by channel: the (random) loads of all G rows are issued before the first store, so a lane
// has G requests in flight instead of one - what a shuffled probe side is bound by
template <int G, typename T>
PG_D void gather_rows_col(const T* __restrict__ src, const uint8_t* __restrict__ src_nulls, T* __restrict__ out, uint8_t* __restrict__ out_nulls,
                          const int64_t (&row)[G], const int (&cd)[G], unsigned inr, T zero) {
    T v[G];
    uint8_t nl[G];
#pragma unroll
    for (int g = 0; g < G; g++) {
        const bool hit = ((inr >> g) & 1u) && cd[g] >= 0;
        v[g] = hit ? __ldg(src + cd[g]) : zero;
        nl[g] = hit ? (src_nulls ? __ldg(src_nulls + cd[g]) : (uint8_t)0) : (uint8_t)1;
    }
#pragma unroll
    for (int g = 0; g < G; g++) {
        if (!((inr >> g) & 1u)) continue;
        out[row[g]] = nl[g] ? zero : v[g];
        if (out_nulls) out_nulls[row[g]] = nl[g] ? 1 : 0;
    }
}
template <int G>
PG_D void gather_rows(const GatherTab& t, const int64_t (&row)[G], const int (&cd)[G], unsigned inr) {
    for (int c = 0; c < t.n; c++) {
        switch (t.width[c]) {
            case 1: gather_rows_col<G>((const uint8_t*)t.src[c], t.src_nulls[c], (uint8_t*)t.out[c], t.out_nulls[c], row, cd, inr, (uint8_t)0); break;
            case 4: gather_rows_col<G>((const int32_t*)t.src[c], t.src_nulls[c], (int32_t*)t.out[c], t.out_nulls[c], row, cd, inr, (int32_t)0); break;
            case 8: gather_rows_col<G>((const int64_t*)t.src[c], t.src_nulls[c], (int64_t*)t.out[c], t.out_nulls[c], row, cd, inr, (int64_t)0); break;
            default: gather_rows_col<G>((const int4*)t.src[c], t.src_nulls[c], (int4*)t.out[c], t.out_nulls[c], row, cd, inr, make_int4(0, 0, 0, 0));
        }
    }
}
template <bool GATHER>
__global__ void __launch_bounds__(SCAN_THREADS) j_probe_dense_kernel(ProbeArgs a, const int64_t* __restrict__ keys, const uint8_t* __restrict__ nulls,
                                                                     Pair64* tile_sums, unsigned long long* total_out,
                                                                     int32_t* out_probe, int64_t* out_build, GatherTab gt) {
    __shared__ int s_cnt[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    int emitted = 0;
    const int G = 4;
#pragma unroll
    for (int j0 = 0; j0 < SCAN_ITEMS; j0 += G) {
        unsigned long long key[G];
        int p[G], left[G], leader[G], code[G];
        unsigned probing = 0, validm = 0, inr = 0;
#pragma unroll
        for (int g = 0; g < G; g++) {
            const int64_t row = tile_base + (int64_t)(j0 + g) * SCAN_THREADS + threadIdx.x;
            const bool in_range = row < a.n;
            const bool valid = in_range && !a.build_empty && !(nulls != nullptr && nulls[row] != 0);   // JoinProbe.java:109-116
            key[g] = valid ? (unsigned long long)keys[row] : 0ULL;
            if (valid) validm |= 1u << g;
            if (in_range) inr |= 1u << g;
        }
#pragma unroll
        for (int g = 0; g < G; g++) {
            const bool valid = (validm >> g) & 1u;
            const unsigned long long prev = __shfl_up_sync(0xffffffffu, key[g], 1);
            const int pv = __shfl_up_sync(0xffffffffu, valid ? 1 : 0, 1);
            const bool start = lane == 0 || pv != (valid ? 1 : 0) || prev != key[g];
            const unsigned starts = __ballot_sync(0xffffffffu, start);
            leader[g] = 31 - __clz(starts & (0xffffffffu >> (31 - lane)));
            code[g] = -1;
            p[g] = (int)a.L.total; left[g] = 0;
            if (valid && lane == leader[g]) {
                probing |= 1u << g;
                if (key[g] != EMPTY_KEY) { int64_t pp; jl_start(a.L, key[g], fmix64(key[g]), pp, left[g]); p[g] = (int)pp; }
            }
        }
        while (true) {
            int4 sv[G];
#pragma unroll
            for (int g = 0; g < G; g++) if ((probing >> g) & 1u) sv[g] = __ldg((const int4*)(a.slots + p[g]));        // {word lo, word hi, count, head}
#pragma unroll
            for (int g = 0; g < G; g++) {
                if (!((probing >> g) & 1u)) continue;
                const unsigned long long w = slot_word(sv[g]);
                if (w == key[g]) { code[g] = sv[g].z >= 1 ? sv[g].w : -1; probing &= ~(1u << g); }
                else if (w == EMPTY_KEY || p[g] == (int)a.L.total) probing &= ~(1u << g);
                else { int64_t pp = p[g]; jl_next(a.L, fmix64(key[g]), pp, left[g]); p[g] = (int)pp; }
            }
            if (!__any_sync(0xffffffffu, probing != 0)) break;
        }
        int64_t rows[G];
        int cds[G];
#pragma unroll
        for (int g = 0; g < G; g++) {
            int cd = __shfl_sync(0xffffffffu, code[g], leader[g]);
            if (!((validm >> g) & 1u)) cd = -1;
            cds[g] = cd;
            rows[g] = tile_base + (int64_t)(j0 + g) * SCAN_THREADS + threadIdx.x;
            if ((inr >> g) & 1u) {
                const int64_t row = rows[g];
                if (!GATHER) a.probe_slot[row] = cd;        // (the fused form never compacts: a page with misses is redone by the two-step path)
                out_probe[row] = (int32_t)row;
                if (out_build) out_build[row] = cd;
                if (cd >= 0 || a.join_type == 1) emitted++;    // PROBE_OUTER: appendNullForBuild (LookupJoinOperator.java:354-361)
            }
        }
        if (GATHER) gather_rows<G>(gt, rows, cds, inr);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) emitted += __shfl_down_sync(0xffffffffu, emitted, d);
    if (lane == 0) s_cnt[warp] = emitted;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < SCAN_THREADS / 32; w++) t += s_cnt[w];
        tile_sums[blockIdx.x] = Pair64{t, 0};
        if (t) atomicAdd(total_out, (unsigned long long)t);
    }
}

// ------------------------------------------------------------------ gather (LookupSource.appendTo / PagesHashStrategy.appendTo)
__global__ void j_gather_fixed_kernel(DCol src, const int64_t* pos, int64_t n, void* out_values, uint8_t* out_nulls) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int64_t p = pos[i];
        bool isnull = p < 0 || dcol_is_null(src, p);
        if (out_nulls) out_nulls[i] = isnull ? 1 : 0;
        switch (src.type) {
            case PG_INTEGER: case PG_DATE: ((int32_t*)out_values)[i] = isnull ? 0 : ((const int32_t*)src.values)[p]; break;
            case PG_BOOLEAN: ((uint8_t*)out_values)[i] = isnull ? 0 : ((const uint8_t*)src.values)[p]; break;
            case PG_LONG_DECIMAL: {
                const int64_t* q = (const int64_t*)src.values;
                ((int64_t*)out_values)[2 * i] = isnull ? 0 : q[2 * p]; ((int64_t*)out_values)[2 * i + 1] = isnull ? 0 : q[2 * p + 1];
                break;
            }
            default: ((int64_t*)out_values)[i] = isnull ? 0 : ((const int64_t*)src.values)[p];
        }
    }
}
PG_D int64_t gather_len(const DCol& src, int64_t p) {
    if (p < 0 || dcol_is_null(src, p)) return 0;
    return src.offsets[p + 1] - src.offsets[p];
}
__global__ void __launch_bounds__(SCAN_THREADS) j_gather_var_count_kernel(DCol src, const int64_t* pos, int64_t n, Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i < n) acc.a += gather_len(src, pos[i]);
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_THREADS) j_gather_var_kernel(DCol src, const int64_t* pos, int64_t n, const Pair64* tile_prefix,
                                                                    uint8_t* out_bytes, int32_t* out_offsets, uint8_t* out_nulls) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = tile_prefix[blockIdx.x];
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        Pair64 v{0, 0};
        int64_t p = -1;
        if (i < n) { p = pos[i]; v.a = gather_len(src, p); }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (i < n) {
            int64_t dst = run.a + ex.a;
            out_offsets[i] = (int32_t)dst;
            if (i == n - 1) out_offsets[n] = (int32_t)(dst + v.a);
            if (out_nulls) out_nulls[i] = (p < 0 || dcol_is_null(src, p)) ? 1 : 0;
            if (v.a) {
                const uint8_t* b = (const uint8_t*)src.values + src.offsets[p];
                for (int64_t t = 0; t < v.a; t++) out_bytes[dst + t] = b[t];
            }
        }
        run = run + total;
    }
}

// gather rows `pos` of a device column into fresh device buffers (also used by the PageProcessor's general path, project.cu)
void gather_device_col(pg_ctx* ctx, const DCol& src, const int64_t* d_pos, int64_t n, DevBuf& values, DevBuf& nulls, DevBuf& offsets, DCol& out) {
    out.type = src.type; out.pad = 0; out.offsets = nullptr;
    nulls.alloc(ctx, n > 0 ? n : 1);
    out.nulls = src.nulls ? nulls.as<uint8_t>() : nullptr;
    if (!type_is_var(src.type)) {
        values.alloc(ctx, (n > 0 ? n : 1) * (int64_t)type_width(src.type));
        out.values = values.p;
        if (n > 0) PG_LAUNCH(ctx, j_gather_fixed_kernel, grid_for(ctx, n, 256), 256, 0, src, d_pos, n, values.p, (uint8_t*)out.nulls);
        return;
    }
    const int64_t ntiles = num_tiles(n > 0 ? n : 1);
    DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
    offsets.alloc(ctx, (n + 1) * 4);
    int64_t tot[2] = {0, 0};
    if (n > 0) {
        PG_LAUNCH(ctx, j_gather_var_count_kernel, (int)ntiles, SCAN_THREADS, 0, src, d_pos, n, tiles.as<Pair64>());
        scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, tiles.as<Pair64>() + ntiles);
        read_scalars(ctx, tiles.as<Pair64>() + ntiles, 2, tot);
    } else memset_async(ctx, offsets.p, 0, 4);
    values.alloc(ctx, tot[0] > 0 ? tot[0] : 8);
    out.values = values.p; out.offsets = offsets.as<int32_t>();
    if (n > 0) PG_LAUNCH(ctx, j_gather_var_kernel, (int)ntiles, SCAN_THREADS, 0, src, d_pos, n, tiles.as<Pair64>(), (uint8_t*)values.p, offsets.as<int32_t>(),
                         (uint8_t*)out.nulls);
}

void gather_device_col_into(pg_ctx* ctx, const DCol& src, const int64_t* d_pos, int64_t n, void* values, uint8_t* nulls) {
    if (n > 0) PG_LAUNCH(ctx, j_gather_fixed_kernel, grid_for(ctx, n, 256), 256, 0, src, d_pos, n, values, nulls);
}

// ------------------------------------------------------------------ outer side (LookupOuterOperator / OuterLookupSource)
__global__ void j_mark_visited_kernel(const int64_t* __restrict__ build_pos, int64_t n, unsigned int* visited) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int64_t p = build_pos[i];
        if (p >= 0) atomicOr(&visited[p >> 5], 1u << (p & 31));
    }
}
__global__ void __launch_bounds__(SCAN_THREADS) j_outer_count_kernel(const unsigned int* __restrict__ visited, int64_t rows, Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i < rows && !((visited[i >> 5] >> (i & 31)) & 1u)) acc.a += 1;
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
// positions in ascending order (SharedLookupOuterPositionIterator walks currentPosition upwards): a tile's rows are visited in
// SCAN_ITEMS rounds of SCAN_THREADS consecutive rows, each round ranked by a block scan
__global__ void __launch_bounds__(SCAN_THREADS) j_outer_emit_kernel(const unsigned int* __restrict__ visited, int64_t rows, const Pair64* __restrict__ tile_prefix,
                                                                    int64_t* out_pos, int64_t capacity) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = tile_prefix[blockIdx.x];
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        Pair64 v{0, 0};
        if (i < rows && !((visited[i >> 5] >> (i & 31)) & 1u)) v.a = 1;
        Pair64 total;
        const Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (v.a && run.a + ex.a < capacity) out_pos[run.a + ex.a] = i;
        run = run + total;
    }
}

// ------------------------------------------------------------------ host side
static JoinCols build_join_cols(const std::vector<ColStore>& cols, const std::vector<int32_t>& jch) {
    JoinCols jc;
    memset(&jc, 0, sizeof(jc));
    jc.n = (int32_t)jch.size();
    for (int k = 0; k < jc.n; k++) jc.c[k] = cols[jch[k]].view();
    jc.bigint_fast = (jc.n == 1 && jc.c[0].type == PG_BIGINT) ? 1 : 0;
    return jc;
}

// slots of the hashed region behind a range region: a quarter of the rows (about 2 % of uniformly spread keys overflow their line)
static int64_t overflow_slots(int64_t rows) {
    int64_t c = 1024;
    while (c < rows / 4) c <<= 1;
    return c;
}

static void lookup_build(pg_lookup* ls) {
    pg_ctx* ctx = ls->ctx;
    int64_t rows = ls->rows;

    JoinCols build = build_join_cols(ls->cols, ls->join_channels);
    ls->bigint_fast = build.bigint_fast != 0;
    int64_t plain_cap = array_size_075(rows > 0 ? rows : 1);   // PagesHash.java:63
    if (plain_cap < 1024) plain_cap = 1024;
    ls->meta.alloc(ctx, 128);
    memset_async(ctx, ls->meta.p, 0, 128);
    const bool fast = ls->bigint_fast && plain_cap < (1LL << 30) && !getenv("PRESTOGPU_NO_FAST_BUILD");
    JLayout& L = ls->layout;
    memset(&L, 0, sizeof(L));
    L.omask = plain_cap - 1; L.total = plain_cap;
    // ---- plan the range region from the build keys' range (single BIGINT key, enough rows to matter)
    const char* renv = getenv("PRESTOGPU_JOIN_RANGE");
    if (fast && rows >= 4096 && !(renv && renv[0] == '0')) {
        const DCol& kc = build.c[0];
        const long long init[2] = {0x7fffffffffffffffLL, -0x7fffffffffffffffLL - 1};
        PG_CUDA(cudaMemcpyAsync(ls->meta.as<long long>() + 6, init, 16, cudaMemcpyHostToDevice, ctx->stream));
        PG_LAUNCH(ctx, j_key_range_kernel, grid_for(ctx, rows, 256, 8, 2), 256, 0, (const int64_t*)kc.values, kc.nulls, rows, ls->meta.as<long long>());
        int64_t mm[2];
        read_scalars(ctx, ls->meta.as<long long>() + 6, 2, mm);
        if (mm[0] <= mm[1]) {
            const double span = (double)((unsigned long long)mm[1] - (unsigned long long)mm[0]) + 1.0;
            const double width = 16.0 * span / (double)rows;               // key range holding 16 keys at the average density
            int shift = 0;
            while (shift < 62 && (double)(2ULL << shift) <= width) shift++;     // largest power of two <= width: 8..16 keys per chunk
            const int64_t nchunks = (int64_t)(((unsigned long long)mm[1] - (unsigned long long)mm[0]) >> shift) + 1;
            const int64_t R = nchunks * 32, ocap = overflow_slots(rows);
            if (R <= 5 * rows + 4096 && R + ocap < (1LL << 31) - 4) {
                L.range = 1; L.fib = getenv("PRESTOGPU_JOIN_NOFIB") ? 0 : 1; L.shift = shift; L.kmin = (unsigned long long)mm[0]; L.nchunks = nchunks; L.R = R;
                L.omask = ocap - 1; L.total = R + ocap;
            }
        }
    }
    DevBuf row_slot;
    if (rows > 0) row_slot.alloc(ctx, rows * 4);
    DCol hc; memset(&hc, 0, sizeof(hc));
    int has_hash = 0;
    if (ls->hash_channel >= 0 && !ls->bigint_fast) { hc = ls->cols[ls->hash_channel].view(); has_hash = 1; }
    int* d_has_dups = ls->meta.as<int>();
    while (true) {
        ls->cap = L.total;
        const int64_t nslots = ls->cap + 1;
        ls->slots.alloc(ctx, nslots * (int64_t)sizeof(JSlot));
        {   // every slot = {word: empty marker, count 0, head -1}
            const unsigned long long empty = ls->bigint_fast ? EMPTY_KEY : EMPTY_WORD;
            fill16_async(ctx, ls->slots.p, empty, 0xFFFFFFFF00000000ULL, nslots * (int64_t)sizeof(JSlot));
        }
            if (rows == 0) return;
        if (fast) {
            const DCol& kc = build.c[0];
            PG_LAUNCH(ctx, j_build_fast_kernel, grid_for(ctx, rows, J_THREADS, 4), J_THREADS, 0, ls->slots.as<JSlot>(), L, (const int64_t*)kc.values,
                      kc.nulls, rows, row_slot.as<int32_t>(), ls->meta.as<unsigned long long>());
            int64_t m[10];
            read_scalars(ctx, ls->meta.p, 10, m);
            // load-factor bound of the hashed region: probing it for an absent key ends at a free slot, so it may never fill up
            // (a row that needed more than 2048 steps already set m[8]); past 3/4 the plain table is built instead
            if (L.range && m[9] > (L.omask + 1) / 4 * 3) m[8] = 1;
                    if (L.range && m[8] != 0) {
                // the keys are spread too unevenly for the range region (its hashed overflow filled up): build the plain table instead
                memset(&L, 0, sizeof(L));
                L.omask = plain_cap - 1; L.total = plain_cap;
                memset_async(ctx, ls->meta.p, 0, 128);
                ls->slots.release();
                continue;
            }
            ls->has_dups = (m[0] & 0xffffffff) != 0 || m[4] != m[5];   // fewer claimed slots than inserted rows: some key repeats
        } else {
            PG_LAUNCH(ctx, j_build_kernel, grid_for(ctx, rows, J_THREADS), J_THREADS, 0, ls->slots.as<JSlot>(), ls->cap, build, hc, has_hash, rows,
                      row_slot.as<int32_t>(), d_has_dups);
            int64_t flag;
            read_scalars(ctx, d_has_dups, 1, &flag);
            ls->has_dups = (flag & 0xffffffff) != 0;
        }
        break;
    }
    const int64_t nslots = ls->cap + 1;
    if (!ls->has_dups) {         // JoinHashSupplier.java:72: positionLinks == Optional.empty()
        row_slot.release();
            return;
    }
    // CSR segments for repeated keys
    int64_t ntiles = num_tiles(nslots);
    DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
    Pair64* d_total = tiles.as<Pair64>() + ntiles;
    PG_LAUNCH(ctx, j_seg_count_kernel, (int)ntiles, SCAN_THREADS, 0, ls->slots.as<JSlot>(), nslots, tiles.as<Pair64>());
    scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, d_total);
    int64_t tot[2];
    read_scalars(ctx, d_total, 2, tot);
    ls->dup_rows = tot[0];
    ls->seg_off.alloc(ctx, nslots * 4);
    ls->seg_rows.alloc(ctx, (ls->dup_rows > 0 ? ls->dup_rows : 1) * 4);
    PG_LAUNCH(ctx, j_seg_offsets_kernel, (int)ntiles, SCAN_THREADS, 0, ls->slots.as<JSlot>(), nslots, tiles.as<Pair64>(),
              ls->seg_off.as<int32_t>());
    DevBuf fill(ctx, nslots * 4);
    memset_async(ctx, fill.p, 0, nslots * 4);
    PG_LAUNCH(ctx, j_seg_fill_kernel, grid_for(ctx, rows, 256), 256, 0, ls->slots.as<JSlot>(), row_slot.as<int32_t>(), rows,
              ls->seg_off.as<int32_t>(), fill.as<int32_t>(), ls->seg_rows.as<int32_t>());
    // reuse `fill` as the large-segment queue
    int* d_large_count = ls->meta.as<int>() + 2;
    memset_async(ctx, fill.p, 0, 4);
    PG_LAUNCH(ctx, j_seg_sort_small_kernel, grid_for(ctx, nslots, 256), 256, 0, ls->slots.as<JSlot>(), nslots, ls->seg_off.as<int32_t>(),
              ls->seg_rows.as<int32_t>(), fill.as<int32_t>(), d_large_count);
    int64_t lc;
    read_scalars(ctx, d_large_count, 1, &lc);
    int nlarge = (int)(lc & 0xffffffff);
    if (nlarge > 0)
        PG_LAUNCH(ctx, j_seg_sort_large_kernel, nlarge, 256, 0, ls->slots.as<JSlot>(), ls->seg_off.as<int32_t>(), ls->seg_rows.as<int32_t>(),
                  fill.as<int32_t>());
}

static int64_t lookup_device_bytes(const pg_lookup* ls) {
    int64_t b = ls->slots.bytes + ls->seg_off.bytes + ls->seg_rows.bytes;
    for (const ColStore& c : ls->cols) b += c.values.bytes + c.nulls.bytes + c.offsets.bytes;
    return b;
}

}  // namespace pg

using namespace pg;

extern "C" {

int32_t pg_joinbuild_create(pg_ctx* ctx, const int32_t* types, int32_t ncols, const int32_t* join_channels, int32_t njoin,
                            int32_t hash_channel, int64_t expected_rows, pg_joinbuild** out) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        PG_CHECK_ARG(out && types && join_channels, "null argument");
        PG_CHECK_ARG(ncols >= 1 && ncols <= MAX_JOIN_COLS, "between 1 and %d build channels are supported", MAX_JOIN_COLS);
        PG_CHECK_ARG(njoin >= 1 && njoin <= MAX_KEYS, "between 1 and %d join channels are supported", MAX_KEYS);
        PG_CHECK_ARG(hash_channel < ncols, "hash channel out of range");
        for (int k = 0; k < ncols; k++) PG_CHECK_ARG(type_valid(types[k]), "unknown type %d", types[k]);
        for (int k = 0; k < njoin; k++) PG_CHECK_ARG(join_channels[k] >= 0 && join_channels[k] < ncols, "join channel out of range");
        if (hash_channel >= 0) PG_CHECK_ARG(types[hash_channel] == PG_BIGINT, "hash channel must be BIGINT");
        pg_joinbuild* b = new pg_joinbuild();
        b->ctx = ctx;
        b->types.assign(types, types + ncols);
        b->join_channels.assign(join_channels, join_channels + njoin);
        b->hash_channel = hash_channel;
        b->cols.resize(ncols);
        for (int k = 0; k < ncols; k++) b->cols[k].type = types[k];
        (void)expected_rows;
        *out = b;
    });
}

int32_t pg_joinbuild_destroy(pg_joinbuild* b) {
    if (!b) return PG_OK;
    return guard(b->ctx, [&] { delete b; });
}

int32_t pg_joinbuild_add_page(pg_joinbuild* b, const pg_col* cols, int64_t n) {
    if (!b) return PG_E_INVALID_ARGUMENT;
    return guard(b->ctx, [&] {
        pg_ctx* ctx = b->ctx;
        PG_CHECK_ARG(cols != nullptr && n >= 0, "bad arguments");
        // PagesIndex.addPage :234-236
        if (b->rows + n > 0x7fffffffLL - 1) fail(PG_E_INSUFFICIENT_RESOURCES, "Size of pages index cannot exceed 2 billion entries");
        if (n == 0) return;
        for (size_t k = 0; k < b->cols.size(); k++) {
            PG_CHECK_ARG(cols[k].type == b->types[k], "build channel %d has type %d, expected %d", (int)k, cols[k].type, b->types[k]);
            ColImport in;
            import_col(ctx, cols[k], n, in);
            store_append(ctx, b->cols[k], in, b->rows, n);
        }
        b->rows += n;
        sync(ctx);
    });
}

int32_t pg_joinbuild_finish(pg_joinbuild* b, pg_lookup** out) {
    if (!b || !out) return PG_E_INVALID_ARGUMENT;
    return guard(b->ctx, [&] {
        pg_lookup* ls = new pg_lookup();
        try {
            ls->ctx = b->ctx; ls->types = b->types; ls->join_channels = b->join_channels; ls->hash_channel = b->hash_channel;
            ls->cols = std::move(b->cols);
            ls->rows = b->rows;
            b->cols.clear(); b->cols.resize(b->types.size());
            for (size_t k = 0; k < b->types.size(); k++) b->cols[k].type = b->types[k];
            b->rows = 0;
            lookup_build(ls);
            sync(ls->ctx);
        } catch (...) { delete ls; throw; }
        *out = ls;
    });
}

int32_t pg_lookup_destroy(pg_lookup* ls) {
    if (!ls) return PG_OK;
    return guard(ls->ctx, [&] { delete ls; });
}

int32_t pg_lookup_stats(pg_lookup* ls, int64_t* rows, int64_t* hash_size, int64_t* device_bytes, int32_t* has_duplicates) {
    if (!ls) return PG_E_INVALID_ARGUMENT;
    if (rows) *rows = ls->rows;
    if (hash_size) *hash_size = array_size_075(ls->rows);
    if (device_bytes) *device_bytes = lookup_device_bytes(ls);
    if (has_duplicates) *has_duplicates = ls->has_dups ? 1 : 0;
    return PG_OK;
}

// set by pg_lookup_join_page around its internal pg_lookup_probe call: the page's samples had many misses (count, then emit)
static thread_local bool t_sparse_probe_hint = false;

int32_t pg_lookup_probe(pg_lookup* ls, const pg_col* probe_keys, int32_t nkeys, const pg_col* hash_col, int64_t n, int32_t join_type,
                        int32_t* out_probe_idx, int64_t* out_build_pos, int32_t out_mem, int64_t capacity, int64_t* n_out) {
    if (!ls) return PG_E_INVALID_ARGUMENT;
    return guard(ls->ctx, [&] {
        pg_ctx* ctx = ls->ctx;
        PG_CHECK_ARG(probe_keys && n_out && n >= 0 && capacity >= 0, "bad arguments");
        PG_CHECK_ARG(nkeys == (int32_t)ls->join_channels.size(), "probe has %d join channels, build has %d", nkeys,
                     (int)ls->join_channels.size());
        PG_CHECK_ARG(join_type == 0 || join_type == 1, "join type must be INNER (0) or PROBE_OUTER (1)");
        *n_out = 0;
        if (n == 0) return;
        PG_CHECK_ARG(n <= MAX_CHUNK_ROWS, "probe page too large");
        JoinCols build = build_join_cols(ls->cols, ls->join_channels);
        JoinCols probe;
        memset(&probe, 0, sizeof(probe));
        probe.n = nkeys; probe.bigint_fast = build.bigint_fast;
        std::vector<ColImport> imp(nkeys);
        for (int k = 0; k < nkeys; k++) {
            PG_CHECK_ARG(probe_keys[k].type == build.c[k].type, "probe join channel %d has type %d, build side has %d", k, probe_keys[k].type,
                         build.c[k].type);
            import_col(ctx, probe_keys[k], n, imp[k]);
            probe.c[k] = imp[k].d;
        }
        ColImport himp;
        DCol hc; memset(&hc, 0, sizeof(hc));
        int has_hash = 0;
        if (hash_col && !ls->bigint_fast) {
            PG_CHECK_ARG(hash_col->type == PG_BIGINT, "hash channel must be BIGINT");
            import_col(ctx, *hash_col, n, himp);
            hc = himp.d; has_hash = 1;
        }
        int64_t ntiles = num_tiles(n);
        ProbeArgs a;
        a.slots = ls->slots.as<JSlot>(); a.cap = ls->cap; a.seg_off = ls->seg_off.as<int32_t>(); a.seg_rows = ls->seg_rows.as<int32_t>();
        a.probe_slot = nullptr; a.n = n; a.join_type = join_type; a.build_empty = ls->rows == 0 ? 1 : 0;
        a.L = ls->layout;
        const bool fast_table = ls->bigint_fast && (ls->layout.range || (ls->cap < (1LL << 30) && !getenv("PRESTOGPU_NO_FAST_PROBE")));
        if (fast_table && !ls->has_dups && capacity >= n && out_probe_idx && out_build_pos && !t_sparse_probe_hint && !getenv("PRESTOGPU_NO_FUSED_PROBE")) {
            // unique build keys and room for one output row per probe row: resolve, and write the output as if every row emitted
            // scratch belongs to the call, not to the lookup source: several probe operators share one source (LookupJoinOperatorFactory)
            DevBuf probe_slot(ctx, n * 4), tile_sums(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64)), total_buf(ctx, 64);
            Pair64* tiles = tile_sums.as<Pair64>();
            a.probe_slot = probe_slot.as<int32_t>();
            unsigned long long* d_total = total_buf.as<unsigned long long>();
            memset_async(ctx, d_total, 0, 8);
            OutVec<int32_t> op; OutVec<int64_t> ob;
            op.begin(ctx, out_probe_idx, out_mem, n);
            ob.begin(ctx, out_build_pos, out_mem, n);
            GatherTab none;
            memset(&none, 0, sizeof(none));
            PG_LAUNCH(ctx, j_probe_dense_kernel<false>, (int)ntiles, SCAN_THREADS, 0, a, (const int64_t*)probe.c[0].values, probe.c[0].nulls, tiles, d_total,
                      op.d, ob.d, none);
            int64_t tot;
            read_scalars(ctx, d_total, 1, &tot);
            *n_out = tot;
            if (tot == 0) return;
            if (tot < n) {   // some rows do not emit: compact with the per-row codes the kernel stored
                scan_tile_totals(ctx, tiles, ntiles, tiles + ntiles);
                PG_LAUNCH(ctx, j_probe_emit_unique_kernel, (int)ntiles, SCAN_THREADS, 0, a, tiles, op.d, ob.d);
            }
            if (ls->track_visited) PG_LAUNCH(ctx, j_mark_visited_kernel, grid_for(ctx, tot, 256), 256, 0, ob.d, tot, ls->visited.as<unsigned int>());
            op.finish(tot); ob.finish(tot);
            sync(ctx);
            return;
        }
        DevBuf probe_slot(ctx, n * 4), tile_sums(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
        Pair64* tiles = tile_sums.as<Pair64>();
        a.probe_slot = probe_slot.as<int32_t>();
        const bool fast = fast_table;
        if (fast)
            PG_LAUNCH(ctx, j_probe_count_fast_kernel, (int)ntiles, SCAN_THREADS, 0, a, (const int64_t*)probe.c[0].values, probe.c[0].nulls, tiles);
        else
            PG_LAUNCH(ctx, j_probe_count_kernel, (int)ntiles, SCAN_THREADS, 0, a, build, probe, hc, has_hash, tiles);
        scan_tile_totals(ctx, tiles, ntiles, tiles + ntiles);
        int64_t tot[2];
        read_scalars(ctx, tiles + ntiles, 2, tot);
        *n_out = tot[0];
        if (tot[0] > capacity) fail(PG_NEED_MEMORY, "join output needs %lld rows, buffers hold %lld", (long long)tot[0], (long long)capacity);
        if (tot[0] == 0) return;
        PG_CHECK_ARG(out_probe_idx && out_build_pos, "null output buffers");
        OutVec<int32_t> op; OutVec<int64_t> ob;
        op.begin(ctx, out_probe_idx, out_mem, tot[0]);
        ob.begin(ctx, out_build_pos, out_mem, tot[0]);
        if (fast && !ls->has_dups) PG_LAUNCH(ctx, j_probe_emit_unique_kernel, (int)ntiles, SCAN_THREADS, 0, a, tiles, op.d, ob.d);
        else if (fast) PG_LAUNCH(ctx, j_probe_emit_fast_kernel, (int)ntiles, SCAN_THREADS, 0, a, tiles, op.d, ob.d);
        else PG_LAUNCH(ctx, j_probe_emit_kernel, (int)ntiles, SCAN_THREADS, 0, a, tiles, op.d, ob.d);
        if (ls->track_visited) PG_LAUNCH(ctx, j_mark_visited_kernel, grid_for(ctx, tot[0], 256), 256, 0, ob.d, tot[0], ls->visited.as<unsigned int>());
        op.finish(); ob.finish();
        sync(ctx);
    });
}

int32_t pg_lookup_gather(pg_lookup* ls, const int64_t* build_pos, int32_t pos_mem, int64_t n, const int32_t* channels, int32_t nchannels,
                         pg_col_out* out) {
    if (!ls) return PG_E_INVALID_ARGUMENT;
    return guard(ls->ctx, [&] {
        pg_ctx* ctx = ls->ctx;
        PG_CHECK_ARG(n >= 0 && nchannels >= 0 && (n == 0 || build_pos) && (nchannels == 0 || (channels && out)), "bad arguments");
        if (n == 0 || nchannels == 0) return;
        DevBuf pbuf;
        const int64_t* d_pos = build_pos;
        if (pos_mem != PG_MEM_DEVICE) { pbuf.alloc(ctx, n * 8); copy_async(ctx, pbuf.p, build_pos, n * 8, cudaMemcpyHostToDevice); d_pos = pbuf.as<int64_t>(); }
        std::vector<OutStage> st(nchannels);
        std::vector<int64_t> vbytes(nchannels, 0);
        for (int k = 0; k < nchannels; k++) {
            int ch = channels[k];
            PG_CHECK_ARG(ch >= 0 && ch < (int)ls->cols.size(), "build channel %d out of range", ch);
            PG_CHECK_ARG(out[k].type == ls->types[ch], "output %d type mismatch", k);
            DCol src = ls->cols[ch].view();
            if (!type_is_var(src.type)) {
                vbytes[k] = n * type_width(src.type);
                st[k].begin(ctx, &out[k], n, vbytes[k]);
                PG_LAUNCH(ctx, j_gather_fixed_kernel, grid_for(ctx, n, 256), 256, 0, src, d_pos, n, st[k].d_values, st[k].d_nulls);
            } else {
                int64_t ntiles = num_tiles(n);
                DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
                PG_LAUNCH(ctx, j_gather_var_count_kernel, (int)ntiles, SCAN_THREADS, 0, src, d_pos, n, tiles.as<Pair64>());
                scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, tiles.as<Pair64>() + ntiles);
                int64_t tot[2];
                read_scalars(ctx, tiles.as<Pair64>() + ntiles, 2, tot);
                vbytes[k] = tot[0];
                st[k].begin(ctx, &out[k], n, vbytes[k]);
                PG_LAUNCH(ctx, j_gather_var_kernel, (int)ntiles, SCAN_THREADS, 0, src, d_pos, n, tiles.as<Pair64>(), (uint8_t*)st[k].d_values,
                          st[k].d_offsets, st[k].d_nulls);
            }
        }
        for (int k = 0; k < nchannels; k++) st[k].finish(vbytes[k]);
        sync(ctx);
    });
}

/* LookupJoinOperator's step for one probe page in one call: JoinProbe + LookupSource.appendTo of `build_channels`
 * (OP/LookupJoinOperator.java:333-361, OP/JoinProbe.java, PagesHashStrategy.appendTo).  out_probe_idx[capacity] receives the probe
 * positions, out_build[k] (capacity >= `capacity` rows) the build channels of the matched rows (null for PROBE_OUTER misses).
 * Unique single-BIGINT-key build sides with device outputs run one kernel (no 64-bit build positions are materialised); every
 * other case runs pg_lookup_probe + pg_lookup_gather internally.  PG_NEED_MEMORY as in pg_lookup_probe. */
int32_t pg_lookup_join_page(pg_lookup* ls, const pg_col* probe_keys, int32_t nkeys, const pg_col* hash_col, int64_t n, int32_t join_type,
                            const int32_t* build_channels, int32_t nchannels, int32_t* out_probe_idx, int32_t out_mem, pg_col_out* out_build,
                            int64_t capacity, int64_t* n_out) {
    if (!ls) return PG_E_INVALID_ARGUMENT;
    return guard(ls->ctx, [&] {
        pg_ctx* ctx = ls->ctx;
        PG_CHECK_ARG(probe_keys && n_out && n >= 0 && capacity >= 0 && nchannels >= 0 && (nchannels == 0 || (build_channels && out_build)), "bad arguments");
        PG_CHECK_ARG(nkeys == (int32_t)ls->join_channels.size(), "probe has %d join channels, build has %d", nkeys, (int)ls->join_channels.size());
        PG_CHECK_ARG(join_type == 0 || join_type == 1, "join type must be INNER (0) or PROBE_OUTER (1)");
        *n_out = 0;
        if (n == 0) return;
        PG_CHECK_ARG(n <= MAX_CHUNK_ROWS, "probe page too large");
        bool fused = ls->bigint_fast && !ls->has_dups && capacity >= n && out_probe_idx && out_mem == PG_MEM_DEVICE && nchannels <= JG_MAX_COLS &&
                     (ls->layout.range || ls->cap < (1LL << 30)) && probe_keys[0].type == PG_BIGINT && !getenv("PRESTOGPU_NO_FUSED_PROBE") &&
                     !ls->track_visited;      // (the visited tracker works on the materialised build positions)
        for (int k = 0; k < nchannels; k++) {
            const int ch = build_channels[k];
            PG_CHECK_ARG(ch >= 0 && ch < (int)ls->cols.size(), "build channel %d out of range", ch);
            PG_CHECK_ARG(out_build[k].type == ls->types[ch], "output %d type mismatch", k);
            if (type_is_var(ls->types[ch]) || out_build[k].mem != PG_MEM_DEVICE || out_build[k].capacity < n) fused = false;
            if (join_type == 1 && !out_build[k].nulls) fused = false;      // misses need a null mask
        }
        ColImport imp;
        bool sparse_hint = false;
        if (fused) import_col(ctx, probe_keys[0], n, imp);
        if (fused && join_type == 0 && n >= 64 * (int64_t)SCAN_TILE && !getenv("PRESTOGPU_NO_PROBE_SAMPLE")) {
            // an INNER probe with misses has to compact its output, which the one-kernel path cannot do: look at three small samples of
            // the page first (start, middle, end) and go straight to the compacting path when any of their rows finds no match
            const int64_t srows = 2 * (int64_t)SCAN_TILE;
            DevBuf scodes(ctx, srows * 4), stiles(ctx, 6 * (int64_t)sizeof(Pair64));
            ProbeArgs sa;
            sa.slots = ls->slots.as<JSlot>(); sa.cap = ls->cap; sa.seg_off = ls->seg_off.as<int32_t>(); sa.seg_rows = ls->seg_rows.as<int32_t>();
            sa.n = srows; sa.join_type = 0; sa.build_empty = ls->rows == 0 ? 1 : 0; sa.L = ls->layout; sa.probe_slot = scodes.as<int32_t>();
            const int64_t at[3] = {0, (n / 2) & ~(int64_t)31, (n - srows) & ~(int64_t)31};
            for (int r = 0; r < 3; r++)
                PG_LAUNCH(ctx, j_probe_count_fast_kernel, 2, SCAN_THREADS, 0, sa, (const int64_t*)imp.d.values + at[r],
                          imp.d.nulls ? imp.d.nulls + at[r] : nullptr, stiles.as<Pair64>() + 2 * r);
            int64_t got[12];
            read_scalars(ctx, stiles.p, 12, got);
            int64_t matched = 0;
            for (int r = 0; r < 6; r++) matched += got[2 * r];
            if (matched < 3 * srows) fused = false;
            // with many misses even the two-step path should not write an output row per probe row first: count, then emit
            sparse_hint = matched < 3 * srows / 4 * 3;
        }
        if (fused) {
            const int64_t ntiles = num_tiles(n);
            ProbeArgs a;
            a.slots = ls->slots.as<JSlot>(); a.cap = ls->cap; a.seg_off = ls->seg_off.as<int32_t>(); a.seg_rows = ls->seg_rows.as<int32_t>();
            a.n = n; a.join_type = join_type; a.build_empty = ls->rows == 0 ? 1 : 0; a.L = ls->layout;
            DevBuf tile_sums(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64)), total_buf(ctx, 64);
            a.probe_slot = nullptr;
            unsigned long long* d_total = total_buf.as<unsigned long long>();
            memset_async(ctx, d_total, 0, 8);
            GatherTab gt;
            memset(&gt, 0, sizeof(gt));
            gt.n = nchannels;
            for (int k = 0; k < nchannels; k++) {
                const DCol src = ls->cols[build_channels[k]].view();
                gt.src[k] = src.values; gt.src_nulls[k] = src.nulls; gt.out[k] = out_build[k].values; gt.out_nulls[k] = out_build[k].nulls;
                gt.width[k] = type_width(src.type);
            }
            PG_LAUNCH(ctx, j_probe_dense_kernel<true>, (int)ntiles, SCAN_THREADS, 0, a, (const int64_t*)imp.d.values, imp.d.nulls, tile_sums.as<Pair64>(),
                      d_total, out_probe_idx, (int64_t*)nullptr, gt);
            int64_t tot;
            read_scalars(ctx, d_total, 1, &tot);
            if (tot == n) { *n_out = n; return; }        // every probe row emitted exactly one row: the output is complete
            // some INNER rows found no match: the compacting path below redoes the page
        }
        DevBuf tp(ctx, std::max<int64_t>(capacity, 1) * 4), tb(ctx, std::max<int64_t>(capacity, 1) * 8);
        t_sparse_probe_hint = sparse_hint;           // per thread, not per lookup source: probers share the source
        int32_t rc = pg_lookup_probe(ls, probe_keys, nkeys, hash_col, n, join_type, tp.as<int32_t>(), tb.as<int64_t>(), PG_MEM_DEVICE, capacity, n_out);
        t_sparse_probe_hint = false;
        if (rc != PG_OK) { const std::string msg = ctx->err; fail(rc, "%s", msg.c_str()); }
        if (*n_out == 0) return;
        if (out_probe_idx)
            copy_async(ctx, out_probe_idx, tp.p, *n_out * 4, out_mem == PG_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost);
        if (nchannels > 0) {
            rc = pg_lookup_gather(ls, tb.as<int64_t>(), PG_MEM_DEVICE, *n_out, build_channels, nchannels, out_build);
            if (rc != PG_OK) { const std::string msg = ctx->err; fail(rc, "%s", msg.c_str()); }
        }
        sync(ctx);
    });
}

/* OuterLookupSource: start tracking which build positions a probe emits (RIGHT / FULL joins; OP/OuterLookupSource.java:176-205).
 * Call before the first probe. */
int32_t pg_lookup_track_visited(pg_lookup* ls) {
    if (!ls) return PG_E_INVALID_ARGUMENT;
    return guard(ls->ctx, [&] {
        if (ls->track_visited) return;
        const int64_t words = (ls->rows + 31) / 32 + 1;
        ls->visited.alloc(ls->ctx, words * 4);
        memset_async(ls->ctx, ls->visited.p, 0, words * 4);
        ls->track_visited = true;
    });
}

/* OuterPositionIterator (OP/OuterLookupSource.java:118-160): the build positions no probe has emitted, ascending.  *n_out = how many
 * there are; at most `capacity` are written (PG_NEED_MEMORY when more exist: call again with room for *n_out). */
int32_t pg_lookup_outer_positions(pg_lookup* ls, int64_t* out_pos, int32_t out_mem, int64_t capacity, int64_t* n_out) {
    if (!ls || !n_out) return PG_E_INVALID_ARGUMENT;
    return guard(ls->ctx, [&] {
        pg_ctx* ctx = ls->ctx;
        PG_CHECK_STATE(ls->track_visited, "pg_lookup_track_visited was not called on this lookup source");
        PG_CHECK_ARG(capacity >= 0 && (capacity == 0 || out_pos), "bad arguments");
        *n_out = 0;
        if (ls->rows == 0) return;
        const int64_t ntiles = num_tiles(ls->rows);
        DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
        PG_LAUNCH(ctx, j_outer_count_kernel, (int)ntiles, SCAN_THREADS, 0, ls->visited.as<unsigned int>(), ls->rows, tiles.as<Pair64>());
        scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, tiles.as<Pair64>() + ntiles);
        int64_t tot[2];
        read_scalars(ctx, tiles.as<Pair64>() + ntiles, 2, tot);
        *n_out = tot[0];
        if (tot[0] > capacity) fail(PG_NEED_MEMORY, "%lld outer positions, the buffer holds %lld", (long long)tot[0], (long long)capacity);
        if (tot[0] == 0) return;
        OutVec<int64_t> o;
        o.begin(ctx, out_pos, out_mem, tot[0]);
        PG_LAUNCH(ctx, j_outer_emit_kernel, (int)ntiles, SCAN_THREADS, 0, ls->visited.as<unsigned int>(), ls->rows, tiles.as<Pair64>(), o.d, capacity);
        o.finish();
        sync(ctx);
    });
}

int32_t pg_lookup_gather_var_bytes(pg_lookup* ls, const int64_t* build_pos, int32_t pos_mem, int64_t n, int32_t channel, int64_t* out_bytes) {
    if (!ls || !out_bytes) return PG_E_INVALID_ARGUMENT;
    return guard(ls->ctx, [&] {
        pg_ctx* ctx = ls->ctx;
        *out_bytes = 0;
        PG_CHECK_ARG(channel >= 0 && channel < (int)ls->cols.size(), "build channel out of range");
        if (n <= 0 || !type_is_var(ls->types[channel])) return;
        DevBuf pbuf;
        const int64_t* d_pos = build_pos;
        if (pos_mem != PG_MEM_DEVICE) { pbuf.alloc(ctx, n * 8); copy_async(ctx, pbuf.p, build_pos, n * 8, cudaMemcpyHostToDevice); d_pos = pbuf.as<int64_t>(); }
        int64_t ntiles = num_tiles(n);
        DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
        PG_LAUNCH(ctx, j_gather_var_count_kernel, (int)ntiles, SCAN_THREADS, 0, ls->cols[channel].view(), d_pos, n, tiles.as<Pair64>());
        scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, tiles.as<Pair64>() + ntiles);
        int64_t tot[2];
        read_scalars(ctx, tiles.as<Pair64>() + ntiles, 2, tot);
        *out_bytes = tot[0];
    });
}

}  // extern "C"
