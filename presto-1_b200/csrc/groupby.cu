// groupby.cu — GroupByHash kernels and host logic.  See groupby.h for the HBM layout.
#include "groupby.h"
#include "groupby_dev.cuh"

namespace pg {

static const int GB_THREADS = 256;

// ------------------------------------------------------------------ table init / rebuild
// re-insert every assigned group into a fresh table (BigintGroupByHash.tryRehash :287-303 walks valuesByGroupId
// the same way; MultiChannelGroupByHash.tryRehash :384-404 walks the old slots)
__global__ void gb_rebuild_bigint_kernel(Slot* slots, int64_t mask, const int64_t* values_by_group, int64_t groups,
                                         const Slot* old_special) {
    const int32_t magic_gid = old_special[0].gid, null_gid = old_special[1].gid;
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; g < groups; g += stride) {
        if (g == null_gid || g == magic_gid) continue;
        unsigned long long key = (unsigned long long)values_by_group[g];
        int64_t p = (int64_t)(fmix64(key) & (uint64_t)mask);
        while (true) {
            unsigned long long old = atomicCAS(&slots[p].word, EMPTY_KEY, key);
            if (old == EMPTY_KEY) { slots[p].gid = (int32_t)g; break; }
            p = (p + 1) & mask;
        }
    }
}
__global__ void gb_rebuild_multi_kernel(Slot* slots, int64_t mask, const int64_t* hash_by_group, int64_t groups) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; g < groups; g += stride) {
        uint64_t f = fmix64((uint64_t)hash_by_group[g]);
        unsigned long long word = (f & 0xFFFFFFFF00000000ULL) | (unsigned long long)(uint32_t)g;
        int64_t p = (int64_t)(f & (uint64_t)mask);
        while (true) {
            unsigned long long old = atomicCAS(&slots[p].word, EMPTY_WORD, word);
            if (old == EMPTY_WORD) { slots[p].gid = (int32_t)g; break; }
            p = (p + 1) & mask;
        }
    }
}
// carry the two special slots (magic key, null group) of bigint mode across a rebuild; pending ones are reset
__global__ void gb_copy_special_kernel(Slot* dst, int64_t dst_cap, const Slot* src, int64_t src_cap) {
    int t = threadIdx.x;
    if (t < 2) {
        Slot s = src[src_cap + t];
        if (s.gid < 0) s.first_row = 0x7fffffff;
        dst[dst_cap + t] = s;
    }
}

// ------------------------------------------------------------------ K1: insert
// Each row finds or claims its slot; pending slots get first_row = min(row).  row_slot[row] = slot index.
// A block stops claiming when it has used its quota (the host then rebuilds a larger table and redoes the chunk).
struct InsertArgs {
    Slot* slots;
    int64_t cap;
    int64_t mask;
    int32_t* row_slot;
    GbMeta* meta;
    int64_t n;
    int64_t quota;   // max new slots per block
    const uint64_t* key_rows;
    const uint8_t* arena;
};

// finds or claims the slot of `row` (claimed = this row created the slot); the body of K1, shared with the small-page kernel
template <bool BIGINT>
PG_D int64_t gb_find_or_claim(const InsertArgs& a, const KeyCols& kc, const KeyLayout& L, const DCol& hash_col, int has_hash, int64_t row,
                              bool& claimed) {
    // Slot words are read through L1 (plain loads): a key / claim never changes once written, so a cached
    // non-empty word is always right; a stale cached EMPTY is resolved by the CAS, which returns the live value.
    if (BIGINT) {
        const DCol& c = kc.c[0];
        if (dcol_is_null(c, row)) return a.cap + 1;
        unsigned long long key = (unsigned long long)((const int64_t*)c.values)[row];
        if (key == EMPTY_KEY) return a.cap;
        int64_t p = (int64_t)(fmix64(key) & (uint64_t)a.mask);
        while (true) {
            unsigned long long k = a.slots[p].word;
            if (k == key) break;
            if (k == EMPTY_KEY) {
                unsigned long long old = atomicCAS(&a.slots[p].word, EMPTY_KEY, key);
                if (old == EMPTY_KEY) { claimed = true; break; }
                if (old == key) break;
            }
            p = (p + 1) & a.mask;
        }
        return p;
    }
    int64_t raw = has_hash ? ((const int64_t*)hash_col.values)[row] : keys_row_hash(kc, L.nkeys, row);
    uint64_t f = fmix64((uint64_t)raw);
    uint32_t tag = (uint32_t)(f >> 32);
    unsigned long long mine = ((unsigned long long)tag << 32) | (unsigned long long)(REF_PENDING | (uint32_t)row);
    int64_t p = (int64_t)(f & (uint64_t)a.mask);
    while (true) {
        unsigned long long w = a.slots[p].word;
        if (w == EMPTY_WORD) {
            unsigned long long old = atomicCAS(&a.slots[p].word, EMPTY_WORD, mine);
            if (old == EMPTY_WORD) { claimed = true; break; }
            w = old;
        }
        if ((uint32_t)(w >> 32) == tag) {
            uint32_t ref = (uint32_t)w;
            bool eq = (ref & REF_PENDING) ? keys_equal_rows(kc, L, row, (int64_t)(ref & ~REF_PENDING))
                                          : keys_equal_store(kc, L, row, a.key_rows, a.arena, (int64_t)ref);
            if (eq) break;
        }
        p = (p + 1) & a.mask;
    }
    return p;
}

// K1.  Four rows per thread, and every probe round is warp-uniform: all lanes load the slot words of their probing rows first (four
// independent loads in flight per lane), then compare / claim, and the warp leaves the loop together (__any_sync) - lanes that leave a
// data-dependent loop at different iterations are not brought back together by the hardware, which left the one-row-per-lane loop
// running at a few active lanes per instruction on long probe sequences.
template <bool BIGINT>
__global__ void __launch_bounds__(GB_THREADS) gb_insert_kernel(InsertArgs a, KeyCols kc, KeyLayout L, DCol hash_col, int has_hash) {
    __shared__ int s_inserted;
    __shared__ int s_stop;
    if (threadIdx.x == 0) { s_inserted = 0; s_stop = 0; }
    __syncthreads();
    constexpr int R = 4;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * R;
    int iter = 0;
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x * R; base < a.n; base += stride, iter++) {
        // warp-uniform stop test (no block barrier inside the loop: warps of a block run independently)
        int stop = 0;
        if (lane == 0) {
            stop = *((volatile int*)&s_stop);
            if (!stop && (iter & 1) == 1) stop = *((volatile int32_t*)&a.meta->abort);
        }
        stop = __shfl_sync(0xffffffffu, stop, 0);
        if (stop) break;
        int64_t p[R];
        unsigned long long want[R];      // BIGINT: the key; otherwise the claim word {tag, PENDING | row}
        bool probing[R], claimed[R];
#pragma unroll
        for (int k = 0; k < R; k++) {
            const int64_t row = base + (int64_t)k * blockDim.x + threadIdx.x;
            probing[k] = false; claimed[k] = false; p[k] = -1; want[k] = 0;
            if (row >= a.n) continue;
            if (BIGINT) {
                const DCol& c = kc.c[0];
                if (dcol_is_null(c, row)) { p[k] = a.cap + 1; continue; }
                want[k] = (unsigned long long)((const int64_t*)c.values)[row];
                if (want[k] == EMPTY_KEY) { p[k] = a.cap; continue; }
                p[k] = (int64_t)(fmix64(want[k]) & (uint64_t)a.mask);
            } else {
                const int64_t raw = has_hash ? ((const int64_t*)hash_col.values)[row] : keys_row_hash(kc, L.nkeys, row);
                const uint64_t f = fmix64((uint64_t)raw);
                want[k] = ((unsigned long long)(uint32_t)(f >> 32) << 32) | (unsigned long long)(REF_PENDING | (uint32_t)row);
                p[k] = (int64_t)(f & (uint64_t)a.mask);
            }
            probing[k] = true;
        }
        while (true) {
            // Slot words are read through L1 (plain loads): a key / claim never changes once written, so a cached non-empty word is
            // always right; a stale cached EMPTY is resolved by the CAS, which returns the live value.
            unsigned long long w[R];
#pragma unroll
            for (int k = 0; k < R; k++) if (probing[k]) w[k] = a.slots[p[k]].word;
            bool more = false;
#pragma unroll
            for (int k = 0; k < R; k++) {
                if (!probing[k]) continue;
                if (BIGINT) {
                    if (w[k] == want[k]) probing[k] = false;
                    else {
                        if (w[k] == EMPTY_KEY) {
                            const unsigned long long old = atomicCAS(&a.slots[p[k]].word, EMPTY_KEY, want[k]);
                            if (old == EMPTY_KEY) { claimed[k] = true; probing[k] = false; }
                            else if (old == want[k]) probing[k] = false;
                        }
                        if (probing[k]) p[k] = (p[k] + 1) & a.mask;
                    }
                } else {
                    if (w[k] == EMPTY_WORD) {
                        const unsigned long long old = atomicCAS(&a.slots[p[k]].word, EMPTY_WORD, want[k]);
                        if (old == EMPTY_WORD) { claimed[k] = true; probing[k] = false; }
                        else w[k] = old;
                    }
                    if (probing[k] && (uint32_t)(w[k] >> 32) == (uint32_t)(want[k] >> 32)) {
                        const int64_t row = base + (int64_t)k * blockDim.x + threadIdx.x;
                        const uint32_t ref = (uint32_t)w[k];
                        const bool eq = (ref & REF_PENDING) ? keys_equal_rows(kc, L, row, (int64_t)(ref & ~REF_PENDING))
                                                            : keys_equal_store(kc, L, row, a.key_rows, a.arena, (int64_t)ref);
                        if (eq) probing[k] = false;
                    }
                    if (probing[k]) p[k] = (p[k] + 1) & a.mask;
                }
                more |= probing[k];
            }
            if (!__any_sync(0xffffffffu, more)) break;
        }
        int nclaimed = 0;
#pragma unroll
        for (int k = 0; k < R; k++) {
            const int64_t row = base + (int64_t)k * blockDim.x + threadIdx.x;
            const bool valid = row < a.n;
            const int64_t s = p[k];
            bool pending = false;
            if (valid) {
                pending = a.slots[s].gid < 0;   // group ids are only assigned between launches: the cached value is current
                a.row_slot[row] = (int32_t)s;
            }
            // rows of still-unnumbered slots compete for first_row; lanes of a warp that share a slot send one atomic
            const unsigned pend = __ballot_sync(0xffffffffu, pending);
            if (pend) {
                const unsigned peers = __match_any_sync(0xffffffffu, pending ? s : (int64_t)(-1 - lane));
                if (pending && (__ffs(peers) - 1) == lane) atomicMin(&a.slots[s].first_row, (int32_t)row);
            }
            nclaimed += claimed[k] ? 1 : 0;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) nclaimed += __shfl_down_sync(0xffffffffu, nclaimed, d);
        if (lane == 0 && nclaimed) {
            const int before = atomicAdd(&s_inserted, nclaimed);
            if ((int64_t)before + nclaimed >= a.quota) *((volatile int*)&s_stop) = 1;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_inserted) atomicAdd((unsigned long long*)&a.meta->inserted, (unsigned long long)s_inserted);
        // stopping before the last row means the chunk is incomplete
        if (s_stop) a.meta->abort = 1;   // (a block that stops on its very last row also reports abort: harmless redo)
    }
}

// ------------------------------------------------------------------ K2: first-seen group-id assignment
// A row is the representative of a new group iff its slot is still pending and it is the slot's first_row.
PG_D bool row_is_new_group(const Slot* slots, const int32_t* row_slot, int64_t row, int32_t* slot_out) {
    int32_t s = row_slot[row];
    *slot_out = s;
    const Slot& sl = slots[s];
    return sl.gid < 0 && sl.first_row == (int32_t)row;
}
PG_D int64_t row_var_bytes(const KeyCols& kc, const KeyLayout& L, int64_t row) {
    int64_t b = 0;
    for (int k = 0; k < L.nkeys; k++) {
        const DCol& c = kc.c[k];
        if ((c.type == PG_VARCHAR || c.type == PG_VARBINARY) && !dcol_is_null(c, row)) b += c.offsets[row + 1] - c.offsets[row];
    }
    return b;
}

__global__ void __launch_bounds__(SCAN_THREADS) gb_count_new_kernel(const Slot* slots, const int32_t* row_slot, int64_t n,
                                                                    KeyCols kc, KeyLayout L, int has_var, Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t row = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (row < n) {
            int32_t s;
            if (row_is_new_group(slots, row_slot, row, &s)) {
                acc.a += 1;
                if (has_var) acc.b += row_var_bytes(kc, L, row);
            }
        }
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

struct AssignArgs {
    Slot* slots;
    int64_t cap;
    const int32_t* row_slot;
    int64_t n;
    const Pair64* tile_prefix;   // exclusive prefix per tile
    const GbMeta* meta;          // group_count / arena_bytes before this chunk
    int64_t* values_by_group;    // bigint mode
    uint64_t* key_rows;          // multi mode
    int64_t* hash_by_group;
    uint8_t* arena;
};

// stores the key of new group `gid` (first seen at `row`, slot s; its variable width bytes start at boff) and numbers the slot
template <bool BIGINT>
PG_D void gb_assign_row(const AssignArgs& a, const KeyCols& kc, const KeyLayout& L, const DCol& hash_col, int has_hash, int64_t row, int32_t s,
                        int64_t gid, int64_t boff) {
    if (BIGINT) {
        int64_t key = 0;
        if (s == a.cap) key = (int64_t)EMPTY_KEY;
        else if (s < a.cap) key = ((const int64_t*)kc.c[0].values)[row];
        a.values_by_group[gid] = key;
    } else {
        uint64_t* out = a.key_rows + gid * L.row_words;
        uint64_t nullbits = 0;
        for (int k = 0; k < L.nkeys; k++) {
            const DCol& c = kc.c[k];
            uint64_t* w = out + L.word_off[k];
            if (dcol_is_null(c, row)) {
                nullbits |= 1ULL << k;
                w[0] = 0;
                if (c.type == PG_LONG_DECIMAL) w[1] = 0;
                if (c.type == PG_VARCHAR || c.type == PG_VARBINARY) w[0] = (uint64_t)boff << 24;
                continue;
            }
            switch (c.type) {
                case PG_VARCHAR: case PG_VARBINARY: {
                    int32_t o = c.offsets[row], len = c.offsets[row + 1] - o;
                    const uint8_t* b = (const uint8_t*)c.values;
                    for (int32_t t = 0; t < len; t++) a.arena[boff + t] = b[o + t];
                    w[0] = ((uint64_t)boff << 24) | (uint64_t)(uint32_t)len;
                    boff += len;
                    break;
                }
                case PG_LONG_DECIMAL: {
                    const uint64_t* p = (const uint64_t*)c.values;
                    w[0] = p[2 * row]; w[1] = p[2 * row + 1];
                    break;
                }
                default: w[0] = (uint64_t)dcol_long(c, row);
            }
        }
        out[0] = nullbits;
        int64_t raw = has_hash ? ((const int64_t*)hash_col.values)[row] : keys_row_hash(kc, L.nkeys, row);
        a.hash_by_group[gid] = raw;
        // the slot keeps its tag (= upper half of fmix64(rawHash), what the claim wrote) and now refers to the stored key
        a.slots[s].word = (fmix64((uint64_t)raw) & 0xFFFFFFFF00000000ULL) | (unsigned long long)(uint32_t)gid;
    }
    a.slots[s].gid = (int32_t)gid;
}

template <bool BIGINT>
__global__ void __launch_bounds__(SCAN_THREADS) gb_assign_kernel(AssignArgs a, KeyCols kc, KeyLayout L, DCol hash_col, int has_hash,
                                                                 int has_var) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = a.tile_prefix[blockIdx.x];
    run.a += a.meta->group_count;
    run.b += a.meta->arena_bytes;
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t row = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        int32_t s = 0;
        Pair64 v{0, 0};
        if (row < a.n && row_is_new_group(a.slots, a.row_slot, row, &s)) {
            v.a = 1;
            if (has_var) v.b = row_var_bytes(kc, L, row);
        }
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (v.a) gb_assign_row<BIGINT>(a, kc, L, hash_col, has_hash, row, s, run.a + ex.a, run.b + ex.b);
        run = run + total;
    }
}

__global__ void gb_commit_kernel(GbMeta* meta) {
    meta->group_count += meta->chunk_total.a;
    meta->arena_bytes += meta->chunk_total.b;
}

// K3: per-row group id (GroupByIdBlock)
__global__ void gb_lookup_ids_kernel(const Slot* slots, const int32_t* row_slot, int64_t n, int32_t* out_ids) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out_ids[i] = slots[row_slot[i]].gid;
}

// ------------------------------------------------------------------ small pages: K1 + K2 + K3 in ONE launch of ONE block
// Trino pages are small (<= 8192 rows, PageProcessor.java:56); six launches and two read-backs per page cost far more than the
// work.  One block does the whole getGroupIds: every row finds or claims its slot, the block synchronises, new groups are numbered
// in row order with block-wide scans, keys are stored, ids are emitted.  The host guarantees that the table (and the arena, for
// variable width keys) has room for n new groups, so there is no quota / abort protocol; meta->{group_count, arena_bytes,
// inserted} are what the host reads back afterwards.
static const int SMALL_THREADS = 1024;
static const int64_t SMALL_PAGE_ROWS = 16384;

template <bool BIGINT>
__global__ void __launch_bounds__(SMALL_THREADS) gb_small_page_kernel(InsertArgs a, AssignArgs as, KeyCols kc, KeyLayout L, DCol hash_col, int has_hash,
                                                                      int has_var, int32_t* row_slot, int32_t* out_ids) {
    __shared__ Pair64 warp_tot[SMALL_THREADS / 32];
    __shared__ Pair64 s_run;
    __shared__ int s_claimed;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t n = a.n;
    const int64_t base_groups = a.meta->group_count, base_bytes = a.meta->arena_bytes;
    if (tid == 0) { s_run = Pair64{0, 0}; s_claimed = 0; }
    __syncthreads();
    // ---- K1
    for (int64_t row = tid; row < n; row += SMALL_THREADS) {
        bool claimed = false;
        const int64_t s = gb_find_or_claim<BIGINT>(a, kc, L, hash_col, has_hash, row, claimed);
        row_slot[row] = (int32_t)s;
        if (__ldcg(&a.slots[s].gid) < 0) atomicMin(&a.slots[s].first_row, (int32_t)row);
        if (claimed) atomicAdd(&s_claimed, 1);
    }
    __syncthreads();
    // ---- K2: rows in page order, SMALL_THREADS at a time
    for (int64_t chunk = 0; chunk < n; chunk += SMALL_THREADS) {
        const int64_t row = chunk + tid;
        int32_t s = 0;
        Pair64 v{0, 0};
        if (row < n) {
            // first_row was settled with atomics (L2) while this SM's L1 may still hold the slot as it first saw it: read L2
            s = row_slot[row];
            const int4 sl = __ldcg((const int4*)&a.slots[s]);
            if (sl.z < 0 && sl.w == (int32_t)row) {
                v.a = 1;
                if (has_var) v.b = row_var_bytes(kc, L, row);
            }
        }
        Pair64 inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            Pair64 o = shfl_up_pair(inc, d);
            if (lane >= d) inc = inc + o;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        Pair64 before = s_run;
        for (int w = 0; w < warp; w++) before = before + warp_tot[w];
        if (v.a) gb_assign_row<BIGINT>(as, kc, L, hash_col, has_hash, row, s, base_groups + before.a + inc.a - 1, base_bytes + before.b + inc.b - v.b);
        __syncthreads();
        if (tid == 0) { Pair64 t = s_run; for (int w = 0; w < SMALL_THREADS / 32; w++) t = t + warp_tot[w]; s_run = t; }
        __syncthreads();
    }
    // ---- K3
    if (out_ids) for (int64_t row = tid; row < n; row += SMALL_THREADS) out_ids[row] = __ldcg(&a.slots[row_slot[row]].gid);
    if (tid == 0) {
        a.meta->group_count = base_groups + s_run.a;
        a.meta->arena_bytes = base_bytes + s_run.b;
        a.meta->inserted = s_claimed;
    }
}

// ------------------------------------------------------------------ contains (read-only probe)
template <bool BIGINT>
__global__ void gb_contains_kernel(const Slot* slots, int64_t cap, int64_t mask, const uint64_t* key_rows, const uint8_t* arena,
                                   KeyCols kc, KeyLayout L, DCol hash_col, int has_hash, int64_t n, uint8_t* out) {
    int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; row < n; row += stride) {
        bool found = false;
        if (BIGINT) {
            const DCol& c = kc.c[0];
            if (dcol_is_null(c, row)) found = slots[cap + 1].gid >= 0;
            else {
                unsigned long long key = (unsigned long long)((const int64_t*)c.values)[row];
                if (key == EMPTY_KEY) found = slots[cap].gid >= 0;
                else {
                    int64_t p = (int64_t)(fmix64(key) & (uint64_t)mask);
                    while (true) {
                        unsigned long long k = slots[p].word;
                        if (k == key) { found = slots[p].gid >= 0; break; }
                        if (k == EMPTY_KEY) break;
                        p = (p + 1) & mask;
                    }
                }
            }
        } else {
            int64_t raw = has_hash ? ((const int64_t*)hash_col.values)[row] : keys_row_hash(kc, L.nkeys, row);
            uint64_t f = fmix64((uint64_t)raw);
            uint32_t tag = (uint32_t)(f >> 32);
            int64_t p = (int64_t)(f & (uint64_t)mask);
            while (true) {
                unsigned long long w = slots[p].word;
                if (w == EMPTY_WORD) break;
                uint32_t ref = (uint32_t)w;
                if ((uint32_t)(w >> 32) == tag && !(ref & REF_PENDING) && keys_equal_store(kc, L, row, key_rows, arena, (int64_t)ref)) {
                    found = true; break;
                }
                p = (p + 1) & mask;
            }
        }
        out[row] = found ? 1 : 0;
    }
}

// ------------------------------------------------------------------ export (appendValuesTo / getRawHash)
__global__ void gb_export_bigint_kernel(const int64_t* values_by_group, const Slot* special, int64_t first, int64_t n,
                                        int64_t* out_values, uint8_t* out_nulls, int64_t* out_hash) {
    const int32_t null_gid = special[1].gid;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int64_t g = first + i;
        bool isnull = g == null_gid;
        int64_t v = isnull ? 0 : values_by_group[g];
        if (out_values) out_values[i] = v;
        if (out_nulls) out_nulls[i] = isnull ? 1 : 0;
        if (out_hash) out_hash[i] = isnull ? 0 : long_type_hash(v);   // BigintGroupByHash.getRawHash :200-203
    }
}

__global__ void gb_export_fixed_kernel(const uint64_t* key_rows, KeyLayout L, int k, int64_t first, int64_t n, void* out_values,
                                       uint8_t* out_nulls) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int type = L.types[k];
    for (; i < n; i += stride) {
        const uint64_t* row = key_rows + (first + i) * L.row_words;
        const uint64_t* w = row + L.word_off[k];
        if (out_nulls) out_nulls[i] = (row[0] >> k) & 1;
        switch (type) {
            case PG_INTEGER: case PG_DATE: ((int32_t*)out_values)[i] = (int32_t)w[0]; break;
            case PG_BOOLEAN: ((uint8_t*)out_values)[i] = (uint8_t)w[0]; break;
            case PG_LONG_DECIMAL: ((uint64_t*)out_values)[2 * i] = w[0]; ((uint64_t*)out_values)[2 * i + 1] = w[1]; break;
            default: ((uint64_t*)out_values)[i] = w[0];
        }
    }
}
__global__ void gb_export_hash_kernel(const int64_t* hash_by_group, int64_t first, int64_t n, int64_t* out_hash) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out_hash[i] = hash_by_group[first + i];
}

PG_D int64_t group_var_len(const uint64_t* key_rows, const KeyLayout& L, int k, int64_t g) {
    const uint64_t* row = key_rows + g * L.row_words;
    if ((row[0] >> k) & 1) return 0;
    return (int64_t)(row[L.word_off[k]] & 0xFFFFFF);
}
__global__ void __launch_bounds__(SCAN_THREADS) gb_export_var_count_kernel(const uint64_t* key_rows, KeyLayout L, int k, int64_t first,
                                                                           int64_t n, Pair64* tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i < n) acc.a += group_var_len(key_rows, L, k, first + i);
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_THREADS) gb_export_var_kernel(const uint64_t* key_rows, const uint8_t* arena, KeyLayout L, int k,
                                                                     int64_t first, int64_t n, const Pair64* tile_prefix,
                                                                     uint8_t* out_bytes, int32_t* out_offsets, uint8_t* out_nulls) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run{0, 0};
    if (tile_prefix) run = tile_prefix[blockIdx.x];   // null: a single tile, nothing before it
    for (int j = 0; j < SCAN_ITEMS; j++) {
        int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (tile_base + (int64_t)j * SCAN_THREADS >= n) break;
        Pair64 v{0, 0};
        if (i < n) v.a = group_var_len(key_rows, L, k, first + i);
        Pair64 total;
        Pair64 ex = block_exclusive_scan(v, smem, &total);
        if (i < n) {
            const uint64_t* row = key_rows + (first + i) * L.row_words;
            int64_t src = (int64_t)(row[L.word_off[k]] >> 24);
            int64_t dst = run.a + ex.a;
            out_offsets[i] = (int32_t)dst;
            if (i == n - 1) out_offsets[n] = (int32_t)(dst + v.a);
            if (out_nulls) out_nulls[i] = (row[0] >> k) & 1;
            for (int64_t t = 0; t < v.a; t++) out_bytes[dst + t] = arena[src + t];
        }
        run = run + total;
    }
}
__global__ void gb_sum_var_bytes_kernel(const uint64_t* key_rows, KeyLayout L, int k, int64_t groups, unsigned long long* out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned long long acc = 0;
    for (; i < groups; i += stride) acc += (unsigned long long)group_var_len(key_rows, L, k, i);
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

// ------------------------------------------------------------------ host logic
static const int64_t MIN_PHYS_CAP = 1LL << 20;

static void alloc_table(pg_groupby* gb, int64_t cap, DevBuf& slots) {
    pg_ctx* ctx = gb->ctx;
    slots.alloc(ctx, (cap + 2) * (int64_t)sizeof(Slot));
    // every slot = {word: empty marker, gid -1, first_row INT_MAX}
    fill16_async(ctx, slots.p, gb->bigint_mode ? EMPTY_KEY : EMPTY_WORD, 0x7FFFFFFFFFFFFFFFULL, (cap + 2) * (int64_t)sizeof(Slot));
}

static void ensure_group_storage(pg_groupby* gb) {
    pg_ctx* ctx = gb->ctx;
    int64_t need = gb->max_fill + 2 + 256 * 1024;   // slack: blocks may overshoot their quota by one iteration
    if (gb->bigint_mode) gb->values_by_group.grow(ctx, need * 8, gb->groups * 8);
    else {
        gb->key_rows.grow(ctx, need * gb->layout.row_words * 8, gb->groups * gb->layout.row_words * 8);
        gb->hash_by_group.grow(ctx, need * 8, gb->groups * 8);
    }
}

// replace the table by one with new_cap slots, re-inserting the assigned groups (pending claims are dropped)
static void rebuild_table(pg_groupby* gb, int64_t new_cap) {
    pg_ctx* ctx = gb->ctx;
    if (new_cap > (1LL << 30)) fail(PG_E_INSUFFICIENT_RESOURCES, "Size of hash table cannot exceed 1 billion entries");
    DevBuf fresh;
    alloc_table(gb, new_cap, fresh);
    if (gb->slots.p) {
        if (gb->groups > 0) {
            int grid = grid_for(ctx, gb->groups, 256);
            if (gb->bigint_mode)
                PG_LAUNCH(ctx, gb_rebuild_bigint_kernel, grid, 256, 0, fresh.as<Slot>(), new_cap - 1, gb->values_by_group.as<int64_t>(),
                          gb->groups, gb->slots.as<Slot>() + gb->cap);
            else
                PG_LAUNCH(ctx, gb_rebuild_multi_kernel, grid, 256, 0, fresh.as<Slot>(), new_cap - 1, gb->hash_by_group.as<int64_t>(),
                          gb->groups);
        }
        PG_LAUNCH(ctx, gb_copy_special_kernel, 1, 32, 0, fresh.as<Slot>(), new_cap, gb->slots.as<Slot>(), gb->cap);
        gb->rehashes++;
    }
    gb->slots = std::move(fresh);
    gb->cap = new_cap;
    gb->max_fill = max_fill_075(new_cap);
    gb->occupied = gb->groups;   // special slots are not counted against the fill
    ensure_group_storage(gb);
}

void groupby_create(pg_ctx* ctx, const int32_t* key_types, int32_t nkeys, int32_t has_hash_channel, int64_t expected,
                    pg_groupby** out, bool lazy_table) {
    PG_CHECK_ARG(out != nullptr && key_types != nullptr, "null argument");
    PG_CHECK_ARG(nkeys >= 1 && nkeys <= MAX_KEYS, "between 1 and %d group-by channels are supported, got %d", MAX_KEYS, nkeys);
    PG_CHECK_ARG(expected > 0, "expectedSize must be greater than zero");   // BigintGroupByHash.java:82
    for (int k = 0; k < nkeys; k++) PG_CHECK_ARG(type_valid(key_types[k]), "unknown key type %d", key_types[k]);
    pg_groupby* gb = new pg_groupby();
    try {
        gb->ctx = ctx;
        gb->bigint_mode = (nkeys == 1 && key_types[0] == PG_BIGINT);   // GroupByHash.createGroupByHash :45-59
        gb->has_hash_channel = has_hash_channel != 0;
        gb->expected = expected;
        KeyLayout& L = gb->layout;
        memset(&L, 0, sizeof(L));
        L.nkeys = nkeys;
        int w = 1;
        for (int k = 0; k < nkeys; k++) {
            L.types[k] = key_types[k];
            L.word_off[k] = w;
            w += key_types[k] == PG_LONG_DECIMAL ? 2 : 1;
            if (type_is_var(key_types[k])) gb->has_var = true;
        }
        L.row_words = w;
        gb->ref_capacity = array_size_075(expected);
        gb->meta.alloc(ctx, sizeof(GbMeta));
        memset_async(ctx, gb->meta.p, 0, sizeof(GbMeta));
        int64_t cap = gb->ref_capacity < MIN_PHYS_CAP ? MIN_PHYS_CAP : gb->ref_capacity;
        // lazy_table: the fused operator may never use this table (its many-groups path keeps its own): start with the smallest
        // one and build the planned one when the first rows are inserted
        if (lazy_table && cap > MIN_PHYS_CAP) { gb->deferred_cap = cap; cap = MIN_PHYS_CAP; }
        rebuild_table(gb, cap);
        if (gb->has_var) gb->arena.alloc(ctx, 1 << 20);
    } catch (...) {
        delete gb;
        throw;
    }
    *out = gb;
}

void groupby_reserve(pg_groupby* gb, int64_t n) {
    if (gb->memory_limit <= 0 || n <= 0) return;
    int64_t cap = gb->cap > gb->deferred_cap ? gb->cap : gb->deferred_cap;
    while (max_fill_075(cap) - gb->occupied < n && cap < (1LL << 30)) cap *= 2;
    if (cap == gb->cap) return;                       // no growth possible with this page
    const int64_t per_group = gb->bigint_mode ? 8 : (int64_t)gb->layout.row_words * 8 + 8;
    const int64_t after = groupby_device_bytes(gb) - gb->slots.bytes - gb->values_by_group.bytes - gb->key_rows.bytes - gb->hash_by_group.bytes +
                          (cap + 2) * (int64_t)sizeof(Slot) + (max_fill_075(cap) + 2 + 256 * 1024) * per_group;
    if (after > gb->memory_limit)
        fail(PG_NEED_MEMORY, "the hash table may have to grow to %lld bytes for this page, the memory limit is %lld; nothing was consumed",
             (long long)after, (long long)gb->memory_limit);
}

int64_t groupby_device_bytes(const pg_groupby* gb) {
    return gb->slots.bytes + gb->values_by_group.bytes + gb->key_rows.bytes + gb->hash_by_group.bytes + gb->arena.bytes +
           gb->row_slot.bytes + gb->tile_sums.bytes + gb->meta.bytes;
}

static void process_chunk(pg_groupby* gb, const KeyCols& kc, const DCol* hash_col, int64_t n, int32_t* d_out_ids) {
    pg_ctx* ctx = gb->ctx;
    if (n <= 0) return;
    if (gb->deferred_cap > gb->cap) {
        const int64_t planned = gb->deferred_cap, before = gb->rehashes;
        gb->deferred_cap = 0;
        rebuild_table(gb, planned);
        gb->rehashes = before;          // not a rehash the reference would have done
    }
    const KeyLayout& L = gb->layout;
    DCol hc; memset(&hc, 0, sizeof(hc));
    int has_hash = 0;
    if (!gb->bigint_mode && hash_col) { hc = *hash_col; has_hash = 1; }
    gb->row_slot.grow(ctx, n * 4, 0);
    int64_t ntiles = num_tiles(n);
    gb->tile_sums.grow(ctx, ntiles * (int64_t)sizeof(Pair64), 0);
    GbMeta* meta = gb->meta.as<GbMeta>();

    // ---- small page with room for every row to be a new group: one launch, one read-back
    {
        const bool small_on = !(getenv("PRESTOGPU_GB_SMALL") && getenv("PRESTOGPU_GB_SMALL")[0] == '0');
        const int64_t hint = gb->var_bound_hint;
        if (small_on && n <= SMALL_PAGE_ROWS && (!gb->has_var || (hint >= 0 && hint <= (64LL << 20)))) {
            while (gb->max_fill - gb->occupied < n) rebuild_table(gb, gb->cap * 2);
            if (gb->has_var && gb->arena_used + hint > gb->arena.bytes) gb->arena.grow(ctx, (gb->arena_used + hint) * 2, gb->arena_used);
            InsertArgs a;
            a.slots = gb->slots.as<Slot>(); a.cap = gb->cap; a.mask = gb->cap - 1; a.row_slot = gb->row_slot.as<int32_t>();
            a.meta = meta; a.n = n; a.quota = INT64_MAX / 2; a.key_rows = gb->key_rows.as<uint64_t>(); a.arena = gb->arena.as<uint8_t>();
            AssignArgs as;
            as.slots = a.slots; as.cap = gb->cap; as.row_slot = a.row_slot; as.n = n; as.tile_prefix = nullptr; as.meta = meta;
            as.values_by_group = gb->values_by_group.as<int64_t>(); as.key_rows = gb->key_rows.as<uint64_t>();
            as.hash_by_group = gb->hash_by_group.as<int64_t>(); as.arena = gb->arena.as<uint8_t>();
            if (gb->bigint_mode) PG_LAUNCH(ctx, gb_small_page_kernel<true>, 1, SMALL_THREADS, 0, a, as, kc, L, hc, has_hash, 0, a.row_slot, d_out_ids);
            else PG_LAUNCH(ctx, gb_small_page_kernel<false>, 1, SMALL_THREADS, 0, a, as, kc, L, hc, has_hash, gb->has_var ? 1 : 0, a.row_slot, d_out_ids);
            int64_t fin[3];
            read_scalars(ctx, &meta->group_count, 3, fin);
            gb->groups = fin[0];
            gb->arena_used = fin[1];
            gb->occupied += fin[2];
            return;
        }
    }

    // ---- K1 with quota / rebuild loop
    const int grid = grid_for(ctx, n, GB_THREADS, 4);       // (gb_insert_kernel: four rows per thread)
    bool inserted_pending = false, arena_pending = false;
    while (true) {
        int64_t room = gb->max_fill - gb->occupied;
        int64_t quota;
        if (room >= n) quota = INT64_MAX / 2;
        else {
            quota = room / grid - 4 * GB_THREADS;       // a block may overshoot by the rows its warps had in flight when it stopped
            if (quota < 32) { rebuild_table(gb, gb->cap * 2); continue; }
        }
        memset_async(ctx, &meta->inserted, 0, 8);
        memset_async(ctx, &meta->abort, 0, 4);
        InsertArgs a;
        a.slots = gb->slots.as<Slot>(); a.cap = gb->cap; a.mask = gb->cap - 1; a.row_slot = gb->row_slot.as<int32_t>();
        a.meta = meta; a.n = n; a.quota = quota; a.key_rows = gb->key_rows.as<uint64_t>(); a.arena = gb->arena.as<uint8_t>();
        if (gb->bigint_mode) PG_LAUNCH(ctx, gb_insert_kernel<true>, grid, GB_THREADS, 0, a, kc, L, hc, has_hash);
        else PG_LAUNCH(ctx, gb_insert_kernel<false>, grid, GB_THREADS, 0, a, kc, L, hc, has_hash);
        if (room >= n) { inserted_pending = true; break; }   // no quota, so no abort: the count is read with the chunk's last read-back
        // one read-back: {inserted, abort}
        copy_async(ctx, ctx->h_scalar, &meta->inserted, 8, cudaMemcpyDeviceToHost);
        copy_async(ctx, ctx->h_scalar + 1, &meta->abort, 4, cudaMemcpyDeviceToHost);
        sync(ctx);
        int64_t inserted = ctx->h_scalar[0];
        int32_t aborted; memcpy(&aborted, ctx->h_scalar + 1, 4);
        gb->occupied += inserted;
        if (!aborted) break;
        rebuild_table(gb, gb->cap * 2);   // drops this chunk's pending claims; the chunk is redone
    }

    // ---- K2: flag new groups, scan, assign ids in first-seen order
    PG_LAUNCH(ctx, gb_count_new_kernel, (int)ntiles, SCAN_THREADS, 0, gb->slots.as<Slot>(), gb->row_slot.as<int32_t>(), n, kc, L,
              gb->has_var ? 1 : 0, gb->tile_sums.as<Pair64>());
    scan_tile_totals(ctx, gb->tile_sums.as<Pair64>(), ntiles, &meta->chunk_total);
    if (gb->has_var) {
        const int64_t hint = gb->var_bound_hint;
        if (hint >= 0 && hint <= (64LL << 20)) {   // a cheap bound from the caller: make room for it, learn the exact use later
            if (gb->arena_used + hint > gb->arena.bytes) gb->arena.grow(ctx, (gb->arena_used + hint) * 2, gb->arena_used);
            arena_pending = true;
        } else {
            int64_t tot[2];
            read_scalars(ctx, &meta->chunk_total, 2, tot);
            int64_t need = gb->arena_used + tot[1];
            if (need > gb->arena.bytes) gb->arena.grow(ctx, need * 2, gb->arena_used);
            gb->arena_used = need;
        }
    }
    AssignArgs as;
    as.slots = gb->slots.as<Slot>(); as.cap = gb->cap; as.row_slot = gb->row_slot.as<int32_t>(); as.n = n;
    as.tile_prefix = gb->tile_sums.as<Pair64>(); as.meta = meta; as.values_by_group = gb->values_by_group.as<int64_t>();
    as.key_rows = gb->key_rows.as<uint64_t>(); as.hash_by_group = gb->hash_by_group.as<int64_t>(); as.arena = gb->arena.as<uint8_t>();
    if (gb->bigint_mode) PG_LAUNCH(ctx, gb_assign_kernel<true>, (int)ntiles, SCAN_THREADS, 0, as, kc, L, hc, has_hash, 0);
    else PG_LAUNCH(ctx, gb_assign_kernel<false>, (int)ntiles, SCAN_THREADS, 0, as, kc, L, hc, has_hash, gb->has_var ? 1 : 0);
    PG_LAUNCH(ctx, gb_commit_kernel, 1, 1, 0, meta);

    // ---- K3
    if (d_out_ids) {
        int g3 = grid_for(ctx, n, 256, 4);
        PG_LAUNCH(ctx, gb_lookup_ids_kernel, g3, 256, 0, gb->slots.as<Slot>(), gb->row_slot.as<int32_t>(), n, d_out_ids);
    }
    int64_t fin[3];   // {group_count, arena_bytes, inserted}: adjacent in GbMeta
    read_scalars(ctx, &meta->group_count, 3, fin);
    gb->groups = fin[0];
    if (arena_pending) gb->arena_used = fin[1];
    if (inserted_pending) gb->occupied += fin[2];
}

static KeyCols offset_keys(const KeyCols& kc, const KeyLayout& L, int64_t off) {
    KeyCols r = kc;
    for (int k = 0; k < L.nkeys; k++) {
        DCol& c = r.c[k];
        if (c.nulls) c.nulls += off;
        if (type_is_var(c.type)) c.offsets += off;
        else c.values = (const char*)c.values + off * type_width(c.type);
    }
    return r;
}

void groupby_process(pg_groupby* gb, const KeyCols& keys, const DCol* hash_col, int64_t n, int32_t* d_out_ids) {
    pg_ctx* ctx = gb->ctx;
    // A large page is split into a short prefix and the rest: the prefix numbers the groups a low-cardinality input
    // consists of, so the bulk of the rows only reads finished slots (no first_row atomics on a handful of addresses).
    const int64_t PREFIX = 128 * 1024;
    int64_t off = 0;
    while (off < n) {
        int64_t m = n - off;
        if (off == 0 && n > 4 * PREFIX) m = PREFIX;
        if (m > MAX_CHUNK_ROWS) m = MAX_CHUNK_ROWS;
        KeyCols kc = offset_keys(keys, gb->layout, off);
        DCol hc;
        if (hash_col) { hc = *hash_col; hc.values = (const int64_t*)hc.values + off; }
        process_chunk(gb, kc, hash_col ? &hc : nullptr, m, d_out_ids ? d_out_ids + off : nullptr);
        off += m;
    }
    (void)ctx;
    // the reference's logical capacity schedule (needRehash/tryRehash, BigintGroupByHash.java:262-334)
    while (gb->groups >= max_fill_075(gb->ref_capacity) && gb->ref_capacity < (1LL << 30)) gb->ref_capacity *= 2;
}


// ------------------------------------------------------------------ MarkDistinct / semi join (f4)
// MarkDistinctHash.markDistinctRows (OP/MarkDistinctHash.java:52-69): a row is marked when its group id equals nextDistinctId, i.e.
// when it is the first row (in page order) of a group this page introduced.
__global__ void gb_first_row_of_new_groups_kernel(const int32_t* __restrict__ ids, int64_t n, int64_t base, int32_t* first_row) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) if (ids[i] >= base) atomicMin(&first_row[ids[i] - base], (int32_t)i);
}
__global__ void gb_mark_distinct_kernel(const int32_t* __restrict__ ids, int64_t n, int64_t base, const int32_t* __restrict__ first_row, uint8_t* marks) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) marks[i] = (ids[i] >= base && first_row[ids[i] - base] == (int32_t)i) ? 1 : 0;
}
// HashSemiJoinOperator's output column (OP/HashSemiJoinOperator.java:190-218): NULL probe key -> false on an empty set, else NULL;
// found -> true; not found -> NULL when the set holds a NULL, else false.
__global__ void gb_semi_join_kernel(const uint8_t* __restrict__ found, const uint8_t* __restrict__ probe_nulls, int64_t n, int set_empty, int set_has_null,
                                    uint8_t* values, uint8_t* nulls) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        uint8_t v = 0, isnull = 0;
        if (probe_nulls && probe_nulls[i]) isnull = set_empty ? 0 : 1;
        else if (found[i]) v = 1;
        else if (set_has_null) isnull = 1;
        values[i] = v; nulls[i] = isnull;
    }
}

}  // namespace pg

// ------------------------------------------------------------------ C ABI
using namespace pg;

namespace {
struct ImportedKeys {
    std::vector<ColImport> cols;
    ColImport hash;
    KeyCols kc;
    DCol hc;
    bool has_hash = false;
};
void import_keys(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n, ImportedKeys& out) {
    PG_CHECK_ARG(keys != nullptr, "null key columns");
    const KeyLayout& L = gb->layout;
    out.cols.resize(L.nkeys);
    memset(&out.kc, 0, sizeof(out.kc));
    for (int k = 0; k < L.nkeys; k++) {
        PG_CHECK_ARG(keys[k].type == L.types[k], "key channel %d has type %d, expected %d", k, keys[k].type, L.types[k]);
        PG_CHECK_ARG(keys[k].n == n, "key channel %d has %lld rows, expected %lld", k, (long long)keys[k].n, (long long)n);
        import_col(gb->ctx, keys[k], n, out.cols[k]);
        out.kc.c[k] = out.cols[k].d;
    }
    out.has_hash = false;
    if (hash_col && gb->has_hash_channel) {
        PG_CHECK_ARG(hash_col->type == PG_BIGINT && hash_col->n == n, "hash channel must be BIGINT with %lld rows", (long long)n);
        import_col(gb->ctx, *hash_col, n, out.hash);
        out.hc = out.hash.d;
        out.has_hash = true;
    } else {
        PG_CHECK_ARG(!gb->has_hash_channel || gb->bigint_mode, "this GroupByHash was created with a precomputed hash channel");
    }
}
// small page: a bound on its variable width key bytes (known for host columns) lets process_chunk take its one-launch path
struct VarHint {
    pg_groupby* gb;
    VarHint(pg_groupby* g, const ImportedKeys& ik, int64_t n) : gb(g) {
        if (!g->has_var || n > 16384) return;
        int64_t bound = 0;
        for (size_t k = 0; k < ik.cols.size() && bound >= 0; k++)
            if (type_is_var(ik.cols[k].d.type)) bound = ik.cols[k].value_bytes >= 0 ? bound + ik.cols[k].value_bytes : -1;
        g->var_bound_hint = bound;
    }
    ~VarHint() { gb->var_bound_hint = -1; }
};
}  // namespace

extern "C" {

int32_t pg_groupby_create(pg_ctx* ctx, const int32_t* key_types, int32_t nkeys, int32_t has_hash_channel, int64_t expected_size,
                          pg_groupby** out) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] { groupby_create(ctx, key_types, nkeys, has_hash_channel, expected_size, out); });
}

int32_t pg_groupby_destroy(pg_groupby* gb) {
    if (!gb) return PG_OK;
    pg_ctx* ctx = gb->ctx;
    return guard(ctx, [&] { delete gb; });
}

int32_t pg_groupby_get_group_ids(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n, int32_t* out_ids,
                                 int32_t ids_mem, int64_t* out_group_count) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        PG_CHECK_ARG(n >= 0 && (out_ids != nullptr || n == 0), "bad arguments");
        ImportedKeys ik;
        import_keys(gb, keys, hash_col, n, ik);
        OutVec<int32_t> ids;
        ids.begin(gb->ctx, out_ids, ids_mem, n);
        VarHint hint(gb, ik, n);
        groupby_reserve(gb, n);
        groupby_process(gb, ik.kc, ik.has_hash ? &ik.hc : nullptr, n, ids.d);
        ids.finish();
        sync(gb->ctx);
        if (out_group_count) *out_group_count = gb->groups;
    });
}

int32_t pg_groupby_add_page(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        PG_CHECK_ARG(n >= 0, "bad arguments");
        ImportedKeys ik;
        import_keys(gb, keys, hash_col, n, ik);
        VarHint hint(gb, ik, n);
        groupby_reserve(gb, n);
        groupby_process(gb, ik.kc, ik.has_hash ? &ik.hc : nullptr, n, nullptr);
        sync(gb->ctx);
    });
}

int32_t pg_groupby_set_memory_limit(pg_groupby* gb, int64_t bytes) {
    if (!gb || bytes < 0) return PG_E_INVALID_ARGUMENT;
    gb->memory_limit = bytes;
    return PG_OK;
}

int32_t pg_groupby_contains(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n, uint8_t* out, int32_t out_mem) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        PG_CHECK_ARG(n >= 0 && (out != nullptr || n == 0), "bad arguments");
        if (n == 0) return;
        pg_ctx* ctx = gb->ctx;
        ImportedKeys ik;
        import_keys(gb, keys, hash_col, n, ik);
        OutVec<uint8_t> o;
        o.begin(ctx, out, out_mem, n);
        DCol hc; memset(&hc, 0, sizeof(hc));
        if (ik.has_hash) hc = ik.hc;
        int grid = grid_for(ctx, n, 256);
        if (gb->bigint_mode)
            PG_LAUNCH(ctx, gb_contains_kernel<true>, grid, 256, 0, gb->slots.as<Slot>(), gb->cap, gb->cap - 1, gb->key_rows.as<uint64_t>(),
                      gb->arena.as<uint8_t>(), ik.kc, gb->layout, hc, ik.has_hash ? 1 : 0, n, o.d);
        else
            PG_LAUNCH(ctx, gb_contains_kernel<false>, grid, 256, 0, gb->slots.as<Slot>(), gb->cap, gb->cap - 1, gb->key_rows.as<uint64_t>(),
                      gb->arena.as<uint8_t>(), ik.kc, gb->layout, hc, ik.has_hash ? 1 : 0, n, o.d);
        o.finish();
        sync(ctx);
    });
}

/* MarkDistinctHash.markDistinctRows(page): getGroupIds + "first row of a group this page introduced" */
int32_t pg_groupby_mark_distinct(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n, uint8_t* out_marks, int32_t out_mem) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        PG_CHECK_ARG(n >= 0 && (out_marks != nullptr || n == 0), "bad arguments");
        if (n == 0) return;
        pg_ctx* ctx = gb->ctx;
        ImportedKeys ik;
        import_keys(gb, keys, hash_col, n, ik);
        const int64_t base = gb->groups;
        DevBuf ids(ctx, n * 4);
        VarHint hint(gb, ik, n);
        groupby_reserve(gb, n);
        groupby_process(gb, ik.kc, ik.has_hash ? &ik.hc : nullptr, n, ids.as<int32_t>());
        const int64_t fresh = gb->groups - base;
        OutVec<uint8_t> o;
        o.begin(ctx, out_marks, out_mem, n);
        DevBuf first(ctx, (fresh > 0 ? fresh : 1) * 4);
        memset_async(ctx, first.p, 0x7f, (fresh > 0 ? fresh : 1) * 4);
        const int grid = grid_for(ctx, n, 256);
        if (fresh > 0) PG_LAUNCH(ctx, gb_first_row_of_new_groups_kernel, grid, 256, 0, ids.as<int32_t>(), n, base, first.as<int32_t>());
        PG_LAUNCH(ctx, gb_mark_distinct_kernel, grid, 256, 0, ids.as<int32_t>(), n, base, first.as<int32_t>(), o.d);
        o.finish();
        sync(ctx);
    });
}

/* HashSemiJoinOperator: the BOOLEAN semi-join column for one probe page against the set this GroupByHash holds (ChannelSet).
 * set_contains_null = ChannelSet.containsNull() (the caller asked pg_groupby_contains with a one-row NULL page when the set was built). */
int32_t pg_groupby_semi_join(pg_groupby* gb, const pg_col* probe_key, const pg_col* hash_col, int64_t n, int32_t set_contains_null, uint8_t* out_values,
                             uint8_t* out_nulls, int32_t out_mem) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        PG_CHECK_ARG(n >= 0 && ((out_values && out_nulls) || n == 0) && probe_key, "bad arguments");
        PG_CHECK_ARG(gb->layout.nkeys == 1, "a semi join has one join channel (HashSemiJoinOperator.probeJoinChannel)");
        if (n == 0) return;
        pg_ctx* ctx = gb->ctx;
        ImportedKeys ik;
        import_keys(gb, probe_key, hash_col, n, ik);
        DevBuf found(ctx, n);
        DCol hc; memset(&hc, 0, sizeof(hc));
        if (ik.has_hash) hc = ik.hc;
        const int grid = grid_for(ctx, n, 256);
        if (gb->bigint_mode)
            PG_LAUNCH(ctx, gb_contains_kernel<true>, grid, 256, 0, gb->slots.as<Slot>(), gb->cap, gb->cap - 1, gb->key_rows.as<uint64_t>(),
                      gb->arena.as<uint8_t>(), ik.kc, gb->layout, hc, ik.has_hash ? 1 : 0, n, found.as<uint8_t>());
        else
            PG_LAUNCH(ctx, gb_contains_kernel<false>, grid, 256, 0, gb->slots.as<Slot>(), gb->cap, gb->cap - 1, gb->key_rows.as<uint64_t>(),
                      gb->arena.as<uint8_t>(), ik.kc, gb->layout, hc, ik.has_hash ? 1 : 0, n, found.as<uint8_t>());
        OutVec<uint8_t> ov, on;
        ov.begin(ctx, out_values, out_mem, n);
        on.begin(ctx, out_nulls, out_mem, n);
        PG_LAUNCH(ctx, gb_semi_join_kernel, grid, 256, 0, found.as<uint8_t>(), ik.kc.c[0].nulls, n, gb->groups == 0 ? 1 : 0, set_contains_null ? 1 : 0, ov.d, on.d);
        ov.finish(); on.finish();
        sync(ctx);
    });
}

static const int64_t VAR_BOUND_BYTES = 64 * 1024;

int32_t pg_groupby_varwidth_bytes(pg_groupby* gb, int32_t key, int64_t* out_bytes) {
    if (!gb || !out_bytes) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        pg_ctx* ctx = gb->ctx;
        PG_CHECK_ARG(key >= 0 && key < gb->layout.nkeys, "bad key channel");
        *out_bytes = 0;
        if (!type_is_var(gb->layout.types[key]) || gb->groups == 0) return;
        // small stores: the bytes held for ALL variable width keys bound this one's, and the host already knows them
        if (gb->arena_used <= VAR_BOUND_BYTES) { *out_bytes = gb->arena_used; return; }
        DevBuf acc(ctx, 8);
        memset_async(ctx, acc.p, 0, 8);
        PG_LAUNCH(ctx, gb_sum_var_bytes_kernel, grid_for(ctx, gb->groups, 256), 256, 0, gb->key_rows.as<uint64_t>(), gb->layout, key,
                  gb->groups, acc.as<unsigned long long>());
        read_scalars(ctx, acc.p, 1, out_bytes);
    });
}

int32_t pg_groupby_export_keys(pg_groupby* gb, int64_t first, int64_t n, pg_col_out* out_keys, int64_t* out_hash, int32_t hash_mem) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    return guard(gb->ctx, [&] {
        pg_ctx* ctx = gb->ctx;
        PG_CHECK_ARG(first >= 0 && n >= 0 && first + n <= gb->groups, "group range [%lld, %lld) outside [0, %lld)", (long long)first,
                     (long long)(first + n), (long long)gb->groups);
        if (n == 0) return;
        const KeyLayout& L = gb->layout;
        OutVec<int64_t> oh;
        oh.begin(ctx, out_hash, hash_mem, n);
        int grid = grid_for(ctx, n, 256);
        std::vector<OutStage> st(out_keys ? L.nkeys : 0);
        std::vector<int64_t> vbytes(L.nkeys, 0);
        if (gb->bigint_mode) {
            int64_t* dv = nullptr; uint8_t* dn = nullptr;
            if (out_keys) {
                PG_CHECK_ARG(out_keys[0].type == PG_BIGINT, "output key type mismatch");
                st[0].begin(ctx, &out_keys[0], n, n * 8);
                dv = (int64_t*)st[0].d_values; dn = st[0].d_nulls; vbytes[0] = n * 8;
            }
            PG_LAUNCH(ctx, gb_export_bigint_kernel, grid, 256, 0, gb->values_by_group.as<int64_t>(), gb->slots.as<Slot>() + gb->cap, first,
                      n, dv, dn, oh.d);
        } else {
            if (oh.d) PG_LAUNCH(ctx, gb_export_hash_kernel, grid, 256, 0, gb->hash_by_group.as<int64_t>(), first, n, oh.d);
            for (int k = 0; out_keys && k < L.nkeys; k++) {
                PG_CHECK_ARG(out_keys[k].type == L.types[k], "output key %d type mismatch", k);
                if (!type_is_var(L.types[k])) {
                    vbytes[k] = n * type_width(L.types[k]);
                    st[k].begin(ctx, &out_keys[k], n, vbytes[k]);
                    PG_LAUNCH(ctx, gb_export_fixed_kernel, grid, 256, 0, gb->key_rows.as<uint64_t>(), L, k, first, n, st[k].d_values,
                              st[k].d_nulls);
                } else {
                    int64_t ntiles = num_tiles(n);
                    const bool bounded = gb->arena_used <= VAR_BOUND_BYTES && out_keys[k].values_capacity_bytes >= gb->arena_used;
                    if (ntiles == 1 && bounded) {   // few groups: offsets come out of the one block's own scan
                        vbytes[k] = gb->arena_used;
                        st[k].begin(ctx, &out_keys[k], n, vbytes[k]);
                        PG_LAUNCH(ctx, gb_export_var_kernel, 1, SCAN_THREADS, 0, gb->key_rows.as<uint64_t>(), gb->arena.as<uint8_t>(), L, k, first,
                                  n, (const Pair64*)nullptr, (uint8_t*)st[k].d_values, st[k].d_offsets, st[k].d_nulls);
                        continue;
                    }
                    DevBuf tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
                    Pair64* d_total = tiles.as<Pair64>() + ntiles;
                    PG_LAUNCH(ctx, gb_export_var_count_kernel, (int)ntiles, SCAN_THREADS, 0, gb->key_rows.as<uint64_t>(), L, k, first, n,
                              tiles.as<Pair64>());
                    scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, d_total);
                    if (bounded) {
                        vbytes[k] = gb->arena_used;   // upper bound known on the host: no read-back of the exact total
                    } else {
                        int64_t tot[2];
                        read_scalars(ctx, d_total, 2, tot);
                        vbytes[k] = tot[0];
                    }
                    st[k].begin(ctx, &out_keys[k], n, vbytes[k]);
                    PG_LAUNCH(ctx, gb_export_var_kernel, (int)ntiles, SCAN_THREADS, 0, gb->key_rows.as<uint64_t>(), gb->arena.as<uint8_t>(), L,
                              k, first, n, tiles.as<Pair64>(), (uint8_t*)st[k].d_values, st[k].d_offsets, st[k].d_nulls);
                }
            }
        }
        for (size_t k = 0; k < st.size(); k++) st[k].finish(vbytes[k]);
        oh.finish();
        if (ctx->guard_depth <= 1) sync(ctx);   // inside the fused operator's getOutput the caller's own read-back completes the copies
    });
}

int32_t pg_groupby_stats(pg_groupby* gb, int64_t* group_count, int64_t* capacity, int64_t* collisions, int64_t* device_bytes) {
    if (!gb) return PG_E_INVALID_ARGUMENT;
    if (group_count) *group_count = gb->groups;
    if (capacity) *capacity = gb->ref_capacity;
    if (collisions) *collisions = 0;   // probe-length statistics are not collected on the device path
    if (device_bytes) *device_bytes = groupby_device_bytes(gb);
    return PG_OK;
}

}  // extern "C"
