// lowcard_body.cuh — body of the run-time specialised fused GROUP BY kernel for few groups ("hash-then-update with
// shared-memory pre-aggregation").  NVRTC compiles  <generated prelude> + decimal_ops.cuh + lowcard_args.h + this file  once per
// plan shape, the way the reference generates one accumulator class per aggregation (AccumulatorCompiler.java:261-330).
//
// The prelude defines
//   PG_T            threads per block            PG_VEC       consecutive rows per thread and trip (1, 2 or 4)
//   PG_REC_WORDS    u64 words per group/thread   PG_MIN_BLOCKS  __launch_bounds__ second argument
//   PG_NSUMS        number of SUM slots          PG_PREFETCH  trips ahead whose rows are prefetched into L2 (0 = off)
//   PG_FAST_BITS    magnitude bound (bits) of addends the private sums take
//   PG_KEYS(X)      X(k, CLASS, WORD, SHIFT, HASNULL, WOFF)   CLASS in {K64, KDBL, K32, K8, KVAR}
//   PG_LOADS(X)     X(l, LKIND, HASNULL)                      LKIND in {U64, I32, U8, U128}: input column l of the row loop
//   PG_SLOTS(X)     X(s, KIND, WORD)                          KIND in {SUM, F64, MIN, MAX, CNT}: private state, one word each
//   PG_ROW_FILTER   statements for row v that end by defining `const bool pgx_pass` (a fused filter; `true` without one)
//   PG_ROW_UPDATE   statements for row v: fused projections, the aggregates' conditions and the updates of `rec` / `wrec`
// The two statement macros are generated text over ld<l>[v] / ldh<l>[v] (value, upper half of 128-bit values) and ldn<l>[v]
// (null byte) — see jit_lowcard.cu and project_codegen.  Everything structural is a literal, so the row loop is straight-line
// code; pointers and sizes arrive in PgjArgs.
//
// State.  Every thread owns one 64-bit word per (group, slot) in shared memory, table[(group * PG_REC_WORDS + word) * PG_T +
// thread]: a warp touching one word of 32 different (or equal) groups hits 32 different banks, and nobody else writes a thread's
// column, so there are no atomics in the row loop.  Integer and decimal sums are exact although the private word is narrow: an
// addend v with |v| < 2^PG_FAST_BITS goes into the private signed 64-bit sum, which cannot wrap within one launch (the host derives
// the bound from the rows a thread can see); any other addend (large magnitudes, 128-bit values) is added to a 192-bit
// accumulator per (group, slot) that the whole block shares (`wide`, shared-memory atomics with carry propagation).  The flush
// sign-extends and adds up the private sums of a warp, warp 0 adds the block's wide accumulators, and the totals are applied to
// the global state with carry-chain atomics.
//
// Groups.  The key -> group id table (lut) is either filled from the operator's known groups (a.known_groups > 0: packed keys in
// a.lc_groups) or discovered by the kernel itself from the first a.discover_rows rows of the page: every block runs the same
// deterministic discovery (fingerprints claimed with CAS, first row per fingerprint by atomicMin, keys written by the first
// row's owner, every row verified against the stored key, ids by ascending first row = the reference's first-seen order), so all
// blocks agree without talking to each other, and block 0 publishes the groups for the host.  Rows whose key is not in the
// table are flagged in a.miss_flag and left to the general path.

typedef unsigned long long u64;
typedef long long i64;
typedef unsigned int u32;
typedef unsigned char u8;

#define PGJ_OP_COUNT 0
#define PGJ_OP_SUM_I64 1
#define PGJ_OP_SUM_F64 2
#define PGJ_OP_SUM_DEC 3
#define PGJ_OP_MIN64 4
#define PGJ_OP_MAX64 5

#define PGJ_CAT2(a, b) a##b
#define PGJ_CAT(a, b) PGJ_CAT2(a, b)

// shared-memory key -> group id table: one 32-byte entry per slot, read with two independent 16-byte loads
struct __align__(16) LutEntry { u64 w0, w1; int gid; int first_row; u64 fp; };

// ------------------------------------------------------------------ vector loads (read-only path)
template <int V> __device__ __forceinline__ void load_u64(const void* base, i64 row, u64 (&o)[V]) {
    const u64* p = (const u64*)base + row;
    if constexpr (V == 1) o[0] = __ldg(p);
    else {
#pragma unroll
        for (int i = 0; i < V; i += 2) { ulonglong2 t = __ldg((const ulonglong2*)(p + i)); o[i] = t.x; o[i + 1] = t.y; }
    }
}
template <int V> __device__ __forceinline__ void load_u128(const void* base, i64 row, u64 (&lo)[V], u64 (&hi)[V]) {
    const ulonglong2* p = (const ulonglong2*)base + row;
#pragma unroll
    for (int i = 0; i < V; i++) { ulonglong2 t = __ldg(p + i); lo[i] = t.x; hi[i] = t.y; }
}
template <int V> __device__ __forceinline__ void load_i32(const void* base, i64 row, int (&o)[V]) {
    const int* p = (const int*)base + row;
    if constexpr (V == 1) o[0] = __ldg(p);
    else if constexpr (V == 2) { int2 t = __ldg((const int2*)p); o[0] = t.x; o[1] = t.y; }
    else { int4 t = __ldg((const int4*)p); o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w; }
}
template <int V> __device__ __forceinline__ void load_u8(const u8* base, i64 row, u32 (&o)[V]) {
    const u8* p = base + row;
    if constexpr (V == 1) o[0] = __ldg(p);
    else if constexpr (V == 2) { unsigned short t = __ldg((const unsigned short*)p); o[0] = t & 0xffu; o[1] = t >> 8; }
    else { u32 t = __ldg((const u32*)p); o[0] = t & 0xffu; o[1] = (t >> 8) & 0xffu; o[2] = (t >> 16) & 0xffu; o[3] = t >> 24; }
}
template <int V> __device__ __forceinline__ void load_offsets(const int* off, i64 row, int (&o)[V + 1]) {
    int t[V];
    load_i32<V>(off, row, t);
#pragma unroll
    for (int i = 0; i < V; i++) o[i] = t[i];
    o[V] = __ldg(off + row + V);
}

// ------------------------------------------------------------------ key helpers (same packed form for rows and stored groups)
__device__ __forceinline__ u64 norm_double_key(u64 b) {
    if ((b & 0x7fffffffffffffffULL) == 0) return 0;                                             // -0.0 == +0.0
    if ((b & 0x7ff0000000000000ULL) == 0x7ff0000000000000ULL && (b & 0x000fffffffffffffULL)) return 0x7ff8000000000000ULL;
    return b;
}
__device__ __forceinline__ u32 lut_hash(u64 w0, u64 w1) {
    u32 x = (u32)w0 ^ (u32)(w0 >> 32) ^ (u32)w1 ^ (u32)(w1 >> 29);
    return (x * 0x9E3779B1u) >> 24;
}
__device__ __forceinline__ u64 key_fingerprint(u64 w0, u64 w1) {      // 64-bit, never zero (zero = unclaimed table entry)
    u64 h = (w0 ^ (w1 * 0x9E3779B97F4A7C15ULL)) * 0xff51afd7ed558ccdULL;
    h ^= h >> 33; h = (h ^ w1) * 0xc4ceb9fe1a85ec53ULL; h ^= h >> 29;
    return h | 1ULL;
}
// strings of up to 3 bytes pack as len | b0<<8 | b1<<16 | b2<<24
__device__ __forceinline__ u64 pack_short_string(const u8* b, int len) {
    u64 v = (u64)len;
    if (len > 0) v |= (u64)__ldg(b) << 8;
    if (len > 1) v |= (u64)__ldg(b + 1) << 16;
    if (len > 2) v |= (u64)__ldg(b + 2) << 24;
    return v;
}

// ------------------------------------------------------------------ state updates
// block-wide 192-bit accumulator (shared memory): the exact home of every addend the narrow private sums cannot take
__device__ __noinline__ void wide_add(u64* w, u64 a0, u64 a1, u64 a2) {
    const u64 old0 = atomicAdd(&w[0], a0);
    const u64 c0 = (old0 + a0) < a0 ? 1ULL : 0ULL;
    const u64 add1 = a1 + c0;
    u64 c1 = add1 < c0 ? 1ULL : 0ULL;
    if (add1) {
        const u64 old1 = atomicAdd(&w[1], add1);
        c1 += (old1 + add1) < add1 ? 1ULL : 0ULL;
    }
    const u64 add2 = a2 + c1;
    if (add2) atomicAdd(&w[2], add2);
}
// |v| < 2^PG_FAST_BITS as a signed 64-bit value: such addends go into the thread's private sum
__device__ __forceinline__ bool fits_fast(u64 v) { return ((v + (1ULL << PG_FAST_BITS)) >> (PG_FAST_BITS + 1)) == 0; }
__device__ __forceinline__ void upd_f64(u64* t, u64 bits, int conv) {
    double d = conv ? (double)(i64)bits : __longlong_as_double((i64)bits);
    t[0] = (u64)__double_as_longlong(__longlong_as_double((i64)t[0]) + d);
}
__device__ __forceinline__ i64 double_order_key(i64 b) {
    b = (i64)(((u64)b & 0x7ff0000000000000ULL) == 0x7ff0000000000000ULL && ((u64)b & 0x000fffffffffffffULL) ? 0x7ff8000000000000ULL : (u64)b);
    return b ^ ((b >> 63) & 0x7fffffffffffffffLL);
}
__device__ __forceinline__ bool bits_is_nan(i64 b) {
    return ((u64)b & 0x7ff0000000000000ULL) == 0x7ff0000000000000ULL && ((u64)b & 0x000fffffffffffffULL);
}
__device__ __forceinline__ void upd_min(u64* t, u64 bits, int conv) {
    i64 k = conv ? double_order_key((i64)bits) : (i64)bits;
    if (k < (i64)t[0]) t[0] = (u64)k;
}
__device__ __forceinline__ void upd_max(u64* t, u64 bits, int conv) {   // max(double): NaN loses (MinMaxCompare.java:100-106)
    i64 k = conv ? (bits_is_nan((i64)bits) ? (i64)0x8000000000000000LL : double_order_key((i64)bits)) : (i64)bits;
    if (k > (i64)t[0]) t[0] = (u64)k;
}

// ------------------------------------------------------------------ global state updates at flush (carry-chain atomics)
__device__ __forceinline__ void atomic_add_chain2(u64* w, u64 a0, u64 a1) {
    u64 old0 = atomicAdd(&w[0], a0);
    u64 add1 = a1 + ((old0 + a0) < a0 ? 1ULL : 0ULL);
    if (add1) atomicAdd(&w[1], add1);
}
__device__ __forceinline__ void atomic_add_chain3(u64* w, u64 a0, u64 a1, u64 a2) {
    u64 old0 = atomicAdd(&w[0], a0);
    u64 c0 = (old0 + a0) < a0 ? 1ULL : 0ULL;
    u64 add1 = a1 + c0;
    u64 c1 = add1 < c0 ? 1ULL : 0ULL;
    if (add1) {
        u64 old1 = atomicAdd(&w[1], add1);
        c1 += (old1 + add1) < add1 ? 1ULL : 0ULL;
    }
    u64 add2 = a2 + c1;
    if (add2) atomicAdd(&w[2], add2);
}
__device__ __forceinline__ void wreduce_sum_carry(u64& s, u64& carry) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        u64 os = __shfl_down_sync(0xffffffffu, s, d);
        u64 oc = __shfl_down_sync(0xffffffffu, carry, d);
        u64 ns = s + os;
        carry += oc + (ns < s ? 1ULL : 0ULL);
        s = ns;
    }
}

// ------------------------------------------------------------------ per-kind macro pieces
#define PGJ_LD_U64(l) u64 PGJ_CAT(ld, l)[V]; load_u64<V>(a.load_ptr[l], row, PGJ_CAT(ld, l));
#define PGJ_LD_U128(l) u64 PGJ_CAT(ld, l)[V], PGJ_CAT(ldh, l)[V]; load_u128<V>(a.load_ptr[l], row, PGJ_CAT(ld, l), PGJ_CAT(ldh, l));
#define PGJ_LD_I32(l) u64 PGJ_CAT(ld, l)[V]; { int t32[V]; load_i32<V>(a.load_ptr[l], row, t32); _Pragma("unroll") for (int q = 0; q < V; q++) PGJ_CAT(ld, l)[q] = (u64)(i64)t32[q]; }
#define PGJ_LD_U8(l) u64 PGJ_CAT(ld, l)[V]; { u32 t8[V]; load_u8<V>((const u8*)a.load_ptr[l], row, t8); _Pragma("unroll") for (int q = 0; q < V; q++) PGJ_CAT(ld, l)[q] = t8[q]; }
#define PGJ_ISKIND_SUM 0
#define PGJ_ISKIND_F64 3
#define PGJ_ISKIND_MIN 4
#define PGJ_ISKIND_MAX 5
#define PGJ_ISKIND_CNT 6

// key classes: stage 1 = loads that depend on the row only, stage 2 = value (strings: bytes behind the offsets)
#define PGJ_K1_K64(k) u64 PGJ_CAT(kv, k)[V]; load_u64<V>(a.key_values[k], row, PGJ_CAT(kv, k));
#define PGJ_K1_KDBL(k) PGJ_K1_K64(k)
#define PGJ_K1_K32(k) u64 PGJ_CAT(kv, k)[V]; { int t32[V]; load_i32<V>(a.key_values[k], row, t32); _Pragma("unroll") for (int q = 0; q < V; q++) PGJ_CAT(kv, k)[q] = (u64)(u32)t32[q]; }
#define PGJ_K1_K8(k) u64 PGJ_CAT(kv, k)[V]; { u32 t8[V]; load_u8<V>((const u8*)a.key_values[k], row, t8); _Pragma("unroll") for (int q = 0; q < V; q++) PGJ_CAT(kv, k)[q] = t8[q]; }
#define PGJ_K1_KVAR(k) u64 PGJ_CAT(kv, k)[V]; int PGJ_CAT(ko, k)[V + 1]; load_offsets<V>(a.key_offsets[k], row, PGJ_CAT(ko, k)); \
        u32 PGJ_CAT(ks, k)[V]; const bool PGJ_CAT(kspec, k) = a.row_base + row + V <= kvb[k];                                               \
        if (PGJ_CAT(kspec, k)) load_u8<V>((const u8*)a.key_values[k] + a.row_base, row, PGJ_CAT(ks, k));
#define PGJ_K2_K64(k)
#define PGJ_K2_KDBL(k) _Pragma("unroll") for (int q = 0; q < V; q++) PGJ_CAT(kv, k)[q] = norm_double_key(PGJ_CAT(kv, k)[q]);
#define PGJ_K2_K32(k)
#define PGJ_K2_K8(k)
#define PGJ_K2_KVAR(k) {                                                                                                  \
        bool dense = PGJ_CAT(kspec, k);   /* offsets[q] == first row of the page + q: one byte per row, bytes already loaded */         \
        _Pragma("unroll") for (int q = 0; q <= V; q++) dense = dense && PGJ_CAT(ko, k)[q] == (int)(a.row_base + row) + q;                   \
        if (dense) { _Pragma("unroll") for (int q = 0; q < V; q++) PGJ_CAT(kv, k)[q] = 1ULL | ((u64)PGJ_CAT(ks, k)[q] << 8); } \
        else { _Pragma("unroll") for (int q = 0; q < V; q++) { int o = PGJ_CAT(ko, k)[q], len = PGJ_CAT(ko, k)[q + 1] - o;     \
            if (len > 3) { fits[q] = false; PGJ_CAT(kv, k)[q] = 0; } else PGJ_CAT(kv, k)[q] = pack_short_string((const u8*)a.key_values[k] + o, len); } } }

// what one pass over V consecutive rows does with them
#define PGJ_MODE_AGG 0       // look the groups up and update the thread's private records
#define PGJ_MODE_CLAIM 1     // discovery pass 1: claim a table entry per key fingerprint, remember the first row that carries it
#define PGJ_MODE_WRITE 2     // discovery pass 2: the first row's owner stores the key
#define PGJ_MODE_VERIFY 3    // discovery pass 3: every row's key must equal the stored key of its entry (fingerprint collisions)

// ------------------------------------------------------------------ V consecutive rows of one thread
template <int V, int MODE>
__device__ __forceinline__ void process_rows(const PgjArgs& a, const i64* kvb, i64 row, u64* table, u64* wide, LutEntry* lut, int tid, u64& misses,
                                             int& dflags) {
    bool fits[V];
    int pgx_err = 0;
#pragma unroll
    for (int q = 0; q < V; q++) fits[q] = true;
    // ---- stage 1: every load whose address depends on the row number only
#define PG_X(k, CLASS, WORD, SHIFT, HASNULL, WOFF) PGJ_K1_##CLASS(k) u32 PGJ_CAT(kn, k)[V]; (void)PGJ_CAT(kn, k); if (HASNULL) load_u8<V>(a.key_nulls[k], row, PGJ_CAT(kn, k));
    PG_KEYS(PG_X)
#undef PG_X
#define PG_X(l, LKIND, HASNULL) PGJ_LD_##LKIND(l) u32 PGJ_CAT(ldn, l)[V]; (void)PGJ_CAT(ldn, l); if (HASNULL) load_u8<V>(a.load_nulls[l], row, PGJ_CAT(ldn, l));
    PG_LOADS(PG_X)
#undef PG_X
    // ---- stage 2: key values (string bytes sit behind the offsets just loaded)
#define PG_X(k, CLASS, WORD, SHIFT, HASNULL, WOFF) PGJ_K2_##CLASS(k)
    PG_KEYS(PG_X)
#undef PG_X
    // ---- phase A: pack the keys and find the groups of all V rows (independent look-ups overlap each other)
    int g[V];
    u64 kw0[V], kw1[V];
#pragma unroll
    for (int v = 0; v < V; v++) {
        u64 w0 = 0, w1 = 0;
#define PG_X(k, CLASS, WORD, SHIFT, HASNULL, WOFF)                                          \
        if (HASNULL && PGJ_CAT(kn, k)[v]) w1 |= 1ULL << (56 + k);                           \
        else if (WORD == 0) w0 |= PGJ_CAT(kv, k)[v] << SHIFT;                               \
        else w1 |= PGJ_CAT(kv, k)[v] << SHIFT;
        PG_KEYS(PG_X)
#undef PG_X
        kw0[v] = w0; kw1[v] = w1;
        g[v] = -1;
        if (MODE == PGJ_MODE_AGG && fits[v]) {
            u32 h = lut_hash(w0, w1);
            while (true) {
                const LutEntry e = lut[h];
                if (e.w0 == w0 && e.w1 == w1) { g[v] = e.gid; break; }
                if (e.gid < 0) break;
                h = (h + 1) & (PGJ_LUT_SIZE - 1);
            }
        }
    }
    // ---- phase B, row by row (two rows may share a group).  A row the fused filter rejects does not exist for the operator:
    // it neither counts nor misses nor founds a group.
#pragma unroll
    for (int v = 0; v < V; v++) {
        PG_ROW_FILTER
        if (!pgx_pass) continue;
        if (MODE == PGJ_MODE_AGG) {
            if (g[v] < 0) { misses++; atomicOr(&a.miss_flag[(row + v) >> 5], 1u << ((row + v) & 31)); continue; }
            u64* rec = table + (g[v] * PG_REC_WORDS) * PG_T + tid;
            u64* wrec = wide + g[v] * (PG_NSUMS * 3);
            u32 pg_slow = 0;
            (void)wrec; (void)pg_slow;
            PG_ROW_UPDATE
        } else {
            if (!fits[v]) { dflags |= PGJ_DISC_UNPACKABLE; continue; }     // the host takes the general path
            const u64 fp = key_fingerprint(kw0[v], kw1[v]);
            u32 h = lut_hash(kw0[v], kw1[v]);
            int probes = 0;
            bool found = false;
            while (probes < PGJ_LUT_SIZE) {
                u64 cur = lut[h].fp;
                if (MODE == PGJ_MODE_CLAIM && cur == 0) { cur = atomicCAS(&lut[h].fp, 0ULL, fp); if (cur == 0) cur = fp; }
                if (cur == fp) { found = true; break; }
                if (MODE != PGJ_MODE_CLAIM && cur == 0) break;
                h = (h + 1) & (PGJ_LUT_SIZE - 1);
                probes++;
            }
            if (!found) { dflags |= PGJ_DISC_TABLE_FULL; continue; }      // more distinct keys than the table holds
            if (MODE == PGJ_MODE_CLAIM) atomicMin(&lut[h].first_row, (int)(row + v));
            else if (MODE == PGJ_MODE_WRITE) { if (lut[h].first_row == (int)(row + v)) { lut[h].w0 = kw0[v]; lut[h].w1 = kw1[v]; } }
            else if (lut[h].w0 != kw0[v] || lut[h].w1 != kw1[v]) dflags |= PGJ_DISC_COLLISION;
        }
    }
    if (MODE == PGJ_MODE_AGG && pgx_err) atomicOr((unsigned int*)(a.d_miss + 1), 1u);
}

// L2 prefetch of the V rows a later trip of this thread will read (PG_PREFETCH trips ahead): keeps more of HBM's
// latency covered than the registers of one trip can
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
template <int V>
__device__ __forceinline__ void prefetch_rows(const PgjArgs& a, i64 row) {
#define PGJ_PF_K64(k) prefetch_l2((const u64*)a.key_values[k] + row);
#define PGJ_PF_KDBL(k) PGJ_PF_K64(k)
#define PGJ_PF_K32(k) prefetch_l2((const int*)a.key_values[k] + row);
#define PGJ_PF_K8(k) prefetch_l2((const u8*)a.key_values[k] + row);
#define PGJ_PF_KVAR(k) prefetch_l2(a.key_offsets[k] + row); prefetch_l2((const u8*)a.key_values[k] + a.row_base + row);
#define PG_X(k, CLASS, WORD, SHIFT, HASNULL, WOFF) PGJ_PF_##CLASS(k) if (HASNULL) prefetch_l2(a.key_nulls[k] + row);
    PG_KEYS(PG_X)
#undef PG_X
#define PGJ_PFL_U64(l) prefetch_l2((const u64*)a.load_ptr[l] + row);
#define PGJ_PFL_U128(l) _Pragma("unroll") for (int q = 0; q < V; q += 2) prefetch_l2((const ulonglong2*)a.load_ptr[l] + row + q);
#define PGJ_PFL_I32(l) prefetch_l2((const int*)a.load_ptr[l] + row);
#define PGJ_PFL_U8(l) prefetch_l2((const u8*)a.load_ptr[l] + row);
#define PG_X(l, LKIND, HASNULL) PGJ_PFL_##LKIND(l) if (HASNULL) prefetch_l2(a.load_nulls[l] + row);
    PG_LOADS(PG_X)
#undef PG_X
}

// one discovery pass of the whole block over the first a.discover_rows rows
template <int MODE>
__device__ __forceinline__ void discover_pass(const PgjArgs& a, const i64* kvb, LutEntry* lut, int tid, int& dflags) {
    u64 dummy = 0;
    const i64 m = a.discover_rows;
    for (i64 row = (i64)tid * PG_VEC; row + PG_VEC <= m; row += (i64)PG_T * PG_VEC) process_rows<PG_VEC, MODE>(a, kvb, row, nullptr, nullptr, lut, tid, dummy, dflags);
    if (PG_VEC > 1 && tid == 0)
        for (i64 row = m - m % PG_VEC; row < m; row++) process_rows<1, MODE>(a, kvb, row, nullptr, nullptr, lut, tid, dummy, dflags);
}

extern "C" __global__ void __launch_bounds__(PG_T, PG_MIN_BLOCKS) pg_lowcard_agg(const PgjArgs a) {
    extern __shared__ u64 table[];
    __shared__ LutEntry lut[PGJ_LUT_SIZE];
    __shared__ int s_groups, s_flags;
    __shared__ i64 kvb[PGJ_MAX_KEYS];                       // bytes in each variable width key's value buffer
    const int tid = threadIdx.x;
    if (tid < PGJ_MAX_KEYS) kvb[tid] = (a.key_value_bytes[tid] < 0 && a.key_offsets[tid]) ? (i64)a.key_offsets[tid][a.n] : a.key_value_bytes[tid];
    const int GCAP = a.G;                                   // groups the private tables are sized for
    u64* wide = table + (i64)GCAP * PG_REC_WORDS * PG_T;     // [group][sum slot][3]
    // ---- own column: zeros, then min/max identities; the block's wide accumulators
    for (int i = 0; i < GCAP * PG_REC_WORDS; i++) table[i * PG_T + tid] = 0;
#define PG_X(s, KIND, WORD)                                                                                               \
    if (PGJ_ISKIND_##KIND == PGJ_KIND_MIN) for (int g = 0; g < GCAP; g++) table[(g * PG_REC_WORDS + WORD) * PG_T + tid] = 0x7fffffffffffffffULL; \
    if (PGJ_ISKIND_##KIND == PGJ_KIND_MAX) for (int g = 0; g < GCAP; g++) table[(g * PG_REC_WORDS + WORD) * PG_T + tid] = 0x8000000000000000ULL;
    PG_SLOTS(PG_X)
#undef PG_X
    for (int i = tid; i < GCAP * PG_NSUMS * 3; i += PG_T) wide[i] = 0;
    for (int i = tid; i < PGJ_LUT_SIZE; i += PG_T) { lut[i].w0 = ~0ULL; lut[i].w1 = ~0ULL; lut[i].gid = -1; lut[i].first_row = 0x7fffffff; lut[i].fp = 0; }
    if (tid == 0) { s_groups = 0; s_flags = 0; }
    __syncthreads();

    if (a.known_groups > 0) {
        // ---- the operator's groups: packed keys in id order
        if (tid == 0) {
            const int G = a.known_groups;
            for (int g = 0; g < G; g++) {
                const u64 w0 = a.lc_groups[4 * g], w1 = a.lc_groups[4 * g + 1];
                if (a.lc_groups[4 * g + 3] == 0) continue;            // a group whose key the packed form cannot hold: its rows miss
                u32 h = lut_hash(w0, w1);
                while (lut[h].gid >= 0) h = (h + 1) & (PGJ_LUT_SIZE - 1);
                lut[h].w0 = w0; lut[h].w1 = w1; lut[h].gid = g;
            }
            s_groups = G;
        }
        __syncthreads();
    } else {
        // ---- discovery (every block computes the same table)
        int dflags = 0;
        discover_pass<PGJ_MODE_CLAIM>(a, kvb, lut, tid, dflags);
        __syncthreads();
        discover_pass<PGJ_MODE_WRITE>(a, kvb, lut, tid, dflags);
        __syncthreads();
        discover_pass<PGJ_MODE_VERIFY>(a, kvb, lut, tid, dflags);
        if (dflags) atomicOr(&s_flags, dflags);
        __syncthreads();
        // ids by ascending first row
        for (int i = tid; i < PGJ_LUT_SIZE; i += PG_T) {
            if (lut[i].fp == 0) continue;
            const int mine = lut[i].first_row;
            int rank = 0;
            for (int j = 0; j < PGJ_LUT_SIZE; j++) rank += (lut[j].fp != 0 && lut[j].first_row < mine) ? 1 : 0;
            lut[i].gid = rank;
            atomicAdd(&s_groups, 1);
        }
        __syncthreads();
        if (s_groups > GCAP && tid == 0) atomicOr(&s_flags, PGJ_DISC_TOO_MANY);
        __syncthreads();
        if (blockIdx.x == 0) {      // publish: group table for the host, group count, discovery verdict
            for (int i = tid; i < PGJ_LUT_SIZE; i += PG_T) {
                if (lut[i].fp == 0 || lut[i].gid >= PGJ_MAX_GROUPS) continue;
                u64* o = a.lc_groups_out + 4 * lut[i].gid;
                o[0] = lut[i].w0; o[1] = lut[i].w1; o[2] = (u64)lut[i].first_row; o[3] = 1;
            }
            if (tid == 0) { a.d_miss[2] = (u64)s_groups; a.d_miss[3] = (u64)s_flags; }
        }
        if (s_flags) return;        // block-uniform: nothing has been accumulated, the host retries or takes the general path
    }

    // ---- row loop: thread t of block b takes rows [(b*T + t)*VEC, +VEC), then strides by the whole grid
    u64 misses = 0;
    int unused = 0;
    const i64 n = a.n;
    const i64 stride = (i64)gridDim.x * PG_T * PG_VEC;
    for (i64 row = ((i64)blockIdx.x * PG_T + tid) * PG_VEC; row + PG_VEC <= n; row += stride) {
#if PG_PREFETCH > 0
        if (row + PG_PREFETCH * stride + PG_VEC <= n) prefetch_rows<PG_VEC>(a, row + PG_PREFETCH * stride);
#endif
        process_rows<PG_VEC, PGJ_MODE_AGG>(a, kvb, row, table, wide, lut, tid, misses, unused);
    }
    if (PG_VEC > 1 && blockIdx.x == 0 && tid == 0)
        for (i64 row = n - n % PG_VEC; row < n; row++) process_rows<1, PGJ_MODE_AGG>(a, kvb, row, table, wide, lut, tid, misses, unused);
    __syncthreads();                // the block's wide accumulators are complete

    // ---- flush: per aggregate and group, reduce the warp's 32 columns and fold them into the global state
    const int lane = tid & 31;
    const int G = s_groups;
    {
        u64 z = 0;
        wreduce_sum_carry(misses, z);
        if (lane == 0 && misses) atomicAdd(a.d_miss, misses);
    }
    for (int ai = 0; ai < a.naggs; ai++) {
        const int op = a.agg_op[ai], kind = a.agg_val_kind[ai];
        for (int g = 0; g < G; g++) {
            u64* st = a.agg_state[ai] + (i64)g * a.agg_words[ai];
            if (kind == PGJ_KIND_SUM && tid == 0) {
                // addends that went to the block's wide accumulator (their rows are counted in the private count words)
                const u64* w = wide + (g * PG_NSUMS + a.agg_wide_idx[ai]) * 3;
                if (w[0] | w[1] | w[2]) {
                    if (op == PGJ_OP_SUM_I64) atomic_add_chain2(st + 1, w[0], w[1]);
                    else atomic_add_chain3(st + 1, w[0], w[1], w[2]);
                }
            }
            const u64* rec = table + (g * PG_REC_WORDS) * PG_T + tid;
            u64 cnt = rec[a.agg_cnt_word[ai] * PG_T], zero = 0;
            wreduce_sum_carry(cnt, zero);
            cnt = __shfl_sync(0xffffffffu, cnt, 0);
            if (cnt == 0) continue;   // warp-uniform: none of this warp's rows reached the aggregate in this group
            if (op == PGJ_OP_COUNT) { if (lane == 0) atomicAdd(&st[0], cnt); continue; }
            const u64* col = rec + a.agg_val_word[ai] * PG_T;
            if (kind == PGJ_KIND_SUM) {
                // the thread's signed 64-bit sum, sign-extended and added up over the warp
                const i64 mine = (i64)col[0];
                u64 s0 = (u64)mine, c0 = 0;
                wreduce_sum_carry(s0, c0);
                u64 s1 = (u64)(mine >> 63);     // 0 or all ones
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) s1 += __shfl_down_sync(0xffffffffu, s1, d);
                if (lane == 0) {
                    const u64 a1 = s1 + c0;
                    atomicAdd(&st[0], cnt);
                    if (op == PGJ_OP_SUM_I64) atomic_add_chain2(st + 1, s0, a1);
                    else atomic_add_chain3(st + 1, s0, a1, (u64)((i64)a1 >> 63));
                }
            } else if (kind == PGJ_KIND_F64) {
                double d = __longlong_as_double((i64)col[0]);
#pragma unroll
                for (int k = 16; k > 0; k >>= 1) d += __shfl_down_sync(0xffffffffu, d, k);
                if (lane == 0) { atomicAdd(&st[0], cnt); atomicAdd((double*)&st[1], d); }
            } else if (kind == PGJ_KIND_MIN || kind == PGJ_KIND_MAX) {
                i64 v = (i64)col[0];
#pragma unroll
                for (int k = 16; k > 0; k >>= 1) {
                    i64 o = __shfl_down_sync(0xffffffffu, v, k);
                    v = kind == PGJ_KIND_MIN ? (o < v ? o : v) : (o > v ? o : v);
                }
                if (lane == 0) {
                    st[0] = 1;
                    if (kind == PGJ_KIND_MIN) atomicMin((i64*)&st[1], v); else atomicMax((i64*)&st[1], v);
                }
            }
        }
    }
}
