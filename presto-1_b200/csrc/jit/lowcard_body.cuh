// lowcard_body.cuh — body of the run-time specialised fused GROUP BY kernel for few groups ("hash-then-update with
// shared-memory pre-aggregation").  NVRTC compiles  <generated prelude> + lowcard_args.h + this file  once per plan shape,
// the way the reference generates one accumulator class per aggregation (AccumulatorCompiler.java:261-330).
//
// The prelude defines
//   PG_T            threads per block            PG_VEC       consecutive rows per thread and trip (1, 2 or 4)
//   PG_REC_WORDS    u64 words per group/thread   PG_MIN_BLOCKS  __launch_bounds__ second argument
//   PG_PREFETCH     trips ahead whose rows are prefetched into L2 (0 = off)
//   PG_KEYS(X)      X(k, CLASS, WORD, SHIFT, HASNULL, WOFF)   CLASS in {K64, KDBL, K32, K8, KVAR}
//   PG_LOADS(X)     X(l, LKIND, HASNULL)                      LKIND in {U64, I32, U8, U128}: input column l of the row loop
//   PG_SLOTS(X)     X(s, KIND, WORD)                          KIND in {DEC, I64, I32, F64, MIN, MAX, CNT}: private state layout
//   PG_ROW_FILTER   statements for row v that end by defining `const bool pgx_pass` (a fused filter; `true` without one)
//   PG_ROW_UPDATE   statements for row v: fused projections, the aggregates' conditions and the updates of `rec`
// The two statement macros are generated text over ld<l>[v] / ldh<l>[v] (value, upper half of 128-bit values) and ldn<l>[v]
// (null byte) — see jit_lowcard.cu and project_codegen.  Everything structural is a literal, so the row loop is straight-line
// code; pointers and sizes arrive in PgjArgs.
//
// Layout of the private tables in shared memory: table[(group * PG_REC_WORDS + word) * PG_T + thread] — a warp touching
// one word of 32 different (or equal) groups hits 32 different banks; every thread owns its column, no atomics.

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
struct __align__(16) LutEntry { u64 w0, w1; int gid; int pad[3]; };

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
// strings of up to 3 bytes pack as len | b0<<8 | b1<<16 | b2<<24
__device__ __forceinline__ u64 pack_short_string(const u8* b, int len) {
    u64 v = (u64)len;
    if (len > 0) v |= (u64)__ldg(b) << 8;
    if (len > 1) v |= (u64)__ldg(b + 1) << 16;
    if (len > 2) v |= (u64)__ldg(b + 2) << 24;
    return v;
}

// ------------------------------------------------------------------ private state updates (thread-owned shared memory words)
// 64-bit input into a 128-bit sum.  The upper word is touched only when a carry or a negative value reaches it.
__device__ __forceinline__ void upd_i64(u64* t, u64 v) {
    u64 lo = t[0] + v;
    t[0] = lo;
    if (lo < v || (i64)v < 0) t[PG_T] += (u64)((i64)v >> 63) + (lo < v ? 1ULL : 0ULL);
}
// 128-bit sign-magnitude input into a 192-bit two's complement sum; negatives are negated first (x = ~mag + 1, sign word ~0),
// words 1 and 2 are touched only when the addend or a carry reaches them.
__device__ __forceinline__ void upd_dec(u64* t, u64 lo, u64 hi) {
    u64 a0 = lo, a1 = hi & 0x7fffffffffffffffULL, a2 = 0;
    if ((i64)hi < 0 && (a0 | a1)) {
        a0 = ~a0; a1 = ~a1; a2 = ~0ULL;
        asm("add.cc.u64 %0, %0, 1;\n\taddc.u64 %1, %1, 0;" : "+l"(a0), "+l"(a1));
    }
    u64 s0 = t[0] + a0;
    t[0] = s0;
    u64 c0 = s0 < a0 ? 1ULL : 0ULL;
    if (a1 | a2 | c0) {
        u64 s1 = t[PG_T], s2 = t[2 * PG_T];
        // addend words 1..2 first absorb the carry of word 0, then join the sum: two independent two-word chains
        asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(a1), "+l"(a2) : "l"(c0));
        asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(s1), "+l"(s2) : "l"(a1), "l"(a2));
        t[PG_T] = s1; t[2 * PG_T] = s2;
    }
}
// the same with a two's complement addend (a derived decimal of a fused projection): sign-extend into the third word
__device__ __forceinline__ void upd_dec2c(u64* t, u64 w0, i64 w1) {
    u64 s0 = t[0] + w0;
    t[0] = s0;
    u64 c0 = s0 < w0 ? 1ULL : 0ULL;
    if (w1 | (i64)c0) {
        u64 a1 = (u64)w1, a2 = (u64)(w1 >> 63);
        u64 s1 = t[PG_T], s2 = t[2 * PG_T];
        asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(a1), "+l"(a2) : "l"(c0));
        asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(s1), "+l"(s2) : "l"(a1), "l"(a2));
        t[PG_T] = s1; t[2 * PG_T] = s2;
    }
}
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
#define PGJ_ISKIND_DEC 0
#define PGJ_ISKIND_I64 1
#define PGJ_ISKIND_I32 2
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
        u32 PGJ_CAT(ks, k)[V]; const bool PGJ_CAT(kspec, k) = a.row_base + row + V <= a.key_value_bytes[k];                               \
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

// ------------------------------------------------------------------ V consecutive rows of one thread
template <int V>
__device__ __forceinline__ void process_rows(const PgjArgs& a, i64 row, u64* table, const LutEntry* lut,
                                             int tid, u64& misses) {
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
#pragma unroll
    for (int v = 0; v < V; v++) {
        u64 w0 = 0, w1 = 0;
#define PG_X(k, CLASS, WORD, SHIFT, HASNULL, WOFF)                                          \
        if (HASNULL && PGJ_CAT(kn, k)[v]) w1 |= 1ULL << (56 + k);                           \
        else if (WORD == 0) w0 |= PGJ_CAT(kv, k)[v] << SHIFT;                               \
        else w1 |= PGJ_CAT(kv, k)[v] << SHIFT;
        PG_KEYS(PG_X)
#undef PG_X
        g[v] = -1;
        if (fits[v]) {
            u32 h = lut_hash(w0, w1);
            while (true) {
                const LutEntry e = lut[h];
                if (e.w0 == w0 && e.w1 == w1) { g[v] = e.gid; break; }
                if (e.gid < 0) break;
                h = (h + 1) & (PGJ_LUT_SIZE - 1);
            }
        }
    }
    // ---- phase B: update the thread's private records, row by row (two rows may share a group).  A row the fused filter
    // rejects does not exist for the operator: it neither counts nor misses.
#pragma unroll
    for (int v = 0; v < V; v++) {
        PG_ROW_FILTER
        if (!pgx_pass) continue;
        if (g[v] < 0) { misses++; atomicOr(&a.miss_flag[(row + v) >> 5], 1u << ((row + v) & 31)); continue; }
        u64* rec = table + (g[v] * PG_REC_WORDS) * PG_T + tid;
        PG_ROW_UPDATE
    }
    if (pgx_err) atomicOr((unsigned int*)(a.d_miss + 1), 1u);
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

// stored key of group g in the packed form (thread 0 builds the look-up table once per block)
__device__ bool pack_group(const PgjArgs& a, int g, u64& w0, u64& w1) {
    w0 = 0; w1 = 0;
    if (a.bigint_mode) {
        if (g == a.null_gid) w1 |= 1ULL << 56; else w0 = (u64)a.values_by_group[g];
        return true;
    }
    const u64* rowp = a.key_rows + (i64)g * a.row_words;
    bool ok = true;
#define PGJ_G_K64(w) v = (w);
#define PGJ_G_KDBL(w) v = norm_double_key(w);
#define PGJ_G_K32(w) v = (u64)(u32)(w);
#define PGJ_G_K8(w) v = (w) & 0xff;
#define PGJ_G_KVAR(w) { int len = (int)((w) & 0xFFFFFF); if (len > 3) ok = false; else { const u8* b = a.arena + ((w) >> 24); v = (u64)len; \
        for (int t = 0; t < len; t++) v |= (u64)b[t] << (8 + 8 * t); } }
#define PG_X(k, CLASS, WORD, SHIFT, HASNULL, WOFF)                                 \
    if ((rowp[0] >> k) & 1) w1 |= 1ULL << (56 + k);                                \
    else { u64 v = 0; PGJ_G_##CLASS(rowp[WOFF]) if (WORD == 0) w0 |= v << SHIFT; else w1 |= v << SHIFT; }
    PG_KEYS(PG_X)
#undef PG_X
    return ok;
}

extern "C" __global__ void __launch_bounds__(PG_T, PG_MIN_BLOCKS) pg_lowcard_agg(const PgjArgs a) {
    extern __shared__ u64 table[];
    __shared__ LutEntry lut[PGJ_LUT_SIZE];
    const int tid = threadIdx.x;
    const int G = a.G;
    // ---- own column: zeros, then min/max identities
    for (int i = 0; i < G * PG_REC_WORDS; i++) table[i * PG_T + tid] = 0;
#define PG_X(s, KIND, WORD)                                                                                               \
    if (PGJ_ISKIND_##KIND == PGJ_KIND_MIN) for (int g = 0; g < G; g++) table[(g * PG_REC_WORDS + WORD) * PG_T + tid] = 0x7fffffffffffffffULL; \
    if (PGJ_ISKIND_##KIND == PGJ_KIND_MAX) for (int g = 0; g < G; g++) table[(g * PG_REC_WORDS + WORD) * PG_T + tid] = 0x8000000000000000ULL;
    PG_SLOTS(PG_X)
#undef PG_X
    for (int i = tid; i < PGJ_LUT_SIZE; i += PG_T) { lut[i].w0 = ~0ULL; lut[i].w1 = ~0ULL; lut[i].gid = -1; }
    __syncthreads();
    if (tid == 0) {
        for (int g = 0; g < G; g++) {
            u64 w0, w1;
            if (!pack_group(a, g, w0, w1)) continue;
            u32 h = lut_hash(w0, w1);
            while (lut[h].gid >= 0) h = (h + 1) & (PGJ_LUT_SIZE - 1);
            lut[h].w0 = w0; lut[h].w1 = w1; lut[h].gid = g;
        }
    }
    __syncthreads();

    // ---- row loop: thread t of block b takes rows [(b*T + t)*VEC, +VEC), then strides by the whole grid
    u64 misses = 0;
    const i64 n = a.n;
    const i64 stride = (i64)gridDim.x * PG_T * PG_VEC;
    for (i64 row = ((i64)blockIdx.x * PG_T + tid) * PG_VEC; row + PG_VEC <= n; row += stride) {
#if PG_PREFETCH > 0
        if (row + PG_PREFETCH * stride + PG_VEC <= n) prefetch_rows<PG_VEC>(a, row + PG_PREFETCH * stride);
#endif
        process_rows<PG_VEC>(a, row, table, lut, tid, misses);
    }
    if (PG_VEC > 1 && blockIdx.x == 0 && tid == 0)
        for (i64 row = n - n % PG_VEC; row < n; row++) process_rows<1>(a, row, table, lut, tid, misses);

    // ---- flush: per aggregate and group, reduce the warp's 32 columns and fold them into the global state
    const int lane = tid & 31;
    {
        u64 z = 0;
        wreduce_sum_carry(misses, z);
        if (lane == 0 && misses) atomicAdd(a.d_miss, misses);
    }
    for (int ai = 0; ai < a.naggs; ai++) {
        const int op = a.agg_op[ai], kind = a.agg_val_kind[ai];
        for (int g = 0; g < G; g++) {
            const u64* rec = table + (g * PG_REC_WORDS) * PG_T + tid;
            u64 cnt = rec[a.agg_cnt_word[ai] * PG_T], zero = 0;
            wreduce_sum_carry(cnt, zero);
            cnt = __shfl_sync(0xffffffffu, cnt, 0);
            if (cnt == 0) continue;   // warp-uniform
            u64* st = a.agg_state[ai] + (i64)g * a.agg_words[ai];
            if (op == PGJ_OP_COUNT) { if (lane == 0) atomicAdd(&st[0], cnt); continue; }
            const u64* col = rec + a.agg_val_word[ai] * PG_T;
            if (kind == PGJ_KIND_I64 || kind == PGJ_KIND_I32) {
                u64 s0 = col[0], c0 = 0, s1 = col[PG_T], c1 = 0;
                wreduce_sum_carry(s0, c0);
                wreduce_sum_carry(s1, c1);
                if (lane == 0) {
                    u64 hi = s1 + c0;   // 128-bit signed partial: cannot overflow for < 2^63 rows of 64-bit values
                    atomicAdd(&st[0], cnt);
                    if (op == PGJ_OP_SUM_I64) atomic_add_chain2(st + 1, s0, hi);
                    else atomic_add_chain3(st + 1, s0, hi, (u64)((i64)hi >> 63));
                }
            } else if (kind == PGJ_KIND_DEC) {
                u64 s0 = col[0], c0 = 0, s1 = col[PG_T], c1 = 0, s2 = col[2 * PG_T], c2 = 0;
                wreduce_sum_carry(s0, c0);
                wreduce_sum_carry(s1, c1);
                wreduce_sum_carry(s2, c2);
                if (lane == 0) {
                    u64 a1 = s1 + c0;
                    u64 k1 = c1 + (a1 < s1 ? 1ULL : 0ULL);
                    atomicAdd(&st[0], cnt);
                    atomic_add_chain3(st + 1, s0, a1, s2 + k1);
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
