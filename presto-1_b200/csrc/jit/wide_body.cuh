// wide_body.cuh — body of the run-time specialised fused GROUP BY kernel for MANY groups on one BIGINT key
// ("hash-then-update", BigintGroupByHash.putIfAbsent :213-260 + the generated GroupedAccumulator.addInput loops in ONE pass).
//
// The hash table holds the partial aggregation state inside the bucket of its key, so a row costs one random access:
//   bucket = PGW_BUCKET_WORDS u64 words (a multiple of 4 = whole 32-byte sectors)
//     word 0  the key (0 = free; key 0 and the null key live in two extra buckets behind the table)
//     word 1  ~(operator-wide index of the first row that carried the key), maintained with atomicMax: first-seen
//             order is recovered when the output is built (hashagg_wide.cu), no group ids exist while rows stream in
//     word 2.. state words of the aggregates (counts, 32-bit-limb integer sums, double sums), zero = identity,
//             so a zero-filled table is an empty table
// Claims are atomicCAS on word 0; every state update is an atomic on the same line (resolved in L2).
// The prelude defines PGW_RANGE (0/1: range-chunked front region, see wide_args.h), PGW_BUCKET_WORDS, PGW_HASKEYNULL, PGW_HASONLY, PGW_QUOTA, PGW_FLAGWORD, PGW_VEC (rows per thread and trip: 1 or 4) and
//   PGW_AGGS(X)  X(a, OP, IN, HASMASK, HASMASKNULLS, HASARGNULLS, CNTMODE, CNTWORD, VALWORD, FLAGBIT)
//     OP in {CNT, SUMI64, SUMF64, SUMDEC}; IN in {NONE, I64, I32, F64, I2F, DEC128}; CNTMODE in {C64, FLAG}

#ifndef PGW_FLAGS_UNCONDITIONAL
#define PGW_FLAGS_UNCONDITIONAL 0
#endif
#ifndef PGW_BIDX
#define PGW_BIDX 0
#endif
#ifndef PGW_LDCA
#define PGW_LDCA 1
#endif
#ifndef PGW_ROWIDS
#define PGW_ROWIDS 0
#endif
#ifndef PGW_DYNAMIC
#define PGW_DYNAMIC 0
#endif
#ifndef PGW_PEEK
#define PGW_PEEK 0      // read (key, ~first row) of the bucket before any atomic: a hit needs no compare-and-swap, and no atomicMax when
#endif                  // the stored first row is already the smaller one (both words only ever move one way, so a stale read is safe)
typedef unsigned long long u64;
typedef long long i64;
typedef unsigned int u32;
typedef unsigned char u8;
#define PGW_CHUNK 32   // buckets per chunk of the front region (hashagg_wide.cu: WIDE_CHUNK)

#define PGW_CAT2(a, b) a##b
#define PGW_CAT(a, b) PGW_CAT2(a, b)

__device__ __forceinline__ u64 fmix64(u64 k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return k;
}

// ---- integer sums are kept as sums of 32-bit limbs, each limb in its own 64-bit word: every update is a fire-and-forget
// red.add (no carry to propagate, so no atomic has to return a value and no thread waits for L2); the top limb is signed.
// A limb sum is exact for up to 2^32 rows per group; the limbs are recombined when the output is built (hashagg_wide.cu).
__device__ __forceinline__ void red_add(u64* p, u64 v) { atomicAdd(p, v); }   // result unused -> RED
__device__ __forceinline__ ulonglong2 ld_ca2(const u64* p) {
    ulonglong2 v; asm volatile("ld.global.ca.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p)); return v;
}
__device__ __forceinline__ ulonglong2 ld_cg2(const u64* p) {
    ulonglong2 v; asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p)); return v;
}
// ---- vector loads of V consecutive rows (read-only path)
template <int V> __device__ __forceinline__ void ld_u64(const void* base, i64 row, u64 (&o)[V]) {
    const u64* p = (const u64*)base + row;
    if constexpr (V == 1) o[0] = __ldg(p);
    else {
#pragma unroll
        for (int i = 0; i < V; i += 2) { const ulonglong2 t = __ldg((const ulonglong2*)(p + i)); o[i] = t.x; o[i + 1] = t.y; }
    }
}
template <int V> __device__ __forceinline__ void ld_i32(const void* base, i64 row, u64 (&o)[V]) {
    const int* p = (const int*)base + row;
    if constexpr (V == 1) o[0] = (u64)(i64)__ldg(p);
    else { const int4 t = __ldg((const int4*)p); o[0] = (u64)(i64)t.x; o[1] = (u64)(i64)t.y; o[2] = (u64)(i64)t.z; o[3] = (u64)(i64)t.w; }
}
template <int V> __device__ __forceinline__ void ld_u8(const u8* base, i64 row, u32 (&o)[V]) {
    const u8* p = base + row;
    if constexpr (V == 1) o[0] = __ldg(p);
    else { const u32 t = __ldg((const u32*)p); o[0] = t & 0xffu; o[1] = (t >> 8) & 0xffu; o[2] = (t >> 16) & 0xffu; o[3] = t >> 24; }
}
template <int V> __device__ __forceinline__ void ld_u128(const void* base, i64 row, u64 (&lo)[V], u64 (&hi)[V]) {
    const ulonglong2* p = (const ulonglong2*)base + row;
#pragma unroll
    for (int i = 0; i < V; i++) { const ulonglong2 t = __ldg(p + i); lo[i] = t.x; hi[i] = t.y; }
}

// ---- per aggregate and input kind.  A thread owns V consecutive rows; rows with the same key (TPC-H lineitem arrives
// clustered by order key) are merged in registers first, so one compare-and-swap and one set of atomics serve the run.
//   LD    load the V arguments                  ACC  per-row partials: 128-bit integer (lo, hi) or double, 0 when the row does not count
//   MERGE fold row j - 1 into row j             EMIT atomics of a run's last row             SOLO not merged: each row's own atomics
#define PGW_LD_NONE(a)
#define PGW_LD_I64(a) u64 PGW_CAT(av, a)[V]; ld_u64<V>(args.arg[a], r0, PGW_CAT(av, a));
#define PGW_LD_I32(a) u64 PGW_CAT(av, a)[V]; ld_i32<V>(args.arg[a], r0, PGW_CAT(av, a));
#define PGW_LD_F64(a) PGW_LD_I64(a)
#define PGW_LD_I2F(a) PGW_LD_I64(a)
#define PGW_LD_DEC128(a) u64 PGW_CAT(av, a)[V]; u64 PGW_CAT(aw, a)[V]; ld_u128<V>(args.arg[a], r0, PGW_CAT(av, a), PGW_CAT(aw, a));

#define PGW_ACC_INT(a) u64 PGW_CAT(slo, a)[V]; i64 PGW_CAT(shi, a)[V]; _Pragma("unroll") for (int j = 0; j < V; j++) {        \
        PGW_CAT(slo, a)[j] = PGW_CAT(ok, a)[j] ? PGW_CAT(av, a)[j] : 0; PGW_CAT(shi, a)[j] = PGW_CAT(ok, a)[j] ? (i64)PGW_CAT(av, a)[j] >> 63 : 0; }
#define PGW_ACC_CNT_NONE(a)
#define PGW_ACC_SUMI64_I64(a) PGW_ACC_INT(a)
#define PGW_ACC_SUMI64_I32(a) PGW_ACC_INT(a)
#define PGW_ACC_SUMDEC_I64(a) PGW_ACC_INT(a)
#define PGW_ACC_SUMF64_F64(a) double PGW_CAT(sd, a)[V]; _Pragma("unroll") for (int j = 0; j < V; j++) PGW_CAT(sd, a)[j] = PGW_CAT(ok, a)[j] ? __longlong_as_double((i64)PGW_CAT(av, a)[j]) : 0.0;
#define PGW_ACC_SUMF64_I2F(a) double PGW_CAT(sd, a)[V]; _Pragma("unroll") for (int j = 0; j < V; j++) PGW_CAT(sd, a)[j] = PGW_CAT(ok, a)[j] ? (double)(i64)PGW_CAT(av, a)[j] : 0.0;
#define PGW_ACC_SUMDEC_DEC128(a)

#define PGW_MERGE_INT(a) { const u64 t_ = PGW_CAT(slo, a)[j] + PGW_CAT(slo, a)[j - 1];                                       \
        PGW_CAT(shi, a)[j] += PGW_CAT(shi, a)[j - 1] + (t_ < PGW_CAT(slo, a)[j - 1] ? 1 : 0); PGW_CAT(slo, a)[j] = t_; }
#define PGW_MERGE_CNT_NONE(a)
#define PGW_MERGE_SUMI64_I64(a) PGW_MERGE_INT(a)
#define PGW_MERGE_SUMI64_I32(a) PGW_MERGE_INT(a)
#define PGW_MERGE_SUMDEC_I64(a) PGW_MERGE_INT(a)
#define PGW_MERGE_SUMF64_F64(a) PGW_CAT(sd, a)[j] += PGW_CAT(sd, a)[j - 1];
#define PGW_MERGE_SUMF64_I2F(a) PGW_MERGE_SUMF64_F64(a)
#define PGW_MERGE_SUMDEC_DEC128(a)

// limbs of a 128-bit two's complement partial sum (lo, hi): its low 32 bits, and the rest as one signed quantity
#define PGW_EMIT_LIMBS(a, W) { const u64 l0_ = PGW_CAT(slo, a)[j] & 0xffffffffULL;                                       \
        const u64 l1_ = (PGW_CAT(slo, a)[j] >> 32) | ((u64)PGW_CAT(shi, a)[j] << 32);                                    \
        if (l0_) red_add(b + (W), l0_); if (l1_) red_add(b + (W) + 1, l1_); }
#define PGW_EMIT_CNT_NONE(a, W)
#define PGW_EMIT_SUMI64_I64(a, W) PGW_EMIT_LIMBS(a, W)
#define PGW_EMIT_SUMI64_I32(a, W) PGW_EMIT_LIMBS(a, W)
#define PGW_EMIT_SUMDEC_I64(a, W) PGW_EMIT_LIMBS(a, W)
#define PGW_EMIT_SUMF64_F64(a, W) atomicAdd((double*)(b + (W)), PGW_CAT(sd, a)[j]);
#define PGW_EMIT_SUMF64_I2F(a, W) PGW_EMIT_SUMF64_F64(a, W)
#define PGW_EMIT_SUMDEC_DEC128(a, W)

#define PGW_SOLO_CNT_NONE(a, W)
#define PGW_SOLO_SUMI64_I64(a, W)
#define PGW_SOLO_SUMI64_I32(a, W)
#define PGW_SOLO_SUMDEC_I64(a, W)
#define PGW_SOLO_SUMF64_F64(a, W)
#define PGW_SOLO_SUMF64_I2F(a, W)
#define PGW_SOLO_SUMDEC_DEC128(a, W) if (PGW_CAT(ok, a)[j]) {                                                \
        u64 a0 = PGW_CAT(av, a)[j], a1 = PGW_CAT(aw, a)[j] & 0x7fffffffffffffffULL;                          \
        if ((i64)PGW_CAT(aw, a)[j] < 0) { a0 = ~a0; a1 = ~a1; a0 += 1; if (a0 == 0) a1 += 1; }               \
        const u64 l0_ = a0 & 0xffffffffULL, l1_ = a0 >> 32, l2_ = a1 & 0xffffffffULL, l3_ = (u64)((i64)a1 >> 32); \
        if (l0_) red_add(b + (W), l0_); if (l1_) red_add(b + (W) + 1, l1_);                                  \
        if (l2_) red_add(b + (W) + 2, l2_); if (l3_) red_add(b + (W) + 3, l3_); }

#define PGW_CNT_C64(a, W, BIT) red_add(b + (W), (u64)PGW_CAT(cn, a)[j]);
#define PGW_CNT_FLAG(a, W, BIT) flag_bits |= 1ULL << (BIT);

// Claim quota.  PGW_QUOTA == 0: the host made sure every row of the launch could claim a bucket of `table`, nothing to check.
// PGW_QUOTA == 1 (the table could fill up): every warp adds its claims to counters[0] after each trip and reads the total before the
// next one; once the total has passed args.room it only looks keys up, and rows whose key is not there yet are left out for the
// re-run after the host has grown the table.  The overshoot is bounded by what the warps in flight claim in one trip (the host
// subtracts that from the room it grants).
// V consecutive rows of one thread
template <int V>
__device__ __forceinline__ void process_rows(const PgwArgs& args, const i64 r0, u64& left_out, u64& claims, u64& fclaims) {
    const i64 cap = args.cap;
    // ---- loads
    u64 key[V];
    u32 kn[V], only[V];
    (void)kn; (void)only;
#if PGW_BIDX
    // split mode: the bucket number of every row replaces its key (equal keys have equal buckets)
    u32 bi[V];
    if constexpr (V == 1) bi[0] = __ldg(args.bidx + r0);
    else { const uint4 t = __ldg((const uint4*)(args.bidx + r0)); bi[0] = t.x; bi[1] = t.y; bi[2] = t.z; bi[3] = t.w; }
#pragma unroll
    for (int j = 0; j < V; j++) key[j] = bi[j];
#else
    ld_u64<V>(args.keys, r0, key);
    if (PGW_HASKEYNULL) ld_u8<V>(args.key_nulls, r0, kn);
    if (PGW_HASONLY) ld_u8<V>(args.only_rows, r0, only);
#endif
#define PG_X(a, OP, IN, HASMASK, HASMASKNULLS, HASARGNULLS, CNTMODE, CNTWORD, VALWORD, FLAGBIT)                               \
    PGW_LD_##IN(a)                                                                                                            \
    u32 PGW_CAT(m0, a)[V], PGW_CAT(m1, a)[V], PGW_CAT(m2, a)[V];                                                               \
    (void)PGW_CAT(m0, a); (void)PGW_CAT(m1, a); (void)PGW_CAT(m2, a);                                                          \
    if (HASMASK) ld_u8<V>(args.mask[a], r0, PGW_CAT(m0, a));                                                                  \
    if (HASMASKNULLS) ld_u8<V>(args.mask_nulls[a], r0, PGW_CAT(m1, a));                                                       \
    if (HASARGNULLS) ld_u8<V>(args.arg_nulls[a], r0, PGW_CAT(m2, a));
    PGW_AGGS(PG_X)
#undef PG_X
    bool valid[V], isnull[V];
#pragma unroll
#if PGW_BIDX
    for (int j = 0; j < V; j++) { valid[j] = bi[j] < 0xFFFFFFFEu; isnull[j] = false; }
#else
    for (int j = 0; j < V; j++) { valid[j] = !PGW_HASONLY || only[j] != 0; isnull[j] = PGW_HASKEYNULL && kn[j] != 0; }
#endif
#define PG_X(a, OP, IN, HASMASK, HASMASKNULLS, HASARGNULLS, CNTMODE, CNTWORD, VALWORD, FLAGBIT)                               \
    bool PGW_CAT(ok, a)[V]; u32 PGW_CAT(cn, a)[V];                                                                            \
    _Pragma("unroll") for (int j = 0; j < V; j++) {                                                                           \
        PGW_CAT(ok, a)[j] = valid[j] && (!HASMASK || PGW_CAT(m0, a)[j] != 0) && (!HASMASKNULLS || PGW_CAT(m1, a)[j] == 0) &&  \
                            (!HASARGNULLS || PGW_CAT(m2, a)[j] == 0);                                                         \
        PGW_CAT(cn, a)[j] = PGW_CAT(ok, a)[j] ? 1u : 0u; }                                                                    \
    PGW_ACC_##OP##_##IN(a)
    PGW_AGGS(PG_X)
#undef PG_X
    // ---- merge runs of equal keys: row j absorbs row j - 1, the last row of a run carries the totals and the run's first row
    i64 first_row[V];
    bool live[V];
#pragma unroll
    for (int j = 0; j < V; j++) { first_row[j] = r0 + j; live[j] = valid[j]; }
#if PGW_ROWIDS
    {
        u32 rid[V];
        if constexpr (V == 1) rid[0] = __ldg(args.row_ids + r0);
        else { const uint4 t = __ldg((const uint4*)(args.row_ids + r0)); rid[0] = t.x; rid[1] = t.y; rid[2] = t.z; rid[3] = t.w; }
#pragma unroll
        for (int j = 0; j < V; j++) first_row[j] = (i64)rid[j];
    }
#endif
#pragma unroll
    for (int j = 1; j < V; j++) {
        if (valid[j] && valid[j - 1] && isnull[j] == isnull[j - 1] && (isnull[j] || key[j] == key[j - 1])) {
            live[j - 1] = false;
            first_row[j] = PGW_ROWIDS ? (first_row[j - 1] < first_row[j] ? first_row[j - 1] : first_row[j]) : first_row[j - 1];
#define PG_X(a, OP, IN, HASMASK, HASMASKNULLS, HASARGNULLS, CNTMODE, CNTWORD, VALWORD, FLAGBIT) \
            PGW_CAT(cn, a)[j] += PGW_CAT(cn, a)[j - 1]; PGW_MERGE_##OP##_##IN(a)
            PGW_AGGS(PG_X)
#undef PG_X
        }
    }
#if PGW_BIDX
    u64* bp[V];
#pragma unroll
    for (int j = 0; j < V; j++) {
        bp[j] = nullptr;
        if (live[j]) bp[j] = bi[j] < args.front_buckets ? args.front + (i64)bi[j] * PGW_BUCKET_WORDS
                                                        : args.table + ((i64)bi[j] - (i64)args.front_buckets) * PGW_BUCKET_WORDS;
    }
#else
    // ---- locate or claim the bucket of every run.  All loops below run a warp-uniform number of rounds (__any_sync) with predicated
    // bodies: lanes that leave a data-dependent probe loop at different iterations are NOT brought back together by the hardware, and
    // the rest of the kernel would then run a few lanes at a time (measured: 3.9 of 32 lanes active per instruction, 5.9 ms against 1.x).
    const unsigned wm = __activemask();
    u64* bp[V];
    u32 idx[V];
    u32 st[V];                       // 0 settled (bucket found, or left out), 1 walking its front chunk, 2 at a free front bucket, 3 probing `table`
    u64 old[V];
    bool first_known[V];             // PGW_PEEK: the bucket already records a first row at or before this run's
#pragma unroll
    for (int j = 0; j < V; j++) {
        bp[j] = nullptr; idx[j] = 0; old[j] = 1; st[j] = 0; first_known[j] = false;
        if (!live[j]) continue;
        if (isnull[j] || key[j] == 0) bp[j] = args.table + (cap + (isnull[j] ? 0 : 1)) * PGW_BUCKET_WORDS;
        else st[j] = 3;
    }
#if PGW_RANGE
    // ---- front region.  A chunk's packed keys are probed a 32-byte sector (4 keys) at a time: the sector a key hashes to is read from
    // L2 (never L1: claims by other SMs must be seen), a match or the first free slot is found in registers, and the claims of all
    // lanes and runs are issued together, so a warp pays about two round trips to L2 per trip however its keys collide.
    {
        u32 fleft[V];                // sectors left to visit inside the chunk
        u32 usecg = 0;               // runs whose sector must be re-read from L2 (their L1 copy is known to be stale)
#pragma unroll
        for (int j = 0; j < V; j++) {
            fleft[j] = PGW_CHUNK / 4;
            if (st[j] != 3) continue;
            const u64 seg = (key[j] - (u64)args.k0) >> args.shift;      // keys below k0 wrap around to a huge segment number
            if (seg < args.nseg) { st[j] = 1; idx[j] = (u32)(seg * PGW_CHUNK + ((fmix64(key[j]) >> 40) & (PGW_CHUNK - 4))); }
        }
        while (true) {
            while (true) {
                ulonglong2 ka[V], kb[V];
#pragma unroll
                for (int j = 0; j < V; j++) {
                    if (st[j] != 1) continue;
                    const u64* sp = args.front_keys + (idx[j] & ~3u);
                    if (PGW_LDCA && !((usecg >> j) & 1u)) { ka[j] = ld_ca2(sp); kb[j] = ld_ca2(sp + 2); }
                    else { ka[j] = ld_cg2(sp); kb[j] = ld_cg2(sp + 2); }
                }
                bool walking = false;
#pragma unroll
                for (int j = 0; j < V; j++) {
                    if (st[j] != 1) continue;
                    const u64 k0 = ka[j].x, k1 = ka[j].y, k2 = kb[j].x, k3 = kb[j].y, kk = key[j];
                    const int hit = k0 == kk ? 0 : k1 == kk ? 1 : k2 == kk ? 2 : k3 == kk ? 3 : -1;
                    const int zero = k0 == 0 ? 0 : k1 == 0 ? 1 : k2 == 0 ? 2 : k3 == 0 ? 3 : -1;
                    if (hit >= 0) { idx[j] = (idx[j] & ~3u) | (u32)hit; bp[j] = args.front + (i64)idx[j] * PGW_BUCKET_WORDS; st[j] = 0; }
                    else if (zero >= 0) { idx[j] = (idx[j] & ~3u) | (u32)zero; st[j] = 2; }
                    else {
                        idx[j] = (idx[j] & ~(u32)(PGW_CHUNK - 1)) | ((idx[j] + 4) & (PGW_CHUNK - 4));
                        if (--fleft[j] == 0) st[j] = 3; else walking = true;        // a full chunk: the key lives in `table`
                    }
                }
                if (!__any_sync(wm, walking)) break;
            }
#pragma unroll
            for (int j = 0; j < V; j++) if (st[j] == 2) old[j] = atomicCAS(args.front_keys + idx[j], 0ULL, key[j]);
            bool again = false;
#pragma unroll
            for (int j = 0; j < V; j++) {
                if (st[j] != 2) continue;
                if (old[j] == 0 || old[j] == key[j]) {
                    bp[j] = args.front + (i64)idx[j] * PGW_BUCKET_WORDS; st[j] = 0;
                    if (old[j] == 0) { fclaims++; bp[j][0] = key[j]; }
                } else { st[j] = 1; again = true; usecg |= 1u << j; }     // taken by another key in the meantime: read the sector again, from L2
            }
            if (!__any_sync(wm, again)) break;
        }
    }
#endif
    // ---- the hashed table (every key without a front region; else keys outside its range or whose chunk is full): one
    // compare-and-swap claims a free bucket or recognises the key's bucket in ONE round trip to L2, all of a thread's runs in flight together
    {
        bool allowed = true;
        u32 mine = 0;
#if PGW_QUOTA
        allowed = __ldcg(args.counters) < args.room;
#endif
#pragma unroll
        for (int j = 0; j < V; j++) if (st[j] == 3) idx[j] = (u32)(((fmix64(key[j]) >> 32) * (u64)cap) >> 32);
        while (true) {
#pragma unroll
            for (int j = 0; j < V; j++) {
                if (st[j] != 3) continue;
                u64* bk = args.table + (i64)idx[j] * PGW_BUCKET_WORDS;
#if PGW_PEEK
                const ulonglong2 seen = ld_cg2(bk);
                old[j] = seen.x;
                if (seen.x == key[j]) { if (seen.y >= ~(u64)(args.row_base + first_row[j])) first_known[j] = true; }
                else if (seen.x == 0 && allowed) old[j] = atomicCAS(bk, 0ULL, key[j]);
#else
                old[j] = allowed ? atomicCAS(bk, 0ULL, key[j]) : __ldcg(bk);
#endif
            }
            bool more = false;
#pragma unroll
            for (int j = 0; j < V; j++) {
                if (st[j] != 3) continue;
                if (allowed && old[j] == 0) mine++;
                if (old[j] == key[j] || (allowed && old[j] == 0)) { bp[j] = args.table + (i64)idx[j] * PGW_BUCKET_WORDS; st[j] = 0; }
                else if (old[j] == 0) st[j] = 0;                         // a free bucket but no room left: the run is left out
                else { idx[j] = idx[j] + 1 == (u32)cap ? 0u : idx[j] + 1; more = true; }
            }
            if (!__any_sync(wm, more)) break;
        }
#if PGW_QUOTA
        const u32 all = __reduce_add_sync(wm, mine);
        if (all && (threadIdx.x & 31) == __ffs(wm) - 1) atomicAdd(args.counters, (u64)all);
#else
        claims += mine;
#endif
    }
#endif
    // rows absorbed by a later row share its bucket
#pragma unroll
    for (int j = V - 2; j >= 0; j--) if (valid[j] && !live[j]) bp[j] = bp[j + 1];
    // ---- atomics
#pragma unroll
    for (int j = 0; j < V; j++) {
        u64* b = bp[j];
#if !PGW_BIDX
        if (valid[j] && b == nullptr) { left_out++; args.overflow_flag[r0 + j] = 1; }
#endif
        if (b == nullptr) continue;
        u64 flag_bits = 0;
#if PGW_PEEK && !PGW_BIDX
        const bool skip_first = first_known[j];
#else
        const bool skip_first = false;
#endif
        if (live[j] && !skip_first) atomicMax(b + 1, ~(u64)(args.row_base + first_row[j]));   // smallest row index wins (stored inverted, zero = never seen)
#define PG_X(a, OP, IN, HASMASK, HASMASKNULLS, HASARGNULLS, CNTMODE, CNTWORD, VALWORD, FLAGBIT)                  \
        if (live[j] && PGW_CAT(cn, a)[j]) { PGW_CNT_##CNTMODE(a, CNTWORD, FLAGBIT) PGW_EMIT_##OP##_##IN(a, VALWORD) }    \
        if (valid[j]) { PGW_SOLO_##OP##_##IN(a, VALWORD) }
        PGW_AGGS(PG_X)
#undef PG_X
        // (when no flagged aggregate has a mask or nullable argument every run sets the same bits: a plain store of that constant does)
        if (PGW_FLAGWORD >= 0 && flag_bits) {
            if (PGW_FLAGS_UNCONDITIONAL) b[PGW_FLAGWORD >= 0 ? PGW_FLAGWORD : 0] = flag_bits;
            else atomicOr(b + (PGW_FLAGWORD >= 0 ? PGW_FLAGWORD : 0), flag_bits);
        }
    }
}

// L2 prefetch of the lines a thread's NEXT trip will read (one lane per 128-byte line issues it)
__device__ __forceinline__ void pf_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }
#define PGW_PF_NONE(a)
#define PGW_PF_I64(a) if ((threadIdx.x & 3) == 0) pf_l2((const u64*)args.arg[a] + rn);
#define PGW_PF_F64(a) PGW_PF_I64(a)
#define PGW_PF_I2F(a) PGW_PF_I64(a)
#define PGW_PF_I32(a) if ((threadIdx.x & 7) == 0) pf_l2((const int*)args.arg[a] + rn);
#define PGW_PF_DEC128(a) if ((threadIdx.x & 1) == 0) pf_l2((const ulonglong2*)args.arg[a] + rn);
#ifndef PGW_MINB
#define PGW_MINB 2
#endif
#ifndef PGW_PREFETCH
#define PGW_PREFETCH 1
#endif

extern "C" __global__ void __launch_bounds__(256, PGW_MINB) pg_wide_agg(const PgwArgs args) {
    u64 left_out = 0, claims = 0, fclaims = 0;
    const i64 n = args.n;
#if PGW_DYNAMIC
    // tiles of 4 trips are handed out in page order through counters[3] (zeroed by the host before the launch)
    __shared__ i64 s_tile;
    const i64 tile_rows = (i64)blockDim.x * PGW_VEC * 4;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_tile = (i64)atomicAdd(args.counters + 3, 1ULL);
        __syncthreads();
        const i64 t0 = s_tile * tile_rows;
        if (t0 >= n) break;
#pragma unroll 1
        for (int k = 0; k < 4; k++) {
            const i64 r0 = t0 + ((i64)k * blockDim.x + threadIdx.x) * PGW_VEC;
            if (r0 + PGW_VEC <= n) process_rows<PGW_VEC>(args, r0, left_out, claims, fclaims);
        }
    }
#else
    const i64 stride = (i64)gridDim.x * blockDim.x * PGW_VEC;
    for (i64 r0 = ((i64)blockIdx.x * blockDim.x + threadIdx.x) * PGW_VEC; r0 + PGW_VEC <= n; r0 += stride) {
        if (PGW_PREFETCH && PGW_VEC == 4 && r0 + stride + PGW_VEC <= n) {
            const i64 rn = r0 + stride;
            if (!PGW_BIDX && (threadIdx.x & 3) == 0) pf_l2(args.keys + rn);
            if (PGW_BIDX && (threadIdx.x & 7) == 0) pf_l2(args.bidx + rn);
#define PG_X(a, OP, IN, HASMASK, HASMASKNULLS, HASARGNULLS, CNTMODE, CNTWORD, VALWORD, FLAGBIT) PGW_PF_##IN(a)
            PGW_AGGS(PG_X)
#undef PG_X
        }
        process_rows<PGW_VEC>(args, r0, left_out, claims, fclaims);
    }
#endif
    if (PGW_VEC > 1 && blockIdx.x == 0 && threadIdx.x == 0)
        for (i64 r0 = n - n % PGW_VEC; r0 < n; r0++) process_rows<1>(args, r0, left_out, claims, fclaims);
    // ---- per-block totals
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        left_out += __shfl_down_sync(0xffffffffu, left_out, d); claims += __shfl_down_sync(0xffffffffu, claims, d);
        fclaims += __shfl_down_sync(0xffffffffu, fclaims, d);
    }
    if ((threadIdx.x & 31) == 0 && fclaims) atomicAdd(args.counters + 2, fclaims);
    if ((threadIdx.x & 31) == 0 && left_out) atomicAdd(args.counters + 1, left_out);
    if ((threadIdx.x & 31) == 0 && claims) atomicAdd(args.counters, claims);
}
