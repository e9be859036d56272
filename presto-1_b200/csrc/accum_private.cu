// accum_private.cu — low-cardinality grouped accumulation (the "shared-memory pre-aggregation" path of north_star).
//
// Every thread owns a private copy of the (few) groups' partial states in shared memory, laid out
// table[(word * G + group) * T + thread] so that a warp's accesses to one word hit 32 different banks: updates are
// plain conflict-free LDS/ADD/STS, no atomics in the row loop.  Identical pieces of state are shared between
// aggregates (sum(x) and avg(x) share one sum, all aggregates under the same mask share one count), and partial sums use
// the narrowest exact width (128 bit for 64-bit inputs, 192 bit for 128-bit inputs).  At the end each warp tree-reduces
// its 32 columns with shuffles (sum + carry count per word) and lane 0 folds the result into the global states with the
// carry-chain atomics of accum_dev.cuh.
#include "accum_dev.cuh"
#include "groupby_dev.cuh"

namespace pg {

// slot kinds; the kernel walks one fixed-trip, switch-free loop per kind (slots are sorted by kind on the host)
enum { PS_DEC = 0, PS_I64 = 1, PS_I32 = 2, PS_F64 = 3, PS_MIN = 4, PS_MAX = 5, PS_CNT = 6, PS_KINDS = 7 };
static const int MAX_DEC = 3, MAX_I64 = 6, MAX_I32 = 2, MAX_F64 = 3, MAX_MIN = 2, MAX_MAX = 2, MAX_CNT = 3;
static const int MAX_PSLOTS = MAX_DEC + MAX_I64 + MAX_I32 + MAX_F64 + MAX_MIN + MAX_MAX + MAX_CNT;
static const int MAX_PCONDS = 4;
static const int LUT_SIZE = 256;     // shared-memory key -> group id table of the fused variant

struct FSlot {
    const void* ptr;      // argument values (pre-offset)
    int32_t word_off;     // first private word
    uint32_t cmask;       // bit of the condition under which a row contributes (0 = unused slot)
    int32_t conv;         // F64: 1 = convert int64 -> double; MIN/MAX: 1 = double order key
    int32_t nwords;
};
struct FCond {
    const uint8_t* mask;        // BOOLEAN mask values or null
    const uint8_t* mask_nulls;  // its null flags or null
    const uint8_t* arg_nulls;   // null flags of the argument or null
};
struct PrivAgg {
    int32_t op, words, cnt_word, val_word, val_kind, pad;
    unsigned long long* state;
};
struct PrivPlan {
    int32_t n[PS_KINDS];
    int32_t nconds, words, naggs;
    FSlot dec[MAX_DEC], i64[MAX_I64], i32[MAX_I32], f64[MAX_F64], mn[MAX_MIN], mx[MAX_MAX], cnt[MAX_CNT];
    FCond conds[MAX_PCONDS];
    PrivAgg aggs[MAX_AGGS];
};

// ---- fused group-id look-up: short keys are packed into two 64-bit words (a key never straddles the words) and searched
// in a shared-memory table built from the groups the GroupByHash already numbered.  Bits 56.. of word 1 are null flags.
struct PackKey {
    const void* values;
    const uint8_t* nulls;
    const int32_t* offsets;
    int32_t type, word, shift, word_off;
};
struct PackSpec {
    int32_t nkeys, bigint_mode, null_gid, row_words;
    PackKey k[MAX_PACK_KEYS];
    const uint64_t* key_rows;         // stored groups (GroupByHash key store)
    const uint8_t* arena;
    const int64_t* values_by_group;   // bigint mode
    unsigned int* miss_flag;          // bitmap: bit set for rows the fused pass could not resolve
    unsigned long long* d_miss;
};

static inline int pack_width(int32_t t) {
    switch (t) {
        case PG_BOOLEAN: return 8;
        case PG_INTEGER: case PG_DATE: case PG_VARCHAR: case PG_VARBINARY: return 32;
        case PG_BIGINT: case PG_DOUBLE: case PG_SHORT_DECIMAL: return 64;
        default: return -1;
    }
}
PG_D unsigned long long norm_double_key(long long b) {
    if ((b & 0x7fffffffffffffffLL) == 0) return 0;
    return (unsigned long long)canon_double_bits(b);
}
// pack the key of input row `row`; false when a value does not fit the packed form (string longer than 3 bytes)
PG_D bool pack_row(const PackSpec& ps, int64_t row, unsigned long long& w0, unsigned long long& w1) {
    w0 = 0; w1 = 0;
    bool ok = true;
#pragma unroll 1
    for (int k = 0; k < ps.nkeys; k++) {
        {
            const PackKey& pk = ps.k[k];
            unsigned long long v = 0;
            bool isnull = pk.nulls != nullptr && pk.nulls[row] != 0;
            if (pk.type == PG_VARCHAR || pk.type == PG_VARBINARY) {
                int32_t o = pk.offsets[row], len = pk.offsets[row + 1] - o;
                if (len > 3) ok = false;
                else {
                    const uint8_t* b = (const uint8_t*)pk.values + o;
                    v = (unsigned long long)len;
                    if (len > 0) v |= (unsigned long long)b[0] << 8;
                    if (len > 1) v |= (unsigned long long)b[1] << 16;
                    if (len > 2) v |= (unsigned long long)b[2] << 24;
                }
            } else if (pk.type == PG_INTEGER || pk.type == PG_DATE) v = (unsigned long long)(uint32_t)((const int32_t*)pk.values)[row];
            else if (pk.type == PG_BOOLEAN) v = ((const uint8_t*)pk.values)[row];
            else if (pk.type == PG_DOUBLE) v = norm_double_key(((const long long*)pk.values)[row]);
            else v = ((const unsigned long long*)pk.values)[row];
            if (isnull) w1 |= 1ULL << (56 + k);
            else if (pk.word == 0) w0 |= v << pk.shift;
            else w1 |= v << pk.shift;
        }
    }
    return ok;
}
// pack the stored key of group g (same bit layout)
PG_D bool pack_group(const PackSpec& ps, int g, unsigned long long& w0, unsigned long long& w1) {
    w0 = 0; w1 = 0;
    if (ps.bigint_mode) {
        if (g == ps.null_gid) w1 |= 1ULL << 56; else w0 = (unsigned long long)ps.values_by_group[g];
        return true;
    }
    const uint64_t* rowp = ps.key_rows + (int64_t)g * ps.row_words;
    for (int k = 0; k < ps.nkeys; k++) {
        const PackKey& pk = ps.k[k];
        if ((rowp[0] >> k) & 1) { w1 |= 1ULL << (56 + k); continue; }
        unsigned long long w = rowp[pk.word_off], v;
        switch (pk.type) {
            case PG_VARCHAR: case PG_VARBINARY: {
                int len = (int)(w & 0xFFFFFF);
                if (len > 3) return false;
                const uint8_t* b = ps.arena + (w >> 24);
                v = (unsigned long long)len;
                for (int t = 0; t < len; t++) v |= (unsigned long long)b[t] << (8 + 8 * t);
                break;
            }
            case PG_INTEGER: case PG_DATE: v = (unsigned long long)(uint32_t)w; break;
            case PG_BOOLEAN: v = w & 0xff; break;
            case PG_DOUBLE: v = norm_double_key((long long)w); break;
            default: v = w;
        }
        if (pk.word == 0) w0 |= v << pk.shift; else w1 |= v << pk.shift;
    }
    return true;
}
PG_D unsigned lut_hash(unsigned long long w0, unsigned long long w1) {
    unsigned x = (unsigned)w0 ^ (unsigned)(w0 >> 32) ^ (unsigned)w1 ^ (unsigned)(w1 >> 29);
    return (x * 0x9E3779B1u) >> 24;
}

PG_D void wreduce_sum_carry(unsigned long long& s, unsigned long long& carry) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        unsigned long long os = __shfl_down_sync(0xffffffffu, s, d);
        unsigned long long oc = __shfl_down_sync(0xffffffffu, carry, d);
        unsigned long long ns = s + os;
        carry += oc + (ns < s ? 1ULL : 0ULL);
        s = ns;
    }
}

// per-row values between the load and the update phase; every index is static after unrolling (registers)
struct RowData {
    int g;               // group id, -1: skip the row
    unsigned condbits;
    ulonglong2 dec[MAX_DEC];
    unsigned long long i64[MAX_I64];
    int i32[MAX_I32];
    unsigned long long f64[MAX_F64];
    unsigned long long mn[MAX_MIN], mx[MAX_MAX];
};

// FUSED = true: the group id of a row is resolved inside this kernel from the packed key (no id column in HBM); rows that
// cannot be resolved are flagged in the miss_flag bitmap and counted in *d_miss — the caller runs the general path for exactly those.
template <int T, bool FUSED>
__global__ void __launch_bounds__(T) acc_private2_kernel(const PrivPlan plan, const int32_t* __restrict__ gids, int64_t n, int G,
                                                         const uint8_t* __restrict__ only_rows, const PackSpec pk) {
    extern __shared__ unsigned long long table[];
    __shared__ unsigned long long lut_w0[FUSED ? LUT_SIZE : 1], lut_w1[FUSED ? LUT_SIZE : 1];
    __shared__ int lut_gid[FUSED ? LUT_SIZE : 1];
    const int tid = threadIdx.x;
    const int GT = G * T;
    // ---- init own columns: zeros, then min/max identities
    for (int w = 0; w < plan.words; w++)
        for (int g = 0; g < G; g++) table[w * GT + g * T + tid] = 0;
    for (int s = 0; s < plan.n[PS_MIN]; s++)
        for (int g = 0; g < G; g++) table[plan.mn[s].word_off * GT + g * T + tid] = 0x7fffffffffffffffULL;
    for (int s = 0; s < plan.n[PS_MAX]; s++)
        for (int g = 0; g < G; g++) table[plan.mx[s].word_off * GT + g * T + tid] = 0x8000000000000000ULL;
    if (FUSED) {
        for (int i = tid; i < LUT_SIZE; i += T) lut_gid[i] = -1;
        __syncthreads();
        if (tid == 0) {
            for (int g = 0; g < G; g++) {
                unsigned long long w0, w1;
                if (!pack_group(pk, g, w0, w1)) continue;
                unsigned h = lut_hash(w0, w1);
                while (lut_gid[h] >= 0) h = (h + 1) & (LUT_SIZE - 1);
                lut_w0[h] = w0; lut_w1[h] = w1; lut_gid[h] = g;
            }
        }
        __syncthreads();
    }
    unsigned long long misses = 0;

    auto load_row = [&](int64_t row, RowData& r) {
        r.condbits = 0;
        if (FUSED) {
            unsigned long long w0, w1;
            r.g = -1;
            if (pack_row(pk, row, w0, w1)) {
                unsigned h = lut_hash(w0, w1);
                while (true) {
                    int g = lut_gid[h];
                    if (g < 0) break;
                    if (lut_w0[h] == w0 && lut_w1[h] == w1) { r.g = g; break; }
                    h = (h + 1) & (LUT_SIZE - 1);
                }
            }
            if (r.g < 0) { misses++; atomicOr(&pk.miss_flag[row >> 5], 1u << (row & 31)); return; }
        } else {
            r.g = gids[row];
            if (only_rows != nullptr && only_rows[row] == 0) { r.g = -1; return; }
        }
#pragma unroll
        for (int c = 0; c < MAX_PCONDS; c++) {
            if (c < plan.nconds) {
                const FCond& pc = plan.conds[c];
                bool ok = true;
                if (pc.mask != nullptr) ok = pc.mask[row] != 0;
                if (pc.mask_nulls != nullptr) ok = ok && pc.mask_nulls[row] == 0;
                if (pc.arg_nulls != nullptr) ok = ok && pc.arg_nulls[row] == 0;
                r.condbits |= (ok ? 1u : 0u) << c;
            }
        }
        const unsigned cb = r.condbits;
        if (plan.n[PS_DEC]) {
#pragma unroll
            for (int s = 0; s < MAX_DEC; s++) if (cb & plan.dec[s].cmask) r.dec[s] = ((const ulonglong2*)plan.dec[s].ptr)[row];
        }
        if (plan.n[PS_I64]) {
#pragma unroll
            for (int s = 0; s < MAX_I64; s++) if (cb & plan.i64[s].cmask) r.i64[s] = ((const unsigned long long*)plan.i64[s].ptr)[row];
        }
        if (plan.n[PS_I32]) {
#pragma unroll
            for (int s = 0; s < MAX_I32; s++) if (cb & plan.i32[s].cmask) r.i32[s] = ((const int*)plan.i32[s].ptr)[row];
        }
        if (plan.n[PS_F64]) {
#pragma unroll
            for (int s = 0; s < MAX_F64; s++) if (cb & plan.f64[s].cmask) r.f64[s] = ((const unsigned long long*)plan.f64[s].ptr)[row];
        }
        if (plan.n[PS_MIN]) {
#pragma unroll
            for (int s = 0; s < MAX_MIN; s++) if (cb & plan.mn[s].cmask) r.mn[s] = ((const unsigned long long*)plan.mn[s].ptr)[row];
        }
        if (plan.n[PS_MAX]) {
#pragma unroll
            for (int s = 0; s < MAX_MAX; s++) if (cb & plan.mx[s].cmask) r.mx[s] = ((const unsigned long long*)plan.mx[s].ptr)[row];
        }
    };

    auto add2 = [&](unsigned long long* t, unsigned long long a0) {
        unsigned long long a1 = (unsigned long long)((long long)a0 >> 63);
        unsigned long long s0 = t[0] + a0;
        t[0] = s0;
        t[GT] += a1 + (s0 < a0 ? 1ULL : 0ULL);
    };

    auto apply_row = [&](const RowData& r) {
        if (r.g < 0) return;
        unsigned long long* base = table + r.g * T + tid;
        const unsigned cb = r.condbits;
        if (plan.n[PS_DEC]) {
#pragma unroll
            for (int s = 0; s < MAX_DEC; s++) {
                if (cb & plan.dec[s].cmask) {
                    unsigned long long* t = base + plan.dec[s].word_off * GT;
                    unsigned long long a0, a1, a2;
                    signmag_to_twos(r.dec[s].x, r.dec[s].y, a0, a1, a2);
                    unsigned long long s0 = t[0] + a0;
                    t[0] = s0;
                    unsigned long long c0 = s0 < a0 ? 1ULL : 0ULL;
                    unsigned long long add1 = a1 + c0;
                    unsigned long long c1 = add1 < c0 ? 1ULL : 0ULL;
                    unsigned long long s1 = t[GT] + add1;
                    t[GT] = s1;
                    c1 += s1 < add1 ? 1ULL : 0ULL;
                    t[2 * GT] += a2 + c1;
                }
            }
        }
        if (plan.n[PS_I64]) {
#pragma unroll
            for (int s = 0; s < MAX_I64; s++) if (cb & plan.i64[s].cmask) add2(base + plan.i64[s].word_off * GT, r.i64[s]);
        }
        if (plan.n[PS_I32]) {
#pragma unroll
            for (int s = 0; s < MAX_I32; s++)
                if (cb & plan.i32[s].cmask) add2(base + plan.i32[s].word_off * GT, (unsigned long long)(long long)r.i32[s]);
        }
        if (plan.n[PS_F64]) {
#pragma unroll
            for (int s = 0; s < MAX_F64; s++) {
                if (cb & plan.f64[s].cmask) {
                    unsigned long long* t = base + plan.f64[s].word_off * GT;
                    double d = plan.f64[s].conv ? (double)(long long)r.f64[s] : __longlong_as_double((long long)r.f64[s]);
                    t[0] = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)t[0]) + d);
                }
            }
        }
        if (plan.n[PS_MIN]) {
#pragma unroll
            for (int s = 0; s < MAX_MIN; s++) {
                if (cb & plan.mn[s].cmask) {
                    unsigned long long* t = base + plan.mn[s].word_off * GT;
                    long long k = plan.mn[s].conv ? double_order_key((long long)r.mn[s]) : (long long)r.mn[s];
                    if (k < (long long)t[0]) t[0] = (unsigned long long)k;
                }
            }
        }
        if (plan.n[PS_MAX]) {
#pragma unroll
            for (int s = 0; s < MAX_MAX; s++) {
                if (cb & plan.mx[s].cmask) {
                    unsigned long long* t = base + plan.mx[s].word_off * GT;
                    long long k = plan.mx[s].conv ? max_double_key((long long)r.mx[s]) : (long long)r.mx[s];
                    if (k > (long long)t[0]) t[0] = (unsigned long long)k;
                }
            }
        }
#pragma unroll
        for (int s = 0; s < MAX_CNT; s++)
            if (cb & plan.cnt[s].cmask) base[plan.cnt[s].word_off * GT] += 1;
    };

    // ---- row loop, software-pipelined: the loads of the next row are issued before the shared-memory updates of this one
    const int64_t stride = (int64_t)gridDim.x * T;
    int64_t row = (int64_t)blockIdx.x * T + tid;
    RowData ra, rb;
    ra.g = -1; rb.g = -1;
    if (row < n) load_row(row, ra);
    while (row < n) {                       // two rows per trip, ping-pong registers: no copies
        int64_t r1 = row + stride, r2 = r1 + stride;
        rb.g = -1;
        if (r1 < n) load_row(r1, rb);
        apply_row(ra);
        ra.g = -1;
        if (r2 < n) load_row(r2, ra);
        apply_row(rb);
        row = r2;
    }
    // ---- flush: per aggregate and group, reduce the warp's 32 columns and fold into the global state
    const int lane = tid & 31;
    if (FUSED) {
        unsigned long long z = 0;
        wreduce_sum_carry(misses, z);
        if (lane == 0 && misses) atomicAdd(pk.d_miss, misses);
    }
    for (int a = 0; a < plan.naggs; a++) {
        const PrivAgg& pa = plan.aggs[a];
        for (int g = 0; g < G; g++) {
            unsigned long long cnt = table[pa.cnt_word * GT + g * T + tid], zero = 0;
            wreduce_sum_carry(cnt, zero);
            cnt = __shfl_sync(0xffffffffu, cnt, 0);
            if (cnt == 0) continue;   // warp-uniform
            unsigned long long* st = pa.state + (int64_t)g * pa.words;
            if (pa.op == OP_COUNT) { if (lane == 0) atomicAdd(&st[0], cnt); continue; }
            const unsigned long long* col = table + pa.val_word * GT + g * T + tid;
            switch (pa.val_kind) {
                case PS_I64: case PS_I32: {
                    unsigned long long s0 = col[0], c0 = 0, s1 = col[GT], c1 = 0;
                    wreduce_sum_carry(s0, c0);
                    wreduce_sum_carry(s1, c1);
                    if (lane == 0) {
                        unsigned long long hi = s1 + c0;   // 128-bit signed partial (cannot overflow: < 2^63 rows of 64-bit values)
                        atomicAdd(&st[0], cnt);
                        if (pa.op == OP_SUM_I64) atomic_add_chain2(st + 1, s0, hi);
                        else atomic_add_chain3(st + 1, s0, hi, (unsigned long long)((long long)hi >> 63));
                    }
                    break;
                }
                case PS_DEC: {
                    unsigned long long s0 = col[0], c0 = 0, s1 = col[GT], c1 = 0, s2 = col[2 * GT], c2 = 0;
                    wreduce_sum_carry(s0, c0);
                    wreduce_sum_carry(s1, c1);
                    wreduce_sum_carry(s2, c2);
                    if (lane == 0) {
                        unsigned long long a1 = s1 + c0;
                        unsigned long long k1 = c1 + (a1 < s1 ? 1ULL : 0ULL);
                        atomicAdd(&st[0], cnt);
                        atomic_add_chain3(st + 1, s0, a1, s2 + k1);
                    }
                    break;
                }
                case PS_F64: {
                    double d = __longlong_as_double((long long)col[0]);
#pragma unroll
                    for (int k = 16; k > 0; k >>= 1) d += __shfl_down_sync(0xffffffffu, d, k);
                    if (lane == 0) { atomicAdd(&st[0], cnt); atomicAdd((double*)&st[1], d); }
                    break;
                }
                case PS_MIN: case PS_MAX: {
                    long long v = (long long)col[0];
#pragma unroll
                    for (int k = 16; k > 0; k >>= 1) {
                        long long o = __shfl_down_sync(0xffffffffu, v, k);
                        v = pa.val_kind == PS_MIN ? (o < v ? o : v) : (o > v ? o : v);
                    }
                    if (lane == 0) {
                        st[0] = 1;
                        if (pa.val_kind == PS_MIN) atomicMin((long long*)&st[1], v); else atomicMax((long long*)&st[1], v);
                    }
                    break;
                }
                default: break;
            }
        }
    }
}

// ---- host: build the plan (dedupe conditions and state, sort slots by kind), pick the block size, launch
// keys that fit the two-word packed form can be resolved inside the accumulation kernel
static bool pack_layout(const KeyLayout& L, int32_t* word, int32_t* shift) {
    if (L.nkeys > MAX_PACK_KEYS) return false;
    int used[2] = {0, 0};
    for (int k = 0; k < L.nkeys; k++) {
        int w = pack_width(L.types[k]);
        if (w < 0) return false;
        int target = (used[0] + w <= 64) ? 0 : 1;
        if (target == 1 && used[1] + w > 56) return false;
        if (word) { word[k] = target; shift[k] = used[target]; }
        used[target] += w;
    }
    return true;
}
bool fused_keys_packable(const KeyLayout& L) { return pack_layout(L, nullptr, nullptr); }
bool fused_pack_layout(const KeyLayout& L, int32_t* word, int32_t* shift) { return pack_layout(L, word, shift); }

bool accumulate_private(pg_ctx* ctx, const AggProgram& prog, const int32_t* gids, int64_t n, int64_t group_count, const uint8_t* only_rows,
                        const FusedLookup* fused) {
    if (group_count < 1 || group_count > 64) return false;
    PrivPlan p;
    memset(&p, 0, sizeof(p));
    p.naggs = prog.naggs;
    FSlot* lists[PS_KINDS] = {p.dec, p.i64, p.i32, p.f64, p.mn, p.mx, p.cnt};
    const int maxn[PS_KINDS] = {MAX_DEC, MAX_I64, MAX_I32, MAX_F64, MAX_MIN, MAX_MAX, MAX_CNT};
    const int kwords[PS_KINDS] = {3, 2, 2, 1, 1, 1, 1};
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        if (op.op == OP_MIN128 || op.op == OP_MAX128) return false;
        FCond fc;
        fc.mask = (const uint8_t*)op.mask.values; fc.mask_nulls = op.mask.values ? op.mask.nulls : nullptr;
        fc.arg_nulls = op.skip_null ? op.arg.nulls : nullptr;
        int cond = -1;
        for (int c = 0; c < p.nconds; c++)
            if (p.conds[c].mask == fc.mask && p.conds[c].mask_nulls == fc.mask_nulls && p.conds[c].arg_nulls == fc.arg_nulls) cond = c;
        if (cond < 0) {
            if (p.nconds >= MAX_PCONDS) return false;
            cond = p.nconds++;
            p.conds[cond] = fc;
        }
        auto find_slot = [&](int kind, const void* ptr, int conv) -> int {   // returns the private word offset or -1
            FSlot* l = lists[kind];
            for (int s = 0; s < p.n[kind]; s++)
                if (l[s].ptr == ptr && l[s].cmask == (1u << cond) && l[s].conv == conv) return l[s].word_off;
            if (p.n[kind] >= maxn[kind]) return -1;
            FSlot& f = l[p.n[kind]++];
            f.ptr = ptr; f.cmask = 1u << cond; f.conv = conv; f.nwords = kwords[kind]; f.word_off = p.words;
            p.words += kwords[kind];
            return f.word_off;
        };
        PrivAgg& pa = p.aggs[a];
        pa.op = op.op; pa.words = op.words; pa.state = op.state;
        pa.cnt_word = find_slot(PS_CNT, nullptr, 0);
        if (pa.cnt_word < 0) return false;
        pa.val_word = -1; pa.val_kind = PS_CNT;
        const bool is32 = op.in_type == PG_INTEGER || op.in_type == PG_DATE;
        switch (op.op) {
            case OP_COUNT: break;
            case OP_SUM_I64: pa.val_kind = is32 ? PS_I32 : PS_I64; break;
            case OP_SUM_F64:
                if (is32) return false;
                pa.val_kind = PS_F64;
                break;
            case OP_SUM_DEC: pa.val_kind = op.in_type == PG_LONG_DECIMAL ? PS_DEC : PS_I64; break;
            case OP_MIN64: case OP_MAX64:
                if (is32 || op.in_type == PG_BOOLEAN) return false;
                pa.val_kind = op.op == OP_MIN64 ? PS_MIN : PS_MAX;
                break;
            default: return false;
        }
        if (op.op != OP_COUNT) {
            int conv = 0;
            if (pa.val_kind == PS_F64) conv = op.convert_i2d ? 1 : 0;
            if (pa.val_kind == PS_MIN || pa.val_kind == PS_MAX) conv = op.in_type == PG_DOUBLE ? 1 : 0;
            pa.val_word = find_slot(pa.val_kind, op.arg.values, conv);
            if (pa.val_word < 0) return false;
        }
    }
    PackSpec pk;
    memset(&pk, 0, sizeof(pk));
    if (fused) {
        const KeyLayout& L = *fused->layout;
        int32_t word[MAX_PACK_KEYS], shift[MAX_PACK_KEYS];
        if (!pack_layout(L, word, shift)) return false;
        pk.nkeys = L.nkeys;
        for (int k = 0; k < L.nkeys; k++) {
            const DCol& c = fused->keys->c[k];
            pk.k[k].values = c.values; pk.k[k].nulls = c.nulls; pk.k[k].offsets = c.offsets; pk.k[k].type = c.type;
            pk.k[k].word = word[k]; pk.k[k].shift = shift[k]; pk.k[k].word_off = L.word_off[k];
        }
        pk.key_rows = fused->key_rows; pk.arena = fused->arena; pk.values_by_group = fused->values_by_group;
        pk.bigint_mode = fused->bigint_mode; pk.null_gid = fused->null_gid; pk.row_words = L.row_words;
        pk.miss_flag = fused->miss_flag; pk.d_miss = fused->d_miss;
    }
    const int G = (int)group_count;
    const int64_t per_thread = (int64_t)p.words * G * 8;
    const int64_t budget = 216 * 1024;
    int T = 256;
    if (per_thread * 256 > budget / 2) T = 128;           // keep at least two blocks (or one 256-thread block) per SM
    if (per_thread * T > budget) return false;
    int64_t smem = per_thread * T;
    int blocks_per_sm = (int)(budget / smem);
    if (blocks_per_sm > 2048 / T) blocks_per_sm = 2048 / T;
    static const int min_rows = getenv("PRESTOGPU_PRIV_ROWS") ? atoi(getenv("PRESTOGPU_PRIV_ROWS")) : 1;
    int64_t want = (n + (int64_t)T * min_rows - 1) / ((int64_t)T * min_rows);
    int64_t grid = (int64_t)ctx->sm_count * blocks_per_sm;
    if (grid > want) grid = want;
#define PG_PRIV_LAUNCH(TT, FF)                                                                                               \
    do {                                                                                                                     \
        PG_CUDA(cudaFuncSetAttribute(acc_private2_kernel<TT, FF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget)); \
        PG_LAUNCH(ctx, (acc_private2_kernel<TT, FF>), (int)grid, TT, (size_t)smem, p, gids, n, G, only_rows, pk);             \
    } while (0)
    if (T == 256) { if (fused) PG_PRIV_LAUNCH(256, true); else PG_PRIV_LAUNCH(256, false); }
    else { if (fused) PG_PRIV_LAUNCH(128, true); else PG_PRIV_LAUNCH(128, false); }
#undef PG_PRIV_LAUNCH
    return true;
}

}  // namespace pg
