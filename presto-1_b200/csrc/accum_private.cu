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

enum { PS_CNT = 0, PS_ADD2 = 1, PS_ADD3 = 2, PS_F64 = 3, PS_MIN = 4, PS_MAX = 5 };
enum { CV_INT = 0, CV_I2D = 1, CV_F64 = 2, CV_DEC128 = 3, CV_MINKEY = 4, CV_MAXKEY = 5 };
static const int MAX_PCOLS = 20;
static const int MAX_PCONDS = 8;
static const int MAX_PSLOTS = 10;
static const int MAX_P128 = 3;       // slots fed by 128-bit inputs come first (they carry a second value word per row)
static const int LUT_SIZE = 256;     // shared-memory key -> group id table of the fused variant

struct PrivSlot {
    int32_t kind, conv, word_off, nwords, src, cond, voff, pad;
};
struct PrivCond {
    int32_t mask_col, null_col;   // indices into cols[] or -1
};
struct PrivAgg {
    int32_t op, words, cnt_slot, val_slot;
    unsigned long long* state;
};
struct PrivPlan {
    int32_t ncols, nconds, nslots, words, naggs, pad;
    DCol cols[MAX_PCOLS];
    PrivCond conds[MAX_PCONDS];
    PrivSlot slots[MAX_PSLOTS];
    PrivAgg aggs[MAX_AGGS];
};

// ---- fused group-id look-up: short keys are packed into 128 bits and searched in a shared-memory table built from the
// groups the GroupByHash already numbered.  Bits 120.. of the high word carry one null flag per channel.
static const int MAX_PACK_KEYS = 8;
struct PackSpec {
    int32_t nkeys;
    int32_t shift[MAX_PACK_KEYS];
    DCol cols[MAX_PACK_KEYS];
    // stored groups (GroupByHash key store)
    const uint64_t* key_rows;
    const uint8_t* arena;
    const int64_t* values_by_group;   // bigint mode
    int32_t bigint_mode, null_gid;
    int32_t row_words;
    int32_t word_off[MAX_PACK_KEYS];
    uint8_t* miss_flag;               // set to 1 for rows the fused pass could not resolve
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
PG_D void or_bits(unsigned long long& w0, unsigned long long& w1, unsigned long long v, int shift) {
    if (shift < 64) {
        w0 |= v << shift;
        if (shift > 0) w1 |= v >> (64 - shift);
    } else {
        w1 |= v << (shift - 64);
    }
}
PG_D unsigned long long norm_double_key(long long b) {
    if ((b & 0x7fffffffffffffffLL) == 0) return 0;
    return (unsigned long long)canon_double_bits(b);
}
// pack the key of input row `row`; false when a value does not fit the packed form (string longer than 3 bytes)
PG_D bool pack_row(const PackSpec& ps, int64_t row, unsigned long long& w0, unsigned long long& w1) {
    w0 = 0; w1 = 0;
#pragma unroll
    for (int k = 0; k < MAX_PACK_KEYS; k++) {
        if (k < ps.nkeys) {
            const DCol& c = ps.cols[k];
            if (c.nulls != nullptr && c.nulls[row]) { w1 |= 1ULL << (56 + k); continue; }
            unsigned long long v;
            switch (c.type) {
                case PG_VARCHAR: case PG_VARBINARY: {
                    int32_t o = c.offsets[row], len = c.offsets[row + 1] - o;
                    if (len > 3) return false;
                    const uint8_t* b = (const uint8_t*)c.values + o;
                    v = (unsigned long long)len;
                    for (int t = 0; t < len; t++) v |= (unsigned long long)b[t] << (8 + 8 * t);
                    break;
                }
                case PG_INTEGER: case PG_DATE: v = (unsigned long long)(uint32_t)((const int32_t*)c.values)[row]; break;
                case PG_BOOLEAN: v = ((const uint8_t*)c.values)[row]; break;
                case PG_DOUBLE: v = norm_double_key(((const long long*)c.values)[row]); break;
                default: v = ((const unsigned long long*)c.values)[row];
            }
            or_bits(w0, w1, v, ps.shift[k]);
        }
    }
    return true;
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
        if ((rowp[0] >> k) & 1) { w1 |= 1ULL << (56 + k); continue; }
        unsigned long long w = rowp[ps.word_off[k]], v;
        switch (ps.cols[k].type) {
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
        or_bits(w0, w1, v, ps.shift[k]);
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

// per-row values between the load and the update phase; indexed by slot number only (static after unrolling, so
// the arrays live in registers)
struct RowData {
    int g;               // group id, -1: skip the row
    unsigned condbits;
    unsigned long long v0[MAX_PSLOTS];
    unsigned long long v1[MAX_P128];
};

// FUSED = true: the group id of a row is resolved inside this kernel from the packed key (no id column in HBM); rows that
// cannot be resolved are flagged in miss_flag and counted in *d_miss — the caller runs the general path for exactly those.
template <int T, bool FUSED>
__global__ void __launch_bounds__(T) acc_private2_kernel(const PrivPlan plan, const int32_t* __restrict__ gids, int64_t n, int G,
                                                         const uint8_t* __restrict__ only_rows, const PackSpec pk) {
    extern __shared__ unsigned long long table[];
    __shared__ unsigned long long lut_w0[FUSED ? LUT_SIZE : 1], lut_w1[FUSED ? LUT_SIZE : 1];
    __shared__ int lut_gid[FUSED ? LUT_SIZE : 1];
    const int tid = threadIdx.x;
    const int GT = G * T;
    // ---- init own columns
    for (int s = 0; s < plan.nslots; s++) {
        const PrivSlot& ps = plan.slots[s];
        unsigned long long idv = ps.kind == PS_MIN ? 0x7fffffffffffffffULL : (ps.kind == PS_MAX ? 0x8000000000000000ULL : 0ULL);
        for (int w = 0; w < ps.nwords; w++)
            for (int g = 0; g < G; g++) table[(ps.word_off + w) * GT + g * T + tid] = idv;
    }
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
            if (r.g < 0) { misses++; pk.miss_flag[row] = 1; return; }
        } else {
            r.g = gids[row];
            if (only_rows != nullptr && only_rows[row] == 0) { r.g = -1; return; }
        }
#pragma unroll
        for (int c = 0; c < MAX_PCONDS; c++) {
            if (c < plan.nconds) {
                const PrivCond& pc = plan.conds[c];
                bool ok = true;
                if (pc.mask_col >= 0) {
                    const DCol& m = plan.cols[pc.mask_col];
                    ok = ((const uint8_t*)m.values)[row] != 0 && !(m.nulls != nullptr && m.nulls[row]);
                }
                if (ok && pc.null_col >= 0) ok = plan.cols[pc.null_col].nulls[row] == 0;
                r.condbits |= (ok ? 1u : 0u) << c;
            }
        }
#pragma unroll
        for (int s = 0; s < MAX_PSLOTS; s++) {
            if (s < plan.nslots) {
                const PrivSlot& ps = plan.slots[s];
                if (ps.kind != PS_CNT && ((r.condbits >> ps.cond) & 1)) {
                    const DCol& c = plan.cols[ps.src];
                    if (s < MAX_P128 && ps.conv == CV_DEC128) {
                        const ulonglong2 p = ((const ulonglong2*)c.values)[row];
                        r.v0[s] = p.x; r.v1[s < MAX_P128 ? s : 0] = p.y;
                    } else if (c.type == PG_INTEGER || c.type == PG_DATE) {
                        r.v0[s] = (unsigned long long)(long long)((const int32_t*)c.values)[row];
                    } else if (c.type == PG_BOOLEAN) {
                        r.v0[s] = ((const uint8_t*)c.values)[row];
                    } else {
                        r.v0[s] = ((const unsigned long long*)c.values)[row];
                    }
                }
            }
        }
    };

    auto apply_row = [&](const RowData& r) {
        if (r.g < 0) return;
        unsigned long long* base = table + r.g * T + tid;
#pragma unroll
        for (int s = 0; s < MAX_PSLOTS; s++) {
            if (s < plan.nslots) {
                const PrivSlot& ps = plan.slots[s];
                if ((r.condbits >> ps.cond) & 1) {
                    unsigned long long* t = base + ps.word_off * GT;
                    switch (ps.kind) {
                        case PS_CNT: t[0] += 1; break;
                        case PS_ADD2: {
                            unsigned long long a0 = r.v0[s], a1 = (unsigned long long)((long long)a0 >> 63);
                            unsigned long long s0 = t[0] + a0;
                            t[0] = s0;
                            t[GT] += a1 + (s0 < a0 ? 1ULL : 0ULL);
                            break;
                        }
                        case PS_ADD3: {
                            unsigned long long a0, a1, a2;
                            signmag_to_twos(r.v0[s], r.v1[s < MAX_P128 ? s : 0], a0, a1, a2);
                            unsigned long long s0 = t[0] + a0;
                            t[0] = s0;
                            unsigned long long c0 = s0 < a0 ? 1ULL : 0ULL;
                            unsigned long long add1 = a1 + c0;
                            unsigned long long c1 = add1 < c0 ? 1ULL : 0ULL;
                            unsigned long long s1 = t[GT] + add1;
                            t[GT] = s1;
                            c1 += s1 < add1 ? 1ULL : 0ULL;
                            t[2 * GT] += a2 + c1;
                            break;
                        }
                        case PS_F64: {
                            double d = ps.conv == CV_I2D ? (double)(long long)r.v0[s] : __longlong_as_double((long long)r.v0[s]);
                            t[0] = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)t[0]) + d);
                            break;
                        }
                        case PS_MIN: {
                            long long k = ps.conv == CV_MINKEY ? double_order_key((long long)r.v0[s]) : (long long)r.v0[s];
                            if (k < (long long)t[0]) t[0] = (unsigned long long)k;
                            break;
                        }
                        case PS_MAX: {
                            long long k = ps.conv == CV_MAXKEY ? max_double_key((long long)r.v0[s]) : (long long)r.v0[s];
                            if (k > (long long)t[0]) t[0] = (unsigned long long)k;
                            break;
                        }
                    }
                }
            }
        }
    };

    // ---- row loop, software-pipelined: the loads of the next row are issued before the shared-memory updates of this one
    const int64_t stride = (int64_t)gridDim.x * T;
    int64_t row = (int64_t)blockIdx.x * T + tid;
    RowData cur;
    cur.g = -1;
    if (row < n) load_row(row, cur);
    while (row < n) {
        int64_t nrow = row + stride;
        RowData nxt;
        nxt.g = -1;
        if (nrow < n) load_row(nrow, nxt);
        apply_row(cur);
        cur = nxt;
        row = nrow;
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
        const PrivSlot& cs = plan.slots[pa.cnt_slot];
        for (int g = 0; g < G; g++) {
            unsigned long long cnt = table[cs.word_off * GT + g * T + tid], zero = 0;
            wreduce_sum_carry(cnt, zero);
            cnt = __shfl_sync(0xffffffffu, cnt, 0);
            if (cnt == 0) continue;   // warp-uniform
            unsigned long long* st = pa.state + (int64_t)g * pa.words;
            if (pa.op == OP_COUNT) { if (lane == 0) atomicAdd(&st[0], cnt); continue; }
            const PrivSlot& vs = plan.slots[pa.val_slot];
            const unsigned long long* col = table + vs.word_off * GT + g * T + tid;
            switch (vs.kind) {
                case PS_ADD2: {
                    unsigned long long s0 = col[0], c0 = 0, s1 = col[GT], c1 = 0;
                    wreduce_sum_carry(s0, c0);
                    wreduce_sum_carry(s1, c1);
                    if (lane == 0) {
                        unsigned long long hi = s1 + c0;   // 128-bit signed partial (never overflows: < 2^31 rows of 64-bit values per thread)
                        atomicAdd(&st[0], cnt);
                        if (pa.op == OP_SUM_I64) atomic_add_chain2(st + 1, s0, hi);
                        else atomic_add_chain3(st + 1, s0, hi, (unsigned long long)((long long)hi >> 63));
                    }
                    break;
                }
                case PS_ADD3: {
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
                        v = vs.kind == PS_MIN ? (o < v ? o : v) : (o > v ? o : v);
                    }
                    if (lane == 0) {
                        st[0] = 1;
                        if (vs.kind == PS_MIN) atomicMin((long long*)&st[1], v); else atomicMax((long long*)&st[1], v);
                    }
                    break;
                }
                default: break;
            }
        }
    }
}

// ---- host: build the plan (dedupe columns, conditions and state), pick the block size, launch
static int find_col(PrivPlan& p, const DCol& c) {
    for (int i = 0; i < p.ncols; i++)
        if (p.cols[i].values == c.values && p.cols[i].nulls == c.nulls && p.cols[i].type == c.type) return i;
    if (p.ncols >= MAX_PCOLS) return -1;
    p.cols[p.ncols] = c;
    return p.ncols++;
}

// keys that fit the 120-bit packed form can be resolved inside the accumulation kernel
bool fused_keys_packable(const KeyLayout& L) {
    if (L.nkeys > MAX_PACK_KEYS) return false;
    int bits = 0;
    for (int k = 0; k < L.nkeys; k++) {
        int w = pack_width(L.types[k]);
        if (w < 0) return false;
        bits += w;
    }
    return bits <= 120;
}

bool accumulate_private(pg_ctx* ctx, const AggProgram& prog, const int32_t* gids, int64_t n, int64_t group_count, const uint8_t* only_rows,
                        const FusedLookup* fused) {
    if (group_count < 1 || group_count > 64) return false;
    PrivPlan p;
    memset(&p, 0, sizeof(p));
    p.naggs = prog.naggs;
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        if (op.op == OP_MIN128 || op.op == OP_MAX128) return false;
        int mask_col = -1, null_col = -1;
        if (op.mask.values) { mask_col = find_col(p, op.mask); if (mask_col < 0) return false; }
        int src = -1;
        if (op.op != OP_COUNT || (op.skip_null && op.arg.nulls)) { src = find_col(p, op.arg); if (src < 0) return false; }
        if (op.skip_null && op.arg.nulls) null_col = src;
        int cond = -1;
        for (int c = 0; c < p.nconds; c++) if (p.conds[c].mask_col == mask_col && p.conds[c].null_col == null_col) cond = c;
        if (cond < 0) {
            if (p.nconds >= MAX_PCONDS) return false;
            cond = p.nconds++;
            p.conds[cond].mask_col = mask_col; p.conds[cond].null_col = null_col;
        }
        auto find_slot = [&](int kind, int conv, int src_col, int nwords, int) -> int {
            for (int s = 0; s < p.nslots; s++) {
                const PrivSlot& ps = p.slots[s];
                if (ps.kind == kind && ps.conv == conv && ps.src == src_col && ps.cond == cond) return s;
            }
            if (p.nslots >= MAX_PSLOTS) return -1;
            PrivSlot& ps = p.slots[p.nslots];
            ps.kind = kind; ps.conv = conv; ps.src = src_col; ps.cond = cond; ps.nwords = nwords; ps.word_off = p.words;
            p.words += nwords;
            return p.nslots++;
        };
        PrivAgg& pa = p.aggs[a];
        pa.op = op.op; pa.words = op.words; pa.state = op.state;
        pa.cnt_slot = find_slot(PS_CNT, 0, -1, 1, 0);
        if (pa.cnt_slot < 0) return false;
        pa.val_slot = -1;
        switch (op.op) {
            case OP_COUNT: break;
            case OP_SUM_I64: pa.val_slot = find_slot(PS_ADD2, CV_INT, src, 2, 1); break;
            case OP_SUM_F64: pa.val_slot = find_slot(PS_F64, op.convert_i2d ? CV_I2D : CV_F64, src, 1, 1); break;
            case OP_SUM_DEC:
                pa.val_slot = op.in_type == PG_LONG_DECIMAL ? find_slot(PS_ADD3, CV_DEC128, src, 3, 2) : find_slot(PS_ADD2, CV_INT, src, 2, 1);
                break;
            case OP_MIN64: pa.val_slot = find_slot(PS_MIN, op.in_type == PG_DOUBLE ? CV_MINKEY : CV_INT, src, 1, 1); break;
            case OP_MAX64: pa.val_slot = find_slot(PS_MAX, op.in_type == PG_DOUBLE ? CV_MAXKEY : CV_INT, src, 1, 1); break;
            default: return false;
        }
        if (op.op != OP_COUNT && pa.val_slot < 0) return false;
    }
    // slots fed by 128-bit inputs first (the kernel keeps their second value word in v1[slot], slot < MAX_P128)
    {
        int order[MAX_PSLOTS], remap[MAX_PSLOTS], k = 0;
        for (int s = 0; s < p.nslots; s++) if (p.slots[s].kind == PS_ADD3) order[k++] = s;
        if (k > MAX_P128) return false;
        for (int s = 0; s < p.nslots; s++) if (p.slots[s].kind != PS_ADD3) order[k++] = s;
        PrivSlot tmp[MAX_PSLOTS];
        for (int i = 0; i < p.nslots; i++) { tmp[i] = p.slots[order[i]]; remap[order[i]] = i; }
        for (int i = 0; i < p.nslots; i++) p.slots[i] = tmp[i];
        for (int a = 0; a < p.naggs; a++) {
            p.aggs[a].cnt_slot = remap[p.aggs[a].cnt_slot];
            if (p.aggs[a].val_slot >= 0) p.aggs[a].val_slot = remap[p.aggs[a].val_slot];
        }
    }
    PackSpec pk;
    memset(&pk, 0, sizeof(pk));
    if (fused) {
        const KeyLayout& L = *fused->layout;
        if (!fused_keys_packable(L)) return false;
        pk.nkeys = L.nkeys;
        int bits = 0;
        for (int k = 0; k < L.nkeys; k++) {
            pk.shift[k] = bits; bits += pack_width(L.types[k]);
            pk.cols[k] = fused->keys->c[k];
            pk.word_off[k] = L.word_off[k];
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
    int64_t want = (n + T - 1) / T;
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
