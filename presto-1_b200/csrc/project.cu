// project.cu — PageProcessor (filter + projections): validation and typing of the expression DAG, the interpreting kernels,
// pg_project_process, and C code generation for the run-time specialised kernels.  See project.h / include/prestogpu.h.
#include "project.h"

#include <stdio.h>

#include "jit/decimal_ops.cuh"
#include "scan.cuh"

namespace pg {

namespace {

__constant__ unsigned long long P10_TABLE[39][2];
static unsigned long long h_p10[39][2];
static bool h_p10_ready = false;

void ensure_pow10() {
    if (!h_p10_ready) {
        unsigned __int128 v = 1;
        for (int i = 0; i < 39; i++) { h_p10[i][0] = (unsigned long long)v; h_p10[i][1] = (unsigned long long)(v >> 64); v *= 10; }
        h_p10_ready = true;
    }
}
void upload_pow10() {   // per device; cheap enough to repeat at every create
    ensure_pow10();
    PG_CUDA(cudaMemcpyToSymbol(P10_TABLE, h_p10, sizeof(h_p10)));
}

PG_HD bool is_dec(int t) { return t == PG_SHORT_DECIMAL || t == PG_LONG_DECIMAL; }

struct PV { unsigned long long lo, hi; int null; };

PG_D PgxMag mag_of(const PV& v, int type) { return type == PG_LONG_DECIMAL ? pgx_mag_long(v.lo, v.hi) : pgx_mag_short((long long)v.lo); }
PG_D PgxMag scale10(PgxMag m, int k) {
    if (k <= 0) return m;
    return pgx_mul(m, pgx_mag_const(P10_TABLE[k][0], P10_TABLE[k][1]));
}
PG_D long long value_hash(const PV& v, int type, const uint8_t* base) {   // HASH_CODE per type, null -> 0 (common.cuh dcol_type_hash)
    if (v.null) return 0;
    switch (type) {
        case PG_BIGINT: return long_type_hash((int64_t)v.lo);
        case PG_INTEGER: case PG_DATE: return long_type_hash((int64_t)(int32_t)v.lo);
        case PG_DOUBLE: { int64_t b = (int64_t)v.lo; if ((b & 0x7fffffffffffffffLL) == 0) b = 0; return long_type_hash(canon_double_bits(b)); }
        case PG_SHORT_DECIMAL: return (long long)v.lo;
        case PG_LONG_DECIMAL: return xxh64_long((int64_t)v.lo) ^ xxh64_long((int64_t)v.hi);
        case PG_VARCHAR: case PG_VARBINARY: return xxh64_bytes((const uint8_t*)v.lo, (int64_t)v.hi);
        case PG_BOOLEAN: return v.lo ? 1231 : 1237;
    }
    return 0;
}

// Evaluates the nodes with in_filter == phase_filter for `row`.  err: bit 0 = arithmetic overflow.
PG_D void eval_nodes(const PNode* __restrict__ nodes, int n_nodes, int phase_filter, const ProjCols& in, const uint8_t* __restrict__ cbytes,
                     int64_t row, PV* vals, unsigned int& err) {
    for (int i = 0; i < n_nodes; i++) {
        const PNode e = nodes[i];
        if (e.in_filter != phase_filter) continue;
        PV r; r.lo = 0; r.hi = 0; r.null = 0;
        switch (e.op) {
            case PG_EXPR_INPUT: {
                const DCol& c = in.c[e.channel];
                r.null = dcol_is_null(c, row) ? 1 : 0;
                if (!r.null) {
                    if (c.type == PG_LONG_DECIMAL) { const unsigned long long* p = (const unsigned long long*)c.values + 2 * row; r.lo = p[0]; r.hi = p[1]; }
                    else if (c.type == PG_VARCHAR || c.type == PG_VARBINARY) {
                        const int32_t o = c.offsets[row];
                        r.lo = (unsigned long long)((const uint8_t*)c.values + o); r.hi = (unsigned long long)(c.offsets[row + 1] - o);
                    } else r.lo = (unsigned long long)dcol_long(c, row);
                }
                break;
            }
            case PG_EXPR_CONST:
                r.null = e.is_null;
                if (e.type == PG_VARCHAR || e.type == PG_VARBINARY) { r.lo = (unsigned long long)(cbytes + e.lo); r.hi = (unsigned long long)e.hi; }
                else { r.lo = (unsigned long long)e.lo; r.hi = (unsigned long long)e.hi; }
                break;
            case PG_EXPR_ADD: case PG_EXPR_SUB: case PG_EXPR_MUL: {
                const PV a = vals[e.a0], b = vals[e.a1];
                const int ta = nodes[e.a0].type, tb = nodes[e.a1].type;
                if (a.null || b.null) { r.null = 1; break; }
                if (is_dec(e.type)) {
                    if (e.type == PG_SHORT_DECIMAL) {   // long arithmetic, cannot overflow by construction of the result precision
                        const long long x = (long long)a.lo, y = (long long)b.lo;
                        if (e.op == PG_EXPR_MUL) r.lo = (unsigned long long)(x * y);
                        else {
                            const long long sx = x * (long long)P10_TABLE[e.up0][0], sy = y * (long long)P10_TABLE[e.up1][0];
                            r.lo = (unsigned long long)(e.op == PG_EXPR_ADD ? sx + sy : sx - sy);
                        }
                    } else {
                        PgxMag x = mag_of(a, ta), y = mag_of(b, tb), z;
                        if (e.op == PG_EXPR_MUL) z = pgx_mul(x, y);
                        else {
                            x = scale10(x, e.up0); y = scale10(y, e.up1);
                            if (e.op == PG_EXPR_SUB) y.neg = !y.neg;
                            z = pgx_add(x, y);
                        }
                        if (pgx_overflows(z)) err |= 1u;
                        r.lo = z.lo; r.hi = pgx_long_hi(z);
                    }
                } else if (e.type == PG_DOUBLE) {
                    const double x = __longlong_as_double((long long)a.lo), y = __longlong_as_double((long long)b.lo);
                    const double z = e.op == PG_EXPR_ADD ? x + y : e.op == PG_EXPR_SUB ? x - y : x * y;
                    r.lo = (unsigned long long)__double_as_longlong(z);
                } else if (e.type == PG_BIGINT) {     // Math.addExact / subtractExact / multiplyExact (BigintOperators.java:47-83)
                    const long long x = (long long)a.lo, y = (long long)b.lo;
                    long long z;
                    bool ov;
                    if (e.op == PG_EXPR_ADD) { z = (long long)((unsigned long long)x + (unsigned long long)y); ov = ((x ^ z) & (y ^ z)) < 0; }
                    else if (e.op == PG_EXPR_SUB) { z = (long long)((unsigned long long)x - (unsigned long long)y); ov = ((x ^ y) & (x ^ z)) < 0; }
                    else {
                        z = (long long)((unsigned long long)x * (unsigned long long)y);
                        const long long hi = __mul64hi(x, y);
                        ov = hi != (z >> 63);
                    }
                    if (ov) err |= 1u;
                    r.lo = (unsigned long long)z;
                } else {                               // INTEGER: 32-bit exact arithmetic
                    const long long x = (long long)(int32_t)a.lo, y = (long long)(int32_t)b.lo;
                    const long long z = e.op == PG_EXPR_ADD ? x + y : e.op == PG_EXPR_SUB ? x - y : x * y;
                    if (z != (long long)(int32_t)z) err |= 1u;
                    r.lo = (unsigned long long)(long long)(int32_t)z;
                }
                break;
            }
            case PG_EXPR_LT: case PG_EXPR_LE: case PG_EXPR_GT: case PG_EXPR_GE: case PG_EXPR_EQ: case PG_EXPR_NE: {
                const PV a = vals[e.a0], b = vals[e.a1];
                const int ta = nodes[e.a0].type, tb = nodes[e.a1].type;
                if (a.null || b.null) { r.null = 1; break; }
                int c;   // -1 / 0 / 1, 2 = unordered
                if (ta == PG_DOUBLE) {
                    const double x = __longlong_as_double((long long)a.lo), y = __longlong_as_double((long long)b.lo);
                    c = (x != x || y != y) ? 2 : (x < y ? -1 : x > y ? 1 : 0);
                } else if (ta == PG_VARCHAR || ta == PG_VARBINARY) {
                    const uint8_t *p = (const uint8_t*)a.lo, *q = (const uint8_t*)b.lo;
                    const long long la = (long long)a.hi, lb = (long long)b.hi, m = la < lb ? la : lb;
                    c = 0;
                    for (long long t = 0; t < m && c == 0; t++) c = p[t] < q[t] ? -1 : p[t] > q[t] ? 1 : 0;
                    if (c == 0) c = la < lb ? -1 : la > lb ? 1 : 0;
                } else if (ta == PG_LONG_DECIMAL || tb == PG_LONG_DECIMAL) c = pgx_compare(mag_of(a, ta), mag_of(b, tb));
                else { const long long x = (long long)a.lo, y = (long long)b.lo; c = x < y ? -1 : x > y ? 1 : 0; }
                bool res;
                if (c == 2) res = e.op == PG_EXPR_NE;
                else res = e.op == PG_EXPR_LT ? c < 0 : e.op == PG_EXPR_LE ? c <= 0 : e.op == PG_EXPR_GT ? c > 0 : e.op == PG_EXPR_GE ? c >= 0
                         : e.op == PG_EXPR_EQ ? c == 0 : c != 0;
                r.lo = res ? 1 : 0;
                break;
            }
            case PG_EXPR_AND: {
                const PV a = vals[e.a0], b = vals[e.a1];
                if ((!a.null && !a.lo) || (!b.null && !b.lo)) r.lo = 0;
                else if (a.null || b.null) r.null = 1;
                else r.lo = 1;
                break;
            }
            case PG_EXPR_OR: {
                const PV a = vals[e.a0], b = vals[e.a1];
                if ((!a.null && a.lo) || (!b.null && b.lo)) r.lo = 1;
                else if (a.null || b.null) r.null = 1;
                else r.lo = 0;
                break;
            }
            case PG_EXPR_NOT: r = vals[e.a0]; if (!r.null) r.lo = r.lo ? 0 : 1; break;
            case PG_EXPR_IS_NULL: r.lo = vals[e.a0].null ? 1 : 0; break;
            case PG_EXPR_CAST: {
                const PV a = vals[e.a0];
                const int ta = nodes[e.a0].type;
                if (a.null) { r.null = 1; break; }
                if (e.type == PG_DOUBLE) r.lo = (unsigned long long)__double_as_longlong((double)(long long)a.lo);
                else if (e.type == PG_BIGINT) r.lo = a.lo;
                else {
                    PgxMag z = scale10(mag_of(a, ta), e.up0);
                    const PgxMag lim = pgx_mag_const(P10_TABLE[e.precision][0], P10_TABLE[e.precision][1]);
                    PgxMag zabs = z; zabs.neg = 0;
                    if (z.ovf || (e.precision < 38 ? pgx_compare(zabs, lim) >= 0 : pgx_ge_1e38(z))) err |= 1u;
                    if (e.type == PG_LONG_DECIMAL) { r.lo = z.lo; r.hi = pgx_long_hi(z); }
                    else r.lo = (unsigned long long)pgx_short_value(z);
                }
                break;
            }
            case PG_EXPR_HASH: {
                const unsigned long long prev = e.a0 >= 0 ? vals[e.a0].lo : 0ULL;
                r.lo = 31ULL * prev + (unsigned long long)value_hash(vals[e.a1], nodes[e.a1].type, cbytes);
                break;
            }
        }
        vals[i] = r;
    }
}

// pass 1: filter (and the optional row bitmap) -> one flag byte per row + the number of selected rows per tile
__global__ void __launch_bounds__(SCAN_THREADS) proj_select_kernel(const PNode* __restrict__ nodes, int n_nodes, int filter, ProjCols in,
                                                                   const uint8_t* __restrict__ cbytes, int64_t n, const unsigned int* __restrict__ only_bits,
                                                                   uint8_t* __restrict__ flags, Pair64* __restrict__ tile_sums, unsigned int* __restrict__ d_err) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
    unsigned int err = 0;
    PV vals[PROJ_MAX_NODES];
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i >= n) continue;
        bool pass = only_bits == nullptr || ((only_bits[i >> 5] >> (i & 31)) & 1u);
        if (pass && filter >= 0) {
            eval_nodes(nodes, n_nodes, 1, in, cbytes, i, vals, err);
            pass = !vals[filter].null && vals[filter].lo != 0;
        }
        flags[i] = pass ? 1 : 0;
        acc.a += pass ? 1 : 0;
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
    if (err) atomicOr(d_err, err);
}
// pass 1, common case: the filter is one comparison of a fixed width input column with a constant of the same type (a date or key
// range).  No interpreter: 4 or 8 bytes read per row instead of a walk over the node array with 24-byte values in local memory
// (600 M rows: 10 ms interpreted).  cmp: PG_EXPR_LT .. PG_EXPR_NE; a NULL input never passes.
template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) proj_select_cmp_kernel(const T* __restrict__ values, const uint8_t* __restrict__ nulls, T constant, int cmp,
                                                                       int64_t n, const unsigned int* __restrict__ only_bits, uint8_t* __restrict__ flags,
                                                                       Pair64* __restrict__ tile_sums) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 acc{0, 0};
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t i = tile_base + (int64_t)j * SCAN_THREADS + threadIdx.x;
        if (i >= n) continue;
        bool pass = only_bits == nullptr || ((only_bits[i >> 5] >> (i & 31)) & 1u);
        if (pass) {
            const T v = values[i];
            switch (cmp) {
                case PG_EXPR_LT: pass = v < constant; break;
                case PG_EXPR_LE: pass = v <= constant; break;
                case PG_EXPR_GT: pass = v > constant; break;
                case PG_EXPR_GE: pass = v >= constant; break;
                case PG_EXPR_EQ: pass = v == constant; break;
                default: pass = v != constant;
            }
            if (nulls && nulls[i]) pass = false;
        }
        flags[i] = pass ? 1 : 0;
        acc.a += pass ? 1 : 0;
    }
    Pair64 total;
    block_exclusive_scan(acc, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
// pass 2: selected row numbers in page order
__global__ void __launch_bounds__(SCAN_THREADS) proj_positions_kernel(const uint8_t* __restrict__ flags, int64_t n, const Pair64* __restrict__ tile_prefix,
                                                                      int64_t* __restrict__ sel) {
    __shared__ Pair64 smem[SCAN_THREADS / 32 + 1];
    const int64_t tile_base = (int64_t)blockIdx.x * SCAN_TILE;
    Pair64 run = tile_prefix[blockIdx.x];
    // blocked order inside the tile so that one scan per tile suffices: thread t owns rows [t*ITEMS, (t+1)*ITEMS)
    const int64_t first = tile_base + (int64_t)threadIdx.x * SCAN_ITEMS;
    Pair64 mine{0, 0};
    uint8_t f[SCAN_ITEMS];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) { f[j] = (first + j < n) ? flags[first + j] : 0; mine.a += f[j]; }
    Pair64 total;
    Pair64 ex = block_exclusive_scan(mine, smem, &total);
    int64_t dst = run.a + ex.a;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) if (f[j]) sel[dst++] = first + j;
}

// pass 2 for a filter whose projections only pass fixed-width channels through (the usual ScanFilterAndProject shape): the selected rows
// of every such channel move straight to their place in the output, no list of row numbers in between
struct CompactTab {
    const void* src[PROJ_MAX_OUTPUTS]; const uint8_t* src_nulls[PROJ_MAX_OUTPUTS];
    void* dst[PROJ_MAX_OUTPUTS]; uint8_t* dst_nulls[PROJ_MAX_OUTPUTS];
    int32_t width[PROJ_MAX_OUTPUTS]; int32_t count;
};
// One warp owns 32 * SCAN_ITEMS consecutive rows and walks them 32 at a time, so that loads are coalesced and the selected rows of each
// group of 32 land next to each other (ballot + popc give the place; page order is kept).
template <typename T>
PG_D void compact_column(const T* __restrict__ src, T* __restrict__ dst, const unsigned (&b)[SCAN_ITEMS], int64_t wbase, int64_t d, int lane) {
    // the loads of a batch are issued before its first store: several requests in flight per lane instead of one
    constexpr int BATCH = sizeof(T) > 8 ? 4 : SCAN_ITEMS;
    const T* s0 = src + wbase + lane;
    T* d0 = dst + d;
#pragma unroll
    for (int j0 = 0; j0 < SCAN_ITEMS; j0 += BATCH) {
        T v[BATCH];
#pragma unroll
        for (int j = 0; j < BATCH; j++) if ((b[j0 + j] >> lane) & 1u) v[j] = s0[(j0 + j) * 32];
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
            if ((b[j0 + j] >> lane) & 1u) d0[__popc(b[j0 + j] & ((1u << lane) - 1u))] = v[j];
            d0 += __popc(b[j0 + j]);
        }
    }
}
template <typename T>      // one launch moves the channels of one width (CompactTab.width is all sizeof(T))
__global__ void __launch_bounds__(SCAN_THREADS, 3) proj_compact_kernel(const uint8_t* __restrict__ flags, int64_t n, const Pair64* __restrict__ tile_prefix,
                                                                       CompactTab t) {
    __shared__ int s_warp[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t wbase = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)warp * (32 * SCAN_ITEMS);
    unsigned b[SCAN_ITEMS];
    int mine = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        const int64_t row = wbase + j * 32 + lane;
        b[j] = __ballot_sync(0xffffffffu, row < n && flags[row] != 0);
        mine += __popc(b[j]);
    }
    if (lane == 0) s_warp[warp] = mine;
    __syncthreads();
    int64_t d = tile_prefix[blockIdx.x].a;
    for (int w = 0; w < warp; w++) d += s_warp[w];
    if (mine == 0) return;
    for (int c = 0; c < t.count; c++) {
        compact_column((const T*)t.src[c], (T*)t.dst[c], b, wbase, d, lane);
        if (t.dst_nulls[c]) {
            if (t.src_nulls[c]) compact_column(t.src_nulls[c], t.dst_nulls[c], b, wbase, d, lane);
            else {
                int64_t q = d;
#pragma unroll
                for (int j = 0; j < SCAN_ITEMS; j++) {
                    if ((b[j] >> lane) & 1u) t.dst_nulls[c][q + __popc(b[j] & ((1u << lane) - 1u))] = 0;
                    q += __popc(b[j]);
                }
            }
        }
    }
}

struct ProjOut { void* values[PROJ_MAX_OUTPUTS]; uint8_t* nulls[PROJ_MAX_OUTPUTS]; int32_t node[PROJ_MAX_OUTPUTS]; int32_t count; };

// computed projections over the selected rows (sel == nullptr: every row)
__global__ void proj_eval_kernel(const PNode* __restrict__ nodes, int n_nodes, int has_filter, ProjCols in, const uint8_t* __restrict__ cbytes,
                                 const int64_t* __restrict__ sel, int64_t m, ProjOut out, unsigned int* __restrict__ d_err) {
    PV vals[PROJ_MAX_NODES];
    unsigned int err = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
        const int64_t row = sel ? sel[i] : i;
        unsigned int ferr = 0;
        if (has_filter) eval_nodes(nodes, n_nodes, 1, in, cbytes, row, vals, ferr);    // projections may read nodes of the filter
        eval_nodes(nodes, n_nodes, 0, in, cbytes, row, vals, err);
        for (int k = 0; k < out.count; k++) {
            const PV v = vals[out.node[k]];
            if (out.nulls[k]) out.nulls[k][i] = (uint8_t)v.null;
            switch (nodes[out.node[k]].type) {
                case PG_INTEGER: case PG_DATE: ((int32_t*)out.values[k])[i] = v.null ? 0 : (int32_t)v.lo; break;
                case PG_BOOLEAN: ((uint8_t*)out.values[k])[i] = v.null ? 0 : (uint8_t)v.lo; break;
                case PG_LONG_DECIMAL: ((unsigned long long*)out.values[k])[2 * i] = v.null ? 0 : v.lo; ((unsigned long long*)out.values[k])[2 * i + 1] = v.null ? 0 : v.hi; break;
                default: ((unsigned long long*)out.values[k])[i] = v.null ? 0 : v.lo;
            }
        }
    }
    if (err) atomicOr(d_err, err);
}

}  // namespace

// ------------------------------------------------------------------ validation + typing (the reference's operator signatures)
static void set_decimal(PNode& e, int p, int s) {
    e.precision = p > 38 ? 38 : p; e.scale = s; e.type = e.precision <= 18 ? PG_SHORT_DECIMAL : PG_LONG_DECIMAL;
}

static void project_build(pg_project* p, const pg_expr* in, int32_t n, int32_t filter, const int32_t* proj, int32_t np) {
    PG_CHECK_ARG(in && n >= 1 && n <= PROJ_MAX_NODES, "between 1 and %d expression nodes are supported", PROJ_MAX_NODES);
    PG_CHECK_ARG(np >= 0 && np <= 64 && (np == 0 || proj), "bad projection list");
    PG_CHECK_ARG(filter >= -1 && filter < n, "filter node out of range");
    p->nodes.resize(n);
    p->channel_used.assign(PROJ_MAX_CHANNELS, 0);
    for (int i = 0; i < n; i++) {
        const pg_expr& x = in[i];
        PNode e; memset(&e, 0, sizeof(e));
        e.op = x.op; e.a0 = x.arg0; e.a1 = x.arg1; e.channel = -1;
        auto arg = [&](int a) -> const PNode& {
            PG_CHECK_ARG(a >= 0 && a < i, "node %d: operand %d must be an earlier node", i, a);
            return p->nodes[a];
        };
        switch (x.op) {
            case PG_EXPR_INPUT:
                PG_CHECK_ARG(x.channel >= 0 && x.channel < PROJ_MAX_CHANNELS, "node %d: input channel %d out of range", i, x.channel);
                PG_CHECK_ARG(type_valid(x.type), "node %d: unknown type %d", i, x.type);
                e.type = x.type; e.precision = x.precision; e.scale = x.scale; e.channel = x.channel; e.a0 = e.a1 = -1;
                p->channel_used[x.channel] = 1;
                if (x.channel > p->max_channel) p->max_channel = x.channel;
                break;
            case PG_EXPR_CONST:
                PG_CHECK_ARG(type_valid(x.type), "node %d: unknown type %d", i, x.type);
                e.type = x.type; e.precision = x.precision; e.scale = x.scale; e.is_null = x.is_null ? 1 : 0; e.lo = x.lo; e.hi = x.hi; e.a0 = e.a1 = -1;
                if (type_is_var(x.type)) {
                    PG_CHECK_ARG(x.nbytes >= 0 && (x.nbytes == 0 || x.bytes), "node %d: constant without bytes", i);
                    e.lo = (int64_t)p->const_bytes.size(); e.hi = x.nbytes;
                    p->const_bytes.insert(p->const_bytes.end(), x.bytes, x.bytes + x.nbytes);
                }
                break;
            case PG_EXPR_ADD: case PG_EXPR_SUB: case PG_EXPR_MUL: {
                const PNode &a = arg(x.arg0), &b = arg(x.arg1);
                if (is_dec(a.type) && is_dec(b.type)) {
                    if (x.op == PG_EXPR_MUL) set_decimal(e, a.precision + b.precision, a.scale + b.scale);            // DecimalOperators.java:261-262
                    else {                                                                                           // :84-85, :184-185
                        const int s = a.scale > b.scale ? a.scale : b.scale;
                        const int ip = (a.precision - a.scale) > (b.precision - b.scale) ? (a.precision - a.scale) : (b.precision - b.scale);
                        set_decimal(e, ip + s + 1, s);
                        e.up0 = s - a.scale; e.up1 = s - b.scale;
                    }
                } else {
                    PG_CHECK_ARG(a.type == b.type && (a.type == PG_BIGINT || a.type == PG_INTEGER || a.type == PG_DOUBLE),
                                 "node %d: arithmetic needs two decimals, two bigints, two integers or two doubles (got %d, %d)", i, a.type, b.type);
                    e.type = a.type;
                }
                break;
            }
            case PG_EXPR_LT: case PG_EXPR_LE: case PG_EXPR_GT: case PG_EXPR_GE: case PG_EXPR_EQ: case PG_EXPR_NE: {
                const PNode &a = arg(x.arg0), &b = arg(x.arg1);
                const bool both_dec = is_dec(a.type) && is_dec(b.type);
                PG_CHECK_ARG(both_dec ? a.scale == b.scale : (a.type == b.type && a.type != PG_LONG_DECIMAL) || (type_is_var(a.type) && type_is_var(b.type)),
                             "node %d: comparison of type %d with %d (decimals need equal scales)", i, a.type, b.type);
                e.type = PG_BOOLEAN;
                break;
            }
            case PG_EXPR_AND: case PG_EXPR_OR:
                PG_CHECK_ARG(arg(x.arg0).type == PG_BOOLEAN && arg(x.arg1).type == PG_BOOLEAN, "node %d: AND / OR over non-boolean operands", i);
                e.type = PG_BOOLEAN;
                break;
            case PG_EXPR_NOT:
                PG_CHECK_ARG(arg(x.arg0).type == PG_BOOLEAN, "node %d: NOT over a non-boolean operand", i);
                e.type = PG_BOOLEAN; e.a1 = -1;
                break;
            case PG_EXPR_IS_NULL: (void)arg(x.arg0); e.type = PG_BOOLEAN; e.a1 = -1; break;
            case PG_EXPR_CAST: {
                const PNode& a = arg(x.arg0);
                e.type = x.type; e.precision = x.precision; e.scale = x.scale; e.a1 = -1;
                if (x.type == PG_DOUBLE || x.type == PG_BIGINT) PG_CHECK_ARG(a.type == PG_BIGINT || a.type == PG_INTEGER, "node %d: unsupported cast %d -> %d", i, a.type, x.type);
                else {
                    PG_CHECK_ARG(is_dec(x.type) && (is_dec(a.type) || a.type == PG_BIGINT || a.type == PG_INTEGER), "node %d: unsupported cast %d -> %d", i, a.type, x.type);
                    PG_CHECK_ARG(x.precision >= 1 && x.precision <= 38 && (x.type == PG_SHORT_DECIMAL) == (x.precision <= 18), "node %d: bad decimal precision", i);
                    e.up0 = x.scale - (is_dec(a.type) ? a.scale : 0);
                    PG_CHECK_ARG(e.up0 >= 0 && e.up0 <= 38, "node %d: a cast may not reduce the scale", i);
                }
                break;
            }
            case PG_EXPR_HASH:
                if (x.arg0 >= 0) PG_CHECK_ARG(arg(x.arg0).type == PG_BIGINT, "node %d: the previous hash must be BIGINT", i);
                else e.a0 = -1;
                (void)arg(x.arg1);
                e.type = PG_BIGINT;
                break;
            default: fail(PG_E_INVALID_ARGUMENT, "node %d: unknown expression op %d", i, x.op);
        }
        if (is_dec(e.type)) PG_CHECK_ARG(e.precision >= 1 && e.precision <= 38 && e.scale >= 0 && e.scale <= e.precision && (e.type == PG_SHORT_DECIMAL) == (e.precision <= 18),
                                         "node %d: decimal(%d,%d) does not match its representation", i, e.precision, e.scale);
        p->nodes[i] = e;
    }
    p->filter = filter;
    if (filter >= 0) {
        PG_CHECK_ARG(p->nodes[filter].type == PG_BOOLEAN, "the filter must be a boolean expression");
        p->nodes[filter].in_filter = 1;
        for (int i = n - 1; i >= 0; i--)
            if (p->nodes[i].in_filter) {
                if (p->nodes[i].a0 >= 0) p->nodes[p->nodes[i].a0].in_filter = 1;
                if (p->nodes[i].a1 >= 0) p->nodes[p->nodes[i].a1].in_filter = 1;
            }
    }
    p->proj.assign(proj, proj + np);
    for (int k = 0; k < np; k++) {
        PG_CHECK_ARG(proj[k] >= 0 && proj[k] < n, "projection %d: node out of range", k);
        const PNode& e = p->nodes[proj[k]];
        PG_CHECK_ARG(!type_is_var(e.type) || e.op == PG_EXPR_INPUT, "projection %d: variable width values can only be passed through", k);
    }
    pg_ctx* ctx = p->ctx;
    upload_pow10();
    p->d_nodes.alloc(ctx, (int64_t)n * (int64_t)sizeof(PNode));
    copy_async(ctx, p->d_nodes.p, p->nodes.data(), (int64_t)n * (int64_t)sizeof(PNode), cudaMemcpyHostToDevice);
    p->d_bytes.alloc(ctx, (int64_t)p->const_bytes.size() + 8);
    if (!p->const_bytes.empty()) copy_async(ctx, p->d_bytes.p, p->const_bytes.data(), (int64_t)p->const_bytes.size(), cudaMemcpyHostToDevice);
    sync(ctx);
}

void project_materialize(pg_project* p, const DCol* in, int ncols, int64_t n, const unsigned int* only_bits, const std::vector<char>* want,
                         ProjectedPage& out, const pg_col_out* direct) {
    pg_ctx* ctx = p->ctx;
    PG_CHECK_ARG(p->max_channel < ncols, "the page has %d channels, the processor reads channel %d", ncols, p->max_channel);
    const int np = (int)p->proj.size();
    out.cols.assign(np, DCol{0, 0, nullptr, nullptr, nullptr});
    out.in_place.assign(np, 0);
    out.n = 0;
    ProjCols pc; memset(&pc, 0, sizeof(pc));
    for (int c = 0; c <= p->max_channel; c++) if (p->channel_used[c]) pc.c[c] = in[c];
    for (size_t i = 0; i < p->nodes.size(); i++)
        if (p->nodes[i].op == PG_EXPR_INPUT)
            PG_CHECK_ARG(in[p->nodes[i].channel].type == p->nodes[i].type, "channel %d has type %d, the expression expects %d", p->nodes[i].channel,
                         in[p->nodes[i].channel].type, p->nodes[i].type);
    DevBuf err(ctx, 8);
    memset_async(ctx, err.p, 0, 8);
    const PNode* d_nodes = p->d_nodes.as<PNode>();
    const uint8_t* d_bytes = p->d_bytes.as<uint8_t>();
    const int n_nodes = (int)p->nodes.size();
    const bool selecting = p->filter >= 0 || only_bits != nullptr;
    int64_t m = n;
    DevBuf sel;
    if (selecting && n > 0) {
        const int64_t ntiles = num_tiles(n);
        DevBuf flags(ctx, n), tiles(ctx, (ntiles + 1) * (int64_t)sizeof(Pair64));
        // the filter as "input column <cmp> constant" over integer-like values of one width (exact for BIGINT, INTEGER, DATE and for
        // short decimals of equal scale; doubles and everything else go through the interpreter)
        int simple = 0;
        const PNode* fa = nullptr; const PNode* fb = nullptr;
        if (p->filter >= 0) {
            const PNode& f = p->nodes[p->filter];
            if (f.op >= PG_EXPR_LT && f.op <= PG_EXPR_NE && f.a0 >= 0 && f.a1 >= 0) {
                fa = &p->nodes[f.a0]; fb = &p->nodes[f.a1];
                const bool ints = fa->type == fb->type && (fa->type == PG_BIGINT || fa->type == PG_INTEGER || fa->type == PG_DATE ||
                                                           (fa->type == PG_SHORT_DECIMAL && fa->scale == fb->scale));
                if (ints && fa->op == PG_EXPR_INPUT && fb->op == PG_EXPR_CONST && !fb->is_null) simple = f.op;
            }
        }
        if (simple && !getenv("PRESTOGPU_PROJECT_INTERPRET")) {
            const DCol& c = in[fa->channel];
            if (type_width(fa->type) == 4)
                PG_LAUNCH(ctx, proj_select_cmp_kernel<int32_t>, (int)ntiles, SCAN_THREADS, 0, (const int32_t*)c.values, c.nulls, (int32_t)fb->lo, simple, n, only_bits,
                          flags.as<uint8_t>(), tiles.as<Pair64>());
            else
                PG_LAUNCH(ctx, proj_select_cmp_kernel<int64_t>, (int)ntiles, SCAN_THREADS, 0, (const int64_t*)c.values, c.nulls, (int64_t)fb->lo, simple, n, only_bits,
                          flags.as<uint8_t>(), tiles.as<Pair64>());
        } else
        PG_LAUNCH(ctx, proj_select_kernel, (int)ntiles, SCAN_THREADS, 0, d_nodes, n_nodes, p->filter, pc, d_bytes, n, only_bits, flags.as<uint8_t>(),
                  tiles.as<Pair64>(), err.as<unsigned int>());
        scan_tile_totals(ctx, tiles.as<Pair64>(), ntiles, tiles.as<Pair64>() + ntiles);
        int64_t tot[2];
        read_scalars(ctx, tiles.as<Pair64>() + ntiles, 2, tot);
        m = tot[0];
        bool compactable = m > 0 && m < n && !getenv("PRESTOGPU_PROJECT_NO_COMPACT");
        for (int k = 0; k < np && compactable; k++) {
            if (want && !(*want)[k]) continue;
            const PNode& e = p->nodes[p->proj[k]];
            if (e.op != PG_EXPR_INPUT || type_is_var(e.type)) compactable = false;
        }
        if (compactable) {
            out.n = m;
            for (int width : {1, 4, 8, 16}) {
            CompactTab ct; memset(&ct, 0, sizeof(ct));
            auto flush = [&] {
                if (ct.count > 0) {
                    if (width == 1) PG_LAUNCH(ctx, proj_compact_kernel<uint8_t>, (int)ntiles, SCAN_THREADS, 0, flags.as<uint8_t>(), n, tiles.as<Pair64>(), ct);
                    else if (width == 4) PG_LAUNCH(ctx, proj_compact_kernel<int32_t>, (int)ntiles, SCAN_THREADS, 0, flags.as<uint8_t>(), n, tiles.as<Pair64>(), ct);
                    else if (width == 8) PG_LAUNCH(ctx, proj_compact_kernel<int64_t>, (int)ntiles, SCAN_THREADS, 0, flags.as<uint8_t>(), n, tiles.as<Pair64>(), ct);
                    else PG_LAUNCH(ctx, proj_compact_kernel<int4>, (int)ntiles, SCAN_THREADS, 0, flags.as<uint8_t>(), n, tiles.as<Pair64>(), ct);
                }
                ct.count = 0;
            };
            for (int k = 0; k < np; k++) {
                if (want && !(*want)[k]) continue;
                const PNode& e = p->nodes[p->proj[k]];
                if (type_width(e.type) != width) continue;
                const DCol& src = in[e.channel];
                DCol& oc = out.cols[k];
                const bool into = direct && direct[k].mem == PG_MEM_DEVICE && direct[k].values && direct[k].type == e.type && direct[k].capacity >= m;
                oc.type = e.type; oc.pad = 0; oc.offsets = nullptr;
                if (into) { oc.values = direct[k].values; oc.nulls = direct[k].nulls; out.in_place[k] = 1; }
                else {
                    out.bufs.emplace_back(ctx, m * (int64_t)type_width(e.type));
                    oc.values = out.bufs.back().p;
                    oc.nulls = nullptr;
                    if (src.nulls) { out.bufs.emplace_back(ctx, m); oc.nulls = (uint8_t*)out.bufs.back().p; }
                }
                if (ct.count == PROJ_MAX_OUTPUTS) flush();
                const int c = ct.count++;
                ct.src[c] = src.values; ct.src_nulls[c] = src.nulls; ct.dst[c] = (void*)oc.values; ct.dst_nulls[c] = (uint8_t*)oc.nulls; ct.width[c] = type_width(e.type);
            }
            flush();
            }
            int64_t e = 0;
            read_scalars(ctx, err.p, 1, &e);
            if (e & 1) fail(PG_E_NUMERIC_OVERFLOW, "Decimal overflow / numeric value out of range in a filter or projection expression");
            return;
        }
        sel.alloc(ctx, (m > 0 ? m : 1) * 8);
        if (m > 0 && m < n) PG_LAUNCH(ctx, proj_positions_kernel, (int)ntiles, SCAN_THREADS, 0, flags.as<uint8_t>(), n, tiles.as<Pair64>(), sel.as<int64_t>());
    }
    const bool identity = m == n;     // every row selected: pass-through projections alias the input columns
    out.n = m;
    ProjOut po; memset(&po, 0, sizeof(po));
    for (int k = 0; k < np && m > 0; k++) {
        if (want && !(*want)[k]) continue;
        const PNode& e = p->nodes[p->proj[k]];
        DCol& oc = out.cols[k];
        // `direct`: the caller's own device buffers; fixed-width columns are produced there instead of in a buffer of ours
        const bool into = direct && direct[k].mem == PG_MEM_DEVICE && direct[k].values && direct[k].type == e.type && !type_is_var(e.type) &&
                          direct[k].capacity >= m;
        if (e.op == PG_EXPR_INPUT) {
            if (identity) { oc = in[e.channel]; continue; }
            if (into) {
                gather_device_col_into(ctx, in[e.channel], sel.as<int64_t>(), m, direct[k].values, direct[k].nulls);
                oc.type = e.type; oc.values = direct[k].values; oc.nulls = direct[k].nulls; oc.offsets = nullptr;
                out.in_place[k] = 1;
                continue;
            }
            out.bufs.emplace_back(); out.bufs.emplace_back(); out.bufs.emplace_back();
            const size_t b = out.bufs.size();
            gather_device_col(ctx, in[e.channel], sel.as<int64_t>(), m, out.bufs[b - 3], out.bufs[b - 2], out.bufs[b - 1], oc);
            continue;
        }
        if (po.count == PROJ_MAX_OUTPUTS) {     // flush a full batch
            PG_LAUNCH(ctx, proj_eval_kernel, grid_for(ctx, m, 128), 128, 0, d_nodes, n_nodes, p->filter >= 0 ? 1 : 0, pc, d_bytes, identity ? nullptr : sel.as<int64_t>(), m, po,
                      err.as<unsigned int>());
            po.count = 0;
        }
        void* v; uint8_t* nl;
        if (into && direct[k].nulls) { v = direct[k].values; nl = direct[k].nulls; out.in_place[k] = 1; }
        else {
            out.bufs.emplace_back(ctx, m * (int64_t)type_width(e.type));
            v = out.bufs.back().p;
            out.bufs.emplace_back(ctx, m);
            nl = (uint8_t*)out.bufs.back().p;
        }
        oc.type = e.type; oc.values = v; oc.nulls = nl; oc.offsets = nullptr;
        po.values[po.count] = v; po.nulls[po.count] = nl; po.node[po.count] = p->proj[k]; po.count++;
    }
    if (po.count > 0)
        PG_LAUNCH(ctx, proj_eval_kernel, grid_for(ctx, m, 128), 128, 0, d_nodes, n_nodes, p->filter >= 0 ? 1 : 0, pc, d_bytes, identity ? nullptr : sel.as<int64_t>(), m, po,
                  err.as<unsigned int>());
    int64_t e = 0;
    read_scalars(ctx, err.p, 1, &e);
    if (e & 1) fail(PG_E_NUMERIC_OVERFLOW, "Decimal overflow / numeric value out of range in a filter or projection expression");
    // computed columns whose inputs carry no nulls cannot be null: drop the mask so that downstream kernels skip it
    for (int k = 0; k < np; k++) {
        if (!out.cols[k].values || p->nodes[p->proj[k]].op == PG_EXPR_INPUT) continue;
        bool nullable = false;
        std::vector<int> stack{p->proj[k]};
        while (!stack.empty() && !nullable) {
            const PNode& q = p->nodes[stack.back()]; stack.pop_back();
            if (q.op == PG_EXPR_INPUT) nullable = in[q.channel].nulls != nullptr;
            else if (q.op == PG_EXPR_CONST) nullable = q.is_null != 0;
            else if (q.op != PG_EXPR_IS_NULL) { if (q.a0 >= 0) stack.push_back(q.a0); if (q.a1 >= 0) stack.push_back(q.a1); }
        }
        if (!nullable && !out.in_place[k]) out.cols[k].nulls = nullptr;      // (a caller's mask was written by the kernel: all zero)
    }
}

// ------------------------------------------------------------------ code generation
static std::string u64lit(unsigned long long v) { char b[40]; snprintf(b, sizeof(b), "0x%llxULL", v); return b; }
static std::string i64lit(long long v) {
    if (v == (long long)0x8000000000000000ULL) return "(-0x7fffffffffffffffLL - 1)";
    char b[40]; snprintf(b, sizeof(b), "%lldLL", v); return b;
}

bool project_codegen(const pg_project* p, const std::vector<int>& roots, const std::vector<std::string>& in_lo, const std::vector<std::string>& in_hi,
                     const std::vector<std::string>& in_null, ProjCode& out) {
    ensure_pow10();
    const int n = (int)p->nodes.size();
    std::vector<char> needed(n, 0);
    for (int r : roots) if (r >= 0 && r < n) needed[r] = 1;
    if (p->filter >= 0) needed[p->filter] = 1;
    for (int i = n - 1; i >= 0; i--)
        if (needed[i]) { if (p->nodes[i].a0 >= 0) needed[p->nodes[i].a0] = 1; if (p->nodes[i].a1 >= 0) needed[p->nodes[i].a1] = 1; }
    out.lo.assign(n, ""); out.hi.assign(n, ""); out.null.assign(n, "false");
    out.decl.clear(); out.can_overflow = false;
    char b[512];
    auto var = [&](int i, const char* suf) { snprintf(b, sizeof(b), "x%d%s", i, suf); return std::string(b); };
    // a long decimal lives as Pgx2c (two's complement); this gives the Pgx2c expression of any decimal operand
    auto as2c = [&](int a) -> std::string {
        if (p->nodes[a].type == PG_LONG_DECIMAL) return var(a, "");
        return "pgx2c_from_short((pgx_i64)(" + out.lo[a] + "))";
    };
    auto null_or = [&](int a, int c) -> std::string {
        const std::string &x = out.null[a], &y = out.null[c];
        if (x == "false") return y;
        if (y == "false") return x;
        return "(" + x + " || " + y + ")";
    };
    for (int phase = 1; phase >= 0; phase--) {
        for (int i = 0; i < n; i++) {
            const PNode& e = p->nodes[i];
            if (!needed[i] || e.in_filter != phase) continue;
            std::string& D = out.decl;
            switch (e.op) {
                case PG_EXPR_INPUT:
                    if (type_is_var(e.type)) return false;          // strings are not handled by the generated code
                    if (in_lo[e.channel].empty()) return false;
                    out.null[i] = in_null[e.channel].empty() ? "false" : in_null[e.channel];
                    if (e.type == PG_LONG_DECIMAL) {
                        D += "const Pgx2c " + var(i, "") + " = pgx2c_from_long(" + in_lo[e.channel] + ", " + in_hi[e.channel] + ");\n";
                        out.lo[i] = var(i, "") + ".w0"; out.hi[i] = var(i, "") + ".w1";
                    } else out.lo[i] = in_lo[e.channel];
                    break;
                case PG_EXPR_CONST:
                    if (type_is_var(e.type)) return false;
                    out.null[i] = e.is_null ? "true" : "false";
                    if (e.type == PG_LONG_DECIMAL) {
                        unsigned long long lo = (unsigned long long)e.lo, hi = (unsigned long long)e.hi & 0x7fffffffffffffffULL;
                        if (e.hi < 0) { lo = 0ULL - lo; hi = ~hi + (lo == 0 ? 1ULL : 0ULL); }
                        D += "Pgx2c " + var(i, "") + "; " + var(i, "") + ".w0 = " + u64lit(lo) + "; " + var(i, "") + ".w1 = (pgx_i64)" + u64lit(hi) + ";\n";
                        out.lo[i] = var(i, "") + ".w0"; out.hi[i] = var(i, "") + ".w1";
                    } else out.lo[i] = "((pgx_u64)" + i64lit(e.lo) + ")";
                    break;
                case PG_EXPR_ADD: case PG_EXPR_SUB: case PG_EXPR_MUL: {
                    const PNode &qa = p->nodes[e.a0], &qb = p->nodes[e.a1];
                    out.null[i] = null_or(e.a0, e.a1);
                    const char* sym = e.op == PG_EXPR_ADD ? "+" : e.op == PG_EXPR_SUB ? "-" : "*";
                    if (e.type == PG_SHORT_DECIMAL) {
                        std::string x = "(pgx_i64)(" + out.lo[e.a0] + ")", y = "(pgx_i64)(" + out.lo[e.a1] + ")";
                        if (e.op != PG_EXPR_MUL) {
                            if (e.up0) x = "(" + x + " * " + i64lit((long long)h_p10[e.up0][0]) + ")";
                            if (e.up1) y = "(" + y + " * " + i64lit((long long)h_p10[e.up1][0]) + ")";
                        }
                        D += "const pgx_u64 " + var(i, "") + " = (pgx_u64)(" + x + " " + sym + " " + y + ");\n";
                        out.lo[i] = var(i, "");
                    } else if (e.type == PG_LONG_DECIMAL) {
                        const bool la = qa.type == PG_LONG_DECIMAL, lb = qb.type == PG_LONG_DECIMAL;
                        bool checks = true;
                        D += "int " + var(i, "e") + " = 0; ";
                        if (e.op == PG_EXPR_MUL) {
                            if (!la && !lb) { D += "const Pgx2c " + var(i, "") + " = pgx2c_mul_ss((pgx_i64)(" + out.lo[e.a0] + "), (pgx_i64)(" + out.lo[e.a1] + "));\n"; checks = false; }
                            else if (la && !lb) D += "const Pgx2c " + var(i, "") + " = pgx2c_mul_ls(" + var(e.a0, "") + ", (pgx_i64)(" + out.lo[e.a1] + "), " + var(i, "e") + ");\n";
                            else if (!la && lb) D += "const Pgx2c " + var(i, "") + " = pgx2c_mul_ls(" + var(e.a1, "") + ", (pgx_i64)(" + out.lo[e.a0] + "), " + var(i, "e") + ");\n";
                            else return false;       // long x long: interpreter only
                        } else {
                            if (e.up0 > 18 || e.up1 > 18) return false;
                            std::string x = as2c(e.a0), y = as2c(e.a1);
                            if (e.up0) x = "pgx2c_mul_ls(" + x + ", " + i64lit((long long)h_p10[e.up0][0]) + ", " + var(i, "e") + ")";
                            if (e.up1) y = "pgx2c_mul_ls(" + y + ", " + i64lit((long long)h_p10[e.up1][0]) + ", " + var(i, "e") + ")";
                            if (e.op == PG_EXPR_SUB) y = "pgx2c_neg(" + y + ")";
                            D += "const Pgx2c " + var(i, "") + " = pgx2c_add(" + x + ", " + y + ", " + var(i, "e") + ");\n";
                        }
                        if (checks) {     // an operand that is null makes the result null, not an error
                            if (out.null[i] == "false") D += "pgx_err |= " + var(i, "e") + ";\n";
                            else D += "if (!" + out.null[i] + ") pgx_err |= " + var(i, "e") + ";\n";
                            out.can_overflow = true;
                        }
                        out.lo[i] = var(i, "") + ".w0"; out.hi[i] = var(i, "") + ".w1";
                    } else if (e.type == PG_DOUBLE) {
                        D += "const pgx_u64 " + var(i, "") + " = (pgx_u64)__double_as_longlong(__longlong_as_double((pgx_i64)(" + out.lo[e.a0] + ")) " + sym +
                             " __longlong_as_double((pgx_i64)(" + out.lo[e.a1] + ")));\n";
                        out.lo[i] = var(i, "");
                    } else return false;      // checked bigint / integer arithmetic: interpreter only
                    break;
                }
                case PG_EXPR_LT: case PG_EXPR_LE: case PG_EXPR_GT: case PG_EXPR_GE: case PG_EXPR_EQ: case PG_EXPR_NE: {
                    const PNode &qa = p->nodes[e.a0], &qb = p->nodes[e.a1];
                    if (type_is_var(qa.type)) return false;
                    const char* sym = e.op == PG_EXPR_LT ? "<" : e.op == PG_EXPR_LE ? "<=" : e.op == PG_EXPR_GT ? ">" : e.op == PG_EXPR_GE ? ">=" : e.op == PG_EXPR_EQ ? "==" : "!=";
                    out.null[i] = null_or(e.a0, e.a1);
                    std::string v;
                    if (qa.type == PG_DOUBLE) v = "(__longlong_as_double((pgx_i64)(" + out.lo[e.a0] + ")) " + sym + " __longlong_as_double((pgx_i64)(" + out.lo[e.a1] + ")))";
                    else if (qa.type == PG_LONG_DECIMAL || qb.type == PG_LONG_DECIMAL) v = "(pgx2c_compare(" + as2c(e.a0) + ", " + as2c(e.a1) + ") " + sym + " 0)";
                    else v = "((pgx_i64)(" + out.lo[e.a0] + ") " + sym + " (pgx_i64)(" + out.lo[e.a1] + "))";
                    D += "const pgx_u64 " + var(i, "") + " = " + v + " ? 1ULL : 0ULL;\n";
                    out.lo[i] = var(i, "");
                    break;
                }
                case PG_EXPR_AND: case PG_EXPR_OR: {
                    const std::string &an = out.null[e.a0], &bn = out.null[e.a1], &av = out.lo[e.a0], &bv = out.lo[e.a1];
                    const bool is_and = e.op == PG_EXPR_AND;
                    // decided: one operand is a non-null FALSE (AND) / TRUE (OR)
                    const std::string dec = std::string("((!") + an + " && " + (is_and ? "!" : "") + av + ") || (!" + bn + " && " + (is_and ? "!" : "") + bv + "))";
                    D += "const bool " + var(i, "d") + " = " + dec + ";\n";
                    D += "const bool " + var(i, "n") + " = !" + var(i, "d") + " && (" + an + " || " + bn + ");\n";
                    D += "const pgx_u64 " + var(i, "") + " = " + var(i, "d") + " ? " + (is_and ? "0ULL" : "1ULL") + " : " + (is_and ? "1ULL" : "0ULL") + ";\n";
                    out.lo[i] = var(i, ""); out.null[i] = (an == "false" && bn == "false") ? "false" : var(i, "n");
                    break;
                }
                case PG_EXPR_NOT:
                    D += "const pgx_u64 " + var(i, "") + " = " + out.lo[e.a0] + " ? 0ULL : 1ULL;\n";
                    out.lo[i] = var(i, ""); out.null[i] = out.null[e.a0];
                    break;
                case PG_EXPR_IS_NULL:
                    D += "const pgx_u64 " + var(i, "") + " = " + out.null[e.a0] + " ? 1ULL : 0ULL;\n";
                    out.lo[i] = var(i, "");
                    break;
                default: return false;          // casts and hashes: interpreter only
            }
        }
        if (phase == 1) out.decl += "PGX_FILTER_DONE\n";
    }
    return true;
}

}  // namespace pg

using namespace pg;

extern "C" {

int32_t pg_project_create(pg_ctx* ctx, const pg_expr* nodes, int32_t n_nodes, int32_t filter, const int32_t* projections, int32_t n_projections,
                          pg_project** out) {
    if (!ctx || !out) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        pg_project* p = new pg_project();
        p->ctx = ctx;
        try { project_build(p, nodes, n_nodes, filter, projections, n_projections); } catch (...) { delete p; throw; }
        *out = p;
    });
}

int32_t pg_project_destroy(pg_project* p) {
    if (!p) return PG_OK;
    return guard(p->ctx, [&] { delete p; });
}

int32_t pg_project_output_types(pg_project* p, int32_t* types, int32_t* precisions, int32_t* scales, int32_t capacity, int32_t* ncols) {
    if (!p || !ncols) return PG_E_INVALID_ARGUMENT;
    return guard(p->ctx, [&] {
        *ncols = (int32_t)p->proj.size();
        if (!types) return;
        PG_CHECK_ARG(capacity >= *ncols, "types buffer too small");
        for (size_t k = 0; k < p->proj.size(); k++) {
            const PNode& e = p->nodes[p->proj[k]];
            types[k] = e.type;
            if (precisions) precisions[k] = e.precision;
            if (scales) scales[k] = e.scale;
        }
    });
}

int32_t pg_project_process(pg_project* p, const pg_col* cols, int32_t ncols, int64_t n, pg_col_out* out, int64_t* n_out) {
    if (!p || !n_out) return PG_E_INVALID_ARGUMENT;
    return guard(p->ctx, [&] {
        pg_ctx* ctx = p->ctx;
        PG_CHECK_ARG(cols && ncols > 0 && n >= 0 && (p->proj.empty() || out), "bad arguments");
        *n_out = 0;
        if (n == 0) return;
        std::vector<ColImport> imp(ncols);
        std::vector<DCol> in(ncols, DCol{0, 0, nullptr, nullptr, nullptr});
        for (int c = 0; c < ncols && c < PROJ_MAX_CHANNELS; c++) {
            if (!p->channel_used[c]) continue;
            PG_CHECK_ARG(cols[c].n == n, "channel %d has %lld rows, expected %lld", c, (long long)cols[c].n, (long long)n);
            import_col(ctx, cols[c], n, imp[c]);
            in[c] = imp[c].d;
        }
        ProjectedPage page;
        project_materialize(p, in.data(), ncols, n, nullptr, nullptr, page, out);
        const int64_t m = page.n;
        *n_out = m;
        // copy the projected columns into the caller's buffers (those not produced there already)
        for (size_t k = 0; k < p->proj.size() && m > 0; k++) {
            if (page.in_place[k]) continue;
            const DCol& c = page.cols[k];
            pg_col_out& o = out[k];
            PG_CHECK_ARG(o.type == c.type && o.capacity >= m, "output column %d: type %d / capacity %lld do not fit type %d / %lld rows", (int)k, o.type,
                         (long long)o.capacity, c.type, (long long)m);
            const cudaMemcpyKind kind = o.mem == PG_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
            if (type_is_var(c.type)) {
                PG_CHECK_ARG(o.offsets != nullptr, "output column %d needs an offsets buffer", (int)k);
                int32_t last = 0;
                flush_staged_inputs(ctx);
                PG_CUDA(cudaMemcpyAsync(&last, c.offsets + m, 4, cudaMemcpyDeviceToHost, ctx->stream));
                sync(ctx);
                PG_CHECK_ARG(o.values_capacity_bytes >= last, "output column %d needs %d value bytes", (int)k, last);
                // a pass-through over every row keeps the input's offsets, which need not start at zero: rebase on the host side of the copy
                copy_async(ctx, o.values, c.values, last, kind);
                copy_async(ctx, o.offsets, c.offsets, (m + 1) * 4, kind);
            } else copy_async(ctx, o.values, c.values, m * (int64_t)type_width(c.type), kind);
            if (o.nulls) {
                if (c.nulls) copy_async(ctx, o.nulls, c.nulls, m, kind);
                else if (o.mem == PG_MEM_DEVICE) memset_async(ctx, o.nulls, 0, m);
                else memset(o.nulls, 0, (size_t)m);
            }
        }
        sync(ctx);
    });
}

}  // extern "C"
