// hashagg.cu — the fused GROUP BY operator: HashAggregationOperator + InMemoryHashAggregationBuilder.
// Replaces OP/HashAggregationOperator.java:381-518 (addInput/getOutput) and
// OP/aggregation/builder/InMemoryHashAggregationBuilder.java:139-155 (processPage), :243-298 (buildResult).
// One page = getGroupIds (groupby.cu) -> one accumulate() launch over ALL aggregates (accum.cu); group ids never
// leave the device.
#include <stdlib.h>

#include "accum.h"
#include "groupby.h"
#include "hashagg_wide.h"
#include "jit.h"
#include "project.h"
#include "jit/lowcard_args.h"

struct pg_hashagg {
    pg_ctx* ctx = nullptr;
    int32_t step = 0;
    std::vector<int32_t> key_channels;
    std::vector<int32_t> key_types;
    int32_t hash_channel = -1;
    std::vector<pg_agg_spec> specs;
    int64_t expected = 0;
    pg_groupby* gb = nullptr;
    std::vector<pg_agg*> aggs;
    pg::DevBuf ids;
    int64_t input_rows = 0;
    pg_project* proc = nullptr;      // fused filter + projections in front of the aggregation (pg_hashagg_set_input_processor); not owned
    // few groups: the groups as packed keys in id order (jit/lowcard_args.h: 4 words each).  While gb_current is false the groups
    // exist only here (discovered by the specialised kernel) and the GroupByHash is brought up to date on demand (hashagg_sync_gb).
    struct { pg::DevBuf groups; int64_t count = 0; bool gb_current = true; } lc;
    pg::WideTable* wide = nullptr;   // many-groups path (hashagg_wide.h), chosen on the first page
    int64_t rows_since_reset = 0;    // first-seen order of the wide path is kept by operator-wide row numbers
    // page accumulator (pg_hashagg_set_page_batching): small host pages are appended to pinned buffers and enter the kernels as one
    // page of `batch_rows` rows; anything that looks at the operator's state flushes first
    int64_t batch_rows = 0;
    struct PageBatch* batch = nullptr;
    ~pg_hashagg();
};

// A piece of a caller-pinned page and where it goes in the device-side accumulator.  The pieces of a whole batch are copied by ONE
// kernel that reads the host memory directly (page-locked memory is mapped into the device's address space): one cudaMemcpyAsync per
// page and column made 16 driver threads queue up on the driver's lock (51 000 calls per SF10 pass: 254 ms against 51 ms for one bulk copy).
struct GatherDesc { const char* src; char* dst; int64_t bytes; };
static __global__ void __launch_bounds__(256) batch_gather_kernel(const GatherDesc* __restrict__ descs, int ndesc) {
    for (int d = blockIdx.y; d < ndesc; d += gridDim.y) {
        const GatherDesc g = descs[d];
        const uintptr_t mis = (uintptr_t)g.src | (uintptr_t)g.dst | (uintptr_t)g.bytes;
        const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
        if ((mis & 15) == 0) { for (int64_t i = t; i < g.bytes / 16; i += stride) ((uint4*)g.dst)[i] = ((const uint4*)g.src)[i]; }
        else if ((mis & 7) == 0) { for (int64_t i = t; i < g.bytes / 8; i += stride) ((unsigned long long*)g.dst)[i] = ((const unsigned long long*)g.src)[i]; }
        else if ((mis & 3) == 0) { for (int64_t i = t; i < g.bytes / 4; i += stride) ((unsigned int*)g.dst)[i] = ((const unsigned int*)g.src)[i]; }
        else { for (int64_t i = t; i < g.bytes; i += stride) g.dst[i] = g.src[i]; }
    }
}

// One pinned host copy of the channels the operator reads, `cap_rows` rows deep.  Kept by the context between operators
// (pg_ctx::page_batch_pool): pinning tens of megabytes costs milliseconds.
struct PageBatch {
    struct Col {
        int32_t type = 0;
        bool used = false, has_nulls = false;
        char* values = nullptr; int64_t values_cap = 0, values_used = 0;
        uint8_t* nulls = nullptr;
        int32_t* offsets = nullptr;
        pg::DevBuf dvalues, dnulls, doffsets;      // device mode (pages in caller-pinned memory are copied straight to the device)
    };
    int64_t cap_rows = 0, rows = 0;
    bool device_mode = false;                      // what the rows accumulated so far are: copies in pinned host memory, or device copies
    std::vector<GatherDesc> descs;                 // device mode: the copies the next flush has to make
    std::vector<Col> cols;
    static void* pinned(int64_t bytes) {
        void* p = nullptr;
        if (cudaMallocHost(&p, (size_t)(bytes > 0 ? bytes : 8)) != cudaSuccess) { cudaGetLastError(); pg::fail(PG_E_INSUFFICIENT_RESOURCES, "pinned allocation of %lld bytes failed", (long long)bytes); }
        return p;
    }
    void release() {
        for (Col& c : cols) { if (c.values) cudaFreeHost(c.values); if (c.nulls) cudaFreeHost(c.nulls); if (c.offsets) cudaFreeHost(c.offsets); }
        cols.clear(); descs.clear(); cap_rows = 0; rows = 0; device_mode = false;
    }
    ~PageBatch() { release(); }
};

pg_hashagg::~pg_hashagg() {
    if (batch) { batch->rows = 0; batch->descs.clear(); ctx->page_batch_pool.push_back(batch); batch = nullptr; }
    {
        delete wide;
        for (pg_agg* a : aggs) delete a;
        delete gb;
    }
}

namespace pg {

void page_batch_pool_free(pg_ctx* ctx) {
    for (void* b : ctx->page_batch_pool) delete (PageBatch*)b;
    ctx->page_batch_pool.clear();
}

static DCol dcol_offset(const DCol& c, int64_t off) {
    DCol r = c;
    if (off == 0) return r;
    if (r.nulls) r.nulls += off;
    if (type_is_var(r.type)) r.offsets += off;
    else r.values = (const char*)r.values + off * type_width(r.type);
    return r;
}

static bool step_input_raw(int s) { return s == PG_STEP_SINGLE || s == PG_STEP_PARTIAL; }
static bool step_output_partial(int s) { return s == PG_STEP_PARTIAL || s == PG_STEP_INTERMEDIATE; }


static __global__ void expand_bits_kernel(const unsigned int* __restrict__ bits, int64_t n, uint8_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = (bits[i >> 5] >> (i & 31)) & 1u;
}

void hashagg_build(pg_hashagg* op) {
    groupby_create(op->ctx, op->key_types.data(), (int32_t)op->key_types.size(), op->hash_channel >= 0 ? 1 : 0, op->expected, &op->gb, true);
    for (const pg_agg_spec& s : op->specs) {
        pg_agg* a = nullptr;
        agg_create(op->ctx, s.kind, s.in_type, &a);
        op->aggs.push_back(a);
    }
}

static void hashagg_add_page(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n);

static int64_t hashagg_group_count(const pg_hashagg* op) { return op->lc.gb_current ? op->gb->groups : op->lc.count; }

// key columns of the packed groups [first, first + n) in fresh device buffers
struct UnpackedKeys { KeyCols kc; std::vector<DevBuf> bufs; };
static void lowcard_unpack_to_cols(pg_hashagg* op, int64_t first, int64_t n, UnpackedKeys& u) {
    pg_ctx* ctx = op->ctx;
    const KeyLayout& L = op->gb->layout;
    memset(&u.kc, 0, sizeof(u.kc));
    LowcardKeyOut out;
    memset(&out, 0, sizeof(out));
    u.bufs.reserve(3 * L.nkeys);
    for (int k = 0; k < L.nkeys; k++) {
        const bool var = type_is_var(L.types[k]);
        u.bufs.emplace_back(ctx, var ? 3 * n + 8 : n * (int64_t)type_width(L.types[k]) + 8);
        out.values[k] = u.bufs.back().p;
        u.bufs.emplace_back(ctx, n + 8);
        out.nulls[k] = (uint8_t*)u.bufs.back().p;
        u.bufs.emplace_back(ctx, (n + 1) * 4 + 8);
        out.offsets[k] = (int32_t*)u.bufs.back().p;
        DCol& c = u.kc.c[k];
        c.type = L.types[k]; c.values = out.values[k]; c.nulls = out.nulls[k]; c.offsets = var ? out.offsets[k] : nullptr;
    }
    lowcard_unpack_groups(ctx, L, op->lc.groups.as<unsigned long long>(), first, n, out);
}

// the GroupByHash learns the groups the specialised kernel discovered (same ids: they are inserted in id order)
static void hashagg_sync_gb(pg_hashagg* op) {
    if (op->lc.gb_current) return;
    op->lc.gb_current = true;
    if (op->lc.count == 0) return;
    PG_CHECK_STATE(op->gb->groups == 0, "group tables out of step");
    UnpackedKeys u;
    lowcard_unpack_to_cols(op, 0, op->lc.count, u);
    groupby_process(op->gb, u.kc, nullptr, op->lc.count, nullptr);
    PG_CHECK_STATE(op->gb->groups == op->lc.count, "the GroupByHash numbered %lld groups, the kernel %lld", (long long)op->gb->groups, (long long)op->lc.count);
}

// The few-groups fast path: one specialised kernel over all rows of the page (jit_lowcard.cu).  Groups: the operator's known
// ones, or — first page — discovered by the kernel itself in first-seen order; rows with a key the kernel does not know are
// handed to on_misses(bitmap) after the GroupByHash has been brought up to date.  Returns false when the plan does not fit or
// the discovery gave up; nothing has been accumulated then.
template <typename MissFn>
static bool hashagg_lowcard(pg_hashagg* op, const KeyCols& kc, ColImport** kimp, int nk, const AggProgram& prog, const LowcardExprs* px, int64_t n,
                            MissFn&& on_misses) {
    pg_ctx* ctx = op->ctx;
    if (!jit_available()) return false;
    const int64_t known = hashagg_group_count(op);
    if (known > 64) return false;
    const bool discover = known == 0;
    if (discover) {
        // the packed form must reproduce the stored key exactly (doubles are normalised in it) and there is no raw hash to keep
        if (op->hash_channel >= 0) return false;
        for (int t : op->key_types) if (t == PG_DOUBLE) return false;
        if (const char* e = getenv("PRESTOGPU_LOWCARD_DISCOVER")) if (e[0] == '0') return false;
    }
    if (!op->lc.groups.p) op->lc.groups.alloc(ctx, 64 * 4 * 8);
    if (!discover && op->lc.gb_current && op->lc.count != op->gb->groups) {      // the GroupByHash has grown since the table was packed
        lowcard_pack_groups(ctx, op->gb, op->gb->groups, op->lc.groups.as<unsigned long long>());
        op->lc.count = op->gb->groups;
    }
    for (pg_agg* a : op->aggs) agg_ensure_capacity(a, discover ? 64 : known);
    const int64_t miss_words = (n + 31) / 32;
    DevBuf miss_flag(ctx, 32 + miss_words * 4);            // [misses][error bits][groups discovered][discovery verdict][bitmap]
    memset_async(ctx, miss_flag.p, 0, 32 + miss_words * 4);
    unsigned long long* d_miss = miss_flag.as<unsigned long long>();
    unsigned int* d_miss_bits = (unsigned int*)(d_miss + 4);
    FusedLookup fl;
    memset(&fl, 0, sizeof(fl));
    fl.layout = &op->gb->layout; fl.keys = &kc; fl.miss_flag = d_miss_bits; fl.d_miss = d_miss; fl.row_base = 0;
    for (int k = 0; k < nk && k < 8; k++) fl.key_value_bytes[k] = kimp[k]->value_bytes;     // -1: the kernel reads offsets[n]
    LowcardGroups lg;
    lg.known = discover ? nullptr : op->lc.groups.as<unsigned long long>();
    lg.known_count = discover ? 0 : (int)known;
    lg.discovered_out = op->lc.groups.as<unsigned long long>();
    lg.discover_rows = 8192;
    lg.capacity_hint = ctx->jit ? ctx->jit->lowcard_groups_hint : 0;
    if (!accumulate_lowcard_jit(ctx, prog, n, fl, lg, px)) return false;
    int64_t res[4];
    read_scalars(ctx, d_miss, 4, res);
    if (discover) {
        const int64_t found = res[2], verdict = res[3];
        if (ctx->jit && found > 0 && found <= 64) ctx->jit->lowcard_groups_hint = (int)found;
        if (verdict == PGJ_DISC_TOO_MANY && found <= 64) {
            // the table the kernel published is complete: come back with the groups known (block shape chosen for them)
            memset_async(ctx, miss_flag.p, 0, 32);
            lg.known = op->lc.groups.as<unsigned long long>(); lg.known_count = (int)found;
            if (!accumulate_lowcard_jit(ctx, prog, n, fl, lg, px)) return false;
            read_scalars(ctx, d_miss, 4, res);
        } else if (verdict != 0 || found == 0) return false;
        op->lc.count = found;
        op->lc.gb_current = false;
    }
    if (res[1] & 1) fail(PG_E_NUMERIC_OVERFLOW, "Decimal overflow / numeric value out of range in a filter or projection expression");
    if (res[0] > 0) {
        hashagg_sync_gb(op);
        on_misses(d_miss_bits);
    }
    return true;
}

// channels of the operator's logical input page (= the processor's projections) that the operator reads
static std::vector<char> channels_read(const pg_hashagg* op, int np) {
    std::vector<char> want(np, 0);
    auto mark = [&](int ch) { PG_CHECK_ARG(ch >= 0 && ch < np, "channel %d out of range (the input processor has %d projections)", ch, np); want[ch] = 1; };
    for (int ch : op->key_channels) mark(ch);
    if (op->hash_channel >= 0) mark(op->hash_channel);
    for (size_t i = 0; i < op->specs.size(); i++) {
        const pg_agg_spec& s = op->specs[i];
        if (s.kind != PG_AGG_COUNT) {
            mark(s.input_channel);
            if (!(op->step == PG_STEP_SINGLE || op->step == PG_STEP_PARTIAL) && agg_n_intermediate(op->aggs[i]) > 1) mark(s.input_channel + 1);
        }
        if (s.mask_channel >= 0) mark(s.mask_channel);
    }
    return want;
}

// The general path of a fused processor: materialise the projections over the selected rows (only_bits restricts the rows
// further), then hand the projected device page to the plain operator.
static void hashagg_add_projected(pg_hashagg* op, const DCol* in, int ncols, int64_t n, const unsigned int* only_bits) {
    pg_project* proc = op->proc;
    const int np = (int)proc->proj.size();
    std::vector<char> want = channels_read(op, np);
    ProjectedPage page;
    project_materialize(proc, in, ncols, n, only_bits, &want, page);
    if (page.n == 0) return;
    std::vector<pg_col> pc(np);
    for (int k = 0; k < np; k++) {
        pg_col& c = pc[k];
        memset(&c, 0, sizeof(c));
        const PNode& e = proc->nodes[proc->proj[k]];
        c.type = e.type; c.mem = PG_MEM_DEVICE; c.n = page.n; c.precision = e.precision; c.scale = e.scale;
        c.values = page.cols[k].values; c.nulls = page.cols[k].nulls; c.offsets = page.cols[k].offsets;
    }
    op->proc = nullptr;
    try { hashagg_add_page(op, pc.data(), np, page.n); } catch (...) { op->proc = proc; throw; }
    op->proc = proc;
}

// Fast path of a fused processor with few groups: the filter and the projections run inside the specialised aggregation kernel
// over the raw input columns (derived columns are never written).  Group numbering stays the reference's: the rows of a short
// prefix that pass the filter are projected and go through the GroupByHash alone (first-seen ids); the kernel then handles all
// rows of the page; rows that pass the filter but carry a key the prefix did not show are projected and take the general path
// in page order, so later groups are numbered after the prefix's, as they would be row by row.
// Returns false when the plan does not fit (nothing has been accumulated then; groups found so far keep their ids).
static bool hashagg_try_fused_processor(pg_hashagg* op, const DCol* in, ColImport* imports, int ncols, int64_t n) {
    pg_ctx* ctx = op->ctx;
    pg_project* proc = op->proc;
    const int64_t PREFIX = 16 * 1024;
    if (!step_input_raw(op->step) || op->aggs.empty() || n < 8 * PREFIX || op->hash_channel >= 0 || op->wide) return false;
    if (!jit_available() || !fused_keys_packable(op->gb->layout)) return false;
    if (const char* e = getenv("PRESTOGPU_FUSED_PROJECT")) if (e[0] == '0') return false;
    const int np = (int)proc->proj.size();
    const int nk = (int)op->key_channels.size();
    KeyCols kc;
    memset(&kc, 0, sizeof(kc));
    ColImport* kimp[MAX_KEYS];
    for (int k = 0; k < nk; k++) {                         // group-by channels must be passed through unchanged
        const PNode& e = proc->nodes[proc->proj[op->key_channels[k]]];
        if (e.op != PG_EXPR_INPUT || e.type != op->key_types[k]) return false;
        kc.c[k] = in[e.channel];
        kimp[k] = &imports[e.channel];
    }
    LowcardExprs px;
    px.proc = proc; px.in = in;
    for (size_t i = 0; i < op->aggs.size(); i++) {
        const pg_agg_spec& s = op->specs[i];
        px.arg_proj[i] = s.kind != PG_AGG_COUNT ? s.input_channel : -1;
        px.mask_proj[i] = s.mask_channel;
        if (px.arg_proj[i] >= np || px.mask_proj[i] >= np) return false;
    }
    AggProgram prog;
    memset(&prog, 0, sizeof(prog));
    prog.naggs = (int32_t)op->aggs.size();
    // state arrays move when they grow: the program is built after the capacity is final (64 groups while discovering)
    auto fill_prog = [&] {
        const int64_t g = hashagg_group_count(op);
        for (pg_agg* a : op->aggs) agg_ensure_capacity(a, g > 0 ? g : 64);
        for (size_t i = 0; i < op->aggs.size(); i++) prog.ops[i] = agg_make_op(op->aggs[i], nullptr, nullptr);
    };
    auto on_misses = [&](const unsigned int* bits) {
        const int64_t rows = op->rows_since_reset;
        hashagg_add_projected(op, in, ncols, n, bits);
        op->rows_since_reset = rows;
    };
    fill_prog();
    if (hashagg_lowcard(op, kc, kimp, nk, prog, &px, n, on_misses)) { op->rows_since_reset += n; return true; }
    if (hashagg_group_count(op) > 0) return false;
    // The kernel's own discovery gave up (e.g. a long string among the keys): number the groups of a projected prefix through the
    // GroupByHash, as the plain operator does, and try once more with them known.
    {
        std::vector<char> want(np, 0);
        for (int k = 0; k < nk; k++) want[op->key_channels[k]] = 1;
        ProjectedPage pp;
        project_materialize(proc, in, ncols, PREFIX, nullptr, &want, pp);
        if (pp.n > 0) {
            KeyCols pk;
            memset(&pk, 0, sizeof(pk));
            for (int k = 0; k < nk; k++) pk.c[k] = pp.cols[op->key_channels[k]];
            groupby_process(op->gb, pk, nullptr, pp.n, nullptr);
        }
    }
    const int64_t groups = op->gb->groups;
    if (groups < 1 || groups > 64) return false;
    fill_prog();
    if (!hashagg_lowcard(op, kc, kimp, nk, prog, &px, n, on_misses)) return false;
    op->rows_since_reset += n;
    return true;
}

static void hashagg_add_page(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n) {
    pg_ctx* ctx = op->ctx;
    if (n == 0) return;
    if (op->proc) {
        pg_project* proc = op->proc;
        PG_CHECK_ARG(proc->max_channel < ncols, "the page has %d channels, the input processor reads channel %d", ncols, proc->max_channel);
        std::vector<ColImport> pimp(ncols);
        std::vector<DCol> in(ncols, DCol{0, 0, nullptr, nullptr, nullptr});
        for (int c = 0; c < ncols && c < PROJ_MAX_CHANNELS; c++) {
            if (!proc->channel_used[c]) continue;
            PG_CHECK_ARG(cols[c].n == n, "channel %d has %lld rows, expected %lld", c, (long long)cols[c].n, (long long)n);
            import_col(ctx, cols[c], n, pimp[c]);
            in[c] = pimp[c].d;
        }
        const int64_t before = op->input_rows;
        if (!hashagg_try_fused_processor(op, in.data(), pimp.data(), ncols, n)) hashagg_add_projected(op, in.data(), ncols, n, nullptr);
        op->input_rows = before + n;        // rows entering the fused operator, not rows surviving the filter
        sync(ctx);
        return;
    }
    // import each referenced channel once
    std::vector<ColImport> imp(ncols);
    std::vector<char> loaded(ncols, 0);
    auto need = [&](int ch) -> const DCol& {
        PG_CHECK_ARG(ch >= 0 && ch < ncols, "channel %d out of range (page has %d)", ch, ncols);
        if (!loaded[ch]) {
            PG_CHECK_ARG(cols[ch].n == n, "channel %d has %lld rows, expected %lld", ch, (long long)cols[ch].n, (long long)n);
            import_col(ctx, cols[ch], n, imp[ch]);
            loaded[ch] = 1;
        }
        return imp[ch].d;
    };
    KeyCols kc;
    memset(&kc, 0, sizeof(kc));
    for (size_t k = 0; k < op->key_channels.size(); k++) {
        kc.c[k] = need(op->key_channels[k]);
        PG_CHECK_ARG(kc.c[k].type == op->key_types[k], "group-by channel %d has type %d, expected %d", op->key_channels[k], kc.c[k].type,
                     op->key_types[k]);
    }
    DCol hc; memset(&hc, 0, sizeof(hc));
    bool has_hash = op->hash_channel >= 0;
    if (has_hash) { hc = need(op->hash_channel); PG_CHECK_ARG(hc.type == PG_BIGINT, "hash channel must be BIGINT"); }
    const bool raw = step_input_raw(op->step);
    auto build_prog = [&](int64_t off) {
        AggProgram prog;
        memset(&prog, 0, sizeof(prog));
        prog.naggs = (int32_t)op->aggs.size();
        for (size_t i = 0; i < op->aggs.size(); i++) {
            const pg_agg_spec& s = op->specs[i];
            DCol arg, mask;
            const DCol* parg = nullptr; const DCol* pmask = nullptr;
            if (s.kind != PG_AGG_COUNT) {
                arg = dcol_offset(need(s.input_channel), off);
                PG_CHECK_ARG(arg.type == s.in_type, "aggregate %d input channel has type %d, expected %d", (int)i, arg.type, s.in_type);
                parg = &arg;
            }
            if (s.mask_channel >= 0) { mask = dcol_offset(need(s.mask_channel), off); PG_CHECK_ARG(mask.type == PG_BOOLEAN, "mask must be BOOLEAN"); pmask = &mask; }
            prog.ops[i] = agg_make_op(op->aggs[i], parg, pmask);
        }
        return prog;
    };
    auto generic_rows = [&](int64_t off, int64_t m, const uint8_t* only_rows, bool with_aggregates = true) {
        KeyCols kco = kc;
        for (size_t k = 0; k < op->key_channels.size(); k++) kco.c[k] = dcol_offset(kc.c[k], off);
        DCol hco = hc;
        if (has_hash) hco = dcol_offset(hc, off);
        op->ids.grow(ctx, m * 4, 0);
        groupby_process(op->gb, kco, has_hash ? &hco : nullptr, m, op->ids.as<int32_t>());
        if (op->aggs.empty()) return;
        for (pg_agg* a : op->aggs) agg_ensure_capacity(a, op->gb->groups);
        if (raw && with_aggregates) accumulate(ctx, build_prog(off), op->ids.as<int32_t>(), m, op->gb->groups, only_rows);
    };
    // Large raw pages: a short prefix goes through the general path and numbers the groups; if there are only a few and
    // the keys are short, the rest is resolved and accumulated in ONE kernel (packed-key look-up in shared memory, private
    // partial states), and only rows that kernel could not resolve take the general path afterwards.
    // Many groups on one BIGINT key: state inside the hash buckets, one kernel per page, ids assigned at output time.
    if (raw && op->gb->bigint_mode && !has_hash) {
        // PRESTOGPU_WIDE: 0 = never, 1 = whenever there are many groups, unset = only when the input is not clustered by key.
        // Measured on B200, GROUP BY orderkey over lineitem SF10 (profiles/README.md): rows in random order 13.8 ms against
        // 21.3 ms on the general path; rows clustered by key 7.4 ms against 5.3 ms (the general path's per-page passes then
        // run out of L2, one random access per row no longer pays).
        const char* wide_env = getenv("PRESTOGPU_WIDE");
        const bool wide_on = wide_env ? wide_env[0] != '0' : true;
        const bool wide_always = wide_env && wide_env[0] == '1';
        if (wide_on && !op->wide && op->rows_since_reset == 0 && hashagg_group_count(op) == 0 && n >= 256 * 1024 && wide_supported(op->specs, op->aggs) &&
            jit_available()) {
            int64_t repeats = 0;
            KeySample ks;
            const int64_t d = sample_distinct_keys(ctx, kc.c[0], 2048, &repeats, &ks);
            int64_t est = (int64_t)((double)d / 2048.0 * (double)n);
            if (est > n / 2) est = n / 2;
            const int64_t target = op->expected * 4 >= est ? op->expected : est;
            // rows sorted by key: the range-chunked front region makes the bucket accesses sequential; rows in no key order: one
            // random access per row beats the general path's two; rows clustered by key but not sorted: the general path
            const bool sorted = d > 128 && wide_range_applicable(ks) && !(getenv("PRESTOGPU_WIDE_RANGE") && getenv("PRESTOGPU_WIDE_RANGE")[0] == '0');
            if (d > 128 && (wide_always || sorted || repeats * 4 < 2048))
                op->wide = wide_create(ctx, op->specs, op->aggs, op->step, target, sorted ? &ks : nullptr);
        }
        if (op->wide) {
            std::vector<DCol> wargs(op->aggs.size()), wmasks(op->aggs.size());
            for (size_t i = 0; i < op->aggs.size(); i++) {
                const pg_agg_spec& s = op->specs[i];
                memset(&wargs[i], 0, sizeof(DCol)); memset(&wmasks[i], 0, sizeof(DCol));
                if (s.kind != PG_AGG_COUNT) {
                    wargs[i] = need(s.input_channel);
                    PG_CHECK_ARG(wargs[i].type == s.in_type, "aggregate %d input channel has type %d, expected %d", (int)i, wargs[i].type, s.in_type);
                }
                if (s.mask_channel >= 0) { wmasks[i] = need(s.mask_channel); PG_CHECK_ARG(wmasks[i].type == PG_BOOLEAN, "mask must be BOOLEAN"); }
            }
            if (wide_add_page(op->wide, kc.c[0], wargs.data(), wmasks.data(), n, op->rows_since_reset)) {
                op->input_rows += n;
                op->rows_since_reset += n;
                return;
            }
            PG_CHECK_STATE(op->rows_since_reset == 0, "specialised kernel became unavailable after the first page: %s",
                           ctx->jit ? ctx->jit->last_error.c_str() : "");
            delete op->wide; op->wide = nullptr;   // nothing consumed yet: fall through to the general path
        }
    }
    const int64_t PREFIX = 16 * 1024;
    bool fused_done = false;
    if (raw && !op->aggs.empty() && n >= 8 * PREFIX && fused_keys_packable(op->gb->layout)) {
        ColImport* kimp[MAX_KEYS];
        int nkimp = 0;
        for (size_t k = 0; k < op->key_channels.size() && k < (size_t)MAX_KEYS; k++) kimp[nkimp++] = &imp[op->key_channels[k]];
        auto on_misses = [&](const unsigned int* bits) {     // rows the kernel could not resolve take the general path, in page order
            DevBuf only(ctx, n);
            PG_LAUNCH(ctx, expand_bits_kernel, grid_for(ctx, n, 256, 4), 256, 0, bits, n, only.as<uint8_t>());
            generic_rows(0, n, only.as<uint8_t>());
        };
        auto lowcard = [&]() -> bool {
            const int64_t g = hashagg_group_count(op);
            for (pg_agg* a : op->aggs) agg_ensure_capacity(a, g > 0 ? g : 64);      // final before the program takes the state pointers
            return hashagg_lowcard(op, kc, kimp, nkimp, build_prog(0), nullptr, n, on_misses);
        };
        // first choice: the kernel specialised at run time for this plan over ALL rows of the page; it finds the groups of a first
        // page itself.  When that does not apply (hash channel, double keys, a long string in the sample, ...): number the groups of a
        // short prefix through the GroupByHash and come back with them known; failing that, the ahead-of-time fused kernel.
        fused_done = lowcard();
        if (!fused_done) {
            hashagg_sync_gb(op);
            if (op->gb->has_var) {   // one read-back gives the value bytes of the page and of the prefix (the store's growth bound)
                int64_t pb[MAX_KEYS], bound = 0;
                resolve_value_bytes(ctx, kimp, nkimp, PREFIX, pb);
                for (int k = 0; k < nkimp; k++) bound += pb[k];
                op->gb->var_bound_hint = bound;
            }
            generic_rows(0, PREFIX, nullptr, false);   // numbers the prefix's groups; its rows are accumulated below
            op->gb->var_bound_hint = -1;
            const int64_t groups = op->gb->groups;
            if (groups <= 64) fused_done = lowcard();
            if (!fused_done) {
                accumulate(ctx, build_prog(0), op->ids.as<int32_t>(), PREFIX, groups);      // the prefix by its ids
                if (groups <= 64) {
                    const int64_t m = n - PREFIX, miss_words = (m + 31) / 32;
                    DevBuf miss_flag(ctx, 8 + miss_words * 4);
                    memset_async(ctx, miss_flag.p, 0, 8 + miss_words * 4);
                    unsigned long long* d_miss_cnt = miss_flag.as<unsigned long long>();
                    unsigned int* d_miss_bits = (unsigned int*)(d_miss_cnt + 1);
                    KeyCols kco = kc;
                    for (size_t k = 0; k < op->key_channels.size(); k++) kco.c[k] = dcol_offset(kc.c[k], PREFIX);
                    FusedLookup fl;
                    memset(&fl, 0, sizeof(fl));
                    fl.layout = &op->gb->layout; fl.keys = &kco; fl.key_rows = op->gb->key_rows.as<uint64_t>(); fl.arena = op->gb->arena.as<uint8_t>();
                    fl.values_by_group = op->gb->values_by_group.as<int64_t>(); fl.bigint_mode = op->gb->bigint_mode ? 1 : 0;
                    fl.null_gid = -1;
                    if (op->gb->bigint_mode) {
                        int64_t sp[2];
                        read_scalars(ctx, op->gb->slots.as<Slot>() + op->gb->cap + 1, 2, sp);   // the null group's slot {word, gid|first_row}
                        fl.null_gid = (int32_t)(sp[1] & 0xffffffff);
                    }
                    fl.miss_flag = d_miss_bits; fl.d_miss = d_miss_cnt; fl.row_base = PREFIX;
                    resolve_value_bytes(ctx, kimp, nkimp);
                    for (int k = 0; k < nkimp && k < 8; k++) fl.key_value_bytes[k] = kimp[k]->value_bytes;
                    if (accumulate_private(ctx, build_prog(PREFIX), nullptr, m, groups, nullptr, &fl)) {
                        int64_t misses;
                        read_scalars(ctx, d_miss_cnt, 1, &misses);
                        if (misses > 0) {
                            DevBuf only(ctx, m);
                            PG_LAUNCH(ctx, expand_bits_kernel, grid_for(ctx, m, 256, 4), 256, 0, d_miss_bits, m, only.as<uint8_t>());
                            generic_rows(PREFIX, m, only.as<uint8_t>());
                        }
                        fused_done = true;
                    }
                }
                if (!fused_done) { generic_rows(PREFIX, n - PREFIX, nullptr); fused_done = true; }
            }
        }
    }
    op->rows_since_reset += n;
    if (fused_done) {
        op->input_rows += n;
        sync(ctx);
        return;
    }
    hashagg_sync_gb(op);
    op->ids.grow(ctx, n * 4, 0);
    if (op->gb->has_var && n <= 16384) {     // small page: a bound on its variable width key bytes lets the GroupByHash take its one-launch path
        int64_t bound = 0;
        for (size_t k = 0; k < op->key_channels.size() && bound >= 0; k++) {
            const ColImport& ci = imp[op->key_channels[k]];
            if (type_is_var(ci.d.type)) bound = ci.value_bytes >= 0 ? bound + ci.value_bytes : -1;
        }
        op->gb->var_bound_hint = bound;
    }
    groupby_process(op->gb, kc, has_hash ? &hc : nullptr, n, op->ids.as<int32_t>());
    op->gb->var_bound_hint = -1;
    int64_t groups = op->gb->groups;
    if (op->aggs.empty()) { op->input_rows += n; return; }
    for (pg_agg* a : op->aggs) agg_ensure_capacity(a, groups);
    if (raw) {
        accumulate(ctx, build_prog(0), op->ids.as<int32_t>(), n, groups);
    } else {
        DCol s0[MAX_AGGS], s1[MAX_AGGS];
        memset(s1, 0, sizeof(s1));
        for (size_t i = 0; i < op->aggs.size(); i++) {
            const pg_agg_spec& s = op->specs[i];
            s0[i] = need(s.input_channel);
            if (agg_n_intermediate(op->aggs[i]) > 1) s1[i] = need(s.input_channel + 1);
        }
        accumulate_intermediate_batch(ctx, op->aggs.data(), s0, s1, (int)op->aggs.size(), op->ids.as<int32_t>(), n, groups);
    }
    op->input_rows += n;
    sync(ctx);   // staged host columns are released when `imp` goes out of scope
}


// ------------------------------------------------------------------ page accumulator
static void hashagg_reserve(pg_hashagg* op, int64_t n) {
    // memory yield: checked before the first row is consumed, for whichever table the rows may grow
    if (op->gb->memory_limit <= 0) return;
    if (op->wide && !op->wide->prepared) {
        WideTable* w = op->wide;
        if (w->claimed + n > w->max_fill) {
            int64_t cap = w->cap;
            while ((int64_t)(cap * 0.75) < w->claimed + n) cap *= 2;
            const int64_t after = w->device_bytes() - w->table.bytes + (cap + 2) * (int64_t)w->bucket_words * 8;
            if (after > op->gb->memory_limit)
                fail(PG_NEED_MEMORY, "the aggregation table may have to grow to %lld bytes for this page, the memory limit is %lld; nothing was consumed",
                     (long long)after, (long long)op->gb->memory_limit);
        }
    } else groupby_reserve(op->gb, n);
}

static bool channel_is_read(const pg_hashagg* op, int c) {
    if (op->proc) return c < (int)op->proc->channel_used.size() && op->proc->channel_used[c];
    if (c == op->hash_channel) return true;
    for (int k : op->key_channels) if (k == c) return true;
    for (const pg_agg_spec& s : op->specs) if (s.input_channel == c || s.mask_channel == c) return true;
    return false;
}

// true when the page was taken into the accumulator
static bool batch_append(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n) {
    if (op->batch_rows <= 0 || n <= 0 || n * 4 > op->batch_rows) return false;
    bool pinned = true;        // every channel the operator reads lives in caller-pinned memory: device mode
    for (int c = 0; c < ncols; c++) {
        if (!channel_is_read(op, c)) continue;
        if (cols[c].mem != PG_MEM_HOST && cols[c].mem != PG_MEM_HOST_PINNED) return false;
        if (cols[c].mem != PG_MEM_HOST_PINNED) pinned = false;
    }
    PageBatch* b = op->batch;
    if (!b) {
        pg_ctx* ctx = op->ctx;
        if (!ctx->page_batch_pool.empty()) { b = (PageBatch*)ctx->page_batch_pool.back(); ctx->page_batch_pool.pop_back(); }
        else b = new PageBatch();
        op->batch = b;
    }
    // (re)shape the accumulator when it comes from another operator's plan
    bool same = b->cap_rows >= op->batch_rows && (int)b->cols.size() == ncols;
    for (int c = 0; same && c < ncols; c++) {
        const bool used = channel_is_read(op, c);
        if (b->cols[c].used != used || (used && b->cols[c].type != cols[c].type)) same = false;
    }
    if (!same) {
        PG_CHECK_STATE(b->rows == 0, "page accumulator reshaped while it holds rows");
        b->release();
        b->cap_rows = op->batch_rows;
        b->cols.resize(ncols);
        for (int c = 0; c < ncols; c++) {
            PageBatch::Col& bc = b->cols[c];
            bc.used = channel_is_read(op, c);
            bc.type = cols[c].type;
            if (!bc.used) continue;
            PG_CHECK_ARG(type_valid(bc.type), "unknown column type %d", bc.type);
            if (type_is_var(bc.type)) {
                bc.values_cap = b->cap_rows * 16;
                bc.offsets = (int32_t*)PageBatch::pinned((b->cap_rows + 1) * 4);
            } else bc.values_cap = b->cap_rows * (int64_t)type_width(bc.type);
            bc.values = (char*)PageBatch::pinned(bc.values_cap);
            bc.nulls = (uint8_t*)PageBatch::pinned(b->cap_rows);
        }
    }
    if (b->rows + n > b->cap_rows) return false;           // the caller flushes and tries again
    if (b->rows > 0 && b->device_mode != pinned) return false;      // one kind of copy per batch
    b->device_mode = pinned;
    if (pinned) {
        pg_ctx* ctx = op->ctx;
        for (int c = 0; c < ncols; c++) {                // room first: a page that does not fit must leave the batch untouched
            PageBatch::Col& bc = b->cols[c];
            if (!bc.used || !type_is_var(bc.type)) continue;
            PG_CHECK_ARG(cols[c].offsets != nullptr, "variable width channel %d without offsets", c);
            const int64_t bytes = (int64_t)cols[c].offsets[n] - cols[c].offsets[0];
            if (bc.values_used + bytes > bc.values_cap) return false;
        }
        for (int c = 0; c < ncols; c++) {
            PageBatch::Col& bc = b->cols[c];
            if (!bc.used) continue;
            const pg_col& in = cols[c];
            PG_CHECK_ARG(in.type == bc.type && in.n == n, "channel %d: type %d / %lld rows, the accumulated pages have type %d / this page %lld rows", c, in.type,
                         (long long)in.n, bc.type, (long long)n);
            if (!bc.dnulls.p) {
                bc.dvalues.alloc(ctx, bc.values_cap);
                bc.dnulls.alloc(ctx, b->cap_rows);
                if (type_is_var(bc.type)) bc.doffsets.alloc(ctx, (b->cap_rows + 1) * 4);
            }
            if (type_is_var(bc.type)) {
                const int64_t first = in.offsets[0], bytes = (int64_t)in.offsets[n] - first;
                if (bytes > 0) b->descs.push_back(GatherDesc{(const char*)in.values + first, (char*)bc.dvalues.p + bc.values_used, bytes});
                for (int64_t i = 0; i <= n; i++) bc.offsets[b->rows + i] = (int32_t)(bc.values_used + (in.offsets[i] - first));     // leave with the flush
                bc.values_used += bytes;
            } else {
                const int w = type_width(bc.type);
                b->descs.push_back(GatherDesc{(const char*)in.values, (char*)bc.dvalues.p + b->rows * w, n * w});
            }
            if (in.nulls) {
                if (!bc.has_nulls) { memset_async(ctx, bc.dnulls.p, 0, b->rows); bc.has_nulls = true; }
                b->descs.push_back(GatherDesc{(const char*)in.nulls, (char*)bc.dnulls.p + b->rows, n});
            } else if (bc.has_nulls) memset_async(ctx, (char*)bc.dnulls.p + b->rows, 0, n);
        }
        b->rows += n;
        return true;
    }
    for (int c = 0; c < ncols; c++) {
        PageBatch::Col& bc = b->cols[c];
        if (!bc.used) continue;
        const pg_col& in = cols[c];
        PG_CHECK_ARG(in.type == bc.type && in.n == n, "channel %d: type %d / %lld rows, the accumulated pages have type %d / this page %lld rows", c, in.type,
                     (long long)in.n, bc.type, (long long)n);
        if (type_is_var(bc.type)) {
            PG_CHECK_ARG(in.offsets != nullptr, "variable width channel %d without offsets", c);
            const int64_t first = in.offsets[0], bytes = (int64_t)in.offsets[n] - first;
            if (bc.values_used + bytes > bc.values_cap) {
                int64_t cap = bc.values_cap * 2;
                while (cap < bc.values_used + bytes) cap *= 2;
                PG_CHECK_ARG(cap < 0x7fffffffLL, "variable width channel %d exceeds 2 GB per batch", c);
                char* nv = (char*)PageBatch::pinned(cap);
                memcpy(nv, bc.values, (size_t)bc.values_used);
                cudaFreeHost(bc.values);
                bc.values = nv; bc.values_cap = cap;
            }
            if (bytes > 0) memcpy(bc.values + bc.values_used, (const char*)in.values + first, (size_t)bytes);
            for (int64_t i = 0; i <= n; i++) bc.offsets[b->rows + i] = (int32_t)(bc.values_used + (in.offsets[i] - first));
            bc.values_used += bytes;
        } else {
            const int w = type_width(bc.type);
            memcpy(bc.values + b->rows * w, in.values, (size_t)(n * w));
        }
        if (in.nulls) {
            if (!bc.has_nulls) { memset(bc.nulls, 0, (size_t)b->rows); bc.has_nulls = true; }
            memcpy(bc.nulls + b->rows, in.nulls, (size_t)n);
        } else if (bc.has_nulls) memset(bc.nulls + b->rows, 0, (size_t)n);
    }
    b->rows += n;
    return true;
}

// hands the accumulated rows to the kernels as one page; on PG_NEED_MEMORY they stay accumulated
static void batch_flush(pg_hashagg* op) {
    PageBatch* b = op->batch;
    if (!b || b->rows == 0) return;
    DevBuf d_descs;
    if (b->device_mode && !b->descs.empty()) {
        pg_ctx* ctx = op->ctx;
        const int64_t bytes = (int64_t)b->descs.size() * (int64_t)sizeof(GatherDesc);
        const GatherDesc* dd = (const GatherDesc*)in_arena_stage(ctx, b->descs.data(), bytes);
        if (!dd) { d_descs.alloc(ctx, bytes); copy_async(ctx, d_descs.p, b->descs.data(), bytes, cudaMemcpyHostToDevice); dd = (const GatherDesc*)d_descs.p; }
        const int nd = (int)b->descs.size();
        dim3 grid(8, nd < 16384 ? nd : 16384);
        PG_LAUNCH(ctx, batch_gather_kernel, grid, 256, 0, dd, nd);
    }
    std::vector<pg_col> cols(b->cols.size());
    for (size_t c = 0; c < cols.size(); c++) {
        const PageBatch::Col& bc = b->cols[c];
        memset(&cols[c], 0, sizeof(pg_col));
        cols[c].type = bc.type; cols[c].mem = PG_MEM_HOST; cols[c].n = b->rows;
        if (!bc.used) continue;
        if (b->device_mode) {
            cols[c].mem = PG_MEM_DEVICE;
            cols[c].values = bc.dvalues.p; cols[c].nulls = bc.has_nulls ? (const uint8_t*)bc.dnulls.p : nullptr;
            if (type_is_var(bc.type)) {
                copy_async(op->ctx, bc.doffsets.p, bc.offsets, (b->rows + 1) * 4, cudaMemcpyHostToDevice);
                cols[c].offsets = (const int32_t*)bc.doffsets.p;
            }
            continue;
        }
        cols[c].values = bc.values; cols[c].nulls = bc.has_nulls ? bc.nulls : nullptr; cols[c].offsets = type_is_var(bc.type) ? bc.offsets : nullptr;
    }
    hashagg_reserve(op, b->rows);
    hashagg_add_page(op, cols.data(), (int32_t)cols.size(), b->rows);      // synchronises: the pinned rows have left when it returns
    b->rows = 0;
    b->descs.clear();
    for (PageBatch::Col& bc : b->cols) { bc.values_used = 0; bc.has_nulls = false; }
}

static void hashagg_add_page_batched(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n) {
    if (batch_append(op, cols, ncols, n)) {
        if (op->batch->rows >= op->batch_rows) batch_flush(op);
        return;
    }
    batch_flush(op);                                   // keeps page order: earlier rows first
    if (batch_append(op, cols, ncols, n)) return;      // it did not fit before the flush
    hashagg_reserve(op, n);
    hashagg_add_page(op, cols, ncols, n);
}

}  // namespace pg

using namespace pg;

extern "C" {

int32_t pg_hashagg_create(pg_ctx* ctx, const int32_t* key_channels, const int32_t* key_types, int32_t nkeys, int32_t hash_channel,
                          const pg_agg_spec* aggs, int32_t naggs, int32_t step, int64_t expected_groups, pg_hashagg** out) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        PG_CHECK_ARG(out && key_channels && key_types && nkeys >= 1, "a grouped aggregation needs at least one group-by channel");
        PG_CHECK_ARG(naggs >= 0 && naggs <= MAX_AGGS && (naggs == 0 || aggs), "between 0 and %d aggregates are supported", MAX_AGGS);
        PG_CHECK_ARG(step >= PG_STEP_SINGLE && step <= PG_STEP_INTERMEDIATE, "bad step %d", step);
        pg_hashagg* op = new pg_hashagg();
        try {
            op->ctx = ctx; op->step = step; op->hash_channel = hash_channel; op->expected = expected_groups;
            op->key_channels.assign(key_channels, key_channels + nkeys);
            op->key_types.assign(key_types, key_types + nkeys);
            op->specs.assign(aggs, aggs + naggs);
            hashagg_build(op);
        } catch (...) { delete op; throw; }
        *out = op;
    });
}

int32_t pg_hashagg_destroy(pg_hashagg* op) {
    if (!op) return PG_OK;
    return guard(op->ctx, [&] { delete op; });
}

int32_t pg_hashagg_add_page(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        PG_CHECK_ARG(cols && ncols > 0 && n >= 0, "bad arguments");
        hashagg_add_page_batched(op, cols, ncols, n);
    });
}

int32_t pg_hashagg_set_page_batching(pg_hashagg* op, int64_t rows) {
    if (!op || rows < 0) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        batch_flush(op);
        op->batch_rows = rows;
    });
}

int32_t pg_hashagg_add_pages(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n, int64_t page_rows) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        PG_CHECK_ARG(cols && ncols > 0 && n >= 0 && page_rows > 0, "bad arguments");
        std::vector<pg_col> page(ncols);
        for (int64_t at = 0; at < n; at += page_rows) {
            const int64_t m = n - at < page_rows ? n - at : page_rows;
            for (int c = 0; c < ncols; c++) {
                PG_CHECK_ARG(cols[c].mem == PG_MEM_HOST || cols[c].mem == PG_MEM_HOST_PINNED, "pg_hashagg_add_pages takes host pages");
                page[c] = cols[c];
                page[c].n = m;
                if (page[c].nulls) page[c].nulls += at;
                if (type_is_var(page[c].type)) { if (page[c].offsets) page[c].offsets += at; }
                else if (page[c].values) page[c].values = (const char*)page[c].values + at * type_width(page[c].type);
            }
            hashagg_add_page_batched(op, page.data(), ncols, m);       // the same path as one pg_hashagg_add_page call per page
        }
    });
}

int32_t pg_hashagg_set_memory_limit(pg_hashagg* op, int64_t bytes) {
    if (!op || bytes < 0) return PG_E_INVALID_ARGUMENT;
    op->gb->memory_limit = bytes;
    return PG_OK;
}

int32_t pg_hashagg_stats(pg_hashagg* op, int64_t* group_count, int64_t* capacity, int64_t* device_bytes, int64_t* input_rows) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    if (op->batch && op->batch->rows > 0) {          // accumulated pages count as consumed input: bring the state up to date first
        const int32_t rc = guard(op->ctx, [&] { batch_flush(op); });
        if (rc != PG_OK) return rc;
    }
    if (op->wide && !op->wide->prepared) {
        // ids do not exist yet; the count and the reference's capacity schedule follow the buckets in use
        op->gb->groups = op->wide->groups();
        while (op->gb->groups >= max_fill_075(op->gb->ref_capacity) && op->gb->ref_capacity < (1LL << 30)) op->gb->ref_capacity *= 2;
    }
    if (!op->lc.gb_current) {   // groups known to the few-groups kernel only: same count, same logical capacity schedule
        while (op->lc.count >= max_fill_075(op->gb->ref_capacity) && op->gb->ref_capacity < (1LL << 30)) op->gb->ref_capacity *= 2;
    }
    if (group_count) *group_count = hashagg_group_count(op);
    if (capacity) *capacity = op->gb->ref_capacity;
    if (device_bytes) {
        int64_t b = groupby_device_bytes(op->gb) + op->ids.bytes + (op->wide ? op->wide->device_bytes() : 0);
        for (pg_agg* a : op->aggs) b += a->state.bytes;
        *device_bytes = b;
    }
    if (input_rows) *input_rows = op->input_rows;
    return PG_OK;
}

/* output column types in order: keys..., (hash), then per aggregate its final type or its intermediate columns */
int32_t pg_hashagg_output_types(pg_hashagg* op, int32_t* types, int32_t capacity, int32_t* ncols) {
    if (!op || !ncols) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        std::vector<int32_t> t(op->key_types);
        if (op->hash_channel >= 0) t.push_back(PG_BIGINT);
        bool partial = step_output_partial(op->step);
        for (pg_agg* a : op->aggs) {
            bool dec = a->op == OP_SUM_DEC;
            if (!partial) {
                int32_t ft;
                switch (a->kind) {
                    case PG_AGG_COUNT: case PG_AGG_COUNT_COL: ft = PG_BIGINT; break;
                    case PG_AGG_SUM: ft = dec ? PG_LONG_DECIMAL : (a->op == OP_SUM_I64 ? PG_BIGINT : PG_DOUBLE); break;
                    case PG_AGG_AVG: ft = dec ? a->in_type : PG_DOUBLE; break;
                    default: ft = a->in_type;
                }
                t.push_back(ft);
            } else {
                switch (a->kind) {
                    case PG_AGG_COUNT: case PG_AGG_COUNT_COL: t.push_back(PG_BIGINT); break;
                    case PG_AGG_SUM:
                        if (dec) t.push_back(PG_VARBINARY);
                        else { t.push_back(PG_BIGINT); t.push_back(a->op == OP_SUM_I64 ? PG_BIGINT : PG_DOUBLE); }
                        break;
                    case PG_AGG_AVG:
                        if (dec) t.push_back(PG_VARBINARY);
                        else { t.push_back(PG_DOUBLE); t.push_back(PG_BIGINT); }
                        break;
                    default: t.push_back(a->in_type);
                }
            }
        }
        *ncols = (int32_t)t.size();
        if (types) {
            PG_CHECK_ARG(capacity >= (int32_t)t.size(), "types buffer too small");
            for (size_t i = 0; i < t.size(); i++) types[i] = t[i];
        }
    });
}

/* buildResult for groups [first, first+n): out[] in pg_hashagg_output_types order */
int32_t pg_hashagg_get_output(pg_hashagg* op, int64_t first, int64_t n, pg_col_out* out, int64_t* error_group) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        pg_ctx* ctx = op->ctx;
        batch_flush(op);
        if (op->wide) wide_prepare_output(op->wide, op->gb, op->aggs, op->rows_since_reset);
        const int64_t total = hashagg_group_count(op);
        PG_CHECK_ARG(out && first >= 0 && n >= 0 && first + n <= total, "group range outside [0, %lld)", (long long)total);
        if (error_group) *error_group = -1;
        if (n == 0) return;
        int nk = (int)op->key_types.size();
        bool has_hash = op->hash_channel >= 0;
        std::vector<OutStage> kst;
        if (!op->lc.gb_current && n <= 64) {
            // the groups live in the packed table only: unpack the keys straight into the output columns
            const KeyLayout& L = op->gb->layout;
            kst.resize(nk);
            LowcardKeyOut ko;
            memset(&ko, 0, sizeof(ko));
            for (int k = 0; k < nk; k++) {
                PG_CHECK_ARG(out[k].type == L.types[k], "key output column %d has type %d, expected %d", k, out[k].type, L.types[k]);
                const bool var = type_is_var(L.types[k]);
                kst[k].begin(ctx, &out[k], n, var ? 3 * n : n * (int64_t)type_width(L.types[k]));
                ko.values[k] = kst[k].d_values; ko.nulls[k] = kst[k].d_nulls; ko.offsets[k] = kst[k].d_offsets;
            }
            lowcard_unpack_groups(ctx, L, op->lc.groups.as<unsigned long long>(), first, n, ko);
            for (int k = 0; k < nk; k++) kst[k].finish(type_is_var(L.types[k]) ? 3 * n : n * (int64_t)type_width(L.types[k]));
        } else {
            hashagg_sync_gb(op);
            int64_t* hash_ptr = nullptr; int32_t hash_mem = PG_MEM_DEVICE;
            if (has_hash) {
                PG_CHECK_ARG(out[nk].type == PG_BIGINT && out[nk].capacity >= n, "hash output column must be BIGINT with capacity");
                hash_ptr = (int64_t*)out[nk].values; hash_mem = out[nk].mem;
            }
            int32_t rc = pg_groupby_export_keys(op->gb, first, n, out, hash_ptr, hash_mem);
            if (rc != PG_OK) fail(rc, "%s", ctx->err.c_str());
        }
        if (has_hash && out[nk].nulls) {
            if (out[nk].mem == PG_MEM_DEVICE) memset_async(ctx, out[nk].nulls, 0, n); else memset(out[nk].nulls, 0, n);
        }
        int col = nk + (has_hash ? 1 : 0);
        bool partial = step_output_partial(op->step);
        int64_t first_err = -1;
        const int na = (int)op->aggs.size();
        DevBuf errs(ctx, 8 * (na > 0 ? na : 1));       // one sticky "first overflowing group" word per aggregate
        memset_async(ctx, errs.p, 0x7f, 8 * (na > 0 ? na : 1));
        agg_evaluate_batch(ctx, op->aggs.data(), na, first, n, partial ? 0 : 1, out + col, errs.as<long long>());
        int64_t eg[MAX_AGGS];
        if (na > 0) {
            read_scalars(ctx, errs.p, na, eg);          // also completes the copies into host output columns
            for (int i = 0; i < na; i++) if (eg[i] != 0x7f7f7f7f7f7f7f7fLL && first_err < 0) first_err = eg[i];
        } else sync(ctx);
        if (first_err >= 0) {
            if (error_group) *error_group = first_err;
            fail(PG_E_NUMERIC_OVERFLOW, "aggregate overflow in group %lld", (long long)first_err);
        }
    });
}

/* value bytes needed for output column `col` (variable width key or VARBINARY intermediate state) */
int32_t pg_hashagg_varwidth_bytes(pg_hashagg* op, int32_t col, int64_t* out_bytes) {
    if (!op || !out_bytes) return PG_E_INVALID_ARGUMENT;
    if (op->batch && op->batch->rows > 0) {
        const int32_t rc = guard(op->ctx, [&] { batch_flush(op); });
        if (rc != PG_OK) return rc;
    }
    int nk = (int)op->key_types.size();
    if (col < nk && !op->lc.gb_current) { *out_bytes = 3 * op->lc.count; return PG_OK; }     // packed keys: strings of at most 3 bytes
    if (col < nk) return pg_groupby_varwidth_bytes(op->gb, col, out_bytes);
    *out_bytes = 32 * hashagg_group_count(op);   // decimal intermediate states: at most 32 bytes per group
    return PG_OK;
}

/* ---- packed pages: the PARTIAL -> FINAL wire format (see include/prestogpu.h) */
namespace {
struct PackedHeader { int64_t total_bytes, n; int32_t ncols, magic; };
struct PackedCol { int32_t type, has_nulls; int64_t values_bytes, offsets_bytes, data_offset; };
const int32_t PACKED_MAGIC = 0x31504750;   // "PGP1"
inline int64_t align8(int64_t x) { return (x + 7) & ~(int64_t)7; }
inline bool packed_is_var(int32_t t) { return t == PG_VARCHAR || t == PG_VARBINARY; }
inline int64_t packed_width(int32_t t) {
    switch (t) {
        case PG_INTEGER: case PG_DATE: return 4;
        case PG_BOOLEAN: return 1;
        case PG_LONG_DECIMAL: return 16;
        default: return 8;
    }
}
}  // namespace

static int32_t get_output_packed_impl(pg_hashagg* op, int64_t first, int64_t n, void* buf, int32_t mem, int64_t capacity, int64_t* out_bytes,
                                      int64_t* error_group) {
    if (!op || !out_bytes) return PG_E_INVALID_ARGUMENT;
    int32_t types[256], ncols = 0;
    int32_t rc = pg_hashagg_output_types(op, types, 256, &ncols);
    if (rc != PG_OK) return rc;
    std::vector<PackedCol> cols(ncols);
    const int64_t header_bytes = align8((int64_t)sizeof(PackedHeader) + (int64_t)ncols * (int64_t)sizeof(PackedCol));
    int64_t pos = header_bytes;
    for (int i = 0; i < ncols; i++) {
        PackedCol& c = cols[i];
        c.type = types[i]; c.has_nulls = 1; c.offsets_bytes = 0;
        if (packed_is_var(types[i])) {
            int64_t vb = 0;
            rc = pg_hashagg_varwidth_bytes(op, i, &vb);
            if (rc != PG_OK) return rc;
            c.values_bytes = vb > 0 ? vb : 1;
            c.offsets_bytes = (n + 1) * 4;
        } else c.values_bytes = (n > 0 ? n : 1) * packed_width(types[i]);
        c.data_offset = pos;
        pos += align8(c.values_bytes) + align8(c.offsets_bytes) + align8(n > 0 ? n : 1);
    }
    *out_bytes = pos;
    if (!buf || capacity < pos) {
        return guard(op->ctx, [&] { fail(PG_NEED_MEMORY, "packed page needs %lld bytes, the buffer holds %lld", (long long)pos, (long long)capacity); });
    }
    char* base = (char*)buf;
    std::vector<char> hdr((size_t)header_bytes, 0);
    PackedHeader* h = (PackedHeader*)hdr.data();
    h->total_bytes = pos; h->n = n; h->ncols = ncols; h->magic = PACKED_MAGIC;
    memcpy(hdr.data() + sizeof(PackedHeader), cols.data(), (size_t)ncols * sizeof(PackedCol));
    if (mem == PG_MEM_DEVICE) {
        // zero the page (null flags, padding), then the header; the columns are written in place by the output kernels
        rc = guard(op->ctx, [&] {
            memset_async(op->ctx, base, 0, pos);
            copy_async(op->ctx, base, hdr.data(), header_bytes, cudaMemcpyHostToDevice);
        });
        if (rc != PG_OK) return rc;
    } else memcpy(base, hdr.data(), (size_t)header_bytes);
    std::vector<pg_col_out> outs(ncols);
    for (int i = 0; i < ncols; i++) {
        const PackedCol& c = cols[i];
        pg_col_out& o = outs[i];
        memset(&o, 0, sizeof(o));
        o.type = c.type; o.mem = mem; o.capacity = n;
        o.values = base + c.data_offset;
        o.offsets = c.offsets_bytes ? (int32_t*)(base + c.data_offset + align8(c.values_bytes)) : nullptr;
        o.nulls = (uint8_t*)(base + c.data_offset + align8(c.values_bytes) + align8(c.offsets_bytes));
        o.values_capacity_bytes = packed_is_var(c.type) ? c.values_bytes : 0;
        if (mem != PG_MEM_DEVICE) memset(o.nulls, 0, (size_t)(n > 0 ? n : 1));
    }
    return pg_hashagg_get_output(op, first, n, outs.data(), error_group);
}

int32_t pg_hashagg_get_output_packed(pg_hashagg* op, int64_t first, int64_t n, void* buf, int64_t capacity, int64_t* out_bytes,
                                     int64_t* error_group) {
    return get_output_packed_impl(op, first, n, buf, PG_MEM_HOST, capacity, out_bytes, error_group);
}

int32_t pg_hashagg_get_output_packed_device(pg_hashagg* op, int64_t first, int64_t n, void* device_buf, int64_t capacity, int64_t* out_bytes,
                                            int64_t* error_group) {
    return get_output_packed_impl(op, first, n, device_buf, PG_MEM_DEVICE, capacity, out_bytes, error_group);
}

int32_t pg_hashagg_add_packed_device(pg_hashagg* op, const void* device_bufs, int64_t stride, int32_t nbufs) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    pg_ctx* ctx = op->ctx;
    std::vector<const void*> ptrs;
    int32_t rc = guard(ctx, [&] {
        PG_CHECK_ARG(device_bufs && nbufs > 0 && stride >= (int64_t)sizeof(PackedHeader) && stride % 8 == 0, "bad arguments");
        const int64_t bytes = stride * nbufs;
        PG_CHECK_ARG(bytes <= (256LL << 20), "packed pages of %lld bytes are too large for this entry point", (long long)bytes);
        if (ctx->pk_cap < bytes) {
            if (ctx->pk_pinned) { cudaFreeHost(ctx->pk_pinned); ctx->pk_pinned = nullptr; ctx->pk_cap = 0; }
            int64_t cap = 1 << 20;
            while (cap < bytes) cap <<= 1;
            PG_CUDA(cudaHostAlloc((void**)&ctx->pk_pinned, (size_t)cap, cudaHostAllocDefault));
            ctx->pk_cap = cap;
        }
        // only the bytes each page uses would do, but their count sits in the pages themselves: one copy of everything is cheaper than two trips
        copy_async(ctx, ctx->pk_pinned, device_bufs, bytes, cudaMemcpyDeviceToHost);
        sync(ctx);
        for (int b = 0; b < nbufs; b++) {
            const PackedHeader* h = (const PackedHeader*)(ctx->pk_pinned + (int64_t)b * stride);
            PG_CHECK_ARG(h->magic == PACKED_MAGIC && h->total_bytes <= stride, "buffer %d is not a packed page of at most %lld bytes", b, (long long)stride);
            ptrs.push_back(h);
        }
    });
    if (rc != PG_OK) return rc;
    return pg_hashagg_add_packed(op, ptrs.data(), nbufs);
}

int32_t pg_hashagg_add_packed(pg_hashagg* op, const void* const* bufs, int32_t nbufs) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    if (op->batch && op->batch->rows > 0) {
        const int32_t rc = guard(op->ctx, [&] { batch_flush(op); });
        if (rc != PG_OK) return rc;
    }
    int32_t ncols = -1;
    int64_t total = 0;
    std::vector<std::vector<char>> values, nulls;
    std::vector<std::vector<int32_t>> offsets;
    std::vector<int32_t> types;
    int32_t rc = guard(op->ctx, [&] {
        PG_CHECK_ARG(bufs && nbufs > 0, "bad arguments");
        for (int b = 0; b < nbufs; b++) {
            const char* base = (const char*)bufs[b];
            PG_CHECK_ARG(base != nullptr, "null buffer");
            const PackedHeader* h = (const PackedHeader*)base;
            PG_CHECK_ARG(h->magic == PACKED_MAGIC && h->n >= 0 && h->ncols > 0 && h->ncols <= 256, "buffer %d is not a packed page", b);
            const PackedCol* pc = (const PackedCol*)(base + sizeof(PackedHeader));
            if (ncols < 0) {
                ncols = h->ncols;
                values.resize(ncols); nulls.resize(ncols); offsets.resize(ncols); types.resize(ncols);
                for (int i = 0; i < ncols; i++) { types[i] = pc[i].type; if (packed_is_var(types[i])) offsets[i].push_back(0); }
            }
            PG_CHECK_ARG(h->ncols == ncols, "buffer %d has %d columns, the first one %d", b, h->ncols, ncols);
            const int64_t n = h->n;
            for (int i = 0; i < ncols; i++) {
                const PackedCol& c = pc[i];
                PG_CHECK_ARG(c.type == types[i], "buffer %d column %d has type %d, expected %d", b, i, c.type, types[i]);
                const char* v = base + c.data_offset;
                const char* nl = v + align8(c.values_bytes) + align8(c.offsets_bytes);
                if (packed_is_var(c.type)) {
                    const int32_t* o = (const int32_t*)(v + align8(c.values_bytes));
                    const int32_t used = n > 0 ? o[n] : 0, shift = offsets[i].back();
                    values[i].insert(values[i].end(), v, v + used);
                    for (int64_t r = 1; r <= n; r++) offsets[i].push_back(o[r] + shift);
                } else values[i].insert(values[i].end(), v, v + n * packed_width(c.type));
                if (c.has_nulls) nulls[i].insert(nulls[i].end(), nl, nl + n);
                else nulls[i].insert(nulls[i].end(), (size_t)n, 0);
            }
            total += n;
        }
    });
    if (rc != PG_OK) return rc;
    if (total == 0) return PG_OK;
    std::vector<pg_col> cols(ncols);
    for (int i = 0; i < ncols; i++) {
        pg_col& c = cols[i];
        memset(&c, 0, sizeof(c));
        c.type = types[i]; c.mem = PG_MEM_HOST; c.n = total;
        if (values[i].empty()) values[i].resize(8);
        c.values = values[i].data();
        bool any_null = false;
        for (char x : nulls[i]) if (x) { any_null = true; break; }
        c.nulls = any_null ? (const uint8_t*)nulls[i].data() : nullptr;
        c.offsets = packed_is_var(types[i]) ? offsets[i].data() : nullptr;
    }
    return pg_hashagg_add_page(op, cols.data(), ncols, total);
}

int32_t pg_hashagg_set_input_processor(pg_hashagg* op, pg_project* p) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        PG_CHECK_STATE(op->input_rows == 0, "the input processor must be attached before the first page");
        if (p) {
            PG_CHECK_ARG(p->ctx->device == op->ctx->device, "processor and operator live on different devices");
            (void)channels_read(op, (int)p->proj.size());     // validates the operator's channel numbers against the projections
        }
        op->proc = p;
    });
}

/* partial-aggregation flush: drop all groups and states, keep the configuration (HashAggregationOperator.java:501-517) */
int32_t pg_hashagg_reset(pg_hashagg* op) {
    if (!op) return PG_E_INVALID_ARGUMENT;
    return guard(op->ctx, [&] {
        if (op->batch) { op->batch->rows = 0; op->batch->descs.clear(); for (PageBatch::Col& bc : op->batch->cols) { bc.values_used = 0; bc.has_nulls = false; } }
        for (pg_agg* a : op->aggs) delete a;
        op->aggs.clear();
        delete op->gb; op->gb = nullptr;
        delete op->wide; op->wide = nullptr;
        op->rows_since_reset = 0;
        op->lc.count = 0; op->lc.gb_current = true;
        hashagg_build(op);
    });
}

}  // extern "C"
