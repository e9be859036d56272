// jit_lowcard.cu — plans, generates and launches the run-time specialised low-cardinality GROUP BY kernel
// (jit/lowcard_body.cuh).  The plan mirrors what the reference's generated GroupedAccumulator classes hard-code per
// aggregation (AccumulatorCompiler.java:487-566: which channels are read, which mask applies, which state is updated);
// identical pieces of state are shared (sum(x) and avg(x) keep one sum, every aggregate under one mask keeps one count).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "accum.h"
#include "groupby.h"
#include "jit.h"
#include "project.h"
#include "jit/lowcard_args.h"
#include "jit_embedded_gen.h"   // LOWCARD_ARGS_SRC, LOWCARD_BODY_SRC (written by build.py from csrc/jit/)

namespace pg {

namespace {

struct LoadPlan { const void* ptr; const uint8_t* nulls; int lkind; };     // lkind: 0 U64, 1 I32, 2 U8, 3 U128
struct SlotPlan { int kind, cond, conv, word; std::string lo, hi; int form; int wide_idx; };
const char* LKIND_NAME[] = {"U64", "I32", "U8", "U128"};

const char* kind_name(int kind) {
    static const char* names[] = {"SUM", "?", "?", "F64", "MIN", "MAX", "CNT"};
    return names[kind];
}
// how a SUM slot's addend arrives
enum { SUM_I64 = 0 /* 64-bit signed */, SUM_SM128 = 1 /* sign-magnitude 128 (Int128ArrayBlock) */, SUM_2C128 = 2 /* two's complement 128 (expression) */ };
// registers one row of this load keeps live between the load and its use
int lkind_regs(int lkind) { return lkind == 3 ? 4 : lkind == 0 ? 2 : 1; }
int lkind_of_type(int32_t t) {
    switch (t) {
        case PG_INTEGER: case PG_DATE: return 1;
        case PG_BOOLEAN: return 2;
        case PG_LONG_DECIMAL: return 3;
        case PG_VARCHAR: case PG_VARBINARY: return -1;
        default: return 0;
    }
}

const char* key_class(int32_t type) {
    switch (type) {
        case PG_BIGINT: case PG_SHORT_DECIMAL: return "K64";
        case PG_DOUBLE: return "KDBL";
        case PG_INTEGER: case PG_DATE: return "K32";
        case PG_BOOLEAN: return "K8";
        case PG_VARCHAR: case PG_VARBINARY: return "KVAR";
        default: return nullptr;
    }
}

bool aligned(const void* p, int64_t bytes) { return ((uintptr_t)p % (uintptr_t)bytes) == 0; }

std::string one_line(const std::string& s) { std::string r = s; for (char& c : r) if (c == '\n') c = ' '; return r; }

}  // namespace

int lowcard_max_threads() {
    int t = 512;     // measured on B200 (Q1, 6 words per group): 512 x 4 rows 0.63 ms, 768 x 2 0.66 ms, 1024 x 2 0.68 ms
    if (const char* e = getenv("PRESTOGPU_JIT_T")) { const int v = atoi(e); if (v >= 64 && v <= 1024 && v % 32 == 0) t = v; }
    return t;
}

bool accumulate_lowcard_jit(pg_ctx* ctx, const AggProgram& prog, int64_t n, const FusedLookup& fused, const LowcardGroups& groups, const LowcardExprs* px,
                            int* gcap_used) {
    if (n <= 0 || prog.naggs > PGJ_MAX_AGGS || groups.known_count > PGJ_MAX_GROUPS) return false;
    if (!jit_available()) return false;
    const KeyLayout& L = *fused.layout;
    int32_t kword[MAX_PACK_KEYS_PUBLIC], kshift[MAX_PACK_KEYS_PUBLIC];
    if (L.nkeys > PGJ_MAX_KEYS || !fused_pack_layout(L, kword, kshift)) return false;

    // ---- plan: input columns of the row loop, conditions, private slots, flush table
    std::vector<LoadPlan> loads;
    std::vector<std::string> conds;
    std::vector<SlotPlan> slots;
    PgjArgs args;
    memset(&args, 0, sizeof(args));
    int words = 0, nsums = 0;
    char buf[512];
    bool failed = false;
    auto find_load = [&](const DCol& c) -> int {
        const int lk = lkind_of_type(c.type);
        if (lk < 0 || c.values == nullptr) { failed = true; return 0; }
        for (size_t i = 0; i < loads.size(); i++)
            if (loads[i].ptr == c.values && loads[i].lkind == lk) { if (c.nulls && !loads[i].nulls) loads[i].nulls = c.nulls; return (int)i; }
        if ((int)loads.size() >= PGJ_MAX_LOADS) { failed = true; return 0; }
        loads.push_back(LoadPlan{c.values, c.nulls, lk});
        return (int)loads.size() - 1;
    };
    auto ld = [&](const char* pre, int l) { snprintf(buf, sizeof(buf), "%s%d[v]", pre, l); return std::string(buf); };
    auto find_cond = [&](const std::string& text) -> int {
        for (size_t i = 0; i < conds.size(); i++) if (conds[i] == text) return (int)i;
        conds.push_back(text);
        return (int)conds.size() - 1;
    };
    auto find_slot = [&](int kind, const std::string& lo, const std::string& hi, int cond, int conv, int form) -> int {
        for (size_t i = 0; i < slots.size(); i++)
            if (slots[i].kind == kind && slots[i].lo == lo && slots[i].cond == cond && slots[i].conv == conv && slots[i].form == form) return (int)i;
        if ((int)slots.size() >= PGJ_MAX_SLOTS) return -1;
        SlotPlan s{kind, cond, conv, words, lo, hi, form, kind == PGJ_KIND_SUM ? nsums++ : -1};
        words += 1;
        slots.push_back(s);
        return (int)slots.size() - 1;
    };
    // fused filter + projections: straight-line code over the processor's input columns
    ProjCode code;
    std::string filter_pass = "true";
    if (px) {
        const pg_project* proc = px->proc;
        std::vector<std::string> in_lo(PROJ_MAX_CHANNELS), in_hi(PROJ_MAX_CHANNELS), in_null(PROJ_MAX_CHANNELS);
        std::vector<int> roots;
        for (int a = 0; a < prog.naggs; a++) {
            if (px->arg_proj[a] >= 0) roots.push_back(proc->proj[px->arg_proj[a]]);
            if (px->mask_proj[a] >= 0) roots.push_back(proc->proj[px->mask_proj[a]]);
        }
        // only the channels those expressions (and the filter) reach are loaded
        std::vector<char> needed(proc->nodes.size(), 0);
        for (int r : roots) needed[r] = 1;
        if (proc->filter >= 0) needed[proc->filter] = 1;
        for (int i = (int)proc->nodes.size() - 1; i >= 0; i--)
            if (needed[i]) { if (proc->nodes[i].a0 >= 0) needed[proc->nodes[i].a0] = 1; if (proc->nodes[i].a1 >= 0) needed[proc->nodes[i].a1] = 1; }
        for (size_t i = 0; i < proc->nodes.size(); i++) {
            const PNode& e = proc->nodes[i];
            if (!needed[i] || e.op != PG_EXPR_INPUT || !in_lo[e.channel].empty()) continue;
            const DCol& c = px->in[e.channel];
            if (c.type != e.type) return false;
            const int l = find_load(c);
            if (failed) return false;
            in_lo[e.channel] = ld("ld", l);
            in_hi[e.channel] = ld("ldh", l);
            if (c.nulls) in_null[e.channel] = "(" + ld("ldn", l) + " != 0)";
        }
        if (!project_codegen(proc, roots, in_lo, in_hi, in_null, code)) return false;
        if (proc->filter >= 0) {
            const std::string& fn = code.null[proc->filter];
            filter_pass = "(" + code.lo[proc->filter] + " != 0" + (fn == "false" ? "" : " && !" + fn) + ")";
        }
    }
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        // what the aggregate reads: a column (plain operator) or a projection of the fused processor
        std::string lo, hi, arg_null = "false", cond_text;
        if (px) {
            if (px->mask_proj[a] >= 0) {
                const int node = px->proc->proj[px->mask_proj[a]];
                if (px->proc->nodes[node].type != PG_BOOLEAN) return false;
                cond_text = "(" + code.lo[node] + " != 0)";
                if (code.null[node] != "false") cond_text += " && !" + code.null[node];
            }
            if (px->arg_proj[a] >= 0) {
                const int node = px->proc->proj[px->arg_proj[a]];
                if (px->proc->nodes[node].type != op.in_type) return false;
                lo = code.lo[node]; hi = code.hi[node]; arg_null = code.null[node];
            }
        } else {
            if (op.mask.values) {
                const int l = find_load(op.mask);
                cond_text = "(" + ld("ld", l) + " != 0)";
                if (op.mask.nulls) cond_text += " && !" + ld("ldn", l);
            }
            if (op.op != OP_COUNT || (op.skip_null && op.arg.values && op.arg.nulls)) {      // count(col) only needs the null flags
                const int l = find_load(op.arg);
                lo = ld("ld", l); hi = ld("ldh", l);
                if (op.arg.nulls) arg_null = "(" + ld("ldn", l) + " != 0)";
            }
            if (failed) return false;
        }
        if (op.skip_null && arg_null != "false") cond_text += std::string(cond_text.empty() ? "" : " && ") + "!" + arg_null;
        if (cond_text.empty()) cond_text = "true";
        const int cond = find_cond(cond_text);
        const bool is32 = op.in_type == PG_INTEGER || op.in_type == PG_DATE;
        int kind = PGJ_KIND_CNT, conv = 0, form = SUM_I64;
        switch (op.op) {
            case OP_COUNT: break;
            case OP_SUM_I64: kind = PGJ_KIND_SUM; break;      // 32-bit columns are sign-extended when they are loaded
            case OP_SUM_F64:
                if (is32) return false;
                kind = PGJ_KIND_F64; conv = op.convert_i2d ? 1 : 0;
                break;
            case OP_SUM_DEC: kind = PGJ_KIND_SUM; form = op.in_type == PG_LONG_DECIMAL ? (px ? SUM_2C128 : SUM_SM128) : SUM_I64; break;
            case OP_MIN64: case OP_MAX64:
                if (is32 || op.in_type == PG_BOOLEAN) return false;
                kind = op.op == OP_MIN64 ? PGJ_KIND_MIN : PGJ_KIND_MAX; conv = op.in_type == PG_DOUBLE ? 1 : 0;
                break;
            default: return false;   // 128-bit min/max need two passes
        }
        args.agg_state[a] = op.state; args.agg_op[a] = op.op; args.agg_words[a] = op.words;
        const int cs = find_slot(PGJ_KIND_CNT, "", "", cond, 0, 0);
        if (cs < 0) return false;
        args.agg_cnt_word[a] = slots[cs].word;
        args.agg_val_word[a] = -1; args.agg_val_kind[a] = PGJ_KIND_CNT; args.agg_wide_idx[a] = 0;
        if (op.op != OP_COUNT) {
            const int vs = find_slot(kind, lo, hi, cond, conv, form);
            if (vs < 0) return false;
            args.agg_val_word[a] = slots[vs].word;
            args.agg_val_kind[a] = kind;
            args.agg_wide_idx[a] = slots[vs].wide_idx < 0 ? 0 : slots[vs].wide_idx;
        }
    }
    args.naggs = prog.naggs;

    // ---- block shape: the largest block whose private tables fit; rows per thread and trip limited by alignment and registers.
    // Known groups size the tables exactly; a discovering launch takes the largest block and whatever group capacity that leaves
    // (the kernel reports how many groups it found when they do not fit, and the caller comes back with them as known groups).
    const int64_t budget = 216 * 1024;   // 227 KB per block minus the static look-up table and a little slack
    const int64_t word_bytes = (int64_t)words * 8;
    int T = lowcard_max_threads(), G;
    auto fits_block = [&](int t, int g) { return word_bytes * g * t + (int64_t)g * (nsums > 0 ? nsums : 1) * 24 <= budget; };
    if (groups.known_count > 0) {
        G = groups.known_count;
        while (T > 64 && !fits_block(T, G)) T -= T > 256 ? 128 : 64;
        if (!fits_block(T, G)) return false;
    } else {
        G = groups.capacity_hint > 0 ? groups.capacity_hint : 4;
        if (G > PGJ_MAX_GROUPS) G = PGJ_MAX_GROUPS;
        while (T > 64 && !fits_block(T, G)) T -= T > 256 ? 128 : 64;
        if (!fits_block(T, G)) return false;
        while (G < PGJ_MAX_GROUPS && fits_block(T, G + 1)) G++;          // use what the block has room for
    }
    if (gcap_used) *gcap_used = G;
    const int64_t per_thread = word_bytes * G;
    int row_regs = 0;
    for (const LoadPlan& l : loads) row_regs += lkind_regs(l.lkind) + (l.nulls ? 1 : 0);
    row_regs += 3 * L.nkeys;
    int vec = 4;
    const int regs_for_rows = 65536 / T - (T > 512 ? 30 : 44);       // what the block size leaves beside the loop's fixed registers
    while (vec > 1 && vec * row_regs > regs_for_rows) vec >>= 1;
    auto fits_vec = [&](int v) {
        if (v == 1) return true;
        for (const LoadPlan& l : loads) {
            if (!aligned(l.ptr, l.lkind == 1 ? 4 * v : l.lkind == 2 ? v : 16)) return false;
            if (l.nulls && !aligned(l.nulls, v)) return false;
        }
        for (int k = 0; k < L.nkeys; k++) {
            const DCol& c = fused.keys->c[k];
            if (c.nulls && !aligned(c.nulls, v)) return false;
            switch (c.type) {
                case PG_VARCHAR: case PG_VARBINARY:
                    if (!aligned(c.offsets, 4 * v) || !aligned((const char*)c.values + fused.row_base, v)) return false;
                    break;
                case PG_INTEGER: case PG_DATE: if (!aligned(c.values, 4 * v)) return false; break;
                case PG_BOOLEAN: if (!aligned(c.values, v)) return false; break;
                default: if (!aligned(c.values, 16)) return false;
            }
        }
        return true;
    };
    while (!fits_vec(vec)) vec >>= 1;

    // ---- source = generated prelude + argument block + body; the prelude doubles as the cache signature
    std::string pre;
    int prefetch = 1;
    if (const char* e = getenv("PRESTOGPU_JIT_PREFETCH")) prefetch = atoi(e);
    if (const char* e = getenv("PRESTOGPU_JIT_VEC")) { int v = atoi(e); if ((v == 1 || v == 2 || v == 4) && v <= vec) vec = v; }
    // a thread's private 64-bit sums take addends below 2^PG_FAST_BITS in magnitude: rows per thread x 2^PG_FAST_BITS stays below 2^62
    const int64_t rows_per_thread = (n + (int64_t)ctx->sm_count * T - 1) / ((int64_t)ctx->sm_count * T) + 2 * vec + 64;
    int fast = 61;
    for (int64_t r = 1; r < rows_per_thread; r <<= 1) fast--;
    fast = fast >= 38 ? 38 : fast < 8 ? 8 : fast;      // 38 in all but absurdly long pages: one kernel per plan, not per page size
    snprintf(buf, sizeof(buf), "#define PG_T %d\n#define PG_VEC %d\n#define PG_MIN_BLOCKS 1\n#define PG_REC_WORDS %d\n#define PG_NSUMS %d\n#define PG_PREFETCH %d\n#define PG_FAST_BITS %d\n",
             T, vec, words, nsums > 0 ? nsums : 1, prefetch, fast);
    pre += buf;
    pre += "#define PG_KEYS(X)";
    for (int k = 0; k < L.nkeys; k++) {
        const DCol& c = fused.keys->c[k];
        const char* cls = key_class(c.type);
        if (!cls) return false;
        snprintf(buf, sizeof(buf), " X(%d, %s, %d, %d, %d, %d)", k, cls, kword[k], kshift[k], c.nulls ? 1 : 0, L.word_off[k]);
        pre += buf;
    }
    pre += "\n#define PG_LOADS(X)";
    for (size_t l = 0; l < loads.size(); l++) {
        snprintf(buf, sizeof(buf), " X(%d, %s, %d)", (int)l, LKIND_NAME[loads[l].lkind], loads[l].nulls ? 1 : 0);
        pre += buf;
    }
    pre += "\n#define PG_SLOTS(X)";
    for (size_t s = 0; s < slots.size(); s++) {
        snprintf(buf, sizeof(buf), " X(%d, %s, %d)", (int)s, kind_name(slots[s].kind), slots[s].word);
        pre += buf;
    }
    // the two statement macros: (1) the filter's part of the expressions and the verdict, (2) the rest, the conditions, the updates
    std::string filter_code, rest_code;
    if (px) {
        const std::string text = one_line(code.decl);
        const size_t cut = text.find("PGX_FILTER_DONE");
        filter_code = text.substr(0, cut);
        rest_code = cut == std::string::npos ? "" : text.substr(cut + strlen("PGX_FILTER_DONE"));
    }
    pre += "\n#define PG_ROW_FILTER " + filter_code + " const bool pgx_pass = " + filter_pass + ";";
    pre += "\n#define PG_ROW_UPDATE " + rest_code;
    for (size_t c = 0; c < conds.size(); c++) {
        snprintf(buf, sizeof(buf), " const bool pgc%d = ", (int)c);
        pre += buf + conds[c] + ";";
    }
    // Updates are branch-free selects on the thread's own words; the addends the narrow sums cannot take (rare) are collected in
    // one bit mask per row and handled behind a single branch.
    std::string slow_code;
    bool any_sum = false;
    for (size_t si = 0; si < slots.size(); si++) {
        const SlotPlan& s = slots[si];
        snprintf(buf, sizeof(buf), "rec[%d * PG_T]", s.word);
        const std::string at = buf;
        snprintf(buf, sizeof(buf), "rec + %d * PG_T", s.word);
        const std::string atp = buf;
        snprintf(buf, sizeof(buf), "pgc%d", s.cond);
        const std::string c = buf;
        switch (s.kind) {
            case PGJ_KIND_SUM: {
                any_sum = true;
                snprintf(buf, sizeof(buf), "%d", (int)si);
                const std::string id = buf;
                std::string w0 = s.lo, w1;
                if (s.form == SUM_SM128) {
                    pre += " const Pgx2c pgv" + id + " = pgx2c_from_long(" + s.lo + ", " + s.hi + ");";
                    w0 = "pgv" + id + ".w0"; w1 = "pgv" + id + ".w1";
                } else if (s.form == SUM_2C128) w1 = "(i64)(" + s.hi + ")";
                const std::string ok = w1.empty() ? "fits_fast(" + w0 + ")" : "(fits_fast(" + w0 + ") && " + w1 + " == ((i64)(" + w0 + ") >> 63))";
                pre += " const bool pgo" + id + " = " + ok + "; " + at + " += (" + c + " && pgo" + id + ") ? (u64)(" + w0 + ") : 0ULL;";
                pre += " pg_slow |= (" + c + " && !pgo" + id + ") ? " + "(1u << " + id + ") : 0u;";
                snprintf(buf, sizeof(buf), "wrec + %d", (s.wide_idx < 0 ? 0 : s.wide_idx) * 3);
                const std::string hi_ext = w1.empty() ? "(u64)((i64)(" + w0 + ") >> 63)" : "(u64)(" + w1 + ")";
                const std::string top = w1.empty() ? hi_ext : "(u64)((i64)(" + w1 + ") >> 63)";
                slow_code += " if (pg_slow & (1u << " + id + ")) wide_add(" + std::string(buf) + ", (u64)(" + w0 + "), " + hi_ext + ", " + top + ");";
                break;
            }
            case PGJ_KIND_F64: pre += " if (" + c + ") upd_f64(" + atp + ", " + s.lo + ", " + (s.conv ? "1" : "0") + ");"; break;
            case PGJ_KIND_MIN: pre += " if (" + c + ") upd_min(" + atp + ", " + s.lo + ", " + (s.conv ? "1" : "0") + ");"; break;
            case PGJ_KIND_MAX: pre += " if (" + c + ") upd_max(" + atp + ", " + s.lo + ", " + (s.conv ? "1" : "0") + ");"; break;
            default: pre += " " + at + " += " + c + " ? 1ULL : 0ULL;";
        }
    }
    if (any_sum) pre += " if (pg_slow) {" + slow_code + " }";
    pre += "\n";
    JitKernel* kern = jit_get(ctx, pre, pre + DECIMAL_OPS_SRC + LOWCARD_ARGS_SRC + LOWCARD_BODY_SRC, "pg_lowcard_agg");
    if (!kern) return false;

    // ---- arguments
    for (int k = 0; k < L.nkeys; k++) {
        const DCol& c = fused.keys->c[k];
        args.key_values[k] = c.values; args.key_nulls[k] = c.nulls; args.key_offsets[k] = c.offsets;
        args.key_value_bytes[k] = fused.key_value_bytes[k];
    }
    for (size_t l = 0; l < loads.size(); l++) { args.load_ptr[l] = loads[l].ptr; args.load_nulls[l] = loads[l].nulls; }
    args.lc_groups = groups.known; args.lc_groups_out = groups.discovered_out;
    args.miss_flag = fused.miss_flag; args.d_miss = fused.d_miss;
    args.n = n; args.row_base = fused.row_base;
    args.discover_rows = groups.known_count > 0 ? 0 : (groups.discover_rows < n ? groups.discover_rows : n);
    args.G = G; args.known_groups = groups.known_count;

    const size_t smem = (size_t)(per_thread * T + (int64_t)G * (nsums > 0 ? nsums : 1) * 24);
    if ((int)smem > kern->max_dyn_smem) {
        PG_CUDA(cudaFuncSetAttribute((const void*)kern->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern->max_dyn_smem = (int)smem;
    }
    int blocks_per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, (const void*)kern->fn, T, smem) != cudaSuccess || blocks_per_sm < 1) {
        cudaGetLastError();
        blocks_per_sm = 1;
    }
    int64_t want = (n + (int64_t)T * vec - 1) / ((int64_t)T * vec);
    int64_t grid = (int64_t)ctx->sm_count * blocks_per_sm;
    if (grid > want) grid = want;
    if (grid < ctx->sm_count && grid < want) return false;      // (cannot happen: the bound on rows per thread assumes a full grid)
    args.fast_bits = fast;
    void* kargs[] = {&args};
    jit_launch(ctx, kern, "lowcard_agg_jit", (int)grid, T, smem, kargs);
    return true;
}

// ------------------------------------------------------------------ packed-key group tables
namespace {
struct PackInfo { int32_t nkeys, row_words, bigint_mode, pad; int32_t types[8], word[8], shift[8], woff[8]; };

PG_D unsigned long long lc_norm_double(unsigned long long b) {
    if ((b & 0x7fffffffffffffffULL) == 0) return 0;
    if ((b & 0x7ff0000000000000ULL) == 0x7ff0000000000000ULL && (b & 0x000fffffffffffffULL)) return 0x7ff8000000000000ULL;
    return b;
}
PG_D int lc_width(int32_t t) {
    switch (t) {
        case PG_BOOLEAN: return 8;
        case PG_INTEGER: case PG_DATE: case PG_VARCHAR: case PG_VARBINARY: return 32;
        default: return 64;
    }
}

// stored keys of the GroupByHash (groupby.h layouts) -> packed form, the same the kernel computes for input rows
__global__ void lc_pack_groups_kernel(PackInfo pi, const uint64_t* __restrict__ key_rows, const uint8_t* __restrict__ arena,
                                      const int64_t* __restrict__ values_by_group, const Slot* __restrict__ null_slot, int64_t groups,
                                      unsigned long long* __restrict__ out) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= groups) return;
    unsigned long long w0 = 0, w1 = 0, ok = 1;
    if (pi.bigint_mode) {
        if (null_slot && null_slot->gid == (int32_t)g) w1 |= 1ULL << 56; else w0 = (unsigned long long)values_by_group[g];
    } else {
        const uint64_t* rowp = key_rows + g * pi.row_words;
        for (int k = 0; k < pi.nkeys; k++) {
            if ((rowp[0] >> k) & 1) { w1 |= 1ULL << (56 + k); continue; }
            const unsigned long long w = rowp[pi.woff[k]];
            unsigned long long v = 0;
            switch (pi.types[k]) {
                case PG_DOUBLE: v = lc_norm_double(w); break;
                case PG_INTEGER: case PG_DATE: v = (unsigned long long)(uint32_t)w; break;
                case PG_BOOLEAN: v = w & 0xff; break;
                case PG_VARCHAR: case PG_VARBINARY: {
                    const int len = (int)(w & 0xFFFFFF);
                    if (len > 3) ok = 0;
                    else { const uint8_t* b = arena + (w >> 24); v = (unsigned long long)len; for (int t = 0; t < len; t++) v |= (unsigned long long)b[t] << (8 + 8 * t); }
                    break;
                }
                default: v = w;
            }
            if (pi.word[k] == 0) w0 |= v << pi.shift[k]; else w1 |= v << pi.shift[k];
        }
    }
    out[4 * g] = w0; out[4 * g + 1] = w1; out[4 * g + 2] = 0; out[4 * g + 3] = ok;
}

// packed form -> key columns for groups [first, first + n) (n <= 64; one block)
__global__ void lc_unpack_groups_kernel(PackInfo pi, const unsigned long long* __restrict__ groups, int64_t first, int64_t n, LowcardKeyOut out) {
    const int i = threadIdx.x;
    if (i < n) {
        const unsigned long long w0 = groups[4 * (first + i)], w1 = groups[4 * (first + i) + 1];
        for (int k = 0; k < pi.nkeys; k++) {
            const bool isnull = (w1 >> (56 + k)) & 1;
            const int width = lc_width(pi.types[k]);
            unsigned long long v = (pi.word[k] == 0 ? w0 : w1) >> pi.shift[k];
            if (width < 64) v &= (1ULL << width) - 1;
            if (isnull) v = 0;
            if (out.nulls[k]) out.nulls[k][i] = isnull ? 1 : 0;
            switch (pi.types[k]) {
                case PG_INTEGER: case PG_DATE: ((int32_t*)out.values[k])[i] = (int32_t)v; break;
                case PG_BOOLEAN: ((uint8_t*)out.values[k])[i] = (uint8_t)v; break;
                case PG_VARCHAR: case PG_VARBINARY: break;
                default: ((unsigned long long*)out.values[k])[i] = v;
            }
        }
    }
    if (i == 0) {
        for (int k = 0; k < pi.nkeys; k++) {
            if (pi.types[k] != PG_VARCHAR && pi.types[k] != PG_VARBINARY) continue;
            int32_t off = 0;
            for (int64_t r = 0; r < n; r++) {
                const unsigned long long w0 = groups[4 * (first + r)], w1 = groups[4 * (first + r) + 1];
                const bool isnull = (w1 >> (56 + k)) & 1;
                const unsigned long long v = ((pi.word[k] == 0 ? w0 : w1) >> pi.shift[k]) & 0xffffffffULL;
                const int len = isnull ? 0 : (int)(v & 0xff);
                out.offsets[k][r] = off;
                for (int t = 0; t < len; t++) ((uint8_t*)out.values[k])[off + t] = (uint8_t)(v >> (8 + 8 * t));
                off += len;
            }
            out.offsets[k][n] = off;
        }
    }
}

bool make_pack_info(const KeyLayout& L, bool bigint_mode, PackInfo& pi) {
    memset(&pi, 0, sizeof(pi));
    int32_t kword[MAX_PACK_KEYS_PUBLIC], kshift[MAX_PACK_KEYS_PUBLIC];
    if (L.nkeys > 8 || !fused_pack_layout(L, kword, kshift)) return false;
    pi.nkeys = L.nkeys; pi.row_words = L.row_words; pi.bigint_mode = bigint_mode ? 1 : 0;
    for (int k = 0; k < L.nkeys; k++) { pi.types[k] = L.types[k]; pi.word[k] = kword[k]; pi.shift[k] = kshift[k]; pi.woff[k] = L.word_off[k]; }
    return true;
}
}  // namespace

void lowcard_pack_groups(pg_ctx* ctx, const pg_groupby* gb, int64_t groups, unsigned long long* d_out) {
    PackInfo pi;
    PG_CHECK_STATE(make_pack_info(gb->layout, gb->bigint_mode, pi), "group-by keys do not fit the packed form");
    if (groups <= 0) return;
    const Slot* null_slot = gb->bigint_mode && gb->slots.p ? gb->slots.as<Slot>() + gb->cap + 1 : nullptr;
    PG_LAUNCH(ctx, lc_pack_groups_kernel, (int)((groups + 63) / 64), 64, 0, pi, gb->key_rows.as<uint64_t>(), gb->arena.as<uint8_t>(), gb->values_by_group.as<int64_t>(),
              null_slot, groups, d_out);
}

void lowcard_unpack_groups(pg_ctx* ctx, const KeyLayout& L, const unsigned long long* d_groups, int64_t first, int64_t n, const LowcardKeyOut& out) {
    PackInfo pi;
    PG_CHECK_STATE(make_pack_info(L, L.nkeys == 1 && L.types[0] == PG_BIGINT, pi), "group-by keys do not fit the packed form");
    PG_CHECK_ARG(n >= 0 && n <= PGJ_MAX_GROUPS, "at most %d groups", PGJ_MAX_GROUPS);
    if (n == 0) return;
    PG_LAUNCH(ctx, lc_unpack_groups_kernel, 1, 64, 0, pi, d_groups, first, n, out);
}

}  // namespace pg
