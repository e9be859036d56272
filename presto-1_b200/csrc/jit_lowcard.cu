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
struct SlotPlan { int kind, cond, conv, word; std::string lo, hi; bool twos; };
const char* LKIND_NAME[] = {"U64", "I32", "U8", "U128"};

int kind_words(int kind) { return kind == PGJ_KIND_DEC ? 3 : (kind == PGJ_KIND_I64 || kind == PGJ_KIND_I32) ? 2 : 1; }
const char* kind_name(int kind) {
    static const char* names[] = {"DEC", "I64", "I32", "F64", "MIN", "MAX", "CNT"};
    return names[kind];
}
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

bool accumulate_lowcard_jit(pg_ctx* ctx, const AggProgram& prog, int64_t n, int64_t group_count, const FusedLookup& fused, const LowcardExprs* px) {
    if (n <= 0 || group_count < 1 || group_count > 64 || prog.naggs > PGJ_MAX_AGGS) return false;
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
    int words = 0;
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
    auto find_slot = [&](int kind, const std::string& lo, const std::string& hi, int cond, int conv) -> int {
        for (const SlotPlan& s : slots)
            if (s.kind == kind && s.lo == lo && s.cond == cond && s.conv == conv) return s.word;
        if ((int)slots.size() >= PGJ_MAX_SLOTS) return -1;
        SlotPlan s{kind, cond, conv, words, lo, hi, px != nullptr};
        words += kind_words(kind);
        slots.push_back(s);
        return s.word;
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
            if (op.op != OP_COUNT) {
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
        int kind = PGJ_KIND_CNT, conv = 0;
        switch (op.op) {
            case OP_COUNT: break;
            case OP_SUM_I64: kind = is32 ? PGJ_KIND_I32 : PGJ_KIND_I64; break;
            case OP_SUM_F64:
                if (is32) return false;
                kind = PGJ_KIND_F64; conv = op.convert_i2d ? 1 : 0;
                break;
            case OP_SUM_DEC: kind = op.in_type == PG_LONG_DECIMAL ? PGJ_KIND_DEC : PGJ_KIND_I64; break;
            case OP_MIN64: case OP_MAX64:
                if (is32 || op.in_type == PG_BOOLEAN) return false;
                kind = op.op == OP_MIN64 ? PGJ_KIND_MIN : PGJ_KIND_MAX; conv = op.in_type == PG_DOUBLE ? 1 : 0;
                break;
            default: return false;   // 128-bit min/max need two passes
        }
        args.agg_state[a] = op.state; args.agg_op[a] = op.op; args.agg_words[a] = op.words;
        args.agg_cnt_word[a] = find_slot(PGJ_KIND_CNT, "", "", cond, 0);
        args.agg_val_word[a] = -1; args.agg_val_kind[a] = PGJ_KIND_CNT;
        if (args.agg_cnt_word[a] < 0) return false;
        if (op.op != OP_COUNT) {
            args.agg_val_word[a] = find_slot(kind, lo, hi, cond, conv);
            args.agg_val_kind[a] = kind;
            if (args.agg_val_word[a] < 0) return false;
        }
    }
    args.naggs = prog.naggs;

    // ---- block shape: the largest block whose private tables fit; rows per thread and trip limited by alignment and registers
    const int G = (int)group_count;
    const int64_t per_thread = (int64_t)words * G * 8;
    const int64_t budget = 220 * 1024;   // 227 KB per block minus the static look-up table
    int T = 512;
    while (T > 64 && per_thread * T > budget) T >>= 1;
    if (per_thread * T > budget) return false;
    int row_regs = 0;
    for (const LoadPlan& l : loads) row_regs += lkind_regs(l.lkind) + (l.nulls ? 1 : 0);
    row_regs += 3 * L.nkeys;
    int vec = 4;
    while (vec > 1 && vec * row_regs > 84) vec >>= 1;
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
    snprintf(buf, sizeof(buf), "#define PG_T %d\n#define PG_VEC %d\n#define PG_MIN_BLOCKS 1\n#define PG_REC_WORDS %d\n#define PG_PREFETCH %d\n", T, vec,
             words, prefetch);
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
    for (const SlotPlan& s : slots) {
        snprintf(buf, sizeof(buf), " if (pgc%d) { ", s.cond);
        pre += buf;
        snprintf(buf, sizeof(buf), "rec + %d * PG_T", s.word);
        const std::string at = buf;
        switch (s.kind) {
            case PGJ_KIND_DEC:
                if (s.twos) pre += "upd_dec2c(" + at + ", " + s.lo + ", (i64)(" + s.hi + "));";
                else pre += "upd_dec(" + at + ", " + s.lo + ", " + s.hi + ");";
                break;
            case PGJ_KIND_I64: case PGJ_KIND_I32: pre += "upd_i64(" + at + ", " + s.lo + ");"; break;
            case PGJ_KIND_F64: pre += "upd_f64(" + at + ", " + s.lo + ", " + (s.conv ? "1" : "0") + ");"; break;
            case PGJ_KIND_MIN: pre += "upd_min(" + at + ", " + s.lo + ", " + (s.conv ? "1" : "0") + ");"; break;
            case PGJ_KIND_MAX: pre += "upd_max(" + at + ", " + s.lo + ", " + (s.conv ? "1" : "0") + ");"; break;
            default: snprintf(buf, sizeof(buf), "rec[%d * PG_T] += 1;", s.word); pre += buf;
        }
        pre += " }";
    }
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
    args.key_rows = (const unsigned long long*)fused.key_rows; args.arena = fused.arena;
    args.values_by_group = (const long long*)fused.values_by_group;
    args.miss_flag = fused.miss_flag; args.d_miss = fused.d_miss;
    args.n = n; args.row_base = fused.row_base;
    args.G = G; args.bigint_mode = fused.bigint_mode; args.null_gid = fused.null_gid; args.row_words = L.row_words;

    const size_t smem = (size_t)(per_thread * T);
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
    void* kargs[] = {&args};
    jit_launch(ctx, kern, "lowcard_agg_jit", (int)grid, T, smem, kargs);
    return true;
}

}  // namespace pg
