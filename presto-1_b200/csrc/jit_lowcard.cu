// jit_lowcard.cu — plans, generates and launches the run-time specialised low-cardinality GROUP BY kernel
// (jit/lowcard_body.cuh).  The plan mirrors what the reference's generated GroupedAccumulator classes hard-code per
// aggregation (AccumulatorCompiler.java:487-566: which channels are read, which mask applies, which state is updated);
// identical pieces of state are shared (sum(x) and avg(x) keep one sum, every aggregate under one mask keeps one count).
#include <stdio.h>
#include <stdlib.h>

#include "accum.h"
#include "groupby.h"
#include "jit.h"
#include "jit/lowcard_args.h"
#include "jit_embedded_gen.h"   // LOWCARD_ARGS_SRC, LOWCARD_BODY_SRC (written by build.py from csrc/jit/)

namespace pg {

namespace {

struct SlotPlan { int kind, cond, conv, word; const void* ptr; };
struct CondPlan { const uint8_t* mask; const uint8_t* mask_nulls; const uint8_t* arg_nulls; };

int kind_words(int kind) { return kind == PGJ_KIND_DEC ? 3 : (kind == PGJ_KIND_I64 || kind == PGJ_KIND_I32) ? 2 : 1; }
const char* kind_name(int kind) {
    static const char* names[] = {"DEC", "I64", "I32", "F64", "MIN", "MAX", "CNT"};
    return names[kind];
}
// registers one row of this slot keeps live between its load and its update
int kind_regs(int kind) { return kind == PGJ_KIND_DEC ? 4 : kind == PGJ_KIND_CNT ? 0 : 2; }

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

}  // namespace

bool accumulate_lowcard_jit(pg_ctx* ctx, const AggProgram& prog, int64_t n, int64_t group_count, const FusedLookup& fused) {
    if (n <= 0 || group_count < 1 || group_count > 64 || prog.naggs > PGJ_MAX_AGGS) return false;
    if (!jit_available()) return false;
    const KeyLayout& L = *fused.layout;
    int32_t kword[MAX_PACK_KEYS_PUBLIC], kshift[MAX_PACK_KEYS_PUBLIC];
    if (L.nkeys > PGJ_MAX_KEYS || !fused_pack_layout(L, kword, kshift)) return false;

    // ---- plan: conditions, private slots, flush table
    std::vector<CondPlan> conds;
    std::vector<SlotPlan> slots;
    PgjArgs args;
    memset(&args, 0, sizeof(args));
    int words = 0;
    auto find_slot = [&](int kind, const void* ptr, int cond, int conv) -> int {
        for (const SlotPlan& s : slots)
            if (s.kind == kind && s.ptr == ptr && s.cond == cond && s.conv == conv) return s.word;
        if ((int)slots.size() >= PGJ_MAX_SLOTS) return -1;
        SlotPlan s{kind, cond, conv, words, ptr};
        words += kind_words(kind);
        slots.push_back(s);
        return s.word;
    };
    for (int a = 0; a < prog.naggs; a++) {
        const AggOp& op = prog.ops[a];
        CondPlan c;
        c.mask = (const uint8_t*)op.mask.values;
        c.mask_nulls = op.mask.values ? op.mask.nulls : nullptr;
        c.arg_nulls = op.skip_null ? op.arg.nulls : nullptr;
        int cond = -1;
        for (size_t i = 0; i < conds.size(); i++)
            if (conds[i].mask == c.mask && conds[i].mask_nulls == c.mask_nulls && conds[i].arg_nulls == c.arg_nulls) cond = (int)i;
        if (cond < 0) {
            if ((int)conds.size() >= PGJ_MAX_CONDS) return false;
            cond = (int)conds.size();
            conds.push_back(c);
        }
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
        args.agg_cnt_word[a] = find_slot(PGJ_KIND_CNT, nullptr, cond, 0);
        args.agg_val_word[a] = -1; args.agg_val_kind[a] = PGJ_KIND_CNT;
        if (args.agg_cnt_word[a] < 0) return false;
        if (op.op != OP_COUNT) {
            args.agg_val_word[a] = find_slot(kind, op.arg.values, cond, conv);
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
    for (const SlotPlan& s : slots) row_regs += kind_regs(s.kind);
    row_regs += 3 * L.nkeys + (int)conds.size();
    int vec = 4;
    while (vec > 1 && vec * row_regs > 84) vec >>= 1;
    auto fits_vec = [&](int v) {
        if (v == 1) return true;
        for (const SlotPlan& s : slots) {
            if (s.kind == PGJ_KIND_CNT) continue;
            if (!aligned(s.ptr, s.kind == PGJ_KIND_I32 ? 4 * v : 16)) return false;
        }
        for (const CondPlan& c : conds)
            if ((c.mask && !aligned(c.mask, v)) || (c.mask_nulls && !aligned(c.mask_nulls, v)) || (c.arg_nulls && !aligned(c.arg_nulls, v))) return false;
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
    char buf[256];
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
    pre += "\n#define PG_CONDS(X)";
    for (size_t c = 0; c < conds.size(); c++) {
        snprintf(buf, sizeof(buf), " X(%d, %d, %d, %d)", (int)c, conds[c].mask ? 1 : 0, conds[c].mask_nulls ? 1 : 0, conds[c].arg_nulls ? 1 : 0);
        pre += buf;
    }
    pre += "\n#define PG_SLOTS(X)";
    for (size_t s = 0; s < slots.size(); s++) {
        snprintf(buf, sizeof(buf), " X(%d, %s, %d, %d, %d)", (int)s, kind_name(slots[s].kind), slots[s].cond, slots[s].conv, slots[s].word);
        pre += buf;
    }
    pre += "\n";
    JitKernel* kern = jit_get(ctx, pre, pre + LOWCARD_ARGS_SRC + LOWCARD_BODY_SRC, "pg_lowcard_agg");
    if (!kern) return false;

    // ---- arguments
    for (int k = 0; k < L.nkeys; k++) {
        const DCol& c = fused.keys->c[k];
        args.key_values[k] = c.values; args.key_nulls[k] = c.nulls; args.key_offsets[k] = c.offsets;
        args.key_value_bytes[k] = fused.key_value_bytes[k];
    }
    for (size_t c = 0; c < conds.size(); c++) {
        args.cond_mask[c] = conds[c].mask; args.cond_mask_nulls[c] = conds[c].mask_nulls; args.cond_arg_nulls[c] = conds[c].arg_nulls;
    }
    for (size_t s = 0; s < slots.size(); s++) args.slot_ptr[s] = slots[s].ptr;
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
