// accum.h — grouped accumulators (replaces the classes AccumulatorCompiler.java:261-330,487-566 generates for
// count/sum/avg/min/max and their state classes under MAIN/operator/aggregation/state/).
//
// HBM layout: one state array per aggregate, `words` u64 words per group (array of structs):
//   COUNT, COUNT_COL               [cnt]
//   SUM(bigint|integer)            [cnt, lo, hi]           128-bit two's complement running sum
//   SUM(double), AVG(double|bigint)[cnt, sum(double bits)]
//   SUM/AVG(decimal)               [cnt, w0, w1, w2]       192-bit two's complement running sum; converted to the
//                                                          reference's (overflow, sign-magnitude 127 bit) state on output
//   MIN/MAX (64-bit types)         [cnt, key]              key = value, or the Double.compare order key for doubles
//   MIN/MAX (long decimal)         [cnt, hi, lo]           two's complement, resolved in two passes (hi, then lo)
// Multi-word sums are updated with one atomicAdd per word plus carry propagation; integer addition commutes, so the
// final state is independent of the order in which rows arrive.
#pragma once
#include "context.h"

struct pg_project;
struct pg_groupby;

namespace pg {

enum OpClass : int32_t {
    OP_COUNT = 0,    // cnt += 1
    OP_SUM_I64 = 1,  // cnt, 128-bit sum
    OP_SUM_F64 = 2,  // cnt, double sum
    OP_SUM_DEC = 3,  // cnt, 192-bit sum
    OP_MIN64 = 4,
    OP_MAX64 = 5,
    OP_MIN128 = 6,
    OP_MAX128 = 7,
};

static const int MAX_AGGS = 16;

struct AggOp {
    int32_t op;         // OpClass
    int32_t in_type;    // PG_* of the argument (0 for count(*))
    int32_t words;      // u64 words per group in the global state
    int32_t skip_null;  // 1: rows whose argument is null do not contribute (everything but count(*))
    int32_t priv_off;   // first word inside the block-private table (PRIVATE mode)
    int32_t convert_i2d;  // avg(bigint): accumulate (double)value
    DCol arg;
    DCol mask;          // optional BOOLEAN mask column (values==nullptr: no mask)
    unsigned long long* state;  // global state base
};

struct AggProgram {
    int32_t naggs;
    int32_t priv_words;  // total private words per group
    AggOp ops[MAX_AGGS];
};

static inline int state_words(int op) {
    switch (op) {
        case OP_COUNT: return 1;
        case OP_SUM_I64: return 3;
        case OP_SUM_F64: return 2;
        case OP_SUM_DEC: return 4;
        case OP_MIN64: case OP_MAX64: return 2;
        default: return 3;
    }
}

// launch the grouped accumulation of `prog` over n rows with group ids `gids` (device int32)
// only_rows (optional, one byte per row): rows whose byte is 0 are skipped (used after a fused pass that handled the rest)
void accumulate(pg_ctx* ctx, const AggProgram& prog, const int32_t* gids, int64_t n, int64_t group_count, const uint8_t* only_rows = nullptr);
// (re)initialise state words of groups [first, first+count): zeros, min/max identities
void init_state(pg_ctx* ctx, int op, unsigned long long* state, int64_t first, int64_t count);
// low-cardinality path; returns false when the program / group count does not fit it
struct KeyLayout;
struct KeyCols;
// what the fused (look-up inside the accumulation kernel) variant needs from the GroupByHash
struct FusedLookup {
    const KeyLayout* layout;
    const KeyCols* keys;              // key columns of the rows being processed
    const uint64_t* key_rows;
    const uint8_t* arena;
    const int64_t* values_by_group;
    int32_t bigint_mode, null_gid;
    unsigned int* miss_flag;          // bitmap of n bits, zeroed by the caller
    unsigned long long* d_miss;       // device counter, zeroed by the caller
    int64_t row_base;                 // page row of the first row handed in (variable width offsets are page-absolute)
    int64_t key_value_bytes[8];       // value buffer size of variable width key columns
};
static const int MAX_PACK_KEYS = 8;
static const int MAX_PACK_KEYS_PUBLIC = MAX_PACK_KEYS;
bool fused_keys_packable(const KeyLayout& L);
// word (0/1) and bit shift of every key inside the packed two-word form; false when the keys do not fit
bool fused_pack_layout(const KeyLayout& L, int32_t* word, int32_t* shift);
bool accumulate_private(pg_ctx* ctx, const AggProgram& prog, const int32_t* gids, int64_t n, int64_t group_count, const uint8_t* only_rows,
                        const FusedLookup* fused);
// the same pass through a kernel specialised at run time for this plan (jit_lowcard.cu); false when it cannot be used.
// px (optional): a fused PageProcessor — aggregate arguments and masks are projections of it, evaluated inside the kernel over
// the processor's input columns `in`; rows its filter rejects do not exist for the operator.  fused.d_miss must then be two
// words: [0] rows left to the general path, [1] error bits (1 = arithmetic overflow in an expression).
struct LowcardExprs {
    const pg_project* proc;
    const DCol* in;               // the processor's input columns, by channel
    int arg_proj[MAX_AGGS];       // projection read by aggregate i (-1: count(*))
    int mask_proj[MAX_AGGS];      // projection used as its mask (-1: none)
};
// The groups the kernel resolves keys against: packed keys in id order, 4 words per group {w0, w1, first row, packable}.
// known_count > 0: that many groups are known (device table `known`).  known_count == 0: the kernel discovers the groups itself
// from the first discover_rows rows, in the reference's first-seen order, and publishes them in discovered_out (64 entries);
// d_miss[2] / d_miss[3] then carry the group count and the verdict (jit/lowcard_args.h PGJ_DISC_*; non-zero = nothing accumulated).
struct LowcardGroups {
    const unsigned long long* known;
    int known_count;
    unsigned long long* discovered_out;
    int64_t discover_rows;
    int capacity_hint;            // discovering launch: groups to make room for at least (0: default)
};
// fused.d_miss must be four words {misses, error bits, groups discovered, discovery verdict}, zeroed by the caller.
bool accumulate_lowcard_jit(pg_ctx* ctx, const AggProgram& prog, int64_t n, const FusedLookup& fused, const LowcardGroups& groups,
                            const LowcardExprs* px = nullptr, int* gcap_used = nullptr);
// packed-key group tables (jit_lowcard.cu): from the GroupByHash's stored keys, and back into key columns
struct LowcardKeyOut { void* values[8]; uint8_t* nulls[8]; int32_t* offsets[8]; };
void lowcard_pack_groups(pg_ctx* ctx, const pg_groupby* gb, int64_t groups, unsigned long long* d_out);
void lowcard_unpack_groups(pg_ctx* ctx, const KeyLayout& L, const unsigned long long* d_groups, int64_t first, int64_t n, const LowcardKeyOut& out);

}  // namespace pg

struct pg_agg {
    pg_ctx* ctx = nullptr;
    int32_t kind = 0;      // PG_AGG_*
    int32_t in_type = 0;
    int32_t op = 0;        // OpClass
    int32_t words = 0;
    int64_t capacity = 0;  // groups the state array holds (initialised)
    pg::DevBuf state;
};

namespace pg {
void agg_create(pg_ctx* ctx, int32_t kind, int32_t in_type, pg_agg** out);
void agg_ensure_capacity(pg_agg* agg, int64_t group_count);
AggOp agg_make_op(pg_agg* agg, const DCol* arg, const DCol* mask);
// evaluate groups [first, first+n) into device buffers described by `st` (already begun)
void agg_evaluate(pg_agg* agg, int64_t first, int64_t n, int32_t final_step, pg_col_out* out, int64_t* error_group, bool do_sync);
void agg_evaluate_async(pg_agg* agg, int64_t first, int64_t n, int32_t final_step, pg_col_out* out, long long* d_err);
int agg_n_intermediate(const pg_agg* agg);
void agg_evaluate_batch(pg_ctx* ctx, pg_agg* const* aggs, int na, int64_t first, int64_t n, int32_t final_step, pg_col_out* out,
                        long long* d_errs);
void accumulate_intermediate_batch(pg_ctx* ctx, pg_agg* const* aggs, const DCol* s0, const DCol* s1, int na, const int32_t* d_ids, int64_t n,
                                   int64_t group_count);
void accumulate_intermediate(pg_agg* agg, const DCol& s0, const DCol& s1, const int32_t* d_ids, int64_t n, int64_t group_count);
}  // namespace pg
