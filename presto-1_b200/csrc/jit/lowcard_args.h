// lowcard_args.h — kernel argument block of the run-time specialised low-cardinality aggregation kernel.
// Included by the host (jit_lowcard.cu) and pasted in front of the NVRTC source, so it uses builtin types only.
#ifndef PGJ_LOWCARD_ARGS_H
#define PGJ_LOWCARD_ARGS_H

#define PGJ_MAX_KEYS 8
#define PGJ_MAX_LOADS 24
#define PGJ_MAX_SLOTS 24
#define PGJ_MAX_AGGS 16
#define PGJ_LUT_SIZE 256

// slot kinds (private partial state per group and thread)
#define PGJ_KIND_DEC 0   // 128-bit sign-magnitude input -> 192-bit sum, 3 words
#define PGJ_KIND_I64 1   // 64-bit input -> 128-bit sum, 2 words
#define PGJ_KIND_I32 2   // 32-bit input -> 128-bit sum, 2 words
#define PGJ_KIND_F64 3   // double sum, 1 word
#define PGJ_KIND_MIN 4   // 1 word
#define PGJ_KIND_MAX 5   // 1 word
#define PGJ_KIND_CNT 6   // 1 word

struct PgjArgs {
    const void* key_values[PGJ_MAX_KEYS];
    const unsigned char* key_nulls[PGJ_MAX_KEYS];
    const int* key_offsets[PGJ_MAX_KEYS];
    long long key_value_bytes[PGJ_MAX_KEYS];   // bytes in a variable width key's value buffer (bounds the speculative load)
    // the input columns the row loop reads (aggregate arguments, masks, operands of fused filter / projection expressions)
    const void* load_ptr[PGJ_MAX_LOADS];
    const unsigned char* load_nulls[PGJ_MAX_LOADS];
    // flush table: one entry per aggregate (several aggregates may share one private slot)
    unsigned long long* agg_state[PGJ_MAX_AGGS];
    int agg_op[PGJ_MAX_AGGS];        // OpClass
    int agg_words[PGJ_MAX_AGGS];     // words per group of the global state
    int agg_cnt_word[PGJ_MAX_AGGS];  // private word holding the aggregate's row count
    int agg_val_word[PGJ_MAX_AGGS];  // first private word of its value (-1: count only)
    int agg_val_kind[PGJ_MAX_AGGS];  // PGJ_KIND_*
    // stored groups (GroupByHash key store) for the shared-memory key -> group id table
    const unsigned long long* key_rows;
    const unsigned char* arena;
    const long long* values_by_group;
    unsigned int* miss_flag;         // bitmap: bit set for rows whose key is not in the table (zeroed by the caller)
    unsigned long long* d_miss;      // [0] number of such rows, [1] error bits (1 = arithmetic overflow in a fused expression)
    long long n;
    long long row_base;              // page row of row 0 (variable width offsets are absolute within the page)
    int naggs, G, bigint_mode, null_gid, row_words, pad;
};

#endif
