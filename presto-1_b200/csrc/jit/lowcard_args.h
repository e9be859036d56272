// lowcard_args.h — kernel argument block of the run-time specialised low-cardinality aggregation kernel.
// Included by the host (jit_lowcard.cu) and pasted in front of the NVRTC source, so it uses builtin types only.
#ifndef PGJ_LOWCARD_ARGS_H
#define PGJ_LOWCARD_ARGS_H

#define PGJ_MAX_KEYS 8
#define PGJ_MAX_LOADS 24
#define PGJ_MAX_SLOTS 24
#define PGJ_MAX_AGGS 16
#define PGJ_LUT_SIZE 256
#define PGJ_MAX_GROUPS 64

// slot kinds (private partial state per group and thread, one 64-bit word each)
#define PGJ_KIND_SUM 0   // exact integer / decimal sum: narrow private word + the block's 192-bit accumulator for everything else
#define PGJ_KIND_F64 3   // double sum
#define PGJ_KIND_MIN 4
#define PGJ_KIND_MAX 5
#define PGJ_KIND_CNT 6

// discovery verdict bits (d_miss[3])
#define PGJ_DISC_TOO_MANY 1      // more groups than the private tables were sized for (d_miss[2] says how many)
#define PGJ_DISC_UNPACKABLE 2    // a key of the sampled prefix does not fit the packed form
#define PGJ_DISC_TABLE_FULL 4
#define PGJ_DISC_COLLISION 8     // two different keys share a fingerprint

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
    int agg_val_word[PGJ_MAX_AGGS];  // private word of its value (-1: count only)
    int agg_val_kind[PGJ_MAX_AGGS];  // PGJ_KIND_*
    int agg_wide_idx[PGJ_MAX_AGGS];  // SUM: index of its slot among the SUM slots (the block's wide accumulators)
    // groups: packed keys in id order, 4 words each {w0, w1, first row, packable}; known_groups == 0: discover them from the first
    // discover_rows rows and publish them in lc_groups_out
    const unsigned long long* lc_groups;
    unsigned long long* lc_groups_out;
    unsigned int* miss_flag;         // bitmap: bit set for rows whose key is not in the table (zeroed by the caller)
    unsigned long long* d_miss;      // [0] number of such rows, [1] error bits (1 = arithmetic overflow in a fused expression),
                                     // [2] groups discovered, [3] discovery verdict (PGJ_DISC_*; non-zero: nothing was accumulated)
    long long n;
    long long row_base;              // page row of row 0 (variable width offsets are absolute within the page)
    long long discover_rows;
    int naggs, G, known_groups, fast_bits;
};

#endif
