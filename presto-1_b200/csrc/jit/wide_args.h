// wide_args.h — argument block of the run-time specialised high-cardinality GROUP BY kernel (jit/wide_body.cuh).
// Shared between the host (hashagg_wide.cu) and the NVRTC source: builtin types only.
#ifndef PGW_WIDE_ARGS_H
#define PGW_WIDE_ARGS_H

#define PGW_MAX_AGGS 16

struct PgwArgs {
    const long long* keys;                 // the BIGINT group-by channel
    const unsigned char* key_nulls;        // or null
    const unsigned char* only_rows;        // or null: rows whose byte is 0 are skipped (overflow re-run)
    const void* arg[PGW_MAX_AGGS];
    const unsigned char* arg_nulls[PGW_MAX_AGGS];
    const unsigned char* mask[PGW_MAX_AGGS];
    const unsigned char* mask_nulls[PGW_MAX_AGGS];
    unsigned long long* table;             // (cap + 2) buckets of PGW_BUCKET_WORDS words; bucket cap = null key, cap + 1 = key 0
    long long cap;
    long long n;
    long long row_base;                    // operator-wide index of row 0 (first-seen order is kept across pages)
    unsigned int quota;                    // buckets one block may claim in this launch
    unsigned int pad;
    unsigned long long* counters;          // [0] buckets claimed, [1] rows left out because the quota ran out
    unsigned char* overflow_flag;          // n bytes, zeroed by the caller; 1 = row left out
};

#endif
