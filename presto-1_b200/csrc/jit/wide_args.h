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
    unsigned long long room;               // PGW_QUOTA: buckets of `table` this launch may claim; a warp stops claiming once counters[0] passed it
    unsigned long long* counters;          // [0] buckets claimed in `table`, [1] rows left out because the quota ran out, [2] buckets claimed in `front`
    unsigned char* overflow_flag;          // n bytes, zeroed by the caller; 1 = row left out
    // range-chunked front region (PGW_RANGE): keys of [k0, k0 + (nseg << shift)) live in chunk (key - k0) >> shift of PGW_CHUNK buckets,
    // so rows that arrive sorted by key touch the table sequentially; a key whose chunk is full, or outside the range, uses `table`
    unsigned long long* front;             // nseg * PGW_CHUNK buckets
    unsigned long long* front_keys;        // the front buckets' keys once more, packed (8 bytes per bucket): probing reads these through L1 and
                                           // claims with a compare-and-swap here; the claimer then copies the key into word 0 of the bucket
    long long k0;
    unsigned long long nseg;
    int shift;
    int pad2;
    // split mode (PGW_BIDX): the bucket of every row was resolved by wide_locate_kernel (hashagg_wide.cu); bucket numbers count the
    // front region first, then the table and its two extra buckets; >= 0xFFFFFFFE: the row is not handled by this launch
    const unsigned int* bidx;
    unsigned long long front_buckets;
    // PGW_ROWIDS: the page was reordered for locality (hashagg_wide.cu: reorder_for_locality); row i of the arrays above is row
    // row_ids[i] of the page the caller handed in, and that index is what first-seen order is built from
    const unsigned int* row_ids;
};

#endif
