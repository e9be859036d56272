// hashagg_wide.h — the fused GROUP BY operator's path for MANY groups on one BIGINT key: aggregation state lives inside
// the hash bucket of its key (one random access per row), group ids are assigned only when the output is built.
// See jit/wide_body.cuh for the bucket layout and hashagg_wide.cu for growth and output numbering.
#pragma once
#include <vector>

#include "accum.h"
#include "groupby.h"

struct pg_hashagg;

namespace pg {

struct KeySample { int64_t rows, distinct, repeats, ascending, min_key, max_key; };

struct WideAgg {
    int32_t op;         // OpClass (OP_COUNT, OP_SUM_I64, OP_SUM_F64, OP_SUM_DEC)
    int32_t in_kind;    // 0 none, 1 i64, 2 i32, 3 f64, 4 int64 -> double, 5 decimal128
    int32_t cnt_mode;   // 0: 64-bit count in cnt_word, 1: "seen a row" flag bit in the flag word
    int32_t cnt_word, val_word, flag_bit;
    int32_t words;      // words per group of the dense global state this aggregate is scattered to
    int32_t val_words;  // words of its value inside the bucket (limb sums: 2 for 64-bit inputs, 4 for long decimals)
    int32_t skip_null;  // rows whose argument is null do not contribute (everything but count(*))
};

struct WideTable {
    pg_ctx* ctx = nullptr;
    int32_t bucket_words = 0, flag_word = -1;
    std::vector<WideAgg> aggs;
    int64_t cap = 0, max_fill = 0;
    int64_t claimed = 0;          // regular buckets of `table` in use (host copy)
    // range-chunked front region, planned from the first page's key sample when the rows arrive (nearly) sorted by key:
    // chunk (key - k0) >> shift of WIDE_CHUNK buckets holds the keys of its key range, hashed inside the chunk, so consecutive
    // rows touch consecutive memory.  Keys outside the planned range or whose chunk is full live in `table` as usual.
    bool split = true;            // locate kernel + update kernel instead of the fused one (fixed for the table's life: the two
                                  // place keys differently inside a front chunk)
    bool range = false;
    int64_t k0 = 0, nseg = 0, front_buckets = 0, front_claimed = 0;
    int32_t shift = 0;
    DevBuf front, front_keys;     // buckets, and their keys once more as a packed array (what probing reads)
    bool null_seen = false, zero_seen = false;
    bool prepared = false;        // output numbering done, dense states filled
    DevBuf table, counters;
    int64_t groups() const { return claimed + front_claimed + (null_seen ? 1 : 0) + (zero_seen ? 1 : 0); }
    int64_t device_bytes() const { return table.bytes + front.bytes + front_keys.bytes + counters.bytes; }
};

// true when the operator's aggregates can run on the wide path (count/sum/avg over 64-bit, double and decimal inputs)
bool wide_supported(const std::vector<pg_agg_spec>& specs, const std::vector<pg_agg*>& aggs);
WideTable* wide_create(pg_ctx* ctx, const std::vector<pg_agg_spec>& specs, const std::vector<pg_agg*>& aggs, int32_t step, int64_t expected_groups,
                       const KeySample* sorted_sample = nullptr);
// one page; `cols` resolves an imported channel.  Returns false when no specialised kernel could be built (no row consumed).
bool wide_add_page(WideTable* w, const DCol& key, const DCol* args, const DCol* masks, int64_t n, int64_t row_base);
// numbers the groups in first-seen order and scatters keys and states into the dense structures the output code reads
void wide_prepare_output(WideTable* w, pg_groupby* gb, const std::vector<pg_agg*>& aggs, int64_t total_rows);
// distinct keys among the first `sample` rows, and how many of those rows repeat the key of the row before them
// (decides between the few-groups path, the general path and the bucket-resident-state path)
int64_t sample_distinct_keys(pg_ctx* ctx, const DCol& key, int64_t sample, int64_t* adjacent_repeats, KeySample* more = nullptr);
// true when the sample says the rows arrive (nearly) sorted by key with a usable key density: the front region pays off
bool wide_range_applicable(const KeySample& s);

}  // namespace pg
