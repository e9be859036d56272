// groupby.h — device-resident GroupByHash (replaces MAIN/operator/BigintGroupByHash.java and
// MAIN/operator/MultiChannelGroupByHash.java; dispatch as in MAIN/operator/GroupByHash.java:45-59).
//
// HBM layout
//   slots[cap + 2]   16 B each: { u64 word; i32 gid; i32 first_row }
//       bigint mode : word = the key itself (EMPTY_KEY = unused slot)
//       multi mode  : word = (tag32 << 32) | ref, tag32 = high half of fmix64(rawHash),
//                     ref = PENDING|row (claimed in this chunk, key still only in the input page)
//                           or the group id (key lives in key_rows[gid]); all-ones = unused slot
//       slot cap   = the one key whose bit pattern equals EMPTY_KEY (bigint mode only)
//       slot cap+1 = the null group of bigint mode (BigintGroupByHash.java:215-222)
//   values_by_group[maxFill+2]        bigint mode: key per group id (valuesByGroupId)
//   key_rows[(maxFill+2) * row_words] multi mode: packed key row per group id: word0 = null bits,
//                                     then 1 word per channel (2 for long decimal; (off<<24|len) for varchar)
//   hash_by_group[maxFill+2]          multi mode: raw hash per group id ($hashvalue channel / getRawHash)
//   arena[]                           multi mode: bytes of variable width keys, appended in group-id order
//
// Group ids are dense and in first-seen row order (TestGroupByHash.java:161-163), produced by
// insert (atomicCAS + atomicMin(first_row)) -> flag first rows of new slots -> device scan -> look-up.
#pragma once
#include "context.h"
#include "scan.cuh"

namespace pg {

struct Slot {
    unsigned long long word;
    int32_t gid;
    int32_t first_row;
};
static_assert(sizeof(Slot) == 16, "slot must be 16 bytes");

static const unsigned long long EMPTY_KEY = 0x8000000000000001ULL;   // bigint mode sentinel
static const unsigned long long EMPTY_WORD = 0xFFFFFFFFFFFFFFFFULL;  // multi mode sentinel
static const uint32_t REF_PENDING = 0x80000000u;
static const int64_t MAX_CHUNK_ROWS = 1LL << 30;
static const int MAX_KEYS = 32;

struct GbMeta {
    int64_t group_count;   // groups with an assigned id
    int64_t arena_bytes;   // bytes used in the var-width arena
    int64_t inserted;      // slots claimed by the last insert launch
    int64_t base_groups;   // group_count before the chunk being assigned
    int64_t base_bytes;    // arena_bytes before the chunk being assigned
    Pair64 chunk_total;    // (new groups, new var bytes) of the chunk being assigned
    int32_t abort;         // a block exhausted its insert quota
    int32_t pad;
};

// device view of the key layout, passed by value to kernels
struct KeyLayout {
    int32_t nkeys;
    int32_t row_words;            // words per packed key row (multi mode)
    int32_t types[MAX_KEYS];
    int32_t word_off[MAX_KEYS];   // first word of channel k inside the packed row
};

struct KeyCols {
    DCol c[MAX_KEYS];
};

}  // namespace pg

struct pg_groupby {
    pg_ctx* ctx = nullptr;
    bool bigint_mode = false;
    bool has_hash_channel = false;
    bool has_var = false;
    pg::KeyLayout layout;
    int64_t expected = 0;
    int64_t cap = 0;          // physical slots (power of two), excluding the two special slots
    int64_t deferred_cap = 0; // lazy_table: the planned size, built when the first rows are inserted
    int64_t max_fill = 0;     // physical fill limit
    int64_t occupied = 0;     // claimed regular slots (assigned + pending)
    int64_t groups = 0;       // host copy of meta.group_count
    int64_t ref_capacity = 0; // the reference's logical capacity schedule (getCapacity)
    int64_t arena_used = 0;
    int64_t var_bound_hint = -1;   // caller's upper bound on the variable width bytes of the next page (-1: unknown, read them back)
    int64_t rehashes = 0;
    int64_t memory_limit = 0;      // bytes (0 = none): a page that could grow the table past it is refused with PG_NEED_MEMORY
    pg::DevBuf slots, values_by_group, key_rows, hash_by_group, arena, meta;
    // scratch reused across pages
    pg::DevBuf row_slot, tile_sums;
};

namespace pg {
// Shared entry used by pg_groupby_get_group_ids / add_page and by the fused hash-aggregation operator.
// d_out_ids may be null (addPage).  keys are device views with n rows.
void groupby_process(pg_groupby* gb, const KeyCols& keys, const DCol* hash_col, int64_t n, int32_t* d_out_ids);
void groupby_create(pg_ctx* ctx, const int32_t* key_types, int32_t nkeys, int32_t has_hash_channel, int64_t expected,
                    pg_groupby** out, bool lazy_table = false);
int64_t groupby_device_bytes(const pg_groupby* gb);
// The analogue of UpdateMemory.update() returning false inside tryRehash (OP/BigintGroupByHash.java:295-334): called BEFORE any
// row of a page of n rows is consumed; throws PG_NEED_MEMORY when, with every row a new group, the table would have to grow past
// memory_limit.  Nothing has changed then, and the caller retries the same page after raising the limit (Work.process() == false).
void groupby_reserve(pg_groupby* gb, int64_t n);
}  // namespace pg
