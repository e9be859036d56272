// join.h — device-resident join build side and lookup source.
// Replaces MAIN/operator/PagesIndex.java:209-240 (addPage), PagesHash.java:53-125 (build) / :152-222 (lookup),
// ArrayPositionLinks.java:44-100 (duplicate chains), JoinHash.java:81-125 and JoinProbe.java:87-117.
//
// HBM layout
//   cols[k]            every build channel, appended page by page (flat device columns, no synthetic addresses)
//   slots[cap + 1]     16 B each: { u64 word; i32 count; i32 head }
//        word   single BIGINT join key: the key itself (EMPTY_KEY unused; slot `cap` holds the key equal to EMPTY_KEY)
//               otherwise (tag32 << 32) | representative build row, all-ones = unused
//        count  build rows with this key
//        head   largest build row with this key (the reference's chain head: newest row first)
//   seg_off[cap + 1]   only when some key repeats: first index of the key's rows inside seg_rows
//   seg_rows[rows]     build rows of repeated keys, grouped by key, descending row order (== walking positionLinks)
#pragma once
#include "context.h"
#include "groupby.h"
#include "scan.cuh"

namespace pg {

static const int MAX_JOIN_COLS = 32;

struct JSlot {
    unsigned long long word;
    int32_t count;
    int32_t head;
};
static_assert(sizeof(JSlot) == 16, "join slot must be 16 bytes");

struct ColStore {
    int32_t type = 0;
    DevBuf values, nulls, offsets;
    int64_t rows_cap = 0;     // rows the fixed-width / nulls / offsets buffers hold
    int64_t bytes_used = 0;   // var width
    bool has_nulls = false;
    DCol view() const {
        DCol d;
        d.type = type; d.pad = 0; d.values = values.p;
        d.nulls = has_nulls ? nulls.as<uint8_t>() : nullptr;
        d.offsets = type_is_var(type) ? offsets.as<int32_t>() : nullptr;
        return d;
    }
};

struct JoinCols {
    int32_t n;
    int32_t bigint_fast;   // exactly one BIGINT join key
    DCol c[MAX_KEYS];      // join key columns (build side view or probe side view)
};

// slot placement of the single-BIGINT-key table (see join.cu: range-chunked region + hashed region)
struct JLayout {
    unsigned long long kmin;
    int64_t nchunks;     // chunks of the range region
    int64_t R;           // slots of the range region = nchunks * 32
    int64_t omask;       // slots of the hashed region - 1
    int64_t total;       // R + omask + 1; slot `total` holds the key whose bit pattern equals EMPTY_KEY
    int32_t range, shift;
    int32_t fib, pad;    // slot inside a chunk from multiplicative hashing of the key (1) or from the murmur bits (0)
};

}  // namespace pg

struct pg_joinbuild {
    pg_ctx* ctx = nullptr;
    std::vector<int32_t> types;
    std::vector<int32_t> join_channels;
    int32_t hash_channel = -1;
    std::vector<pg::ColStore> cols;
    int64_t rows = 0;
};

struct pg_lookup {
    pg_ctx* ctx = nullptr;
    std::vector<int32_t> types;
    std::vector<int32_t> join_channels;
    int32_t hash_channel = -1;
    std::vector<pg::ColStore> cols;
    int64_t rows = 0;
    int64_t cap = 0;              // slots (both regions); slot `cap` is the special one
    pg::JLayout layout;           // single BIGINT key: where a key's probe sequence starts (join.cu)
    bool bigint_fast = false;
    bool has_dups = false;
    int64_t dup_rows = 0;
    pg::DevBuf slots, seg_off, seg_rows;
    pg::DevBuf meta;       // build-time counters (probe scratch is allocated per call: probers share the source)
    // OuterLookupSource's visited-positions tracker (OP/OuterLookupSource.java:176-205) as a device bitmap over the build rows,
    // switched on with pg_lookup_track_visited for RIGHT / FULL joins; every emitted build position sets its bit
    pg::DevBuf visited;
    bool track_visited = false;
};
