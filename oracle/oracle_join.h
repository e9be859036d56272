// oracle_join.h — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Restates MAIN/operator/{PagesIndex,PagesHash,ArrayPositionLinks,JoinHash,JoinProbe,
// LookupJoinOperator,LookupJoinPageBuilder,PartitionedLookupSource}.java and
// MAIN/operator/exchange/LocalPartitionGenerator.java for the INNER / PROBE_OUTER paths.
#pragma once
#include "oracle_groupby.h"

namespace orc {

enum { JOIN_INNER = 0, JOIN_PROBE_OUTER = 1 };

struct JoinBuild {
    int ncols = 0;
    std::vector<int32_t> types;
    std::vector<int32_t> join_channels;
    int hash_channel = -1;          // precomputed hash channel index among build columns, or -1
    std::vector<Store> cols;        // PagesIndex.addPage retains blocks; we append
    int64_t rows = 0;
    // PagesHash
    int64_t hash_size = 0, mask = 0;
    std::vector<int32_t> key;             // row index per slot, -1 empty
    std::vector<uint8_t> pos_to_hash;     // positionToHashes
    std::vector<int32_t> links;           // ArrayPositionLinks.positionLinks; empty if no duplicates
    bool has_links = false;
    int64_t hash_collisions = 0;
    bool built = false;

    void init(const int32_t* t, int n, const int32_t* jc, int nj, int hc) {
        ncols = n; types.assign(t, t + n); join_channels.assign(jc, jc + nj); hash_channel = hc;
        cols.resize(n); for (int k = 0; k < n; k++) cols[k].type = t[k];
    }
    void add_page(const pg_col* page, int64_t n) {   // PagesIndex.addPage :209-240
        for (int k = 0; k < ncols; k++) for (int64_t i = 0; i < n; i++) cols[k].append(page[k], i);
        rows += n;
    }
    // a pg_col view of stored column k (so hashing/equality reuse the Block accessors)
    pg_col view(int k) const {
        pg_col c; memset(&c, 0, sizeof(c));
        const Store& s = cols[k];
        c.type = s.type; c.mem = PG_MEM_HOST; c.n = rows; c.nulls = s.nulls.data();
        if (s.type == PG_VARCHAR || s.type == PG_VARBINARY) { c.values = s.bytes.data(); c.offsets = s.offs.data(); }
        else c.values = s.v.data();
        return c;
    }
    // Stored fixed-width values are widened to int64; expose them with a 64-bit integer type of
    // identical hash semantics: INTEGER/DATE hash((int)v) == long_type_hash(sign-extended v).
    static int32_t widened(int32_t t) { return (t == PG_INTEGER || t == PG_DATE) ? PG_BIGINT : t; }
    pg_col wview(int k) const { pg_col c = view(k); c.type = widened(c.type); return c; }

    bool row_has_null_key(int64_t r) const {   // SimplePagesHashStrategy.isPositionNull :268-276
        for (int32_t ch : join_channels) if (cols[ch].nulls[r]) return true;
        return false;
    }
    int64_t hash_build_row(int64_t r) const {  // hashPosition: precomputed channel or fold over join channels
        if (hash_channel >= 0) return cols[hash_channel].v[r];
        uint64_t h = 0;
        for (int32_t ch : join_channels) { pg_col c = wview(ch); h = 31 * h + (uint64_t)type_hash(c, r); }
        return (int64_t)h;
    }
    bool rows_equal_ignore_nulls(int64_t a, int64_t b) const {  // positionEqualsPositionIgnoreNulls
        for (int32_t ch : join_channels) { pg_col c = wview(ch); if (!cols[ch].equal_nonnull(a, c, b, false)) return false; }
        return true;
    }
    // PagesHash constructor :53-125 (serial insert; newest row becomes the chain head)
    void build() {
        hash_size = array_size(rows, 0.75f); mask = hash_size - 1;
        key.assign(hash_size, -1); pos_to_hash.assign(rows, 0);
        links.assign(rows, -1); has_links = false;
        for (int64_t r = 0; r < rows; r++) {
            int64_t h = hash_build_row(r);
            pos_to_hash[r] = (uint8_t)h;
            if (row_has_null_key(r)) continue;
            int64_t pos = (int64_t)(fmix64((uint64_t)h) & (uint64_t)mask);
            while (key[pos] != -1) {
                int32_t cur = key[pos];
                if ((uint8_t)h == pos_to_hash[cur] && rows_equal_ignore_nulls(cur, r)) {
                    links[r] = cur; has_links = true;     // ArrayPositionLinks.link(new, old head) :44-50
                    break;
                }
                pos = (pos + 1) & mask; hash_collisions++;
            }
            key[pos] = (int32_t)r;
        }
        built = true;
    }
    // PagesHash.getAddressIndex :152-169
    int64_t join_position(const pg_col* probe_keys, int nk, int64_t i, int64_t raw) const {
        int64_t pos = (int64_t)(fmix64((uint64_t)raw) & (uint64_t)mask);
        while (key[pos] != -1) {
            int32_t cur = key[pos];
            if (pos_to_hash[cur] == (uint8_t)raw) {
                bool eq = true;
                for (int k = 0; k < nk && eq; k++) eq = cols[join_channels[k]].equal_nonnull(cur, probe_keys[k], i, false);
                if (eq) return cur;
            }
            pos = (pos + 1) & mask;
        }
        return -1;
    }
    // LookupJoinOperator.processProbe :299-378 for one probe page; emits (probe index, build row)
    // pairs in the reference's order.  build row -1 == appendNullForBuild (PROBE_OUTER).
    void probe(const pg_col* probe_keys, int nk, const pg_col* hash_col, int64_t n, int join_type,
               std::vector<int32_t>& out_probe, std::vector<int64_t>& out_build) const {
        for (int64_t i = 0; i < n; i++) {
            int64_t jp = -1;
            bool has_null = false;                                   // JoinProbe.currentRowContainsNull :109-116
            for (int k = 0; k < nk; k++) if (col_is_null(probe_keys[k], i)) has_null = true;
            if (!has_null && rows > 0) {
                int64_t raw = hash_col ? col_long(*hash_col, i) : row_hash(probe_keys, nk, i);
                jp = join_position(probe_keys, nk, i, raw);
            }
            bool produced = false;
            while (jp >= 0) {
                out_probe.push_back((int32_t)i); out_build.push_back(jp); produced = true;
                jp = has_links ? links[jp] : -1;                     // JoinHash.getNextJoinPosition :107-113
            }
            if (!produced && join_type == JOIN_PROBE_OUTER) { out_probe.push_back((int32_t)i); out_build.push_back(-1); }
        }
    }
};

// LocalPartitionGenerator.getPartition :50-63 — (int) XxHash64.hash(Long.reverse(rawHash)) & (N-1)
static inline uint64_t bit_reverse64(uint64_t x) {
    x = ((x >> 1) & 0x5555555555555555ULL) | ((x & 0x5555555555555555ULL) << 1);
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    return __builtin_bswap64(x);
}
static inline int32_t local_partition(int64_t raw_hash, int32_t nparts_pow2) {
    return (int32_t)xxh64_long((int64_t)bit_reverse64((uint64_t)raw_hash)) & (nparts_pow2 - 1);
}
// HashGenerator.getPartition :24-35 — (rawHash & MAX_LONG) % partitionCount
static inline int32_t remote_partition(int64_t raw_hash, int32_t nparts) {
    return (int32_t)((raw_hash & 0x7fffffffffffffffLL) % nparts);
}

}  // namespace orc
