// oracle.cpp — TEST INFRASTRUCTURE ONLY.
// extern "C" surface of the CPU oracle (ctypes-friendly).  See oracle_common.h for the
// pinning statement.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load liboracle.so.
#include "oracle_accum.h"
#include "oracle_join.h"
#include "oracle_project.h"

using namespace orc;

#define ORC_API extern "C" __attribute__((visibility("default")))

// ---------------- scalar helpers (known-answer tests) ----------------
ORC_API int64_t orc_long_hash(int64_t v) { return long_type_hash(v); }
ORC_API int64_t orc_xxh64_long(int64_t v) { return xxh64_long(v); }
ORC_API int64_t orc_xxh64_bytes(const uint8_t* p, int64_t n) { return xxh64_bytes(p, n); }
ORC_API uint64_t orc_fmix64(uint64_t v) { return fmix64(v); }
ORC_API int64_t orc_array_size(int64_t expected) { return array_size(expected, 0.75f); }
ORC_API int64_t orc_max_fill(int64_t cap) { return calculate_max_fill(cap); }
// state (lo,hi,overflow) += value (lo,hi); mirrors DecimalSumAggregation.accumulateValueInState
ORC_API void orc_dec_add(uint64_t* state_lo_hi, int64_t* overflow, const uint64_t* v_lo_hi) {
    Dec128 s{state_lo_hi[0], state_lo_hi[1]}, v{v_lo_hi[0], v_lo_hi[1]}, o;
    *overflow += dec_add_with_overflow(s, v, o);
    state_lo_hi[0] = o.lo; state_lo_hi[1] = o.hi;
}
ORC_API int32_t orc_dec_avg(const uint64_t* state_lo_hi, int64_t overflow, int64_t count, uint64_t* out_lo_hi) {
    Dec128 s{state_lo_hi[0], state_lo_hi[1]}; bool fits;
    Dec128 r = i192_div_half_up(i192_from_state(s, overflow), (uint64_t)count, &fits);
    out_lo_hi[0] = r.lo; out_lo_hi[1] = r.hi; return fits ? 0 : 1;
}
ORC_API int32_t orc_dec_exceeds_1e38(const uint64_t* v) { Dec128 d{v[0], v[1]}; return dec_exceeds_1e38(d) ? 1 : 0; }

ORC_API void orc_hash_rows(const pg_col* cols, int32_t ncols, int64_t n, int64_t* out) {
    for (int64_t i = 0; i < n; i++) out[i] = row_hash(cols, ncols, i);
}
// mode 0: LocalPartitionGenerator (nparts power of two); mode 1: HashGenerator.getPartition
ORC_API void orc_partition(const int64_t* raw_hash, int64_t n, int32_t nparts, int32_t mode, int32_t* out) {
    for (int64_t i = 0; i < n; i++) out[i] = mode == 0 ? local_partition(raw_hash[i], nparts) : remote_partition(raw_hash[i], nparts);
}

// ---------------- GroupByHash ----------------
ORC_API void* orc_groupby_create(const int32_t* types, int32_t nkeys, int32_t has_hash, int64_t expected) {
    GroupBy* g = new GroupBy(); g->init(types, nkeys, has_hash != 0, expected); return g;
}
ORC_API void orc_groupby_destroy(void* h) { delete (GroupBy*)h; }
// getGroupIds (out_ids != NULL) / addPage (out_ids == NULL): the Work loops of
// BigintGroupByHash.java:336-424 / MultiChannelGroupByHash.java:568-734 with updateMemory always true
ORC_API int32_t orc_groupby_get_group_ids(void* h, const pg_col* keys, const pg_col* hash_col, int64_t n, int32_t* out_ids) {
    GroupBy* g = (GroupBy*)h;
    for (int64_t i = 0; i < n; i++) {
        if (g->need_rehash()) { int rc = g->rehash(); if (rc) return rc; }
        int64_t id = g->put(keys, hash_col, i);
        if (out_ids) out_ids[i] = (int32_t)id;
    }
    return PG_OK;
}
ORC_API int64_t orc_groupby_group_count(void* h) { return ((GroupBy*)h)->next_group; }
ORC_API int64_t orc_groupby_capacity(void* h) { return ((GroupBy*)h)->capacity; }
ORC_API int64_t orc_groupby_collisions(void* h) { return ((GroupBy*)h)->hash_collisions; }
ORC_API void orc_groupby_contains(void* h, const pg_col* keys, const pg_col* hash_col, int64_t n, uint8_t* out) {
    GroupBy* g = (GroupBy*)h;
    for (int64_t i = 0; i < n; i++) out[i] = g->contains(keys, hash_col, i) ? 1 : 0;
}
static void store_export(const Store& s, int64_t first, int64_t n, pg_col_out* o) {
    for (int64_t i = 0; i < n; i++) {
        int64_t r = first + i;
        if (o->nulls) o->nulls[i] = s.nulls[r];
        switch (s.type) {
            case PG_INTEGER: case PG_DATE: ((int32_t*)o->values)[i] = (int32_t)s.v[r]; break;
            case PG_BOOLEAN: ((uint8_t*)o->values)[i] = (uint8_t)s.v[r]; break;
            case PG_LONG_DECIMAL: ((int64_t*)o->values)[2 * i] = s.v[2 * r]; ((int64_t*)o->values)[2 * i + 1] = s.v[2 * r + 1]; break;
            case PG_VARCHAR: case PG_VARBINARY: {
                if (i == 0) o->offsets[0] = 0;
                int32_t len = s.offs[r + 1] - s.offs[r];
                memcpy((uint8_t*)o->values + o->offsets[i], s.bytes.data() + s.offs[r], len);
                o->offsets[i + 1] = o->offsets[i] + len; break;
            }
            default: ((int64_t*)o->values)[i] = s.v[r];
        }
    }
}
// appendValuesTo for groups [first, first+n): keys (+ raw hash channel when out_hash != NULL)
ORC_API int32_t orc_groupby_export(void* h, int64_t first, int64_t n, pg_col_out* out_keys, int64_t* out_hash) {
    GroupBy* g = (GroupBy*)h;
    if (first < 0 || first + n > g->next_group) return PG_E_INVALID_ARGUMENT;
    if (g->bigint_mode) {
        for (int64_t i = 0; i < n; i++) {
            int64_t id = first + i;
            ((int64_t*)out_keys[0].values)[i] = id == g->null_group ? 0 : g->values_by_group[id];
            if (out_keys[0].nulls) out_keys[0].nulls[i] = id == g->null_group;
        }
    } else {
        for (int k = 0; k < g->nkeys; k++) store_export(g->keys[k], first, n, &out_keys[k]);
    }
    if (out_hash) for (int64_t i = 0; i < n; i++) out_hash[i] = g->raw_hash(first + i);
    return PG_OK;
}
ORC_API int64_t orc_groupby_varwidth_bytes(void* h, int32_t key) {
    GroupBy* g = (GroupBy*)h; return g->bigint_mode ? 0 : (int64_t)g->keys[key].bytes.size();
}

// ---------------- GroupedAccumulator ----------------
ORC_API void* orc_agg_create(int32_t kind, int32_t in_type) { Accum* a = new Accum(); a->kind = kind; a->in_type = in_type; return a; }
ORC_API void orc_agg_destroy(void* h) { delete (Accum*)h; }
ORC_API int32_t orc_agg_add_input(void* h, const int32_t* ids, int64_t groups, const pg_col* arg, const pg_col* mask, int64_t n) {
    Accum* a = (Accum*)h; a->add_input(ids, groups, arg, mask, n);
    return a->overflow_error ? PG_E_NUMERIC_OVERFLOW : PG_OK;
}
ORC_API int32_t orc_agg_add_intermediate(void* h, const int32_t* ids, int64_t groups, const pg_col* st, int64_t n) {
    Accum* a = (Accum*)h; a->add_intermediate(ids, groups, st, n);
    return a->overflow_error ? PG_E_NUMERIC_OVERFLOW : PG_OK;
}
ORC_API int32_t orc_agg_n_intermediate(void* h) { return ((Accum*)h)->n_inter(); }

static inline void put_i64(pg_col_out* o, int64_t i, int64_t v, bool isnull) {
    ((int64_t*)o->values)[i] = isnull ? 0 : v; if (o->nulls) o->nulls[i] = isnull;
}
// evaluateFinal / evaluateIntermediate for groups [first, first+n).  Returns PG_E_NUMERIC_OVERFLOW
// (and *error_group) where the reference's output function throws.
ORC_API int32_t orc_agg_evaluate(void* h, int64_t first, int64_t n, int32_t final_step, pg_col_out* out, int64_t* error_group) {
    Accum* a = (Accum*)h; a->ensure(first + n);
    int32_t rc = PG_OK;
    for (int64_t i = 0; i < n; i++) {
        int64_t g = first + i;
        switch (a->kind) {
            case AGG_COUNT: case AGG_COUNT_COL: put_i64(&out[0], i, a->cnt[g], false); break;
            case AGG_MIN: case AGG_MAX:
                if (a->in_type == PG_DOUBLE) put_i64(&out[0], i, double_to_long_bits(a->dsum[g]), !a->set[g]);
                else if (a->in_type == PG_LONG_DECIMAL) {
                    ((uint64_t*)out[0].values)[2 * i] = a->set[g] ? a->dec[g].lo : 0; ((uint64_t*)out[0].values)[2 * i + 1] = a->set[g] ? a->dec[g].hi : 0;
                    if (out[0].nulls) out[0].nulls[i] = !a->set[g];
                } else if (a->in_type == PG_INTEGER || a->in_type == PG_DATE) {
                    ((int32_t*)out[0].values)[i] = a->set[g] ? (int32_t)a->lsum[g] : 0; if (out[0].nulls) out[0].nulls[i] = !a->set[g];
                } else put_i64(&out[0], i, a->lsum[g], !a->set[g]);
                break;
            case AGG_SUM: case AGG_AVG:
                if (a->is_dec()) {
                    if (!final_step) {   // VARBINARY [count?][overflow][lo][hi]; null when state untouched
                        pg_col_out* o = &out[0]; if (i == 0) o->offsets[0] = 0;
                        int32_t off = o->offsets[i]; uint8_t* p = (uint8_t*)o->values + off;
                        if (a->set[g]) {
                            if (a->kind == AGG_AVG) { memcpy(p, &a->cnt[g], 8); p += 8; }
                            memcpy(p, &a->ovf[g], 8); memcpy(p + 8, &a->dec[g].lo, 8); memcpy(p + 16, &a->dec[g].hi, 8);
                            off += a->kind == AGG_AVG ? 32 : 24;
                        }
                        o->offsets[i + 1] = off; if (o->nulls) o->nulls[i] = !a->set[g];
                    } else if (a->kind == AGG_SUM) {   // DecimalSumAggregation.outputLongDecimal :178-190
                        bool isnull = !a->set[g];
                        if (!isnull && (a->ovf[g] != 0 || dec_exceeds_1e38(a->dec[g]))) { if (rc == PG_OK) { rc = PG_E_NUMERIC_OVERFLOW; if (error_group) *error_group = g; } }
                        ((uint64_t*)out[0].values)[2 * i] = isnull ? 0 : a->dec[g].lo; ((uint64_t*)out[0].values)[2 * i + 1] = isnull ? 0 : a->dec[g].hi;
                        if (out[0].nulls) out[0].nulls[i] = isnull;
                    } else {                             // DecimalAverageAggregation.output* :188-220
                        bool isnull = a->cnt[g] == 0; bool fits = true; Dec128 r{0, 0};
                        if (!isnull) r = i192_div_half_up(i192_from_state(a->dec[g], a->ovf[g]), (uint64_t)a->cnt[g], &fits);
                        if (a->in_type == PG_SHORT_DECIMAL) {
                            // longValueExact
                            bool ok = fits && dec_hi_mag(r) == 0 && (r.lo <= 0x7fffffffffffffffULL || (dec_is_negative(r) && r.lo == 0x8000000000000000ULL));
                            if (!ok && rc == PG_OK) { rc = PG_E_NUMERIC_OVERFLOW; if (error_group) *error_group = g; }
                            int64_t v = dec_is_negative(r) ? (int64_t)(0 - r.lo) : (int64_t)r.lo;
                            put_i64(&out[0], i, v, isnull);
                        } else {
                            if ((!fits || dec_exceeds_1e38(r)) && rc == PG_OK) { rc = PG_E_NUMERIC_OVERFLOW; if (error_group) *error_group = g; }
                            ((uint64_t*)out[0].values)[2 * i] = r.lo; ((uint64_t*)out[0].values)[2 * i + 1] = r.hi;
                            if (out[0].nulls) out[0].nulls[i] = isnull;
                        }
                    }
                } else if (a->kind == AGG_SUM) {
                    bool dbl = a->in_type == PG_DOUBLE;
                    int64_t sv = dbl ? double_to_long_bits(a->dsum[g]) : a->lsum[g];
                    if (final_step) put_i64(&out[0], i, sv, a->cnt[g] == 0);
                    else { put_i64(&out[0], i, a->cnt[g], false); put_i64(&out[1], i, sv, false); }
                } else {  // avg(double|bigint): AverageAggregations.output :68-80
                    if (final_step) put_i64(&out[0], i, double_to_long_bits(a->dsum[g] / (double)a->cnt[g]), a->cnt[g] == 0);
                    else { put_i64(&out[0], i, double_to_long_bits(a->dsum[g]), false); put_i64(&out[1], i, a->cnt[g], false); }
                }
                break;
        }
    }
    return rc;
}

// ---------------- join ----------------
struct JoinHandle { JoinBuild b; std::vector<int32_t> op; std::vector<int64_t> ob; };
ORC_API void* orc_join_create(const int32_t* types, int32_t ncols, const int32_t* join_channels, int32_t nj, int32_t hash_channel) {
    JoinHandle* j = new JoinHandle(); j->b.init(types, ncols, join_channels, nj, hash_channel); return j;
}
ORC_API void orc_join_destroy(void* h) { delete (JoinHandle*)h; }
ORC_API void orc_join_add_page(void* h, const pg_col* cols, int64_t n) { ((JoinHandle*)h)->b.add_page(cols, n); }
ORC_API void orc_join_build(void* h) { ((JoinHandle*)h)->b.build(); }
ORC_API int64_t orc_join_rows(void* h) { return ((JoinHandle*)h)->b.rows; }
ORC_API int64_t orc_join_hash_size(void* h) { return ((JoinHandle*)h)->b.hash_size; }
ORC_API int64_t orc_join_collisions(void* h) { return ((JoinHandle*)h)->b.hash_collisions; }
ORC_API int64_t orc_join_probe(void* h, const pg_col* probe_keys, int32_t nk, const pg_col* hash_col, int64_t n, int32_t join_type) {
    JoinHandle* j = (JoinHandle*)h; j->op.clear(); j->ob.clear();
    j->b.probe(probe_keys, nk, hash_col, n, join_type, j->op, j->ob);
    return (int64_t)j->op.size();
}
ORC_API void orc_join_fetch(void* h, int32_t* out_probe, int64_t* out_build) {
    JoinHandle* j = (JoinHandle*)h;
    if (!j->op.empty()) { memcpy(out_probe, j->op.data(), j->op.size() * 4); memcpy(out_build, j->ob.data(), j->ob.size() * 8); }
}

// ---------------- PageProcessor (filter + projections) ----------------
ORC_API void* orc_project_create(const pg_expr* nodes, int32_t n, int32_t filter, const int32_t* proj, int32_t np) {
    Project* p = new Project(); p->init(nodes, n, filter, proj, np); return p;
}
ORC_API void orc_project_destroy(void* h) { delete (Project*)h; }
ORC_API void orc_project_types(void* h, int32_t* types, int32_t* precisions, int32_t* scales) {
    Project* p = (Project*)h;
    for (size_t k = 0; k < p->proj.size(); k++) { const PType& t = p->types[p->proj[k]]; types[k] = t.type; precisions[k] = t.precision; scales[k] = t.scale; }
}
// out[] sized for n rows (variable width: the input column's value bytes); returns PG_E_NUMERIC_OVERFLOW where the reference throws
ORC_API int32_t orc_project_process(void* h, const pg_col* cols, int64_t n, pg_col_out* out, int64_t* n_out) {
    Project* p = (Project*)h; p->overflow = false;
    *n_out = p->process(cols, n, out);
    return p->overflow ? PG_E_NUMERIC_OVERFLOW : PG_OK;
}
