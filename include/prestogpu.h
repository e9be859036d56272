/*
 * prestogpu.h — C ABI of the B200-native hash-aggregation / hash-join path.
 *
 * This is the drop-in boundary.  Every entry point is what a
 * JNI binding of the reference's operator interfaces would call; the
 * reference interface each one replaces is cited as file:line relative to
 * /root/reference/core/trino-main/src/main/java/io/trino/ (OP = operator/).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no C++/torch types.
 *  - every function returns int32 status (PG_OK == 0); out-params are last.
 *  - a pg_col describes one Trino Block: values + optional 1-byte-per-row
 *    null mask (boolean[] valueIsNull) + optional int32 offsets (n+1) for
 *    variable width blocks.  `mem` says where the pointers live.
 *  - handles are not thread-safe; one CUDA stream per context.
 *  - input columns are copied/consumed before the call returns; nothing is
 *    retained that points into caller memory (unless mem == PG_MEM_DEVICE
 *    and the call documents that it keeps a reference).
 */
#ifndef PRESTOGPU_H
#define PRESTOGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PG_API __attribute__((visibility("default")))

/* ---- status codes (SURVEY §8b "Error conventions") ---- */
enum {
    PG_OK = 0,
    PG_NEED_MEMORY = 1,              /* Work.process()==false analogue: nothing consumed      */
    PG_E_INSUFFICIENT_RESOURCES = 2, /* GENERIC_INSUFFICIENT_RESOURCES (OP/BigintGroupByHash.java:265-267) */
    PG_E_NUMERIC_OVERFLOW = 3,       /* NUMERIC_VALUE_OUT_OF_RANGE (type/BigintOperators.java:47-55)     */
    PG_E_INVALID_ARGUMENT = 4,       /* checkArgument / IllegalArgumentException             */
    PG_E_CUDA = 5,
    PG_E_NCCL = 6,
    PG_E_INTERNAL = 7,
    PG_E_ILLEGAL_STATE = 8           /* checkState / IllegalStateException                    */
};

/* ---- column element types (spi/type/ *.java) ---- */
enum {
    PG_BIGINT = 1,        /* LongArrayBlock, int64                                   */
    PG_INTEGER = 2,       /* IntArrayBlock, int32                                    */
    PG_DATE = 3,          /* IntArrayBlock, int32 days                               */
    PG_DOUBLE = 4,        /* LongArrayBlock holding doubleToLongBits                 */
    PG_SHORT_DECIMAL = 5, /* LongArrayBlock, unscaled int64 (p <= 18)                */
    PG_LONG_DECIMAL = 6,  /* Int128ArrayBlock, 2 x int64 sign-magnitude (p > 18)     */
    PG_VARCHAR = 7,       /* VariableWidthBlock: bytes + int32 offsets[n+1]          */
    PG_VARBINARY = 8,     /* same layout as VARCHAR (intermediate states)            */
    PG_BOOLEAN = 9        /* ByteArrayBlock, 1 byte                                  */
};

enum { PG_MEM_HOST = 0, PG_MEM_DEVICE = 1,
       /* host memory that is page-locked (cudaHostAlloc / cudaHostRegister) AND that the caller keeps valid and unmodified until the
        * operator has consumed it: accepted wherever PG_MEM_HOST is; the page accumulator of pg_hashagg_add_page then copies such pages
        * to the device straight from the caller's memory (no intermediate host copy), and they count as consumed after the next call
        * that reads the operator's state (stats, get_output*, ...), or as soon as pg_ctx_synchronize returns */
       PG_MEM_HOST_PINNED = 2 };

typedef struct pg_col {
    int32_t type;           /* PG_BIGINT ...                                         */
    int32_t mem;            /* PG_MEM_HOST | PG_MEM_DEVICE | PG_MEM_HOST_PINNED      */
    int64_t n;              /* positionCount                                         */
    const void* values;     /* element array (bytes for VARCHAR/VARBINARY)           */
    const uint8_t* nulls;   /* 1 byte per row, non-zero == null; NULL == no nulls    */
    const int32_t* offsets; /* VARCHAR/VARBINARY only: n+1 offsets into values       */
    int32_t precision;      /* decimals                                              */
    int32_t scale;
} pg_col;

/* mutable twin used for outputs: caller allocates, callee fills */
typedef struct pg_col_out {
    int32_t type;
    int32_t mem;
    int64_t capacity;       /* rows the buffers can hold                             */
    void* values;
    uint8_t* nulls;         /* may be NULL if the caller does not want null flags    */
    int32_t* offsets;       /* VARCHAR/VARBINARY: capacity+1                         */
    int64_t values_capacity_bytes; /* VARCHAR/VARBINARY byte capacity                */
} pg_col_out;

typedef struct pg_ctx pg_ctx;

/* ---- context ---- */
PG_API int32_t pg_ctx_create(int32_t device, pg_ctx** out);
PG_API int32_t pg_ctx_destroy(pg_ctx* ctx);
/* message of the last failing call on this context (thread-unsafe, like the handle) */
PG_API const char* pg_last_error(pg_ctx* ctx);
/* use an external stream (e.g. torch's current stream) for all work of this context */
PG_API int32_t pg_ctx_set_stream(pg_ctx* ctx, void* cuda_stream);
PG_API int32_t pg_ctx_synchronize(pg_ctx* ctx);
/* leave `sms` multiprocessors free for work on other streams: the one-wave persistent kernels of this context size their
   grids to (SM count - sms), so a second pipeline stage (e.g. the FINAL step of the previous page set) is not queued
   behind them.  The device-side analogue of splitting driver threads between pipelines (TaskExecutor.java). */
PG_API int32_t pg_ctx_reserve_sms(pg_ctx* ctx, int32_t sms);
/* number of kernels this context has launched so far (bench.py's gpu_launches)  */
PG_API int64_t pg_ctx_launch_count(pg_ctx* ctx);
/* per-kernel CUDA-event timing on the context stream (off by default): measurement support for bench.py */
PG_API int32_t pg_ctx_set_timing(pg_ctx* ctx, int32_t enabled);
PG_API int32_t pg_ctx_kernel_time(pg_ctx* ctx, const char* name_substr, int32_t reset, double* total_ms, int64_t* launches);
PG_API int32_t pg_ctx_kernel_report(pg_ctx* ctx, char* buf, int64_t cap);
PG_API const char* pg_version(void);

/* =====================================================================================
 * GroupByHash  —  replaces OP/GroupByHash.java:31-88, OP/BigintGroupByHash.java:43-425,
 * OP/MultiChannelGroupByHash.java:55-853.
 * Exactly one BIGINT key -> the bigint table (16-byte slots, key in slot); anything else ->
 * the multi-channel table (16-byte slots holding the row hash + packed key rows).  Dispatch mirrors
 * GroupByHash.createGroupByHash (OP/GroupByHash.java:45-59).
 * Group ids are dense and assigned in first-seen row order, like the reference.
 * ===================================================================================== */
typedef struct pg_groupby pg_groupby;

PG_API int32_t pg_groupby_create(pg_ctx* ctx, const int32_t* key_types, int32_t nkeys,
                                 int32_t has_hash_channel, int64_t expected_size, pg_groupby** out);
PG_API int32_t pg_groupby_destroy(pg_groupby* gb);
/* getGroupIds(Page) (OP/GroupByHash.java:75): out_ids[n] int32 (host or device, `ids_mem`),
 * *out_group_count = GroupByIdBlock.groupCount.  hash_col = precomputed $hashvalue channel or NULL. */
PG_API int32_t pg_groupby_get_group_ids(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n,
                                        int32_t* out_ids, int32_t ids_mem, int64_t* out_group_count);
/* addPage(Page) (OP/GroupByHash.java:73) */
PG_API int32_t pg_groupby_add_page(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n);
/* contains(position, page, hashChannels[, rawHash]) for every row of the page (OP/GroupByHash.java:77-83) */
PG_API int32_t pg_groupby_contains(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n,
                                   uint8_t* out, int32_t out_mem);
/* MarkDistinctHash.markDistinctRows(page) (OP/MarkDistinctHash.java:52-69, driven by MarkDistinctOperator): adds the page like
 * getGroupIds and sets out_marks[i] = 1 for the first row of every group this page introduced (ids[i] == nextDistinctId++). */
PG_API int32_t pg_groupby_mark_distinct(pg_groupby* gb, const pg_col* keys, const pg_col* hash_col, int64_t n, uint8_t* out_marks, int32_t out_mem);
/* HashSemiJoinOperator (OP/HashSemiJoinOperator.java:176-221) against the ChannelSet this GroupByHash holds (SetBuilderOperator =
 * pg_groupby_add_page of the build side; ChannelSet.java:32-130): out_values / out_nulls [n] are the appended BOOLEAN column —
 * NULL probe key: false on an empty set, else NULL; found: true; not found: NULL when the set contains NULL, else false. */
PG_API int32_t pg_groupby_semi_join(pg_groupby* gb, const pg_col* probe_key, const pg_col* hash_col, int64_t n, int32_t set_contains_null,
                                    uint8_t* out_values, uint8_t* out_nulls, int32_t out_mem);
/* appendValuesTo(groupId, pageBuilder, 0) for groupId in [first, first+n) (OP/GroupByHash.java:71);
 * out_hash (int64[n], may be NULL) receives the raw-hash channel / getRawHash(groupId) (:85). */
PG_API int32_t pg_groupby_export_keys(pg_groupby* gb, int64_t first, int64_t n, pg_col_out* out_keys,
                                      int64_t* out_hash, int32_t hash_mem);
/* UpdateMemory (OP/UpdateMemory.java, consulted by tryRehash: OP/BigintGroupByHash.java:295-334, MultiChannelGroupByHash.java:
 * 369-420) as a byte limit: once set (> 0), get_group_ids / add_page return PG_NEED_MEMORY — BEFORE consuming any row, the table
 * unchanged — when the page could force the table to grow past the limit (every row counted as a new group).  The caller gets the
 * memory (raises the limit) and hands the same page in again: Work.process() returning false, at page granularity. */
PG_API int32_t pg_groupby_set_memory_limit(pg_groupby* gb, int64_t bytes);
/* bytes needed for variable-width key `key` over all groups (to size pg_col_out.values) */
PG_API int32_t pg_groupby_varwidth_bytes(pg_groupby* gb, int32_t key, int64_t* out_bytes);
/* getGroupCount / getCapacity / getHashCollisions / getEstimatedSize (OP/GroupByHash.java:61-87).
 * capacity is the reference's logical capacity schedule (arraySize(expected,0.75) doubled while
 * groupCount >= maxFill); device_bytes is what the table really holds in HBM. */
PG_API int32_t pg_groupby_stats(pg_groupby* gb, int64_t* group_count, int64_t* capacity,
                                int64_t* collisions, int64_t* device_bytes);

/* =====================================================================================
 * GroupedAccumulator — replaces OP/aggregation/GroupedAccumulator.java:22-39 as generated by
 * OP/aggregation/AccumulatorCompiler.java:261-330,487-566 for count/sum/avg/min/max.
 *
 * Intermediate (partial -> final) column layouts, per aggregate:
 *   COUNT, COUNT_COL      : [BIGINT count]                                   (CountAggregation.java:33-55)
 *   SUM(BIGINT|INTEGER)   : [BIGINT count, BIGINT sum]                       (LongSumAggregation.java:34-64)
 *   SUM(DOUBLE)           : [BIGINT count, DOUBLE sum]                       (DoubleSumAggregation.java:33-63)
 *   AVG(DOUBLE|BIGINT)    : [DOUBLE sum, BIGINT count]                       (AverageAggregations.java:34-80)
 *   SUM(DECIMAL)          : [VARBINARY 24 B = overflow int64 | lo | hi], NULL if untouched
 *                           (state/LongDecimalWithOverflowStateSerializer.java:38-66)
 *   AVG(DECIMAL)          : [VARBINARY 32 B = count | overflow | lo | hi]
 *                           (state/LongDecimalWithOverflowAndLongStateSerializer.java:38-68)
 *   MIN, MAX              : [value type, nullable]                           (NullableLongStateSerializer.java:33-42)
 * ===================================================================================== */
enum { PG_AGG_COUNT = 1, PG_AGG_COUNT_COL = 2, PG_AGG_SUM = 3, PG_AGG_AVG = 4, PG_AGG_MIN = 5, PG_AGG_MAX = 6 };

typedef struct pg_agg pg_agg;
PG_API int32_t pg_agg_create(pg_ctx* ctx, int32_t kind, int32_t in_type, pg_agg** out);
PG_API int32_t pg_agg_destroy(pg_agg* agg);
/* addInput(GroupByIdBlock, Page): ids int32[n] (`ids_mem`), arg NULL for COUNT, mask NULL or BOOLEAN */
PG_API int32_t pg_agg_add_input(pg_agg* agg, const int32_t* ids, int32_t ids_mem, int64_t group_count,
                                const pg_col* arg, const pg_col* mask, int64_t n);
/* addIntermediate(GroupByIdBlock, Block): `state` = the columns listed in the table above */
PG_API int32_t pg_agg_add_intermediate(pg_agg* agg, const int32_t* ids, int32_t ids_mem, int64_t group_count,
                                       const pg_col* state, int64_t n);
PG_API int32_t pg_agg_n_intermediate(pg_agg* agg, int32_t* out);
/* evaluateFinal / evaluateIntermediate for groupId in [first, first+n).  Returns
 * PG_E_NUMERIC_OVERFLOW and *error_group where the reference's output function throws. */
PG_API int32_t pg_agg_evaluate(pg_agg* agg, int64_t first, int64_t n, int32_t final_step,
                               pg_col_out* out, int64_t* error_group);
PG_API int32_t pg_agg_device_bytes(pg_agg* agg, int64_t* out);

/* =====================================================================================
 * Join build + LookupSource — replaces OP/HashBuilderOperator.java:308-350,428-496 (addInput/finish),
 * OP/PagesIndex.java:209-240,489-533, OP/PagesHash.java:53-125 (build) and :152-222 (lookup),
 * OP/ArrayPositionLinks.java:44-100, OP/JoinHash.java:81-125, OP/LookupSource.java:23-50,
 * OP/JoinProbe.java:87-117 and the emit loop of OP/LookupJoinOperator.java:299-378.
 * ===================================================================================== */
enum { PG_JOIN_INNER = 0, PG_JOIN_PROBE_OUTER = 1 };

typedef struct pg_joinbuild pg_joinbuild;
typedef struct pg_lookup pg_lookup;

/* HashBuilderOperatorFactory(types, outputChannels, hashChannels, preComputedHashChannel, expectedPositions)
 * (OP/HashBuilderOperator.java:54-152).  hash_channel = index of the precomputed $hashvalue BIGINT channel or -1. */
PG_API int32_t pg_joinbuild_create(pg_ctx* ctx, const int32_t* types, int32_t ncols, const int32_t* join_channels, int32_t njoin,
                                   int32_t hash_channel, int64_t expected_rows, pg_joinbuild** out);
PG_API int32_t pg_joinbuild_destroy(pg_joinbuild* b);
/* HashBuilderOperator.addInput -> PagesIndex.addPage: the page's columns are copied into device-resident build columns */
PG_API int32_t pg_joinbuild_add_page(pg_joinbuild* b, const pg_col* cols, int64_t n);
/* HashBuilderOperator.finish -> buildLookupSource: builds the hash table; the builder's storage moves into the lookup */
PG_API int32_t pg_joinbuild_finish(pg_joinbuild* b, pg_lookup** out);
PG_API int32_t pg_lookup_destroy(pg_lookup* ls);
/* getJoinPositionCount / PagesHash hashSize / getInMemorySizeInBytes / positionLinks.isPresent() */
PG_API int32_t pg_lookup_stats(pg_lookup* ls, int64_t* rows, int64_t* hash_size, int64_t* device_bytes, int32_t* has_duplicates);
/* One probe page through JoinProbe + LookupJoinOperator.processProbe: emits (probe position, build position) pairs in the
 * reference's order (probe order; within a probe row the build rows newest first = positionLinks chain order).
 * build position -1 = appendNullForBuild (PROBE_OUTER).  Probe rows with a null join key never match.
 * If the output exceeds `capacity`, returns PG_NEED_MEMORY with *n_out = rows needed and nothing consumed. */
PG_API int32_t pg_lookup_probe(pg_lookup* ls, const pg_col* probe_keys, int32_t nkeys, const pg_col* hash_col, int64_t n,
                               int32_t join_type, int32_t* out_probe_idx, int64_t* out_build_pos, int32_t out_mem,
                               int64_t capacity, int64_t* n_out);
/* LookupSource.appendTo(position, pageBuilder, offset) for a vector of build positions and a list of build channels
 * (LookupJoinPageBuilder.appendRow :76-84): gathers build columns; position -1 -> null. */
PG_API int32_t pg_lookup_gather(pg_lookup* ls, const int64_t* build_pos, int32_t pos_mem, int64_t n, const int32_t* channels,
                                int32_t nchannels, pg_col_out* out);
/* LookupJoinOperator's step for one probe page in ONE call: JoinProbe + LookupSource.appendTo of `build_channels`
 * (OP/LookupJoinOperator.java:333-361, OP/JoinProbe.java:40-116, PagesHashStrategy.appendTo).  out_probe_idx[capacity] (memory kind
 * out_mem) receives the probe positions; out_build[k] (capacity >= `capacity` rows each; a variable width channel additionally
 * needs values_capacity_bytes >= the bytes pg_lookup_gather_var_bytes reports, so callers with such channels use the two-step
 * form) receives build channel build_channels[k] of the matched rows, null for PROBE_OUTER misses.  Unique single-BIGINT-key
 * build sides with device outputs run one kernel and never materialise 64-bit build positions; every other case runs
 * pg_lookup_probe + pg_lookup_gather internally.  PG_NEED_MEMORY (nothing consumed, *n_out = rows needed) as in pg_lookup_probe. */
PG_API int32_t pg_lookup_join_page(pg_lookup* ls, const pg_col* probe_keys, int32_t nkeys, const pg_col* hash_col, int64_t n, int32_t join_type,
                                   const int32_t* build_channels, int32_t nchannels, int32_t* out_probe_idx, int32_t out_mem,
                                   pg_col_out* out_build, int64_t capacity, int64_t* n_out);
/* RIGHT / FULL outer joins (f4).  OuterLookupSource (OP/OuterLookupSource.java:34-205) wraps the lookup source with a tracker of the
 * build positions that were appended to some output row; after build and probe have finished, LookupOuterOperator
 * (OP/LookupOuterOperator.java:140-181) emits the positions never visited, ascending, with NULL probe columns.
 * pg_lookup_track_visited switches the tracker on (before the first probe; the one-kernel join_page form is then not used);
 * pg_lookup_outer_positions returns the unvisited positions (feed them to pg_lookup_gather). */
PG_API int32_t pg_lookup_track_visited(pg_lookup* ls);
PG_API int32_t pg_lookup_outer_positions(pg_lookup* ls, int64_t* out_pos, int32_t out_mem, int64_t capacity, int64_t* n_out);
/* value bytes pg_lookup_gather will write for a variable width build channel */
PG_API int32_t pg_lookup_gather_var_bytes(pg_lookup* ls, const int64_t* build_pos, int32_t pos_mem, int64_t n, int32_t channel,
                                          int64_t* out_bytes);

/* =====================================================================================
 * Fused GROUP BY operator — replaces OP/HashAggregationOperator.java:381-518 (addInput/getOutput) together with
 * OP/aggregation/builder/InMemoryHashAggregationBuilder.java:139-155 (processPage) and :243-298 (buildResult).
 * One call per page: group ids are resolved and every aggregate is updated on the device; ids never cross the ABI.
 * ===================================================================================== */
enum { PG_STEP_SINGLE = 0, PG_STEP_PARTIAL = 1, PG_STEP_FINAL = 2, PG_STEP_INTERMEDIATE = 3 };  /* AggregationNode.Step */

typedef struct pg_agg_spec {
    int32_t kind;           /* PG_AGG_*                                                          */
    int32_t in_type;        /* argument type (0 for count(*))                                    */
    int32_t input_channel;  /* raw steps: argument channel (-1 for count(*));
                               FINAL/INTERMEDIATE: first channel of the intermediate columns      */
    int32_t mask_channel;   /* BOOLEAN mask channel or -1                                        */
} pg_agg_spec;

typedef struct pg_hashagg pg_hashagg;
/* HashAggregationOperatorFactory(groupByTypes, groupByChannels, step, accumulatorFactories, hashChannel, expectedGroups)
 * (OP/HashAggregationOperator.java:54-262) */
PG_API int32_t pg_hashagg_create(pg_ctx* ctx, const int32_t* key_channels, const int32_t* key_types, int32_t nkeys,
                                 int32_t hash_channel, const pg_agg_spec* aggs, int32_t naggs, int32_t step,
                                 int64_t expected_groups, pg_hashagg** out);
PG_API int32_t pg_hashagg_destroy(pg_hashagg* op);
/* addInput(Page) */
PG_API int32_t pg_hashagg_add_page(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n);
/* Reference-sized pages.  The reference hands addInput pages of at most ~1 MiB (a few thousand rows: PageBuilder /
 * ScanFilterAndProjectOperator's merging output, OP/project/MergingPageOutput.java); one kernel pass per such page would be
 * launch- and synchronisation-bound.  With batching on (rows > 0), host pages smaller than rows / 4 are appended to a pinned
 * accumulator and enter the kernels as one page of about `rows` rows; page order is kept (first-seen group ids do not change),
 * and every call that looks at the operator's state (stats, get_output*, varwidth_bytes, add_packed*) first hands the accumulated
 * rows in.  Under a memory limit the accumulated rows stay accumulated when the flush answers PG_NEED_MEMORY.  rows = 0 turns it
 * off (the default). */
PG_API int32_t pg_hashagg_set_page_batching(pg_hashagg* op, int64_t rows);
/* Feeds a host page of n rows as consecutive pages of page_rows rows, each through the path of one pg_hashagg_add_page call
 * (a caller that holds many small pages, and bench.py's e2e_pages record). */
PG_API int32_t pg_hashagg_add_pages(pg_hashagg* op, const pg_col* cols, int32_t ncols, int64_t n, int64_t page_rows);
PG_API int32_t pg_hashagg_stats(pg_hashagg* op, int64_t* group_count, int64_t* capacity, int64_t* device_bytes, int64_t* input_rows);
/* buildTypes(): keys..., (hash), then per aggregate its final type or its intermediate columns */
PG_API int32_t pg_hashagg_output_types(pg_hashagg* op, int32_t* types, int32_t capacity, int32_t* ncols);
/* buildResult() for group ids [first, first+n) */
PG_API int32_t pg_hashagg_get_output(pg_hashagg* op, int64_t first, int64_t n, pg_col_out* out, int64_t* error_group);
PG_API int32_t pg_hashagg_varwidth_bytes(pg_hashagg* op, int32_t col, int64_t* out_bytes);
/* The operator's memory yield (HashAggregationOperator.addInput -> unfinishedWork, OP/HashAggregationOperator.java:398-417;
 * InMemoryHashAggregationBuilder.updateMemoryWithYieldInfo): with a limit set, add_page returns PG_NEED_MEMORY before consuming
 * anything when the page could grow the operator's table past it; same contract as pg_groupby_set_memory_limit. */
PG_API int32_t pg_hashagg_set_memory_limit(pg_hashagg* op, int64_t bytes);
/* partial-aggregation flush: forget all groups (OP/HashAggregationOperator.java:501-517) */
PG_API int32_t pg_hashagg_reset(pg_hashagg* op);
/* The output page of group ids [first, first+n) written as ONE self-describing host buffer, and its counterpart that adds several
 * such buffers (e.g. the PARTIAL pages of all ranks after an all-gather) as one input page, rows in buffer order.  This is the
 * wire format between a PARTIAL and a FINAL operator in different processes - the job PagesSerde does for the reference's
 * exchange (MAIN/execution/buffer/PagesSerde.java), restricted to the block kinds this library produces.  Layout (little endian):
 *   int64 total_bytes, int64 n, int32 ncols, int32 magic 'PGP1',
 *   ncols x { int32 type, int32 has_nulls, int64 values_bytes, int64 offsets_bytes, int64 data_offset },
 *   then per column, 8-byte aligned from data_offset: values, offsets (int32[n+1], var width only), nulls (n bytes).
 * get_output_packed fails with PG_NEED_MEMORY (and the size needed in *out_bytes) when `capacity` is too small. */
PG_API int32_t pg_hashagg_get_output_packed(pg_hashagg* op, int64_t first, int64_t n, void* buf, int64_t capacity, int64_t* out_bytes,
                                            int64_t* error_group);
PG_API int32_t pg_hashagg_add_packed(pg_hashagg* op, const void* const* bufs, int32_t nbufs);
/* The same pages in DEVICE memory, for exchanges that never leave the GPUs (NCCL all-gather of the PARTIAL pages of all ranks):
 * get_output_packed_device writes the page at device_buf; add_packed_device takes nbufs pages that start `stride` bytes apart
 * at device_bufs (the all-gather's receive buffer; stride a multiple of 8, every page at most `stride` bytes). */
PG_API int32_t pg_hashagg_get_output_packed_device(pg_hashagg* op, int64_t first, int64_t n, void* device_buf, int64_t capacity,
                                                   int64_t* out_bytes, int64_t* error_group);
PG_API int32_t pg_hashagg_add_packed_device(pg_hashagg* op, const void* device_bufs, int64_t stride, int32_t nbufs);

/* =====================================================================================
 * PageProcessor (filter + projections) — replaces OP/project/PageProcessor.java:54-57,139-330 as ScanFilterAndProjectOperator
 * (OP/ScanFilterAndProjectOperator.java) drives it, together with the filter/projection classes sql/gen/PageFunctionCompiler.java
 * generates, for the RowExpression subset the configs' upstream expressions use: input references, constants, decimal / bigint /
 * integer / double arithmetic (type/DecimalOperators.java:84-330, BigintOperators.java:47-83, IntegerOperators, DoubleOperators),
 * comparisons, AND / OR / NOT with SQL three-valued logic, IS NULL, a few casts and the $hashvalue projection
 * (sql/planner/optimizations/HashGenerationOptimizer.java:98,872,926 -> operator/scalar/CombineHashFunction.java:25-30).
 * Expressions form a DAG stored as an array: operands are earlier nodes.  Result types of operators follow the reference's
 * signatures (ADD/SUB: precision min(38, max(p1-s1, p2-s2) + max(s1, s2) + 1), scale max(s1, s2); MUL: min(38, p1+p2), s1+s2;
 * a decimal with precision <= 18 is PG_SHORT_DECIMAL, otherwise PG_LONG_DECIMAL) and are computed by the library.
 * A row passes the filter iff it evaluates to TRUE (NULL counts as false, PageProcessor / DictionaryAwarePageFilter).
 * Arithmetic overflow (|decimal| >= 10^38, Math.addExact on bigint / integer) fails the call with PG_E_NUMERIC_OVERFLOW, like
 * the reference's NUMERIC_VALUE_OUT_OF_RANGE — for rows the filter rejects as well only if the expression is also part of the filter.
 * ===================================================================================== */
enum {
    PG_EXPR_INPUT = 1,   /* InputReferenceExpression(channel): type / precision / scale given by the caller          */
    PG_EXPR_CONST = 2,   /* ConstantExpression: lo (int64 value, double bits) or lo/hi (long decimal) or bytes      */
    PG_EXPR_ADD = 3, PG_EXPR_SUB = 4, PG_EXPR_MUL = 5,
    PG_EXPR_LT = 6, PG_EXPR_LE = 7, PG_EXPR_GT = 8, PG_EXPR_GE = 9, PG_EXPR_EQ = 10, PG_EXPR_NE = 11,
    PG_EXPR_AND = 12, PG_EXPR_OR = 13, PG_EXPR_NOT = 14, PG_EXPR_IS_NULL = 15,
    PG_EXPR_CAST = 16,   /* arg0 cast to type / precision / scale: integer->bigint, integer|bigint->double,
                            integer|bigint->decimal, decimal->decimal with a scale that does not shrink                */
    PG_EXPR_HASH = 17    /* combine_hash(arg0 = previous hash node or -1 for INITIAL_HASH_VALUE 0, hash_code(arg1)) */
};
typedef struct pg_expr {
    int32_t op;
    int32_t type;           /* INPUT, CONST, CAST: PG_* ; ignored otherwise                          */
    int32_t precision, scale;
    int32_t channel;        /* INPUT: channel of the page handed to process / addInput               */
    int32_t arg0, arg1;     /* operand node indices (smaller than this node's index); -1 = none      */
    int32_t is_null;        /* CONST: the typed NULL                                                  */
    int64_t lo, hi;         /* CONST value                                                            */
    const uint8_t* bytes;   /* CONST VARCHAR (copied at create)                                       */
    int32_t nbytes;
    int32_t reserved;
} pg_expr;

typedef struct pg_project pg_project;
/* filter = node index or -1 (no filter); projections[i] = node index of output channel i */
PG_API int32_t pg_project_create(pg_ctx* ctx, const pg_expr* nodes, int32_t n_nodes, int32_t filter, const int32_t* projections,
                                 int32_t n_projections, pg_project** out);
PG_API int32_t pg_project_destroy(pg_project* p);
PG_API int32_t pg_project_output_types(pg_project* p, int32_t* types, int32_t* precisions, int32_t* scales, int32_t capacity, int32_t* ncols);
/* PageProcessor.process(page): out[] (caller-allocated, capacity >= n rows each; a variable width pass-through column needs
 * values_capacity_bytes >= the input column's value bytes) receives the projections over the rows that pass the filter, in
 * page order; *n_out = the number of such rows. */
PG_API int32_t pg_project_process(pg_project* p, const pg_col* cols, int32_t ncols, int64_t n, pg_col_out* out, int64_t* n_out);
/* Fuses the processor into the aggregation (ScanFilterAndProject -> HashAggregation in one pass over the raw columns): pages
 * handed to pg_hashagg_add_page are then the processor's INPUT pages and the operator's group-by / aggregate / mask channels
 * index the processor's projections.  Derived columns are never materialised on the fast path.  The processor must outlive
 * the operator; pass NULL to detach. */
PG_API int32_t pg_hashagg_set_input_processor(pg_hashagg* op, pg_project* p);

/* =====================================================================================
 * Hash partitioning for the multi-GPU exchange — replaces the row routing of OP/exchange/LocalPartitionGenerator.java:50-63,
 * OP/exchange/PartitioningExchanger.java:79-95 and OP/HashGenerator.java:24-35 (PartitionedOutputOperator).  The exchange
 * itself is an NCCL all-to-all of the partition-major columns, or direct stores into the receivers' buffers over NVLink.
 * ===================================================================================== */
enum { PG_PARTITION_REMOTE = 0 /* (rawHash & MAX_LONG) % n */, PG_PARTITION_LOCAL = 1 /* xxhash64(reverse(rawHash)) & (n-1) */ };
typedef struct pg_partitioner pg_partitioner;
PG_API int32_t pg_partition_create(pg_ctx* ctx, const pg_col* keys, int32_t nkeys, const pg_col* hash_col, int64_t n, int32_t nparts,
                                   int32_t mode, int64_t* out_counts, pg_partitioner** out);
PG_API int32_t pg_partition_destroy(pg_partitioner* pt);
PG_API int32_t pg_partition_scatter(pg_partitioner* pt, const pg_col* in, pg_col_out* out);
PG_API int32_t pg_partition_scatter_peers(pg_partitioner* pt, const pg_col* in, void* const* peer_values, uint8_t* const* peer_nulls,
                                          const int64_t* peer_row_offset);
/* The whole page in one launch: every fixed width column (at most 16 per call) is staged tile by tile in shared memory and stored
 * into the receivers' buffers.  peer_values / peer_nulls hold ncols x nparts pointers ([c * nparts + p]); a column whose
 * peer_nulls[c * nparts] is NULL travels without a mask.  Stream-ordered: it returns after the launch, and a barrier on the same
 * stream (as in presto-1_b200/exchange.py) must order the stores before the receivers read. */
PG_API int32_t pg_partition_scatter_peers_all(pg_partitioner* pt, const pg_col* cols, int32_t ncols, void* const* peer_values,
                                              uint8_t* const* peer_nulls, const int64_t* peer_row_offset);
PG_API int32_t pg_partition_ids(pg_partitioner* pt, uint8_t* out, int32_t out_mem);

/* pg_partition_positions_noalloc — NO COUNTERPART IN THE REFERENCE; added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The positions pass of pg_partition_create over a precomputed raw-hash channel, without any allocation: it calls no cudaMalloc /
 * cudaFree / allocating library routine, creates no handle and does not synchronise (results are ready in stream order on the
 * context's stream).  Every pointer is a DEVICE pointer supplied by the caller:
 *   raw_hash[n]          BIGINT $hashvalue channel (rawHash of every row)
 *   out_part_of_row[n]   partition of every row (HashGenerator.getPartition / LocalPartitionGenerator, `mode` as above)
 *   out_dest[n]          row's position in partition-major order, stable inside a partition (what pg_partition_scatter applies)
 *   out_counts[nparts]   rows per partition
 * Its scratch (per-tile histograms) is one fixed array in device memory, about 1 MiB, held by every GPU that loads the library until
 * exit.  Hence: n <= PG_NOALLOC_MAX_ROWS (larger pages fail with PG_E_INVALID_ARGUMENT - use pg_partition_create), nparts <= 64;
 * all calls on a device share that array, so calls that OVERLAP on the GPU (different streams or contexts) race - calls on one
 * stream are ordered and safe; the array keeps the last call's contents and no call depends on them.  The first call on a device
 * resolves the array's device address (cudaGetSymbolAddress, no allocation): make one call before starting a graph capture. */
#define PG_NOALLOC_MAX_ROWS 4194304
PG_API int32_t pg_partition_positions_noalloc(pg_ctx* ctx, const int64_t* raw_hash, int64_t n, int32_t nparts, int32_t mode,
                                              uint8_t* out_part_of_row, int32_t* out_dest, int64_t* out_counts);

#ifdef __cplusplus
}
#endif
#endif
