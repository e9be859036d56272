/* prestogpu_jni_ops.c — the natives of io.trino.gpu.PrestoGpu, one per pg_* entry point of include/prestogpu.h.
 * Staging and conventions: prestogpu_jni.c.  Pages arrive as (types, values, nulls, offsets, arrayOffsets, positionCount). */
#include "prestogpu_jni.h"

#define PAGE_PARAMS jintArray types, jobjectArray values, jobjectArray nulls, jobjectArray offsets, jintArray arrayOffsets, jlong positionCount
#define PAGE_ARGS types, values, nulls, offsets, arrayOffsets, positionCount
#define OUT_PARAMS jintArray outTypes, jobjectArray outValues, jobjectArray outNulls, jobjectArray outOffsets, jlongArray outValueBytes

/* ================================================================= context */
NATIVE(jint, ctxCreate)(JNIEnv* env, jclass cls, jint device, jlongArray out) {
    pg_ctx* ctx = NULL;
    const int32_t rc = pg_ctx_create(device, &ctx);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)ctx);
    return rc;
}
NATIVE(jint, ctxDestroy)(JNIEnv* env, jclass cls, jlong ctx) { return pg_ctx_destroy(H(pg_ctx, ctx)); }
NATIVE(jstring, lastError)(JNIEnv* env, jclass cls, jlong ctx) {
    const char* msg = pg_last_error(H(pg_ctx, ctx));
    return (*env)->NewStringUTF(env, msg ? msg : "");
}
NATIVE(jstring, version)(JNIEnv* env, jclass cls) { return (*env)->NewStringUTF(env, pg_version()); }
NATIVE(jint, ctxSetStream)(JNIEnv* env, jclass cls, jlong ctx, jlong cudaStream) { return pg_ctx_set_stream(H(pg_ctx, ctx), (void*)(intptr_t)cudaStream); }
NATIVE(jint, ctxSynchronize)(JNIEnv* env, jclass cls, jlong ctx) { return pg_ctx_synchronize(H(pg_ctx, ctx)); }
NATIVE(jint, ctxReserveSms)(JNIEnv* env, jclass cls, jlong ctx, jint sms) { return pg_ctx_reserve_sms(H(pg_ctx, ctx), sms); }
NATIVE(jlong, ctxLaunchCount)(JNIEnv* env, jclass cls, jlong ctx) { return pg_ctx_launch_count(H(pg_ctx, ctx)); }
NATIVE(jint, ctxSetTiming)(JNIEnv* env, jclass cls, jlong ctx, jboolean enabled) { return pg_ctx_set_timing(H(pg_ctx, ctx), enabled ? 1 : 0); }
NATIVE(jint, ctxKernelTime)(JNIEnv* env, jclass cls, jlong ctx, jstring nameSubstr, jboolean reset, jdoubleArray outMs, jlongArray outLaunches) {
    const char* s = nameSubstr ? (*env)->GetStringUTFChars(env, nameSubstr, NULL) : NULL;
    double ms = 0;
    int64_t launches = 0;
    const int32_t rc = pg_ctx_kernel_time(H(pg_ctx, ctx), s, reset ? 1 : 0, &ms, &launches);
    if (s) (*env)->ReleaseStringUTFChars(env, nameSubstr, s);
    jdouble jms = ms;
    if (outMs) (*env)->SetDoubleArrayRegion(env, outMs, 0, 1, &jms);
    pgj_set_long(env, outLaunches, 0, launches);
    return rc;
}
NATIVE(jstring, ctxKernelReport)(JNIEnv* env, jclass cls, jlong ctx) {
    static __thread char buf[1 << 16];
    buf[0] = 0;
    pg_ctx_kernel_report(H(pg_ctx, ctx), buf, (int64_t)sizeof(buf));
    return (*env)->NewStringUTF(env, buf);
}

/* ================================================================= GroupByHash (OP/GroupByHash.java:31-88) */
NATIVE(jint, groupbyCreate)(JNIEnv* env, jclass cls, jlong ctx, jintArray keyTypes, jboolean hasHashChannel, jlong expectedSize, jlongArray out) {
    const jsize nkeys = (*env)->GetArrayLength(env, keyTypes);
    jint kt[PGJ_MAX_COLS];
    if (nkeys > PGJ_MAX_COLS) return PG_E_INVALID_ARGUMENT;
    (*env)->GetIntArrayRegion(env, keyTypes, 0, nkeys, kt);
    pg_groupby* gb = NULL;
    const int32_t rc = pg_groupby_create(H(pg_ctx, ctx), (const int32_t*)kt, nkeys, hasHashChannel ? 1 : 0, expectedSize, &gb);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)gb);
    return rc;
}
NATIVE(jint, groupbyDestroy)(JNIEnv* env, jclass cls, jlong gb) { return pg_groupby_destroy(H(pg_groupby, gb)); }

/* getGroupIds(Page): the page holds the key channels, then the precomputed hash channel when lastIsHash */
NATIVE(jint, groupbyGetGroupIds)(JNIEnv* env, jclass cls, jlong gb, PAGE_PARAMS, jboolean lastIsHash, jintArray outIds, jlongArray outGroupCount) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    int64_t groups = 0;
    if (rc == PG_OK && (!pgj_scratch_reset((size_t)positionCount * 4) || ncols < 1)) rc = PG_E_INSUFFICIENT_RESOURCES;
    if (rc == PG_OK) {
        int32_t* ids = (int32_t*)pgj_scratch((size_t)positionCount * 4);
        rc = pg_groupby_get_group_ids(H(pg_groupby, gb), cols, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, ids, PG_MEM_HOST, &groups);
        if (rc == PG_OK) (*env)->SetIntArrayRegion(env, outIds, 0, (jsize)positionCount, (const jint*)ids);
    }
    pgj_set_long(env, outGroupCount, 0, groups);
    return rc;
}
NATIVE(jint, groupbyAddPage)(JNIEnv* env, jclass cls, jlong gb, PAGE_PARAMS, jboolean lastIsHash) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK && ncols < 1) rc = PG_E_INVALID_ARGUMENT;
    if (rc == PG_OK) rc = pg_groupby_add_page(H(pg_groupby, gb), cols, lastIsHash ? &cols[ncols - 1] : NULL, positionCount);
    return rc;
}
NATIVE(jint, groupbyContains)(JNIEnv* env, jclass cls, jlong gb, PAGE_PARAMS, jboolean lastIsHash, jbooleanArray out) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK && (!pgj_scratch_reset((size_t)positionCount) || ncols < 1)) rc = PG_E_INSUFFICIENT_RESOURCES;
    if (rc == PG_OK) {
        uint8_t* found = (uint8_t*)pgj_scratch((size_t)positionCount);
        rc = pg_groupby_contains(H(pg_groupby, gb), cols, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, found, PG_MEM_HOST);
        if (rc == PG_OK) (*env)->SetBooleanArrayRegion(env, out, 0, (jsize)positionCount, (const jboolean*)found);
    }
    return rc;
}
/* MarkDistinctHash.markDistinctRows(page) */
NATIVE(jint, groupbyMarkDistinct)(JNIEnv* env, jclass cls, jlong gb, PAGE_PARAMS, jboolean lastIsHash, jbooleanArray outMarks) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK && (!pgj_scratch_reset((size_t)positionCount) || ncols < 1)) rc = PG_E_INSUFFICIENT_RESOURCES;
    if (rc == PG_OK) {
        uint8_t* marks = (uint8_t*)pgj_scratch((size_t)positionCount);
        rc = pg_groupby_mark_distinct(H(pg_groupby, gb), cols, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, marks, PG_MEM_HOST);
        if (rc == PG_OK) (*env)->SetBooleanArrayRegion(env, outMarks, 0, (jsize)positionCount, (const jboolean*)marks);
    }
    return rc;
}
/* HashSemiJoinOperator's output column for one probe page: page = probe join channel [, hash channel] */
NATIVE(jint, groupbySemiJoin)(JNIEnv* env, jclass cls, jlong gb, PAGE_PARAMS, jboolean lastIsHash, jboolean setContainsNull, jbooleanArray outValues, jbooleanArray outNulls) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK && (!pgj_scratch_reset(2 * (size_t)positionCount + 128) || ncols < 1)) rc = PG_E_INSUFFICIENT_RESOURCES;
    if (rc == PG_OK) {
        uint8_t* vals = (uint8_t*)pgj_scratch((size_t)positionCount);
        uint8_t* nulls = (uint8_t*)pgj_scratch((size_t)positionCount);
        rc = pg_groupby_semi_join(H(pg_groupby, gb), &cols[0], lastIsHash ? &cols[ncols - 1] : NULL, positionCount, setContainsNull ? 1 : 0, vals, nulls, PG_MEM_HOST);
        if (rc == PG_OK) {
            (*env)->SetBooleanArrayRegion(env, outValues, 0, (jsize)positionCount, (const jboolean*)vals);
            (*env)->SetBooleanArrayRegion(env, outNulls, 0, (jsize)positionCount, (const jboolean*)nulls);
        }
    }
    return rc;
}
/* appendValuesTo for groupId in [first, first + n): key columns into the out* arrays, getRawHash into outHash (may be null) */
NATIVE(jint, groupbyExportKeys)(JNIEnv* env, jclass cls, jlong gb, jlong first, jlong n, OUT_PARAMS, jlongArray outHash) {
    pgj_outs o;
    int32_t rc = pgj_out_begin(env, outTypes, outValues, outNulls, outOffsets, outValueBytes, n, (size_t)n * 8, &o);
    if (rc != PG_OK) return rc;
    int64_t* hash = outHash ? (int64_t*)pgj_scratch((size_t)n * 8) : NULL;
    rc = pg_groupby_export_keys(H(pg_groupby, gb), first, n, o.cols, hash, PG_MEM_HOST);
    if (rc == PG_OK) {
        pgj_out_end(env, &o, n);
        if (hash) (*env)->SetLongArrayRegion(env, outHash, 0, (jsize)n, (const jlong*)hash);
    }
    return rc;
}
NATIVE(jint, groupbyVarwidthBytes)(JNIEnv* env, jclass cls, jlong gb, jint key, jlongArray outBytes) {
    int64_t b = 0;
    const int32_t rc = pg_groupby_varwidth_bytes(H(pg_groupby, gb), key, &b);
    pgj_set_long(env, outBytes, 0, b);
    return rc;
}
NATIVE(jint, groupbySetMemoryLimit)(JNIEnv* env, jclass cls, jlong gb, jlong bytes) { return pg_groupby_set_memory_limit(H(pg_groupby, gb), bytes); }
/* out4 = groupCount, capacity, hashCollisions, device bytes (getGroupCount / getCapacity / getHashCollisions / getEstimatedSize) */
NATIVE(jint, groupbyStats)(JNIEnv* env, jclass cls, jlong gb, jlongArray out4) {
    int64_t v[4] = {0, 0, 0, 0};
    const int32_t rc = pg_groupby_stats(H(pg_groupby, gb), &v[0], &v[1], &v[2], &v[3]);
    for (int i = 0; i < 4; i++) pgj_set_long(env, out4, i, v[i]);
    return rc;
}

/* ================================================================= GroupedAccumulator (OP/aggregation/GroupedAccumulator.java:22-39) */
NATIVE(jint, aggCreate)(JNIEnv* env, jclass cls, jlong ctx, jint kind, jint inType, jlongArray out) {
    pg_agg* a = NULL;
    const int32_t rc = pg_agg_create(H(pg_ctx, ctx), kind, inType, &a);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)a);
    return rc;
}
NATIVE(jint, aggDestroy)(JNIEnv* env, jclass cls, jlong agg) { return pg_agg_destroy(H(pg_agg, agg)); }

/* ids travel as one more INTEGER channel at the end of the page (GroupByIdBlock is an IntArrayBlock of group ids) */
static int split_ids(pg_col* cols, int ncols, const int32_t** ids) {
    if (ncols < 1 || cols[ncols - 1].type != PG_INTEGER) return PG_E_INVALID_ARGUMENT;
    *ids = (const int32_t*)cols[ncols - 1].values;
    return PG_OK;
}
/* addInput(GroupByIdBlock, Page): page = [argument (absent for count(*))] [mask when hasMask] ids */
NATIVE(jint, aggAddInput)(JNIEnv* env, jclass cls, jlong agg, jlong groupCount, PAGE_PARAMS, jboolean hasArg, jboolean hasMask) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    const int32_t* ids = NULL;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK) rc = split_ids(cols, ncols, &ids);
    if (rc == PG_OK && ncols != 1 + (hasArg ? 1 : 0) + (hasMask ? 1 : 0)) rc = PG_E_INVALID_ARGUMENT;
    if (rc == PG_OK) rc = pg_agg_add_input(H(pg_agg, agg), ids, PG_MEM_HOST, groupCount, hasArg ? &cols[0] : NULL, hasMask ? &cols[hasArg ? 1 : 0] : NULL, positionCount);
    return rc;
}
/* addIntermediate(GroupByIdBlock, Block): page = the intermediate state columns, then ids */
NATIVE(jint, aggAddIntermediate)(JNIEnv* env, jclass cls, jlong agg, jlong groupCount, PAGE_PARAMS) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    const int32_t* ids = NULL;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK) rc = split_ids(cols, ncols, &ids);
    if (rc == PG_OK) rc = pg_agg_add_intermediate(H(pg_agg, agg), ids, PG_MEM_HOST, groupCount, cols, positionCount);
    return rc;
}
NATIVE(jint, aggIntermediateColumns)(JNIEnv* env, jclass cls, jlong agg, jintArray out) {
    int32_t n = 0;
    const int32_t rc = pg_agg_n_intermediate(H(pg_agg, agg), &n);
    pgj_set_int(env, out, 0, n);
    return rc;
}
/* evaluateFinal / evaluateIntermediate for groupId in [first, first + n) */
NATIVE(jint, aggEvaluate)(JNIEnv* env, jclass cls, jlong agg, jlong first, jlong n, jboolean finalStep, OUT_PARAMS, jlongArray errorGroup) {
    pgj_outs o;
    int32_t rc = pgj_out_begin(env, outTypes, outValues, outNulls, outOffsets, outValueBytes, n, 0, &o);
    if (rc != PG_OK) return rc;
    int64_t eg = -1;
    rc = pg_agg_evaluate(H(pg_agg, agg), first, n, finalStep ? 1 : 0, o.cols, &eg);
    if (rc == PG_OK) pgj_out_end(env, &o, n);
    pgj_set_long(env, errorGroup, 0, eg);
    return rc;
}
NATIVE(jint, aggDeviceBytes)(JNIEnv* env, jclass cls, jlong agg, jlongArray out) {
    int64_t b = 0;
    const int32_t rc = pg_agg_device_bytes(H(pg_agg, agg), &b);
    pgj_set_long(env, out, 0, b);
    return rc;
}

/* ================================================================= HashBuilderOperator / LookupSource (OP/HashBuilderOperator.java:56-635) */
NATIVE(jint, joinbuildCreate)(JNIEnv* env, jclass cls, jlong ctx, jintArray buildTypes, jintArray joinChannels, jint hashChannel, jlong expectedRows, jlongArray out) {
    const jsize ncols = (*env)->GetArrayLength(env, buildTypes), njoin = (*env)->GetArrayLength(env, joinChannels);
    jint t[PGJ_MAX_COLS], jc[PGJ_MAX_COLS];
    if (ncols > PGJ_MAX_COLS || njoin > PGJ_MAX_COLS) return PG_E_INVALID_ARGUMENT;
    (*env)->GetIntArrayRegion(env, buildTypes, 0, ncols, t);
    (*env)->GetIntArrayRegion(env, joinChannels, 0, njoin, jc);
    pg_joinbuild* b = NULL;
    const int32_t rc = pg_joinbuild_create(H(pg_ctx, ctx), (const int32_t*)t, ncols, (const int32_t*)jc, njoin, hashChannel, expectedRows, &b);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)b);
    return rc;
}
NATIVE(jint, joinbuildDestroy)(JNIEnv* env, jclass cls, jlong b) { return pg_joinbuild_destroy(H(pg_joinbuild, b)); }
NATIVE(jint, joinbuildAddPage)(JNIEnv* env, jclass cls, jlong b, PAGE_PARAMS) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK) rc = pg_joinbuild_add_page(H(pg_joinbuild, b), cols, positionCount);
    return rc;
}
NATIVE(jint, joinbuildFinish)(JNIEnv* env, jclass cls, jlong b, jlongArray outLookup) {
    pg_lookup* ls = NULL;
    const int32_t rc = pg_joinbuild_finish(H(pg_joinbuild, b), &ls);
    pgj_set_long(env, outLookup, 0, (int64_t)(intptr_t)ls);
    return rc;
}
NATIVE(jint, lookupDestroy)(JNIEnv* env, jclass cls, jlong ls) { return pg_lookup_destroy(H(pg_lookup, ls)); }
/* out4 = join position count, PagesHash size, device bytes, 1 when position links exist */
NATIVE(jint, lookupStats)(JNIEnv* env, jclass cls, jlong ls, jlongArray out4) {
    int64_t v[3] = {0, 0, 0};
    int32_t dups = 0;
    const int32_t rc = pg_lookup_stats(H(pg_lookup, ls), &v[0], &v[1], &v[2], &dups);
    for (int i = 0; i < 3; i++) pgj_set_long(env, out4, i, v[i]);
    pgj_set_long(env, out4, 3, dups);
    return rc;
}
/* one probe page -> (probe position, build position) pairs; PG_NEED_MEMORY: nOut[0] = rows needed, Java grows the arrays and calls again */
NATIVE(jint, lookupProbe)(JNIEnv* env, jclass cls, jlong ls, PAGE_PARAMS, jboolean lastIsHash, jint joinType, jintArray outProbeIdx, jlongArray outBuildPos,
                          jlongArray nOut) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    int64_t produced = 0;
    const jsize cap = (*env)->GetArrayLength(env, outProbeIdx);
    if (rc == PG_OK && (ncols < 1 || (*env)->GetArrayLength(env, outBuildPos) < cap)) rc = PG_E_INVALID_ARGUMENT;
    if (rc == PG_OK && !pgj_scratch_reset((size_t)cap * 12 + 128)) rc = PG_E_INSUFFICIENT_RESOURCES;
    if (rc == PG_OK) {
        int32_t* pi = (int32_t*)pgj_scratch((size_t)cap * 4);
        int64_t* bp = (int64_t*)pgj_scratch((size_t)cap * 8);
        const int nkeys = lastIsHash ? ncols - 1 : ncols;
        rc = pg_lookup_probe(H(pg_lookup, ls), cols, nkeys, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, joinType, pi, bp, PG_MEM_HOST, cap, &produced);
        if (rc == PG_OK) {
            (*env)->SetIntArrayRegion(env, outProbeIdx, 0, (jsize)produced, (const jint*)pi);
            (*env)->SetLongArrayRegion(env, outBuildPos, 0, (jsize)produced, (const jlong*)bp);
        }
    }
    pgj_set_long(env, nOut, 0, produced);
    return rc;
}
/* OuterLookupSource / LookupOuterOperator: visited-positions tracker and the positions no probe emitted */
NATIVE(jint, lookupTrackVisited)(JNIEnv* env, jclass cls, jlong ls) { return pg_lookup_track_visited(H(pg_lookup, ls)); }
NATIVE(jint, lookupOuterPositions)(JNIEnv* env, jclass cls, jlong ls, jlongArray outPositions, jlongArray nOut) {
    const jsize cap = (*env)->GetArrayLength(env, outPositions);
    if (!pgj_scratch_reset((size_t)cap * 8)) return PG_E_INSUFFICIENT_RESOURCES;
    int64_t* pos = (int64_t*)pgj_scratch((size_t)cap * 8);
    int64_t n = 0;
    const int32_t rc = pg_lookup_outer_positions(H(pg_lookup, ls), pos, PG_MEM_HOST, cap, &n);
    if (rc == PG_OK) (*env)->SetLongArrayRegion(env, outPositions, 0, (jsize)n, (const jlong*)pos);
    pgj_set_long(env, nOut, 0, n);
    return rc;
}
/* LookupSource.appendTo for a vector of build positions (-1 -> null) */
NATIVE(jint, lookupGather)(JNIEnv* env, jclass cls, jlong ls, jlongArray buildPos, jlong n, jintArray channels, OUT_PARAMS) {
    const jsize nch = (*env)->GetArrayLength(env, channels);
    jint ch[PGJ_MAX_COLS];
    if (nch > PGJ_MAX_COLS || (*env)->GetArrayLength(env, buildPos) < n) return PG_E_INVALID_ARGUMENT;
    (*env)->GetIntArrayRegion(env, channels, 0, nch, ch);
    pgj_outs o;
    int32_t rc = pgj_out_begin(env, outTypes, outValues, outNulls, outOffsets, outValueBytes, n, (size_t)n * 8, &o);
    if (rc != PG_OK) return rc;
    int64_t* pos = (int64_t*)pgj_scratch((size_t)n * 8);
    (*env)->GetLongArrayRegion(env, buildPos, 0, (jsize)n, (jlong*)pos);
    rc = pg_lookup_gather(H(pg_lookup, ls), pos, PG_MEM_HOST, n, (const int32_t*)ch, nch, o.cols);
    if (rc == PG_OK) pgj_out_end(env, &o, n);
    return rc;
}
NATIVE(jint, lookupGatherVarBytes)(JNIEnv* env, jclass cls, jlong ls, jlongArray buildPos, jlong n, jint channel, jlongArray outBytes) {
    if ((*env)->GetArrayLength(env, buildPos) < n || !pgj_scratch_reset((size_t)n * 8)) return PG_E_INVALID_ARGUMENT;
    int64_t* pos = (int64_t*)pgj_scratch((size_t)n * 8);
    (*env)->GetLongArrayRegion(env, buildPos, 0, (jsize)n, (jlong*)pos);
    int64_t b = 0;
    const int32_t rc = pg_lookup_gather_var_bytes(H(pg_lookup, ls), pos, PG_MEM_HOST, n, channel, &b);
    pgj_set_long(env, outBytes, 0, b);
    return rc;
}
/* LookupJoinOperator's step in one call: probe positions + the build channels of the matched rows (fixed width channels) */
NATIVE(jint, lookupJoinPage)(JNIEnv* env, jclass cls, jlong ls, PAGE_PARAMS, jboolean lastIsHash, jint joinType, jintArray buildChannels, jintArray outProbeIdx,
                             OUT_PARAMS, jlongArray nOut) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    const jsize nch = (*env)->GetArrayLength(env, buildChannels), cap = (*env)->GetArrayLength(env, outProbeIdx);
    jint ch[PGJ_MAX_COLS];
    if (rc == PG_OK && (nch > PGJ_MAX_COLS || ncols < 1)) rc = PG_E_INVALID_ARGUMENT;
    if (rc != PG_OK) return rc;
    (*env)->GetIntArrayRegion(env, buildChannels, 0, nch, ch);
    pgj_outs o;
    rc = pgj_out_begin(env, outTypes, outValues, outNulls, outOffsets, outValueBytes, cap, (size_t)cap * 4, &o);
    if (rc != PG_OK) return rc;
    int32_t* pi = (int32_t*)pgj_scratch((size_t)cap * 4);
    int64_t produced = 0;
    const int nkeys = lastIsHash ? ncols - 1 : ncols;
    rc = pg_lookup_join_page(H(pg_lookup, ls), cols, nkeys, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, joinType, (const int32_t*)ch, nch, pi,
                             PG_MEM_HOST, o.cols, cap, &produced);
    if (rc == PG_OK) {
        (*env)->SetIntArrayRegion(env, outProbeIdx, 0, (jsize)produced, (const jint*)pi);
        pgj_out_end(env, &o, produced);
    }
    pgj_set_long(env, nOut, 0, produced);
    return rc;
}

/* ================================================================= fused operator (OP/HashAggregationOperator.java:381-518) */
NATIVE(jint, hashaggCreate)(JNIEnv* env, jclass cls, jlong ctx, jintArray keyChannels, jintArray keyTypes, jint hashChannel,
                            jintArray aggSpecs /* kind, inType, inputChannel, maskChannel per aggregate */, jint step, jlong expectedGroups, jlongArray out) {
    const jsize nkeys = (*env)->GetArrayLength(env, keyChannels), nspec = (*env)->GetArrayLength(env, aggSpecs) / 4;
    jint kc[PGJ_MAX_COLS], kt[PGJ_MAX_COLS], sp[4 * PGJ_MAX_COLS];
    if (nkeys > PGJ_MAX_COLS || nspec > PGJ_MAX_COLS || (*env)->GetArrayLength(env, keyTypes) != nkeys) return PG_E_INVALID_ARGUMENT;
    (*env)->GetIntArrayRegion(env, keyChannels, 0, nkeys, kc);
    (*env)->GetIntArrayRegion(env, keyTypes, 0, nkeys, kt);
    (*env)->GetIntArrayRegion(env, aggSpecs, 0, nspec * 4, sp);
    pg_agg_spec specs[PGJ_MAX_COLS];
    for (jsize a = 0; a < nspec; a++) {
        specs[a].kind = sp[4 * a]; specs[a].in_type = sp[4 * a + 1];
        specs[a].input_channel = sp[4 * a + 2]; specs[a].mask_channel = sp[4 * a + 3];
    }
    pg_hashagg* op = NULL;
    const int32_t rc = pg_hashagg_create(H(pg_ctx, ctx), (const int32_t*)kc, (const int32_t*)kt, nkeys, hashChannel, specs, nspec, step, expectedGroups, &op);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)op);
    return rc;
}
NATIVE(jint, hashaggDestroy)(JNIEnv* env, jclass cls, jlong op) { return pg_hashagg_destroy(H(pg_hashagg, op)); }
NATIVE(jint, hashaggAddPage)(JNIEnv* env, jclass cls, jlong op, PAGE_PARAMS) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK) rc = pg_hashagg_add_page(H(pg_hashagg, op), cols, ncols, positionCount);
    return rc;      /* PG_NEED_MEMORY: the table has to grow past the limit set with hashaggSetMemoryLimit; nothing consumed (Work.process() == false) */
}
/* out4 = group count, capacity, device bytes, input rows */
NATIVE(jint, hashaggStats)(JNIEnv* env, jclass cls, jlong op, jlongArray out4) {
    int64_t v[4] = {0, 0, 0, 0};
    const int32_t rc = pg_hashagg_stats(H(pg_hashagg, op), &v[0], &v[1], &v[2], &v[3]);
    for (int i = 0; i < 4; i++) pgj_set_long(env, out4, i, v[i]);
    return rc;
}
NATIVE(jint, hashaggOutputTypes)(JNIEnv* env, jclass cls, jlong op, jintArray outTypes, jintArray outCount) {
    int32_t t[PGJ_MAX_COLS], n = 0;
    const int32_t rc = pg_hashagg_output_types(H(pg_hashagg, op), t, PGJ_MAX_COLS, &n);
    if (rc == PG_OK && (*env)->GetArrayLength(env, outTypes) >= n) (*env)->SetIntArrayRegion(env, outTypes, 0, n, (const jint*)t);
    pgj_set_int(env, outCount, 0, n);
    return rc;
}
NATIVE(jint, hashaggGetOutput)(JNIEnv* env, jclass cls, jlong op, jlong first, jlong n, OUT_PARAMS, jlongArray errorGroup) {
    pgj_outs o;
    int32_t rc = pgj_out_begin(env, outTypes, outValues, outNulls, outOffsets, outValueBytes, n, 0, &o);
    if (rc != PG_OK) return rc;
    int64_t eg = -1;
    rc = pg_hashagg_get_output(H(pg_hashagg, op), first, n, o.cols, &eg);
    if (rc == PG_OK) pgj_out_end(env, &o, n);
    pgj_set_long(env, errorGroup, 0, eg);
    return rc;
}
NATIVE(jint, hashaggVarwidthBytes)(JNIEnv* env, jclass cls, jlong op, jint col, jlongArray outBytes) {
    int64_t b = 0;
    const int32_t rc = pg_hashagg_varwidth_bytes(H(pg_hashagg, op), col, &b);
    pgj_set_long(env, outBytes, 0, b);
    return rc;
}
NATIVE(jint, hashaggReset)(JNIEnv* env, jclass cls, jlong op) { return pg_hashagg_reset(H(pg_hashagg, op)); }
NATIVE(jint, hashaggSetPageBatching)(JNIEnv* env, jclass cls, jlong op, jlong rows) { return pg_hashagg_set_page_batching(H(pg_hashagg, op), rows); }
/* a page held as one large array set, fed as pages of pageRows rows (each through the per-page path) */
NATIVE(jint, hashaggAddPages)(JNIEnv* env, jclass cls, jlong op, PAGE_PARAMS, jlong pageRows) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc == PG_OK) rc = pg_hashagg_add_pages(H(pg_hashagg, op), cols, ncols, positionCount, pageRows);
    return rc;
}
NATIVE(jint, hashaggSetMemoryLimit)(JNIEnv* env, jclass cls, jlong op, jlong bytes) { return pg_hashagg_set_memory_limit(H(pg_hashagg, op), bytes); }
NATIVE(jint, hashaggSetInputProcessor)(JNIEnv* env, jclass cls, jlong op, jlong project) { return pg_hashagg_set_input_processor(H(pg_hashagg, op), H(pg_project, project)); }
/* packed pages: PARTIAL output -> FINAL input between processes (the job of PagesSerde in the reference's exchange) */
NATIVE(jint, hashaggGetOutputPacked)(JNIEnv* env, jclass cls, jlong op, jlong first, jlong n, jobject directBuffer, jlongArray outBytes, jlongArray errorGroup) {
    void* buf = (*env)->GetDirectBufferAddress(env, directBuffer);
    const jlong cap = (*env)->GetDirectBufferCapacity(env, directBuffer);
    int64_t nbytes = 0, eg = -1;
    const int32_t rc = pg_hashagg_get_output_packed(H(pg_hashagg, op), first, n, buf, cap, &nbytes, &eg);
    pgj_set_long(env, outBytes, 0, nbytes);      /* on PG_NEED_MEMORY: the size a retry needs */
    pgj_set_long(env, errorGroup, 0, eg);
    return rc;
}
NATIVE(jint, hashaggAddPacked)(JNIEnv* env, jclass cls, jlong op, jobjectArray directBuffers) {
    const jsize nb = (*env)->GetArrayLength(env, directBuffers);
    const void* ptrs[PGJ_MAX_COLS];
    if (nb > PGJ_MAX_COLS) return PG_E_INVALID_ARGUMENT;
    for (jsize i = 0; i < nb; i++) ptrs[i] = (*env)->GetDirectBufferAddress(env, (*env)->GetObjectArrayElement(env, directBuffers, i));
    return pg_hashagg_add_packed(H(pg_hashagg, op), ptrs, nb);
}
/* device-resident packed pages: addresses are CUDA device pointers owned by the caller's exchange */
NATIVE(jint, hashaggGetOutputPackedDevice)(JNIEnv* env, jclass cls, jlong op, jlong first, jlong n, jlong deviceBuf, jlong capacity, jlongArray outBytes,
                                           jlongArray errorGroup) {
    int64_t nbytes = 0, eg = -1;
    const int32_t rc = pg_hashagg_get_output_packed_device(H(pg_hashagg, op), first, n, (void*)(intptr_t)deviceBuf, capacity, &nbytes, &eg);
    pgj_set_long(env, outBytes, 0, nbytes);
    pgj_set_long(env, errorGroup, 0, eg);
    return rc;
}
NATIVE(jint, hashaggAddPackedDevice)(JNIEnv* env, jclass cls, jlong op, jlong deviceBufs, jlong stride, jint nbufs) {
    return pg_hashagg_add_packed_device(H(pg_hashagg, op), (const void*)(intptr_t)deviceBufs, stride, nbufs);
}

/* ================================================================= PageProcessor (OP/project/PageProcessor.java:139-330) */
/* nodeInts: 8 per node (op, type, precision, scale, channel, arg0, arg1, isNull); nodeLongs: (lo, hi) per node;
 * nodeBytes[i]: byte[] of a VARCHAR constant or null */
NATIVE(jint, projectCreate)(JNIEnv* env, jclass cls, jlong ctx, jintArray nodeInts, jlongArray nodeLongs, jobjectArray nodeBytes, jint filter,
                            jintArray projections, jlongArray out) {
    const jsize nnodes = (*env)->GetArrayLength(env, nodeInts) / 8, nproj = (*env)->GetArrayLength(env, projections);
    if (nnodes > 256 || nproj > PGJ_MAX_COLS || (*env)->GetArrayLength(env, nodeLongs) < 2 * nnodes) return PG_E_INVALID_ARGUMENT;
    pg_expr* nodes = (pg_expr*)calloc((size_t)nnodes + 1, sizeof(pg_expr));
    jint* ints = (jint*)malloc(((size_t)nnodes * 8 + 1) * sizeof(jint));
    jlong* longs = (jlong*)malloc(((size_t)nnodes * 2 + 1) * sizeof(jlong));
    jbyte* held[256];
    jbyteArray arrs[256];
    jint proj[PGJ_MAX_COLS];
    int32_t rc = PG_OK;
    if (!nodes || !ints || !longs) rc = PG_E_INSUFFICIENT_RESOURCES;
    if (rc == PG_OK) {
        (*env)->GetIntArrayRegion(env, nodeInts, 0, nnodes * 8, ints);
        (*env)->GetLongArrayRegion(env, nodeLongs, 0, nnodes * 2, longs);
        (*env)->GetIntArrayRegion(env, projections, 0, nproj, proj);
    }
    for (jsize i = 0; i < nnodes; i++) { held[i] = NULL; arrs[i] = NULL; }
    for (jsize i = 0; rc == PG_OK && i < nnodes; i++) {
        pg_expr* e = &nodes[i];
        e->op = ints[8 * i]; e->type = ints[8 * i + 1]; e->precision = ints[8 * i + 2]; e->scale = ints[8 * i + 3];
        e->channel = ints[8 * i + 4]; e->arg0 = ints[8 * i + 5]; e->arg1 = ints[8 * i + 6]; e->is_null = ints[8 * i + 7];
        e->lo = longs[2 * i]; e->hi = longs[2 * i + 1];
        arrs[i] = nodeBytes ? (jbyteArray)(*env)->GetObjectArrayElement(env, nodeBytes, i) : NULL;
        if (arrs[i]) {       /* constants are a few bytes and pg_project_create copies them: a plain element copy, no critical section */
            held[i] = (*env)->GetByteArrayElements(env, arrs[i], NULL);
            e->bytes = (const uint8_t*)held[i];
            e->nbytes = (*env)->GetArrayLength(env, arrs[i]);
        }
    }
    pg_project* p = NULL;
    if (rc == PG_OK) rc = pg_project_create(H(pg_ctx, ctx), nodes, nnodes, filter, (const int32_t*)proj, nproj, &p);
    for (jsize i = 0; i < nnodes; i++) if (held[i]) (*env)->ReleaseByteArrayElements(env, arrs[i], held[i], JNI_ABORT);
    free(nodes); free(ints); free(longs);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)p);
    return rc;
}
NATIVE(jint, projectDestroy)(JNIEnv* env, jclass cls, jlong p) { return pg_project_destroy(H(pg_project, p)); }
NATIVE(jint, projectOutputTypes)(JNIEnv* env, jclass cls, jlong p, jintArray outTypes, jintArray outPrecisions, jintArray outScales, jintArray outCount) {
    int32_t t[PGJ_MAX_COLS], pr[PGJ_MAX_COLS], sc[PGJ_MAX_COLS], n = 0;
    const int32_t rc = pg_project_output_types(H(pg_project, p), t, pr, sc, PGJ_MAX_COLS, &n);
    if (rc == PG_OK && (*env)->GetArrayLength(env, outTypes) >= n) {
        (*env)->SetIntArrayRegion(env, outTypes, 0, n, (const jint*)t);
        (*env)->SetIntArrayRegion(env, outPrecisions, 0, n, (const jint*)pr);
        (*env)->SetIntArrayRegion(env, outScales, 0, n, (const jint*)sc);
    }
    pgj_set_int(env, outCount, 0, n);
    return rc;
}
/* PageProcessor.process(page): the projections over the rows that pass the filter; nOut[0] = selected rows */
NATIVE(jint, projectProcess)(JNIEnv* env, jclass cls, jlong p, PAGE_PARAMS, OUT_PARAMS, jlongArray nOut) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    if (rc != PG_OK) return rc;
    pgj_outs o;
    rc = pgj_out_begin(env, outTypes, outValues, outNulls, outOffsets, outValueBytes, positionCount, 0, &o);
    if (rc != PG_OK) return rc;
    int64_t selected = 0;
    rc = pg_project_process(H(pg_project, p), cols, ncols, positionCount, o.cols, &selected);
    if (rc == PG_OK) pgj_out_end(env, &o, selected);
    pgj_set_long(env, nOut, 0, selected);
    return rc;
}

/* ================================================================= hash partitioning (device-resident exchange)
 * These entry points move rows between GPUs and take CUDA device pointers; a JVM host reaches them only through an exchange that
 * owns device buffers (addresses travel as jlong).  The key page of partitionCreate may be a host page like everywhere else. */
NATIVE(jint, partitionCreate)(JNIEnv* env, jclass cls, jlong ctx, PAGE_PARAMS, jboolean lastIsHash, jint nparts, jint mode, jlongArray outCounts, jlongArray out) {
    pg_col cols[PGJ_MAX_COLS];
    int ncols = 0;
    int32_t rc = pgj_page_in(env, PAGE_ARGS, cols, &ncols);
    int64_t counts[64];
    memset(counts, 0, sizeof(counts));
    pg_partitioner* pt = NULL;
    if (rc == PG_OK && (ncols < 1 || nparts > 64)) rc = PG_E_INVALID_ARGUMENT;
    if (rc == PG_OK) rc = pg_partition_create(H(pg_ctx, ctx), cols, lastIsHash ? ncols - 1 : ncols, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, nparts, mode, counts, &pt);
    for (int i = 0; rc == PG_OK && i < nparts; i++) pgj_set_long(env, outCounts, i, counts[i]);
    pgj_set_long(env, out, 0, (int64_t)(intptr_t)pt);
    return rc;
}
NATIVE(jint, partitionDestroy)(JNIEnv* env, jclass cls, jlong pt) { return pg_partition_destroy(H(pg_partitioner, pt)); }
static void device_col(pg_col* c, jint type, jlong values, jlong nulls, jlong n) {
    memset(c, 0, sizeof(*c));
    c->type = type; c->mem = PG_MEM_DEVICE; c->n = n; c->values = (const void*)(intptr_t)values; c->nulls = (const uint8_t*)(intptr_t)nulls;
}
NATIVE(jint, partitionScatter)(JNIEnv* env, jclass cls, jlong pt, jint type, jlong n, jlong inValues, jlong inNulls, jlong outValues, jlong outNulls) {
    pg_col in;
    pg_col_out out;
    device_col(&in, type, inValues, inNulls, n);
    memset(&out, 0, sizeof(out));
    out.type = type; out.mem = PG_MEM_DEVICE; out.capacity = n; out.values = (void*)(intptr_t)outValues; out.nulls = (uint8_t*)(intptr_t)outNulls;
    return pg_partition_scatter(H(pg_partitioner, pt), &in, &out);
}
NATIVE(jint, partitionScatterPeers)(JNIEnv* env, jclass cls, jlong pt, jint type, jlong n, jlong inValues, jlong inNulls, jlongArray peerValues, jlongArray peerNulls,
                                    jlongArray peerRowOffset) {
    const jsize np = (*env)->GetArrayLength(env, peerValues);
    jlong pv[64], pn[64], po[64];
    if (np > 64) return PG_E_INVALID_ARGUMENT;
    memset(pn, 0, sizeof(pn));
    (*env)->GetLongArrayRegion(env, peerValues, 0, np, pv);
    if (peerNulls) (*env)->GetLongArrayRegion(env, peerNulls, 0, np, pn);
    (*env)->GetLongArrayRegion(env, peerRowOffset, 0, np, po);
    void* v[64];
    uint8_t* nl[64];
    int64_t off[64];
    for (jsize i = 0; i < np; i++) { v[i] = (void*)(intptr_t)pv[i]; nl[i] = (uint8_t*)(intptr_t)pn[i]; off[i] = po[i]; }
    pg_col in;
    device_col(&in, type, inValues, inNulls, n);
    return pg_partition_scatter_peers(H(pg_partitioner, pt), &in, v, peerNulls ? nl : NULL, off);
}
/* every column of the page in one launch; peerValues / peerNulls hold ncols x nparts device addresses ([c * nparts + p]) */
NATIVE(jint, partitionScatterPeersAll)(JNIEnv* env, jclass cls, jlong pt, jintArray colTypes, jlong n, jlongArray inValues, jlongArray inNulls, jlongArray peerValues,
                                       jlongArray peerNulls, jlongArray peerRowOffset) {
    const jsize ncols = (*env)->GetArrayLength(env, colTypes), np = (*env)->GetArrayLength(env, peerRowOffset);
    if (ncols > 16 || np > 64 || (*env)->GetArrayLength(env, peerValues) != ncols * np) return PG_E_INVALID_ARGUMENT;
    jint t[16];
    jlong iv[16], in_[16], po[64];
    static __thread jlong pv[16 * 64], pn[16 * 64];
    static __thread void* v[16 * 64];
    static __thread uint8_t* nl[16 * 64];
    memset(in_, 0, sizeof(in_));
    memset(pn, 0, sizeof(pn));
    (*env)->GetIntArrayRegion(env, colTypes, 0, ncols, t);
    (*env)->GetLongArrayRegion(env, inValues, 0, ncols, iv);
    if (inNulls) (*env)->GetLongArrayRegion(env, inNulls, 0, ncols, in_);
    (*env)->GetLongArrayRegion(env, peerValues, 0, ncols * np, pv);
    if (peerNulls) (*env)->GetLongArrayRegion(env, peerNulls, 0, ncols * np, pn);
    (*env)->GetLongArrayRegion(env, peerRowOffset, 0, np, po);
    pg_col cols[16];
    int64_t off[64];
    for (jsize c = 0; c < ncols; c++) device_col(&cols[c], t[c], iv[c], in_[c], n);
    for (jsize i = 0; i < ncols * np; i++) { v[i] = (void*)(intptr_t)pv[i]; nl[i] = (uint8_t*)(intptr_t)pn[i]; }
    for (jsize i = 0; i < np; i++) off[i] = po[i];
    return pg_partition_scatter_peers_all(H(pg_partitioner, pt), cols, ncols, v, nl, off);
}
NATIVE(jint, partitionIds)(JNIEnv* env, jclass cls, jlong pt, jbyteArray out) {
    const jsize n = (*env)->GetArrayLength(env, out);
    if (!pgj_scratch_reset((size_t)n)) return PG_E_INSUFFICIENT_RESOURCES;
    uint8_t* ids = (uint8_t*)pgj_scratch((size_t)n);
    const int32_t rc = pg_partition_ids(H(pg_partitioner, pt), ids, PG_MEM_HOST);
    if (rc == PG_OK) (*env)->SetByteArrayRegion(env, out, 0, n, (const jbyte*)ids);
    return rc;
}
/* no allocation, no handle, no synchronisation: every address is a device pointer (see include/prestogpu.h) */
NATIVE(jint, partitionPositionsNoalloc)(JNIEnv* env, jclass cls, jlong ctx, jlong rawHash, jlong n, jint nparts, jint mode, jlong outPartOfRow, jlong outDest,
                                        jlong outCounts) {
    return pg_partition_positions_noalloc(H(pg_ctx, ctx), (const int64_t*)(intptr_t)rawHash, n, nparts, mode, (uint8_t*)(intptr_t)outPartOfRow,
                                          (int32_t*)(intptr_t)outDest, (int64_t*)(intptr_t)outCounts);
}
