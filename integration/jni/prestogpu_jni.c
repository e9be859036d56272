/* prestogpu_jni.c — JNI shim between io.trino.gpu.PrestoGpu (integration/java) and libprestogpu.so (include/prestogpu.h).
 *
 * NOT built in this repository: the image has no JDK (no jni.h).  A maintainer builds it next to the reference with
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude \
 *       integration/jni/prestogpu_jni.c -Lpresto-1_b200 -lprestogpu -o libprestogpu_jni.so
 *
 * Conventions
 *  - handles travel as jlong; every native returns the pg status, Java turns non-zero into TrinoException
 *    (PrestoGpu.check): PG_E_INSUFFICIENT_RESOURCES -> GENERIC_INSUFFICIENT_RESOURCES,
 *    PG_E_NUMERIC_OVERFLOW -> NUMERIC_VALUE_OUT_OF_RANGE, PG_E_INVALID_ARGUMENT -> IllegalArgumentException, ...
 *  - a Block crosses as its backing arrays: long[]/int[]/byte[] values, boolean[] valueIsNull (1 byte per element in
 *    HotSpot), int[] offsets, plus arrayOffset and positionCount (LongArrayBlock.java:37-76, VariableWidthBlock.java:38-48).
 *    Arrays are pinned with GetPrimitiveArrayCritical only for the duration of the call; the library copies to the device
 *    before returning (include/prestogpu.h, "ownership").
 */
#include <jni.h>
#include <stdint.h>
#include <string.h>
#include "prestogpu.h"

#define MAX_COLS 64

typedef struct {
    jarray values, nulls, offsets;
    void *pv, *pn, *po;
} pinned_col;

/* Java side passes parallel arrays describing the page's channels: types[], values[] (Object[] of primitive arrays),
 * nulls[] (Object[] of boolean[] or null), offsets[] (Object[] of int[] or null), arrayOffsets[]. */
static int pin_page(JNIEnv* env, jintArray jtypes, jobjectArray jvalues, jobjectArray jnulls, jobjectArray joffsets,
                    jintArray jarray_offsets, jlong n, pg_col* cols, pinned_col* pins, int* ncols_out) {
    jsize ncols = (*env)->GetArrayLength(env, jtypes);
    if (ncols > MAX_COLS) return PG_E_INVALID_ARGUMENT;
    jint types[MAX_COLS], aoff[MAX_COLS];
    (*env)->GetIntArrayRegion(env, jtypes, 0, ncols, types);
    (*env)->GetIntArrayRegion(env, jarray_offsets, 0, ncols, aoff);
    for (jsize c = 0; c < ncols; c++) {
        pinned_col* p = &pins[c];
        memset(p, 0, sizeof(*p));
        p->values = (jarray)(*env)->GetObjectArrayElement(env, jvalues, c);
        p->nulls = (jarray)(*env)->GetObjectArrayElement(env, jnulls, c);
        p->offsets = (jarray)(*env)->GetObjectArrayElement(env, joffsets, c);
    }
    /* no JNI calls other than Get/ReleasePrimitiveArrayCritical between here and unpin_page */
    for (jsize c = 0; c < ncols; c++) {
        pinned_col* p = &pins[c];
        p->pv = p->values ? (*env)->GetPrimitiveArrayCritical(env, p->values, NULL) : NULL;
        p->pn = p->nulls ? (*env)->GetPrimitiveArrayCritical(env, p->nulls, NULL) : NULL;
        p->po = p->offsets ? (*env)->GetPrimitiveArrayCritical(env, p->offsets, NULL) : NULL;
        pg_col* col = &cols[c];
        memset(col, 0, sizeof(*col));
        col->type = types[c];
        col->mem = PG_MEM_HOST;
        col->n = n;
        size_t width = (types[c] == PG_INTEGER || types[c] == PG_DATE) ? 4 : (types[c] == PG_LONG_DECIMAL) ? 16 :
                       (types[c] == PG_BOOLEAN) ? 1 : 8;
        if (types[c] == PG_VARCHAR || types[c] == PG_VARBINARY) {
            col->values = p->pv;                                   /* Slice bytes; offsets are absolute */
            col->offsets = (const int32_t*)p->po + aoff[c];
        } else {
            col->values = (const char*)p->pv + (size_t)aoff[c] * width;
        }
        col->nulls = p->pn ? (const uint8_t*)p->pn + aoff[c] : NULL;
    }
    *ncols_out = (int)ncols;
    return PG_OK;
}

static void unpin_page(JNIEnv* env, pinned_col* pins, int ncols) {
    for (int c = ncols - 1; c >= 0; c--) {
        if (pins[c].po) (*env)->ReleasePrimitiveArrayCritical(env, pins[c].offsets, pins[c].po, JNI_ABORT);
        if (pins[c].pn) (*env)->ReleasePrimitiveArrayCritical(env, pins[c].nulls, pins[c].pn, JNI_ABORT);
        if (pins[c].pv) (*env)->ReleasePrimitiveArrayCritical(env, pins[c].values, pins[c].pv, JNI_ABORT);
    }
}

JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_ctxCreate(JNIEnv* env, jclass cls, jint device, jlongArray out) {
    pg_ctx* ctx = NULL;
    int32_t rc = pg_ctx_create(device, &ctx);
    jlong h = (jlong)(intptr_t)ctx;
    (*env)->SetLongArrayRegion(env, out, 0, 1, &h);
    return rc;
}

JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_ctxDestroy(JNIEnv* env, jclass cls, jlong ctx) {
    return pg_ctx_destroy((pg_ctx*)(intptr_t)ctx);
}

JNIEXPORT jstring JNICALL Java_io_trino_gpu_PrestoGpu_lastError(JNIEnv* env, jclass cls, jlong ctx) {
    const char* msg = pg_last_error((pg_ctx*)(intptr_t)ctx);
    return (*env)->NewStringUTF(env, msg ? msg : "");
}

/* ---- fused operator: HashAggregationOperator.addInput / getOutput (OP/HashAggregationOperator.java:381-518) ---- */
JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggCreate(JNIEnv* env, jclass cls, jlong ctx, jintArray keyChannels,
        jintArray keyTypes, jint hashChannel, jintArray aggSpecs /* kind,inType,inputChannel,maskChannel per aggregate */,
        jint step, jlong expectedGroups, jlongArray out) {
    jsize nkeys = (*env)->GetArrayLength(env, keyChannels);
    jsize nspec = (*env)->GetArrayLength(env, aggSpecs) / 4;
    jint kc[MAX_COLS], kt[MAX_COLS], sp[4 * MAX_COLS];
    if (nkeys > MAX_COLS || nspec > MAX_COLS) return PG_E_INVALID_ARGUMENT;
    (*env)->GetIntArrayRegion(env, keyChannels, 0, nkeys, kc);
    (*env)->GetIntArrayRegion(env, keyTypes, 0, nkeys, kt);
    (*env)->GetIntArrayRegion(env, aggSpecs, 0, nspec * 4, sp);
    pg_agg_spec specs[MAX_COLS];
    for (jsize a = 0; a < nspec; a++) {
        specs[a].kind = sp[4 * a]; specs[a].in_type = sp[4 * a + 1];
        specs[a].input_channel = sp[4 * a + 2]; specs[a].mask_channel = sp[4 * a + 3];
    }
    pg_hashagg* op = NULL;
    int32_t rc = pg_hashagg_create((pg_ctx*)(intptr_t)ctx, kc, kt, nkeys, hashChannel, specs, nspec, step, expectedGroups, &op);
    jlong h = (jlong)(intptr_t)op;
    (*env)->SetLongArrayRegion(env, out, 0, 1, &h);
    return rc;
}

JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggAddPage(JNIEnv* env, jclass cls, jlong op, jintArray types,
        jobjectArray values, jobjectArray nulls, jobjectArray offsets, jintArray arrayOffsets, jlong positionCount) {
    pg_col cols[MAX_COLS];
    pinned_col pins[MAX_COLS];
    int ncols = 0;
    int32_t rc = pin_page(env, types, values, nulls, offsets, arrayOffsets, positionCount, cols, pins, &ncols);
    if (rc == PG_OK) rc = pg_hashagg_add_page((pg_hashagg*)(intptr_t)op, cols, ncols, positionCount);
    unpin_page(env, pins, ncols);
    return rc;
}

JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggStats(JNIEnv* env, jclass cls, jlong op, jlongArray out4) {
    int64_t v[4] = {0, 0, 0, 0};
    int32_t rc = pg_hashagg_stats((pg_hashagg*)(intptr_t)op, &v[0], &v[1], &v[2], &v[3]);
    (*env)->SetLongArrayRegion(env, out4, 0, 4, (const jlong*)v);
    return rc;
}

/* getOutput: Java allocates the output arrays (one per output channel, sized from hashaggStats/hashaggVarwidthBytes)
 * and wraps them into LongArrayBlock/IntArrayBlock/Int128ArrayBlock/VariableWidthBlock afterwards. */
JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggGetOutput(JNIEnv* env, jclass cls, jlong op, jlong first, jlong n,
        jintArray types, jobjectArray values, jobjectArray nulls, jobjectArray offsets, jlongArray valueBytes, jlongArray errorGroup) {
    jsize ncols = (*env)->GetArrayLength(env, types);
    if (ncols > MAX_COLS) return PG_E_INVALID_ARGUMENT;
    jint t[MAX_COLS];
    jlong vb[MAX_COLS];
    (*env)->GetIntArrayRegion(env, types, 0, ncols, t);
    (*env)->GetLongArrayRegion(env, valueBytes, 0, ncols, vb);
    pg_col_out outs[MAX_COLS];
    pinned_col pins[MAX_COLS];
    for (jsize c = 0; c < ncols; c++) {
        memset(&pins[c], 0, sizeof(pins[c]));
        pins[c].values = (jarray)(*env)->GetObjectArrayElement(env, values, c);
        pins[c].nulls = (jarray)(*env)->GetObjectArrayElement(env, nulls, c);
        pins[c].offsets = (jarray)(*env)->GetObjectArrayElement(env, offsets, c);
    }
    for (jsize c = 0; c < ncols; c++) {
        pins[c].pv = (*env)->GetPrimitiveArrayCritical(env, pins[c].values, NULL);
        pins[c].pn = pins[c].nulls ? (*env)->GetPrimitiveArrayCritical(env, pins[c].nulls, NULL) : NULL;
        pins[c].po = pins[c].offsets ? (*env)->GetPrimitiveArrayCritical(env, pins[c].offsets, NULL) : NULL;
        memset(&outs[c], 0, sizeof(outs[c]));
        outs[c].type = t[c];
        outs[c].mem = PG_MEM_HOST;
        outs[c].capacity = n;
        outs[c].values = pins[c].pv;
        outs[c].nulls = (uint8_t*)pins[c].pn;
        outs[c].offsets = (int32_t*)pins[c].po;
        outs[c].values_capacity_bytes = vb[c];
    }
    int64_t eg = -1;
    int32_t rc = pg_hashagg_get_output((pg_hashagg*)(intptr_t)op, first, n, outs, &eg);
    for (int c = ncols - 1; c >= 0; c--) {
        if (pins[c].po) (*env)->ReleasePrimitiveArrayCritical(env, pins[c].offsets, pins[c].po, 0);
        if (pins[c].pn) (*env)->ReleasePrimitiveArrayCritical(env, pins[c].nulls, pins[c].pn, 0);
        (*env)->ReleasePrimitiveArrayCritical(env, pins[c].values, pins[c].pv, 0);
    }
    jlong jeg = eg;
    (*env)->SetLongArrayRegion(env, errorGroup, 0, 1, &jeg);
    return rc;
}

JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggDestroy(JNIEnv* env, jclass cls, jlong op) {
    return pg_hashagg_destroy((pg_hashagg*)(intptr_t)op);
}

/* ---- packed pages: PARTIAL output -> FINAL input between processes (the job of PagesSerde in the reference's exchange) ---- */
JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggGetOutputPacked(JNIEnv* env, jclass cls, jlong op, jlong first, jlong n,
        jobject directBuffer, jlongArray outBytes, jlongArray errorGroup) {
    void* buf = (*env)->GetDirectBufferAddress(env, directBuffer);
    jlong cap = (*env)->GetDirectBufferCapacity(env, directBuffer);
    int64_t nbytes = 0, eg = -1;
    jint rc = pg_hashagg_get_output_packed((pg_hashagg*)(intptr_t)op, first, n, buf, cap, &nbytes, &eg);
    jlong jb = nbytes, jeg = eg;
    (*env)->SetLongArrayRegion(env, outBytes, 0, 1, &jb);      /* on PG_NEED_MEMORY: the size a retry needs */
    (*env)->SetLongArrayRegion(env, errorGroup, 0, 1, &jeg);
    return rc;
}

JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_hashaggAddPacked(JNIEnv* env, jclass cls, jlong op, jobjectArray directBuffers) {
    jsize nb = (*env)->GetArrayLength(env, directBuffers);
    const void* ptrs[64];
    if (nb > 64) return PG_E_INVALID_ARGUMENT;
    for (jsize i = 0; i < nb; i++) ptrs[i] = (*env)->GetDirectBufferAddress(env, (*env)->GetObjectArrayElement(env, directBuffers, i));
    return pg_hashagg_add_packed((pg_hashagg*)(intptr_t)op, ptrs, nb);
}

/* ---- GroupByHash.getGroupIds(Page) (OP/GroupByHash.java:75) ---- */
JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_groupbyGetGroupIds(JNIEnv* env, jclass cls, jlong gb, jintArray types,
        jobjectArray values, jobjectArray nulls, jobjectArray offsets, jintArray arrayOffsets, jboolean lastIsHash,
        jlong positionCount, jintArray outIds, jlongArray outGroupCount) {
    pg_col cols[MAX_COLS];
    pinned_col pins[MAX_COLS];
    int ncols = 0;
    int32_t rc = pin_page(env, types, values, nulls, offsets, arrayOffsets, positionCount, cols, pins, &ncols);
    int64_t groups = 0;
    if (rc == PG_OK) {
        jint* ids = (jint*)(*env)->GetPrimitiveArrayCritical(env, outIds, NULL);
        const pg_col* hash = lastIsHash ? &cols[ncols - 1] : NULL;
        rc = pg_groupby_get_group_ids((pg_groupby*)(intptr_t)gb, cols, hash, positionCount, (int32_t*)ids, PG_MEM_HOST, &groups);
        (*env)->ReleasePrimitiveArrayCritical(env, outIds, ids, 0);
    }
    unpin_page(env, pins, ncols);
    jlong g = groups;
    (*env)->SetLongArrayRegion(env, outGroupCount, 0, 1, &g);
    return rc;
}

/* ---- LookupSource: one probe page -> (probe position, build position) pairs (OP/LookupJoinOperator.java:299-378) ---- */
JNIEXPORT jint JNICALL Java_io_trino_gpu_PrestoGpu_lookupProbe(JNIEnv* env, jclass cls, jlong ls, jintArray types,
        jobjectArray values, jobjectArray nulls, jobjectArray offsets, jintArray arrayOffsets, jboolean lastIsHash,
        jlong positionCount, jint joinType, jintArray outProbeIdx, jlongArray outBuildPos, jlongArray nOut) {
    pg_col cols[MAX_COLS];
    pinned_col pins[MAX_COLS];
    int ncols = 0;
    int32_t rc = pin_page(env, types, values, nulls, offsets, arrayOffsets, positionCount, cols, pins, &ncols);
    int64_t produced = 0;
    if (rc == PG_OK) {
        jsize cap = (*env)->GetArrayLength(env, outProbeIdx);
        jint* pi = (jint*)(*env)->GetPrimitiveArrayCritical(env, outProbeIdx, NULL);
        jlong* bp = (jlong*)(*env)->GetPrimitiveArrayCritical(env, outBuildPos, NULL);
        int nkeys = lastIsHash ? ncols - 1 : ncols;
        rc = pg_lookup_probe((pg_lookup*)(intptr_t)ls, cols, nkeys, lastIsHash ? &cols[ncols - 1] : NULL, positionCount, joinType,
                             (int32_t*)pi, (int64_t*)bp, PG_MEM_HOST, cap, &produced);
        (*env)->ReleasePrimitiveArrayCritical(env, outBuildPos, bp, 0);
        (*env)->ReleasePrimitiveArrayCritical(env, outProbeIdx, pi, 0);
    }
    unpin_page(env, pins, ncols);
    jlong p = produced;
    (*env)->SetLongArrayRegion(env, nOut, 0, 1, &p);
    return rc;   /* PG_NEED_MEMORY: nOut[0] = rows needed; Java grows the arrays and calls again */
}
