/* prestogpu_jni.c — JNI shim between io.trino.gpu.PrestoGpu (integration/java) and libprestogpu.so (include/prestogpu.h).
 * One wrapper per pg_* entry point (tests/test_boundary_cpu.py checks the list against the header).
 *
 * The image has no JDK: tests/test_boundary_cpu.py compiles this file against tests/jni_stub/jni.h (the JNI types and the
 * function table entries used here) so that it cannot rot; a maintainer builds it next to the reference with
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude \
 *       integration/jni/prestogpu_jni.c integration/jni/prestogpu_jni_ops.c -Lpresto-1_b200 -lprestogpu -o libprestogpu_jni.so
 *
 * Conventions
 *  - handles travel as jlong; every native returns the pg status, Java turns non-zero into TrinoException
 *    (PrestoGpu.check): PG_E_INSUFFICIENT_RESOURCES -> GENERIC_INSUFFICIENT_RESOURCES,
 *    PG_E_NUMERIC_OVERFLOW -> NUMERIC_VALUE_OUT_OF_RANGE, PG_E_INVALID_ARGUMENT -> IllegalArgumentException, ...
 *  - a Block crosses as its backing arrays: long[]/int[]/byte[] values, boolean[] valueIsNull (1 byte per element in
 *    HotSpot), int[] offsets, plus arrayOffset and positionCount (LongArrayBlock.java:37-76, VariableWidthBlock.java:38-48).
 *  - NO array is pinned while the library runs.  Inputs are copied into a thread-local native staging buffer with
 *    Get<Type>ArrayRegion (a page is <= 1 MiB) and outputs are written by the library into staging and copied back with
 *    Set<Type>ArrayRegion: GetPrimitiveArrayCritical would hold the garbage collector off for the duration of GPU
 *    synchronisations inside the call.
 */
#include "prestogpu_jni.h"

/* ---------------------------------------------------------------- staging */
static __thread pgj_arena tl_in, tl_out;

static int arena_reserve(pgj_arena* a, size_t bytes) {
    a->used = 0;
    if (bytes <= a->cap) return 1;
    size_t cap = a->cap ? a->cap : (1u << 20);
    while (cap < bytes) cap *= 2;
    char* p = (char*)realloc(a->p, cap);
    if (!p) return 0;
    a->p = p; a->cap = cap;
    return 1;
}
static void* arena_take(pgj_arena* a, size_t bytes) {
    size_t at = (a->used + 63) & ~(size_t)63;
    a->used = at + bytes;
    return a->p + at;
}
static size_t padded(size_t bytes) { return (bytes + 63 + 64) & ~(size_t)63; }

int pgj_type_width(int t) {
    switch (t) {
        case PG_INTEGER: case PG_DATE: return 4;
        case PG_LONG_DECIMAL: return 16;
        case PG_BOOLEAN: return 1;
        case PG_VARCHAR: case PG_VARBINARY: return 0;
        default: return 8;
    }
}
static int is_var(int t) { return t == PG_VARCHAR || t == PG_VARBINARY; }

/* copies `count` elements of `width` bytes starting at element `start` of a primitive array of the matching Java type */
static void copy_in(JNIEnv* env, jarray arr, int width, jsize start, jsize count, void* dst) {
    if (count <= 0) return;
    switch (width) {
        case 1: (*env)->GetByteArrayRegion(env, (jbyteArray)arr, start, count, (jbyte*)dst); break;
        case 4: (*env)->GetIntArrayRegion(env, (jintArray)arr, start, count, (jint*)dst); break;
        case 16: (*env)->GetLongArrayRegion(env, (jlongArray)arr, 2 * start, 2 * count, (jlong*)dst); break;   /* Int128ArrayBlock: long[2n] */
        default: (*env)->GetLongArrayRegion(env, (jlongArray)arr, start, count, (jlong*)dst);
    }
}
static void copy_out(JNIEnv* env, jarray arr, int width, jsize count, const void* src) {
    if (count <= 0) return;
    switch (width) {
        case 1: (*env)->SetByteArrayRegion(env, (jbyteArray)arr, 0, count, (const jbyte*)src); break;
        case 4: (*env)->SetIntArrayRegion(env, (jintArray)arr, 0, count, (const jint*)src); break;
        case 16: (*env)->SetLongArrayRegion(env, (jlongArray)arr, 0, 2 * count, (const jlong*)src); break;
        default: (*env)->SetLongArrayRegion(env, (jlongArray)arr, 0, count, (const jlong*)src);
    }
}

/* A page: parallel arrays types[], values[] (Object[] of primitive arrays), nulls[] (Object[] of boolean[] or null),
 * offsets[] (Object[] of int[] or null), arrayOffsets[].  Everything is copied into tl_in; cols[] point there. */
int pgj_page_in(JNIEnv* env, jintArray jtypes, jobjectArray jvalues, jobjectArray jnulls, jobjectArray joffsets, jintArray jarray_offsets,
                jlong n, pg_col* cols, int* ncols_out) {
    const jsize ncols = (*env)->GetArrayLength(env, jtypes);
    *ncols_out = 0;
    if (ncols > PGJ_MAX_COLS || n < 0 || n > 0x7fffffff) return PG_E_INVALID_ARGUMENT;
    jint types[PGJ_MAX_COLS], aoff[PGJ_MAX_COLS];
    jarray av[PGJ_MAX_COLS], an[PGJ_MAX_COLS], ao[PGJ_MAX_COLS];
    jint first_byte[PGJ_MAX_COLS], end_byte[PGJ_MAX_COLS];
    (*env)->GetIntArrayRegion(env, jtypes, 0, ncols, types);
    (*env)->GetIntArrayRegion(env, jarray_offsets, 0, ncols, aoff);
    size_t total = 64;
    for (jsize c = 0; c < ncols; c++) {
        av[c] = (jarray)(*env)->GetObjectArrayElement(env, jvalues, c);
        an[c] = jnulls ? (jarray)(*env)->GetObjectArrayElement(env, jnulls, c) : NULL;
        ao[c] = joffsets ? (jarray)(*env)->GetObjectArrayElement(env, joffsets, c) : NULL;
        if (!av[c] && n > 0) return PG_E_INVALID_ARGUMENT;
        first_byte[c] = end_byte[c] = 0;
        if (is_var(types[c])) {
            if (!ao[c]) return PG_E_INVALID_ARGUMENT;
            (*env)->GetIntArrayRegion(env, (jintArray)ao[c], aoff[c], 1, &first_byte[c]);
            (*env)->GetIntArrayRegion(env, (jintArray)ao[c], aoff[c] + (jsize)n, 1, &end_byte[c]);
            total += padded((size_t)(end_byte[c] - first_byte[c])) + padded(((size_t)n + 1) * 4);
        } else total += padded((size_t)n * (size_t)pgj_type_width(types[c]));
        if (an[c]) total += padded((size_t)n);
    }
    if (!arena_reserve(&tl_in, total)) return PG_E_INSUFFICIENT_RESOURCES;
    for (jsize c = 0; c < ncols; c++) {
        pg_col* col = &cols[c];
        memset(col, 0, sizeof(*col));
        col->type = types[c]; col->mem = PG_MEM_HOST; col->n = n;
        if (is_var(types[c])) {
            const jsize nbytes = end_byte[c] - first_byte[c];
            void* bytes = arena_take(&tl_in, (size_t)nbytes);
            int32_t* offs = (int32_t*)arena_take(&tl_in, ((size_t)n + 1) * 4);
            if (av[c]) copy_in(env, av[c], 1, first_byte[c], nbytes, bytes);       /* only the Slice range this page refers to */
            (*env)->GetIntArrayRegion(env, (jintArray)ao[c], aoff[c], (jsize)n + 1, (jint*)offs);
            for (jlong i = 0; i <= n; i++) offs[i] -= first_byte[c];
            col->values = bytes; col->offsets = offs;
        } else {
            const int w = pgj_type_width(types[c]);
            void* v = arena_take(&tl_in, (size_t)n * (size_t)w);
            if (av[c]) copy_in(env, av[c], w, aoff[c], (jsize)n, v);
            col->values = v;
        }
        if (an[c]) {
            void* nl = arena_take(&tl_in, (size_t)n);
            (*env)->GetBooleanArrayRegion(env, (jbooleanArray)an[c], aoff[c], (jsize)n, (jboolean*)nl);
            col->nulls = (const uint8_t*)nl;
        }
    }
    *ncols_out = (int)ncols;
    return (*env)->ExceptionCheck(env) ? PG_E_INVALID_ARGUMENT : PG_OK;     /* ArrayIndexOutOfBounds from a Region call */
}

/* Output columns: Java allocates one array set per channel (sized from the *Stats / *VarwidthBytes calls); the library writes
 * into tl_out and pgj_out_end copies the first `rows` rows back. */
int pgj_out_begin(JNIEnv* env, jintArray jtypes, jobjectArray jvalues, jobjectArray jnulls, jobjectArray joffsets, jlongArray jvalue_bytes,
                  jlong capacity, size_t extra, pgj_outs* o) {
    memset(o, 0, sizeof(*o));
    const jsize ncols = jtypes ? (*env)->GetArrayLength(env, jtypes) : 0;
    if (ncols > PGJ_MAX_COLS || capacity < 0) return PG_E_INVALID_ARGUMENT;
    o->ncols = (int)ncols; o->capacity = capacity;
    jlong vb[PGJ_MAX_COLS];
    memset(vb, 0, sizeof(vb));
    (*env)->GetIntArrayRegion(env, jtypes, 0, ncols, o->types);
    if (jvalue_bytes) (*env)->GetLongArrayRegion(env, jvalue_bytes, 0, ncols, vb);
    size_t total = 64 + padded(extra);
    for (jsize c = 0; c < ncols; c++) {
        o->values[c] = (jarray)(*env)->GetObjectArrayElement(env, jvalues, c);
        o->nulls[c] = jnulls ? (jarray)(*env)->GetObjectArrayElement(env, jnulls, c) : NULL;
        o->offsets[c] = joffsets ? (jarray)(*env)->GetObjectArrayElement(env, joffsets, c) : NULL;
        if (!o->values[c] || (is_var(o->types[c]) && !o->offsets[c])) return PG_E_INVALID_ARGUMENT;
        total += is_var(o->types[c]) ? padded((size_t)vb[c]) + padded(((size_t)capacity + 1) * 4) : padded((size_t)capacity * (size_t)pgj_type_width(o->types[c]));
        if (o->nulls[c]) total += padded((size_t)capacity);
    }
    if (!arena_reserve(&tl_out, total)) return PG_E_INSUFFICIENT_RESOURCES;
    for (jsize c = 0; c < ncols; c++) {
        pg_col_out* out = &o->cols[c];
        memset(out, 0, sizeof(*out));
        out->type = o->types[c]; out->mem = PG_MEM_HOST; out->capacity = capacity;
        if (is_var(o->types[c])) {
            out->values = arena_take(&tl_out, (size_t)vb[c]);
            out->offsets = (int32_t*)arena_take(&tl_out, ((size_t)capacity + 1) * 4);
            out->values_capacity_bytes = vb[c];
        } else out->values = arena_take(&tl_out, (size_t)capacity * (size_t)pgj_type_width(o->types[c]));
        if (o->nulls[c]) out->nulls = (uint8_t*)arena_take(&tl_out, (size_t)capacity);
    }
    return PG_OK;
}

void pgj_out_end(JNIEnv* env, pgj_outs* o, jlong rows) {
    if (rows > o->capacity) rows = o->capacity;
    for (int c = 0; c < o->ncols; c++) {
        const pg_col_out* out = &o->cols[c];
        if (is_var(o->types[c])) {
            (*env)->SetIntArrayRegion(env, (jintArray)o->offsets[c], 0, (jsize)rows + 1, (const jint*)out->offsets);
            copy_out(env, o->values[c], 1, rows > 0 ? (jsize)out->offsets[rows] : 0, out->values);
        } else copy_out(env, o->values[c], pgj_type_width(o->types[c]), (jsize)rows, out->values);
        if (o->nulls[c]) (*env)->SetBooleanArrayRegion(env, (jbooleanArray)o->nulls[c], 0, (jsize)rows, (const jboolean*)out->nulls);
    }
}

/* scratch in tl_out for plain result vectors (group ids, probe positions): after pgj_out_begin(.., extra, ..) reserved `extra`
 * bytes for them, or after pgj_scratch_reset(total) when the call has no output columns */
void* pgj_scratch(size_t bytes) {
    const size_t at = (tl_out.used + 63) & ~(size_t)63;
    if (at + bytes > tl_out.cap) return NULL;
    tl_out.used = at + bytes;
    return tl_out.p + at;
}
int pgj_scratch_reset(size_t reserve) { return arena_reserve(&tl_out, padded(reserve) + 256); }

void pgj_set_long(JNIEnv* env, jlongArray arr, int i, int64_t v) {
    if (!arr) return;
    jlong j = (jlong)v;
    (*env)->SetLongArrayRegion(env, arr, i, 1, &j);
}
void pgj_set_int(JNIEnv* env, jintArray arr, int i, int32_t v) {
    if (!arr) return;
    jint j = (jint)v;
    (*env)->SetIntArrayRegion(env, arr, i, 1, &j);
}
