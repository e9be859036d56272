/* prestogpu_jni.h — helpers shared by the two halves of the JNI shim (prestogpu_jni.c: staging, prestogpu_jni_ops.c: the natives). */
#ifndef PRESTOGPU_JNI_H
#define PRESTOGPU_JNI_H
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "prestogpu.h"

#define PGJ_MAX_COLS 64

typedef struct { char* p; size_t cap, used; } pgj_arena;
typedef struct {
    int ncols;
    jlong capacity;
    jint types[PGJ_MAX_COLS];
    jarray values[PGJ_MAX_COLS], nulls[PGJ_MAX_COLS], offsets[PGJ_MAX_COLS];
    pg_col_out cols[PGJ_MAX_COLS];
} pgj_outs;

int pgj_type_width(int type);
int pgj_page_in(JNIEnv* env, jintArray types, jobjectArray values, jobjectArray nulls, jobjectArray offsets, jintArray array_offsets, jlong n,
                pg_col* cols, int* ncols_out);
int pgj_out_begin(JNIEnv* env, jintArray types, jobjectArray values, jobjectArray nulls, jobjectArray offsets, jlongArray value_bytes,
                  jlong capacity, size_t extra, pgj_outs* o);
void pgj_out_end(JNIEnv* env, pgj_outs* o, jlong rows);
void* pgj_scratch(size_t bytes);
int pgj_scratch_reset(size_t reserve);
void pgj_set_long(JNIEnv* env, jlongArray arr, int i, int64_t v);
void pgj_set_int(JNIEnv* env, jintArray arr, int i, int32_t v);

#define H(type, handle) ((type*)(intptr_t)(handle))
#define NATIVE(ret, name) JNIEXPORT ret JNICALL Java_io_trino_gpu_PrestoGpu_##name
#endif
