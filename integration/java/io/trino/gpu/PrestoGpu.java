/*
 * Native entry points of libprestogpu_jni.so (integration/jni/prestogpu_jni.c), which forwards to libprestogpu.so
 * (include/prestogpu.h).  Not compiled in this repository (no JDK in the build image); shown so that a maintainer can
 * see exactly what the reference side binds.
 */
package io.trino.gpu;

import io.trino.spi.TrinoException;

import static io.trino.spi.StandardErrorCode.GENERIC_INSUFFICIENT_RESOURCES;
import static io.trino.spi.StandardErrorCode.GENERIC_INTERNAL_ERROR;
import static io.trino.spi.StandardErrorCode.NUMERIC_VALUE_OUT_OF_RANGE;

public final class PrestoGpu
{
    // include/prestogpu.h status codes
    public static final int PG_OK = 0;
    public static final int PG_NEED_MEMORY = 1;
    public static final int PG_E_INSUFFICIENT_RESOURCES = 2;
    public static final int PG_E_NUMERIC_OVERFLOW = 3;
    public static final int PG_E_INVALID_ARGUMENT = 4;
    public static final int PG_E_ILLEGAL_STATE = 8;

    // pg type tags
    public static final int BIGINT = 1;
    public static final int INTEGER = 2;
    public static final int DATE = 3;
    public static final int DOUBLE = 4;
    public static final int SHORT_DECIMAL = 5;
    public static final int LONG_DECIMAL = 6;
    public static final int VARCHAR = 7;
    public static final int VARBINARY = 8;
    public static final int BOOLEAN = 9;

    static {
        System.loadLibrary("prestogpu_jni");
    }

    private PrestoGpu() {}

    public static native int ctxCreate(int device, long[] outHandle);

    public static native int ctxDestroy(long ctx);

    public static native String lastError(long ctx);

    public static native int hashaggCreate(long ctx, int[] keyChannels, int[] keyTypes, int hashChannel, int[] aggSpecs, int step, long expectedGroups, long[] outHandle);

    public static native int hashaggAddPage(long op, int[] types, Object[] values, Object[] nulls, Object[] offsets, int[] arrayOffsets, long positionCount);

    public static native int hashaggStats(long op, long[] out4);

    public static native int hashaggGetOutput(long op, long first, long n, int[] types, Object[] values, Object[] nulls, Object[] offsets, long[] valueBytes, long[] errorGroup);

    public static native int hashaggDestroy(long op);

    /** Groups [first, first+n) as one self-describing packed page in a direct ByteBuffer (PARTIAL -> FINAL wire format, prestogpu.h). */
    public static native int hashaggGetOutputPacked(long op, long first, long n, java.nio.ByteBuffer directBuffer, long[] outBytes, long[] errorGroup);

    /** Packed pages (direct ByteBuffers, one per producer) as one input page of a FINAL / INTERMEDIATE operator. */
    public static native int hashaggAddPacked(long op, java.nio.ByteBuffer[] directBuffers);

    public static native int groupbyGetGroupIds(long gb, int[] types, Object[] values, Object[] nulls, Object[] offsets, int[] arrayOffsets, boolean lastIsHash, long positionCount, int[] outIds, long[] outGroupCount);

    public static native int lookupProbe(long lookup, int[] types, Object[] values, Object[] nulls, Object[] offsets, int[] arrayOffsets, boolean lastIsHash, long positionCount, int joinType, int[] outProbeIdx, long[] outBuildPos, long[] nOut);

    /** Status → the exception the reference's CPU operator would have thrown (SURVEY.md §8b, error conventions). */
    public static void check(long ctx, int status)
    {
        switch (status) {
            case PG_OK:
                return;
            case PG_E_INSUFFICIENT_RESOURCES:
                throw new TrinoException(GENERIC_INSUFFICIENT_RESOURCES, lastError(ctx));
            case PG_E_NUMERIC_OVERFLOW:
                throw new TrinoException(NUMERIC_VALUE_OUT_OF_RANGE, lastError(ctx));
            case PG_E_INVALID_ARGUMENT:
                throw new IllegalArgumentException(lastError(ctx));
            case PG_E_ILLEGAL_STATE:
                throw new IllegalStateException(lastError(ctx));
            default:
                throw new TrinoException(GENERIC_INTERNAL_ERROR, "prestogpu status " + status + ": " + lastError(ctx));
        }
    }
}
