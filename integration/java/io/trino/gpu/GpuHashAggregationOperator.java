/*
 * Operator-level seam: a HashAggregationOperator whose addInput/getOutput run on the GPU.  It is created by a
 * GpuHashAggregationOperatorFactory with the constructor arguments of HashAggregationOperatorFactory
 * (OP/HashAggregationOperator.java:54-262) and plugged in at LocalExecutionPlanner.java:3035-3054.
 * Not compiled here (no JDK); the Python mirror presto-1_b200/operators.py implements the same state machine and is what
 * the parity tests drive.
 */
package io.trino.gpu;

import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.spi.Page;
import io.trino.spi.type.Type;

import java.util.List;

import static io.trino.gpu.PrestoGpu.check;

public class GpuHashAggregationOperator
        implements Operator
{
    private static final int OUTPUT_PAGE_ROWS = 8192;   // PageBuilder default row budget

    private final OperatorContext operatorContext;
    private final long ctx;
    private final long handle;
    private final List<Type> inputTypes;
    private final int[] allChannels;
    private final int[] outputTypes;
    private final boolean partial;
    private final long maxPartialMemory;
    private PageArrays unfinished;    // unfinishedWork (HashAggregationOperator.java:398-417): a page the memory limit has not let in yet
    private long grantedBytes;
    private boolean finishing;
    private boolean flushing;         // partial aggregation over its memory budget: emit what is there, then reset (:501-517)
    private long outputCursor = -1;   // -1: not producing output yet
    private long groupCount;

    public GpuHashAggregationOperator(OperatorContext operatorContext, long ctx, List<Type> inputTypes, int[] keyChannels, int[] keyTypes, int hashChannel,
            int[] aggSpecs, int step, long expectedGroups, long maxPartialMemory, int[] outputTypes)
    {
        this.operatorContext = operatorContext;
        this.ctx = ctx;
        this.inputTypes = List.copyOf(inputTypes);
        this.allChannels = new int[inputTypes.size()];
        for (int c = 0; c < allChannels.length; c++) {
            allChannels[c] = c;
        }
        this.outputTypes = outputTypes;
        this.partial = step == PrestoGpu.STEP_PARTIAL;
        this.maxPartialMemory = maxPartialMemory;
        long[] out = new long[1];
        check(ctx, PrestoGpu.hashaggCreate(ctx, keyChannels, keyTypes, hashChannel, aggSpecs, step, expectedGroups, out));
        this.handle = out[0];
        this.grantedBytes = stats()[2];
        check(ctx, PrestoGpu.hashaggSetMemoryLimit(handle, grantedBytes));
    }

    private long[] stats()
    {
        long[] s = new long[4];
        check(ctx, PrestoGpu.hashaggStats(handle, s));
        return s;
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return !finishing && !flushing && unfinished == null && outputCursor < 0;   // HashAggregationOperator.java:367-378
    }

    /** unfinishedWork.process(): true when no page is waiting any more.  Memory is asked for where the CPU builder's
     *  updateMemoryWithYieldInfo would (InMemoryHashAggregationBuilder.java:189-214): a refused reservation leaves the page waiting. */
    private boolean processUnfinished()
    {
        while (unfinished != null) {
            PageArrays a = unfinished;
            int status = PrestoGpu.hashaggAddPage(handle, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount);
            if (status == PrestoGpu.PG_NEED_MEMORY) {
                long wanted = grantedBytes * 2;
                if (!operatorContext.localUserMemoryContext().trySetBytes(wanted)) {
                    return false;        // the driver yields; getOutput() / the next addInput attempt tries again
                }
                grantedBytes = wanted;
                check(ctx, PrestoGpu.hashaggSetMemoryLimit(handle, grantedBytes));
                continue;
            }
            check(ctx, status);
            unfinished = null;
            long[] s = stats();
            operatorContext.localSystemMemoryContext().setBytes(s[2]);   // device bytes are not JVM heap
            if (partial && s[2] > maxPartialMemory) {
                flushing = true;
            }
        }
        return true;
    }

    @Override
    public void addInput(Page page)
    {
        unfinished = PageArrays.of(page, allChannels, inputTypes);
        processUnfinished();
    }

    @Override
    public void finish()
    {
        finishing = true;
    }

    @Override
    public Page getOutput()
    {
        if (!processUnfinished()) {
            return null;                  // :473-481
        }
        if (!finishing && !flushing) {
            return null;
        }
        if (outputCursor < 0) {
            groupCount = stats()[0];
            outputCursor = 0;
        }
        if (outputCursor >= groupCount) {
            if (flushing && !finishing) {   // partial flush done: forget the groups and take input again
                check(ctx, PrestoGpu.hashaggReset(handle));
                flushing = false;
                outputCursor = -1;
            }
            return null;
        }
        int n = (int) Math.min(OUTPUT_PAGE_ROWS, groupCount - outputCursor);
        long[] varBytes = new long[outputTypes.length];
        for (int c = 0; c < outputTypes.length; c++) {
            if (outputTypes[c] == PrestoGpu.VARCHAR || outputTypes[c] == PrestoGpu.VARBINARY) {
                long[] b = new long[1];
                check(ctx, PrestoGpu.hashaggVarwidthBytes(handle, c, b));
                varBytes[c] = b[0];
            }
        }
        OutputArrays out = OutputArrays.allocate(outputTypes, n, varBytes);      // long[]/int[]/byte[] + boolean[] per channel
        long[] errorGroup = new long[1];
        check(ctx, PrestoGpu.hashaggGetOutput(handle, outputCursor, n, outputTypes, out.values, out.nulls, out.offsets, out.valueBytes, errorGroup));
        outputCursor += n;
        return new Page(n, PageArrays.toBlocks(outputTypes, out, n));
    }

    @Override
    public boolean isFinished()
    {
        return finishing && unfinished == null && outputCursor >= 0 && outputCursor >= groupCount;
    }

    @Override
    public void close()
    {
        PrestoGpu.hashaggDestroy(handle);
    }
}
