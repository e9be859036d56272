/*
 * HashBuilderOperator on the GPU (OP/HashBuilderOperator.java:56-635, states CONSUMING_INPUT -> LOOKUP_SOURCE_BUILT -> CLOSED; the
 * spill states do not exist: the build side lives in device memory).  addInput copies the page's channels into device-resident
 * build columns (PagesIndex.addPage :234-254), finish() builds the table and publishes the GpuLookupSource through the factory's
 * future, which GpuLookupJoinOperator waits on.  Not compiled in this repository (no JDK); presto-1_b200/operators.py
 * (HashBuilderOperator) is the mirror the parity tests drive.
 */
package io.trino.gpu;

import com.google.common.util.concurrent.ListenableFuture;
import com.google.common.util.concurrent.SettableFuture;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.spi.Page;
import io.trino.spi.type.Type;

import java.util.List;

import static io.trino.gpu.PrestoGpu.check;

public class GpuHashBuilderOperator
        implements Operator
{
    private final OperatorContext operatorContext;
    private final long ctx;
    private final long builder;
    private final List<Type> buildTypes;
    private final int[] allChannels;
    private final SettableFuture<GpuLookupSource> lookupSourceFuture;
    private boolean finishing;
    private GpuLookupSource lookupSource;
    private long rows;

    public GpuHashBuilderOperator(OperatorContext operatorContext, long ctx, List<Type> buildTypes, int[] joinChannels, int hashChannel, long expectedRows,
            SettableFuture<GpuLookupSource> lookupSourceFuture)
    {
        this.operatorContext = operatorContext;
        this.ctx = ctx;
        this.buildTypes = List.copyOf(buildTypes);
        this.allChannels = new int[buildTypes.size()];
        for (int c = 0; c < allChannels.length; c++) {
            allChannels[c] = c;
        }
        this.lookupSourceFuture = lookupSourceFuture;
        long[] out = new long[1];
        check(ctx, PrestoGpu.joinbuildCreate(ctx, PageArrays.typeTags(buildTypes), joinChannels, hashChannel, expectedRows, out));
        this.builder = out[0];
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return !finishing;       // :309-315 (no spilling states)
    }

    @Override
    public void addInput(Page page)
    {
        PageArrays a = PageArrays.of(page, allChannels, buildTypes);
        check(ctx, PrestoGpu.joinbuildAddPage(builder, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount));   // "Size of pages index cannot exceed 2 billion entries" comes back as GENERIC_INSUFFICIENT_RESOURCES
        rows += page.getPositionCount();
        operatorContext.recordOutput(page.getSizeInBytes(), page.getPositionCount());
    }

    @Override
    public Page getOutput()
    {
        return null;
    }

    @Override
    public void finish()
    {
        if (finishing) {
            return;
        }
        finishing = true;
        long[] out = new long[1];
        check(ctx, PrestoGpu.joinbuildFinish(builder, out));       // buildLookupSource :585-602
        lookupSource = new GpuLookupSource(ctx, out[0], buildTypes);
        operatorContext.localSystemMemoryContext().setBytes(lookupSource.getInMemorySizeInBytes());
        lookupSourceFuture.set(lookupSource);
    }

    @Override
    public ListenableFuture<?> isBlocked()
    {
        return NOT_BLOCKED;
    }

    @Override
    public boolean isFinished()
    {
        return lookupSource != null;
    }

    @Override
    public void close()
    {
        PrestoGpu.joinbuildDestroy(builder);
        // the lookup source is closed by the last probe operator (GpuLookupJoinOperatorFactory reference counts it)
    }
}
