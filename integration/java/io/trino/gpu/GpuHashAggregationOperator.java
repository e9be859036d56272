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
import io.trino.spi.block.Block;
import io.trino.spi.block.ByteArrayBlock;
import io.trino.spi.block.Int128ArrayBlock;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.block.LongArrayBlock;
import io.trino.spi.block.VariableWidthBlock;

import java.util.Optional;

import static io.airlift.slice.Slices.wrappedBuffer;
import static io.trino.gpu.PrestoGpu.check;

public class GpuHashAggregationOperator
        implements Operator
{
    private static final int OUTPUT_PAGE_ROWS = 8192;   // PageBuilder default row budget

    private final OperatorContext operatorContext;
    private final long ctx;
    private final long handle;
    private final int[] outputTypes;
    private boolean finishing;
    private long outputCursor = -1;   // -1: not producing output yet
    private long groupCount;

    public GpuHashAggregationOperator(OperatorContext operatorContext, long ctx, int[] keyChannels, int[] keyTypes, int hashChannel, int[] aggSpecs, int step, long expectedGroups, int[] outputTypes)
    {
        this.operatorContext = operatorContext;
        this.ctx = ctx;
        this.outputTypes = outputTypes;
        long[] out = new long[1];
        check(ctx, PrestoGpu.hashaggCreate(ctx, keyChannels, keyTypes, hashChannel, aggSpecs, step, expectedGroups, out));
        this.handle = out[0];
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return !finishing && outputCursor < 0;   // HashAggregationOperator.java:367-378
    }

    @Override
    public void addInput(Page page)
    {
        int channels = page.getChannelCount();
        int[] types = new int[channels];
        Object[] values = new Object[channels];
        Object[] nulls = new Object[channels];
        Object[] offsets = new Object[channels];
        int[] arrayOffsets = new int[channels];
        for (int c = 0; c < channels; c++) {
            // BlockArrays flattens Dictionary/RLE/Lazy blocks (getLoadedBlock + copyPositions) and exposes the backing arrays
            BlockArrays arrays = BlockArrays.of(page.getBlock(c));
            types[c] = arrays.typeTag;
            values[c] = arrays.values;
            nulls[c] = arrays.valueIsNull;
            offsets[c] = arrays.offsets;
            arrayOffsets[c] = arrays.arrayOffset;
        }
        check(ctx, PrestoGpu.hashaggAddPage(handle, types, values, nulls, offsets, arrayOffsets, page.getPositionCount()));
        long[] stats = new long[4];
        check(ctx, PrestoGpu.hashaggStats(handle, stats));
        operatorContext.localSystemMemoryContext().setBytes(stats[2]);   // device bytes are not JVM heap
    }

    @Override
    public void finish()
    {
        finishing = true;
    }

    @Override
    public Page getOutput()
    {
        if (!finishing) {
            return null;
        }
        if (outputCursor < 0) {
            long[] stats = new long[4];
            check(ctx, PrestoGpu.hashaggStats(handle, stats));
            groupCount = stats[0];
            outputCursor = 0;
        }
        if (outputCursor >= groupCount) {
            return null;
        }
        int n = (int) Math.min(OUTPUT_PAGE_ROWS, groupCount - outputCursor);
        OutputArrays out = OutputArrays.allocate(outputTypes, n);      // long[]/int[]/byte[] + boolean[] per channel
        long[] errorGroup = new long[1];
        check(ctx, PrestoGpu.hashaggGetOutput(handle, outputCursor, n, outputTypes, out.values, out.nulls, out.offsets, out.valueBytes, errorGroup));
        outputCursor += n;
        Block[] blocks = new Block[outputTypes.length];
        for (int c = 0; c < blocks.length; c++) {
            Optional<boolean[]> isNull = Optional.ofNullable((boolean[]) out.nulls[c]);
            switch (outputTypes[c]) {
                case PrestoGpu.INTEGER:
                case PrestoGpu.DATE:
                    blocks[c] = new IntArrayBlock(n, isNull, (int[]) out.values[c]);
                    break;
                case PrestoGpu.LONG_DECIMAL:
                    blocks[c] = new Int128ArrayBlock(n, isNull, (long[]) out.values[c]);
                    break;
                case PrestoGpu.VARCHAR:
                case PrestoGpu.VARBINARY:
                    blocks[c] = new VariableWidthBlock(n, wrappedBuffer((byte[]) out.values[c]), (int[]) out.offsets[c], isNull);
                    break;
                case PrestoGpu.BOOLEAN:
                    blocks[c] = new ByteArrayBlock(n, isNull, (byte[]) out.values[c]);
                    break;
                default:
                    blocks[c] = new LongArrayBlock(n, isNull, (long[]) out.values[c]);
            }
        }
        return new Page(n, blocks);
    }

    @Override
    public boolean isFinished()
    {
        return finishing && outputCursor >= 0 && outputCursor >= groupCount;
    }

    @Override
    public void close()
    {
        PrestoGpu.hashaggDestroy(handle);
    }
}
