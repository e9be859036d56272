/*
 * GroupedAccumulator on the GPU for count / sum / avg / min / max: what AccumulatorCompiler generates per aggregation function
 * (OP/aggregation/AccumulatorCompiler.java:261-330, 487-566), with the state arrays in device memory.
 * Not compiled in this repository (no JDK); presto-1_b200/accumulators.py is the mirror the parity tests drive.
 */
package io.trino.gpu;

import io.trino.operator.GroupByIdBlock;
import io.trino.operator.aggregation.GroupedAccumulator;
import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.block.BlockBuilder;
import io.trino.spi.type.Type;

import java.util.List;
import java.util.Optional;

import static io.trino.gpu.PrestoGpu.check;

public class GpuGroupedAccumulator
        implements GroupedAccumulator, AutoCloseable
{
    private static final int WINDOW = 8192;

    private final long ctx;
    private final long handle;
    private final int kind;
    private final Optional<Integer> inputChannel;      // empty for count(*)
    private final Optional<Integer> maskChannel;
    private final Type inputType;
    private final Type intermediateType;               // a RowType for avg (count, sum), the state's own type otherwise
    private final List<Type> intermediateColumns;      // its fields, as the library's intermediate columns
    private final Type finalType;
    private long groupCount;
    private int windowFirst = -1;
    private boolean windowFinal;
    private Block[] window;

    public GpuGroupedAccumulator(long ctx, int kind, Type inputType, Optional<Integer> inputChannel, Optional<Integer> maskChannel,
            Type intermediateType, List<Type> intermediateColumns, Type finalType)
    {
        this.ctx = ctx;
        this.kind = kind;
        this.inputType = inputType;
        this.inputChannel = inputChannel;
        this.maskChannel = maskChannel;
        this.intermediateType = intermediateType;
        this.intermediateColumns = List.copyOf(intermediateColumns);
        this.finalType = finalType;
        long[] out = new long[1];
        check(ctx, PrestoGpu.aggCreate(ctx, kind, inputChannel.isPresent() ? PageArrays.typeTag(inputType) : 0, out));
        this.handle = out[0];
    }

    @Override
    public void addInput(GroupByIdBlock groupIdsBlock, Page page)
    {
        groupCount = Math.max(groupCount, groupIdsBlock.getGroupCount());
        // page layout for the native: [argument] [mask] ids
        int[] channels = inputChannel.map(c -> new int[] {c}).orElse(new int[0]);
        List<Type> channelTypes = inputChannel.isPresent() ? List.of(inputType) : List.of();
        Block ids = toIntIds(groupIdsBlock);
        Block[] extra = maskChannel.isPresent() ? new Block[] {page.getBlock(maskChannel.get()), ids} : new Block[] {ids};
        int[] tags = maskChannel.isPresent() ? new int[] {PrestoGpu.BOOLEAN, PrestoGpu.INTEGER} : new int[] {PrestoGpu.INTEGER};
        PageArrays a = PageArrays.of(page, channels, channelTypes, extra, tags);
        check(ctx, PrestoGpu.aggAddInput(handle, groupCount, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount,
                inputChannel.isPresent(), maskChannel.isPresent()));
        windowFirst = -1;
    }

    @Override
    public void addIntermediate(GroupByIdBlock groupIdsBlock, Block block)
    {
        groupCount = Math.max(groupCount, groupIdsBlock.getGroupCount());
        // a RowBlock's fields are the intermediate columns (avg: count, sum); a plain block is the single column
        Block[] fields = intermediateColumns.size() == 1 ? new Block[] {block} : io.trino.spi.block.ColumnarRow.toColumnarRow(block).getFields();
        Block[] all = new Block[fields.length + 1];
        int[] tags = new int[fields.length + 1];
        for (int i = 0; i < fields.length; i++) {
            all[i] = fields[i];
            tags[i] = PageArrays.typeTag(intermediateColumns.get(i));
        }
        all[fields.length] = toIntIds(groupIdsBlock);
        tags[fields.length] = PrestoGpu.INTEGER;
        PageArrays a = PageArrays.of(new Page(block.getPositionCount()), new int[0], List.of(), all, tags);
        check(ctx, PrestoGpu.aggAddIntermediate(handle, groupCount, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, block.getPositionCount()));
        windowFirst = -1;
    }

    private static Block toIntIds(GroupByIdBlock ids)
    {
        int[] v = new int[ids.getPositionCount()];
        for (int i = 0; i < v.length; i++) {
            v[i] = (int) ids.getGroupId(i);
        }
        return new io.trino.spi.block.IntArrayBlock(v.length, Optional.empty(), v);
    }

    private void loadWindow(int groupId, boolean finalStep)
    {
        int first = groupId - groupId % WINDOW;
        int n = (int) Math.min(WINDOW, groupCount - first);
        int[] tags = finalStep ? new int[] {PageArrays.typeTag(finalType)} : PageArrays.typeTags(intermediateColumns);
        OutputArrays out = OutputArrays.allocate(tags, n);
        long[] errorGroup = new long[] {-1};
        check(ctx, PrestoGpu.aggEvaluate(handle, first, n, finalStep, tags, out.values, out.nulls, out.offsets, out.valueBytes, errorGroup));
        window = PageArrays.toBlocks(tags, out, n);
        windowFirst = first;
        windowFinal = finalStep;
    }

    @Override
    public void evaluateIntermediate(int groupId, BlockBuilder output)
    {
        if (windowFirst < 0 || windowFinal || groupId < windowFirst || groupId >= windowFirst + window[0].getPositionCount()) {
            loadWindow(groupId, false);
        }
        int at = groupId - windowFirst;
        if (intermediateColumns.size() == 1) {
            intermediateColumns.get(0).appendTo(window[0], at, output);
            return;
        }
        BlockBuilder row = output.beginBlockEntry();          // avg: ROW(count, sum)
        for (int i = 0; i < window.length; i++) {
            intermediateColumns.get(i).appendTo(window[i], at, row);
        }
        output.closeEntry();
    }

    @Override
    public void evaluateFinal(int groupId, BlockBuilder output)
    {
        if (windowFirst < 0 || !windowFinal || groupId < windowFirst || groupId >= windowFirst + window[0].getPositionCount()) {
            loadWindow(groupId, true);       // NUMERIC_VALUE_OUT_OF_RANGE surfaces here, as in the generated evaluateFinal
        }
        finalType.appendTo(window[0], groupId - windowFirst, output);
    }

    @Override
    public void prepareFinal() {}

    @Override
    public long getEstimatedSize()
    {
        long[] b = new long[1];
        check(ctx, PrestoGpu.aggDeviceBytes(handle, b));
        return b[0];
    }

    @Override
    public Type getFinalType()
    {
        return finalType;
    }

    @Override
    public Type getIntermediateType()
    {
        return intermediateType;
    }

    @Override
    public void close()
    {
        PrestoGpu.aggDestroy(handle);
    }
}
