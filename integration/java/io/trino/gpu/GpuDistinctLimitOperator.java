/*
 * DistinctLimitOperator on the GPU (OP/DistinctLimitOperator.java:115-220): the first row of every new distinct combination of the
 * distinct channels (plus the precomputed hash channel) passes through until `limit` rows have been produced.  The first-row test is
 * the one MarkDistinct uses, one native call per page.  Not compiled in this repository (no JDK); presto-1_b200/operators.py
 * (DistinctLimitOperator) is the mirror the parity tests drive (tests/test_f4_gpu.py replays TestDistinctLimitOperator.java:92-156).
 */
package io.trino.gpu;

import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.UpdateMemory;
import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.type.Type;

import java.util.List;
import java.util.Optional;

public class GpuDistinctLimitOperator
        implements Operator
{
    private final OperatorContext operatorContext;
    private final GpuGroupByHash hash;
    private final int[] outputChannels;
    private long remainingLimit;
    private Page inputPage;
    private boolean[] marks;
    private boolean finishing;

    public GpuDistinctLimitOperator(OperatorContext operatorContext, long ctx, List<Integer> distinctChannels, List<Type> distinctTypes, long limit, Optional<Integer> hashChannel)
    {
        if (limit < 0) {
            throw new IllegalArgumentException("limit must be at least zero");
        }
        this.operatorContext = operatorContext;
        int[] channels = distinctChannels.stream().mapToInt(Integer::intValue).toArray();
        this.outputChannels = new int[channels.length + (hashChannel.isPresent() ? 1 : 0)];
        System.arraycopy(channels, 0, outputChannels, 0, channels.length);
        if (hashChannel.isPresent()) {
            outputChannels[channels.length] = hashChannel.get();
        }
        this.hash = new GpuGroupByHash(ctx, distinctTypes, channels, hashChannel, (int) Math.min(limit, 10_000), UpdateMemory.NOOP);
        this.remainingLimit = limit;
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return !finishing && remainingLimit > 0 && inputPage == null;
    }

    @Override
    public void addInput(Page page)
    {
        inputPage = page;
        marks = hash.markDistinct(page);
        operatorContext.localUserMemoryContext().setBytes(hash.getEstimatedSize());
    }

    @Override
    public Page getOutput()
    {
        if (inputPage == null) {
            return null;
        }
        int[] positions = new int[marks.length];
        int count = 0;
        for (int i = 0; i < marks.length && remainingLimit > 0; i++) {
            if (marks[i]) {
                positions[count++] = i;
                remainingLimit--;
            }
        }
        Page page = inputPage;
        inputPage = null;
        marks = null;
        if (count == 0) {
            return null;
        }
        Block[] blocks = new Block[outputChannels.length];
        for (int c = 0; c < blocks.length; c++) {
            blocks[c] = page.getBlock(outputChannels[c]).getPositions(positions, 0, count);
        }
        return new Page(count, blocks);
    }

    @Override
    public void finish()
    {
        finishing = true;
    }

    @Override
    public boolean isFinished()
    {
        return inputPage == null && (finishing || remainingLimit == 0);
    }

    @Override
    public void close()
    {
        hash.close();
    }
}
