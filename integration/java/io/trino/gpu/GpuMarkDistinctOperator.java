/*
 * MarkDistinctOperator on the GPU (OP/MarkDistinctOperator.java:104-175): every input page leaves with one more BOOLEAN channel that is
 * true for the first row of each distinct combination of the mark channels seen so far.  The marks come from one native call per page
 * (pg_groupby_mark_distinct: the page is added to the device GroupByHash, then the first row of every group it introduced is marked).
 * Not compiled in this repository (no JDK); presto-1_b200/operators.py (MarkDistinctOperator) is the mirror the parity tests drive
 * (tests/test_f4_gpu.py replays TestMarkDistinctOperator.java:92-110).
 */
package io.trino.gpu;

import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.UpdateMemory;
import io.trino.spi.Page;
import io.trino.spi.block.ByteArrayBlock;
import io.trino.spi.type.Type;

import java.util.ArrayList;
import java.util.List;
import java.util.Optional;

public class GpuMarkDistinctOperator
        implements Operator
{
    private final OperatorContext operatorContext;
    private final GpuGroupByHash hash;
    private Page output;
    private boolean finishing;

    public GpuMarkDistinctOperator(OperatorContext operatorContext, long ctx, List<Type> types, List<Integer> markDistinctChannels, Optional<Integer> hashChannel)
    {
        this.operatorContext = operatorContext;
        List<Type> distinctTypes = new ArrayList<>();
        int[] channels = new int[markDistinctChannels.size()];
        for (int i = 0; i < channels.length; i++) {
            channels[i] = markDistinctChannels.get(i);
            distinctTypes.add(types.get(channels[i]));
        }
        this.hash = new GpuGroupByHash(ctx, distinctTypes, channels, hashChannel, 10_000, UpdateMemory.NOOP);   // expectedSize as MarkDistinctHash.java:40-44
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return !finishing && output == null;
    }

    @Override
    public void addInput(Page page)
    {
        boolean[] marks = hash.markDistinct(page);
        byte[] values = new byte[marks.length];
        for (int i = 0; i < marks.length; i++) {
            values[i] = (byte) (marks[i] ? 1 : 0);
        }
        output = page.appendColumn(new ByteArrayBlock(marks.length, Optional.empty(), values));
        operatorContext.localUserMemoryContext().setBytes(hash.getEstimatedSize());
    }

    @Override
    public Page getOutput()
    {
        Page page = output;
        output = null;
        return page;
    }

    @Override
    public void finish()
    {
        finishing = true;
    }

    @Override
    public boolean isFinished()
    {
        return finishing && output == null;
    }

    @Override
    public void close()
    {
        hash.close();
    }
}
