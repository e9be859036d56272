/*
 * SetBuilderOperator on the GPU (OP/SetBuilderOperator.java:137-210): collects the build side's set channel into a device-resident
 * GroupByHash and publishes it as a GpuChannelSet (ChannelSet.java:101-109: the hash plus "does the set hold a NULL", answered by
 * one contains() of a NULL row).  Not compiled in this repository (no JDK); presto-1_b200/operators.py (SetBuilderOperator) is the
 * mirror the parity tests drive.
 */
package io.trino.gpu;

import com.google.common.util.concurrent.ListenableFuture;
import com.google.common.util.concurrent.SettableFuture;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.UpdateMemory;
import io.trino.operator.Work;
import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.type.Type;

import java.util.List;
import java.util.Optional;

public class GpuSetBuilderOperator
        implements Operator
{
    /** What the probe side waits for: the set's hash, its size and whether it holds a NULL. */
    public static final class GpuChannelSet
    {
        final GpuGroupByHash hash;
        final boolean containsNull;

        GpuChannelSet(GpuGroupByHash hash, boolean containsNull)
        {
            this.hash = hash;
            this.containsNull = containsNull;
        }

        public boolean isEmpty()
        {
            return hash.getGroupCount() == 0;
        }

        public int size()
        {
            return hash.getGroupCount();
        }

        public boolean containsNull()
        {
            return containsNull;
        }
    }

    /** The counterpart of SetBuilderOperator.SetSupplier (:39-64). */
    public static final class GpuSetSupplier
    {
        private final Type type;
        private final SettableFuture<GpuChannelSet> future = SettableFuture.create();

        public GpuSetSupplier(Type type)
        {
            this.type = type;
        }

        public Type getType()
        {
            return type;
        }

        public ListenableFuture<GpuChannelSet> getChannelSet()
        {
            return future;
        }
    }

    private final OperatorContext operatorContext;
    private final GpuSetSupplier setSupplier;
    private final int setChannel;
    private final Optional<Integer> hashChannel;
    private final GpuGroupByHash hash;
    private boolean finished;

    public GpuSetBuilderOperator(OperatorContext operatorContext, long ctx, GpuSetSupplier setSupplier, int setChannel, Optional<Integer> hashChannel, int expectedPositions)
    {
        this.operatorContext = operatorContext;
        this.setSupplier = setSupplier;
        this.setChannel = setChannel;
        this.hashChannel = hashChannel;
        // the set page is (set channel[, hash channel]): channel 0 of what addInput hands to the hash
        this.hash = new GpuGroupByHash(ctx, List.of(setSupplier.getType()), new int[] {0}, hashChannel.map(c -> 1), expectedPositions, UpdateMemory.NOOP);
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return !finished;
    }

    @Override
    public void addInput(Page page)
    {
        Page setPage = hashChannel.isPresent() ? page.getColumns(setChannel, hashChannel.get()) : page.getColumns(setChannel);
        Work<?> work = hash.addPage(setPage);
        boolean done = work.process();        // UpdateMemory.NOOP: never yields
        assert done;
        operatorContext.localUserMemoryContext().setBytes(hash.getEstimatedSize());
    }

    @Override
    public Page getOutput()
    {
        return null;
    }

    @Override
    public void finish()
    {
        if (finished) {
            return;
        }
        finished = true;
        Block nullRow = setSupplier.getType().createBlockBuilder(null, 1).appendNull().build();
        boolean containsNull = hash.contains(0, new Page(nullRow), new int[] {0});
        setSupplier.future.set(new GpuChannelSet(hash, containsNull));
    }

    @Override
    public boolean isFinished()
    {
        return finished;
    }
}
