/*
 * HashSemiJoinOperator on the GPU (OP/HashSemiJoinOperator.java:120-225): every probe page leaves with one more BOOLEAN channel,
 * "probe key IN (set)" under SQL's three-valued logic - NULL probe key: false on an empty set, else NULL; found: true; not found:
 * NULL when the set holds a NULL, else false (:190-218).  The column comes from one native call per page (pg_groupby_semi_join).
 * The reference's class is a WorkProcessorOperator; this one is written against the plain Operator interface and blocks on the
 * set's future.  Not compiled in this repository (no JDK); presto-1_b200/operators.py (HashSemiJoinOperator) is the mirror the
 * parity tests drive (tests/test_f4_gpu.py replays TestHashSemiJoinOperator.java:95-431).
 */
package io.trino.gpu;

import com.google.common.util.concurrent.ListenableFuture;
import io.trino.gpu.GpuSetBuilderOperator.GpuChannelSet;
import io.trino.gpu.GpuSetBuilderOperator.GpuSetSupplier;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.spi.Page;
import io.trino.spi.block.ByteArrayBlock;

import java.util.Optional;

import static com.google.common.util.concurrent.Futures.getUnchecked;

public class GpuHashSemiJoinOperator
        implements Operator
{
    private final OperatorContext operatorContext;
    private final ListenableFuture<GpuChannelSet> channelSetFuture;
    private final int probeJoinChannel;
    private final Optional<Integer> probeHashChannel;
    private GpuChannelSet channelSet;
    private Page output;
    private boolean finishing;

    public GpuHashSemiJoinOperator(OperatorContext operatorContext, GpuSetSupplier setSupplier, int probeJoinChannel, Optional<Integer> probeHashChannel)
    {
        if (probeJoinChannel < 0) {
            throw new IllegalArgumentException("probeJoinChannel is negative");
        }
        this.operatorContext = operatorContext;
        this.channelSetFuture = setSupplier.getChannelSet();
        this.probeJoinChannel = probeJoinChannel;
        this.probeHashChannel = probeHashChannel;
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public ListenableFuture<?> isBlocked()
    {
        return channelSetFuture;       // no probe page is taken before the set is built
    }

    @Override
    public boolean needsInput()
    {
        return !finishing && output == null && channelSetFuture.isDone();
    }

    @Override
    public void addInput(Page page)
    {
        if (channelSet == null) {
            channelSet = getUnchecked(channelSetFuture);
        }
        int n = page.getPositionCount();
        boolean[] values = new boolean[n];
        boolean[] isNull = new boolean[n];
        // the set's hash was created over (set channel[, hash channel]); the probe page is handed over in the same shape
        Page probe = probeHashChannel.isPresent() ? page.getColumns(probeJoinChannel, probeHashChannel.get()) : page.getColumns(probeJoinChannel);
        channelSet.hash.semiJoin(probe, channelSet.containsNull(), values, isNull);
        byte[] bytes = new byte[n];
        boolean anyNull = false;
        for (int i = 0; i < n; i++) {
            bytes[i] = (byte) (values[i] ? 1 : 0);
            anyNull |= isNull[i];
        }
        output = page.appendColumn(new ByteArrayBlock(n, anyNull ? Optional.of(isNull) : Optional.empty(), bytes));
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
}
