/*
 * LookupJoinOperator on the GPU for INNER and PROBE_OUTER joins (OP/LookupJoinOperator.java:299-378: processProbe / joinCurrentPosition /
 * outerJoinCurrentPosition, output through LookupJoinPageBuilder).  One probe page = one native call that returns the probe positions
 * and the build-side channels of every match in the reference's order (probe order; within a probe row the build rows in
 * positionLinks chain order).  LOOKUP_OUTER / FULL_OUTER: the library has the visited-positions tracker (PrestoGpu.lookupTrackVisited /
 * lookupOuterPositions) and GpuLookupOuterOperator emits the unmatched build rows once the probes are done.  Not compiled in this repository (no JDK); presto-1_b200/operators.py (LookupJoinOperator) is the mirror
 * the parity tests drive.
 */
package io.trino.gpu;

import com.google.common.util.concurrent.ListenableFuture;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.type.Type;

import java.util.ArrayDeque;
import java.util.List;
import java.util.Optional;

import static com.google.common.util.concurrent.Futures.getDone;

public class GpuLookupJoinOperator
        implements Operator
{
    private static final int MAX_OUTPUT_ROWS = 8192;     // LookupJoinPageBuilder's page budget: larger results leave as several pages

    private final OperatorContext operatorContext;
    private final ListenableFuture<GpuLookupSource> lookupSourceFuture;
    private final Runnable onClose;
    private final int[] probeJoinChannels;
    private final List<Type> probeJoinTypes;
    private final Optional<Integer> probeHashChannel;
    private final int[] probeOutputChannels;
    private final int[] buildOutputChannels;
    private final int joinType;
    private final ArrayDeque<Page> pending = new ArrayDeque<>();
    private boolean finishing;
    private boolean closed;

    public GpuLookupJoinOperator(OperatorContext operatorContext, ListenableFuture<GpuLookupSource> lookupSourceFuture, Runnable onClose, int[] probeJoinChannels,
            List<Type> probeJoinTypes, Optional<Integer> probeHashChannel, int[] probeOutputChannels, int[] buildOutputChannels, int joinType)
    {
        this.operatorContext = operatorContext;
        this.lookupSourceFuture = lookupSourceFuture;
        this.onClose = onClose;
        this.probeJoinChannels = probeJoinChannels;
        this.probeJoinTypes = List.copyOf(probeJoinTypes);
        this.probeHashChannel = probeHashChannel;
        this.probeOutputChannels = probeOutputChannels;
        this.buildOutputChannels = buildOutputChannels;
        this.joinType = joinType;
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public ListenableFuture<?> isBlocked()
    {
        return lookupSourceFuture.isDone() ? NOT_BLOCKED : lookupSourceFuture;      // :236-245: wait for the build side
    }

    @Override
    public boolean needsInput()
    {
        return !finishing && lookupSourceFuture.isDone() && pending.isEmpty();
    }

    @Override
    public void addInput(Page page)
    {
        GpuLookupSource lookup;
        try {
            lookup = getDone(lookupSourceFuture);
        }
        catch (java.util.concurrent.ExecutionException e) {
            throw new RuntimeException(e.getCause());
        }
        if (lookup.isEmpty() && joinType == PrestoGpu.JOIN_INNER) {
            return;                 // :262-268 an empty build side ends an inner join early
        }
        GpuLookupSource.JoinedPage joined = lookup.joinPage(page, probeJoinChannels, probeJoinTypes, probeHashChannel, joinType, buildOutputChannels);
        if (joined.positionCount == 0) {
            return;
        }
        Block[] probeBlocks = GpuLookupSource.selectProbeRows(page, probeOutputChannels, joined);
        Block[] all = new Block[probeBlocks.length + joined.buildBlocks.length];
        System.arraycopy(probeBlocks, 0, all, 0, probeBlocks.length);
        System.arraycopy(joined.buildBlocks, 0, all, probeBlocks.length, joined.buildBlocks.length);
        Page out = new Page(joined.positionCount, all);
        for (int at = 0; at < out.getPositionCount(); at += MAX_OUTPUT_ROWS) {
            pending.add(out.getRegion(at, Math.min(MAX_OUTPUT_ROWS, out.getPositionCount() - at)));
        }
    }

    @Override
    public Page getOutput()
    {
        return pending.poll();
    }

    @Override
    public void finish()
    {
        finishing = true;
    }

    @Override
    public boolean isFinished()
    {
        return finishing && pending.isEmpty();
    }

    @Override
    public void close()
    {
        if (!closed) {
            closed = true;
            onClose.run();
        }
    }
}
