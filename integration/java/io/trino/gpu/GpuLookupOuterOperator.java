/*
 * LookupOuterOperator on the GPU (OP/LookupOuterOperator.java:115-220; RIGHT and FULL joins): once every probe operator of the join is
 * done, the build rows no probe matched are emitted with NULLs on the probe side.  The lookup source keeps a device bitmap of the
 * build positions it emitted (GpuLookupSource.trackVisitedPositions, switched on by the join's factory before the first probe -
 * OuterLookupSource.java:176-205), and unvisitedRows() is OuterPositionIterator + appendTo in one piece.  Output pages are cut to
 * the page builder's row limit on the host.  Not compiled in this repository (no JDK); presto-1_b200/operators.py
 * (LookupOuterOperator) is the mirror the parity tests drive (tests/test_f4_gpu.py, tests/test_join_gpu.py).
 */
package io.trino.gpu;

import com.google.common.util.concurrent.ListenableFuture;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.block.RunLengthEncodedBlock;
import io.trino.spi.type.Type;

import java.util.List;

import static com.google.common.util.concurrent.Futures.getUnchecked;

public class GpuLookupOuterOperator
        implements Operator
{
    private static final int MAX_OUTPUT_ROWS = 8192;          // the row bound of the reference's page builder

    private final OperatorContext operatorContext;
    private final ListenableFuture<GpuLookupSource> lookupSourceWhenProbesAreDone;
    private final List<Type> probeOutputTypes;
    private final int[] buildOutputChannels;
    private final Runnable onClose;
    private Page unvisited;
    private int emitted;
    private boolean loaded;
    private boolean closed;

    /** `lookupSourceWhenProbesAreDone` completes when the last probe operator of the join has finished (the factory's reference counting). */
    public GpuLookupOuterOperator(OperatorContext operatorContext, ListenableFuture<GpuLookupSource> lookupSourceWhenProbesAreDone, List<Type> probeOutputTypes,
            int[] buildOutputChannels, Runnable onClose)
    {
        this.operatorContext = operatorContext;
        this.lookupSourceWhenProbesAreDone = lookupSourceWhenProbesAreDone;
        this.probeOutputTypes = List.copyOf(probeOutputTypes);
        this.buildOutputChannels = buildOutputChannels.clone();
        this.onClose = onClose;
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public ListenableFuture<?> isBlocked()
    {
        return lookupSourceWhenProbesAreDone;
    }

    @Override
    public boolean needsInput()
    {
        return false;        // a source of pages (:160-163)
    }

    @Override
    public void addInput(Page page)
    {
        throw new UnsupportedOperationException();
    }

    @Override
    public Page getOutput()
    {
        if (closed || !lookupSourceWhenProbesAreDone.isDone()) {
            return null;
        }
        if (!loaded) {
            unvisited = getUnchecked(lookupSourceWhenProbesAreDone).unvisitedRows(buildOutputChannels);
            loaded = true;
        }
        if (unvisited == null || emitted == unvisited.getPositionCount()) {
            close();
            return null;
        }
        int n = Math.min(MAX_OUTPUT_ROWS, unvisited.getPositionCount() - emitted);
        Block[] blocks = new Block[probeOutputTypes.size() + buildOutputChannels.length];
        for (int c = 0; c < probeOutputTypes.size(); c++) {
            Block nullValue = probeOutputTypes.get(c).createBlockBuilder(null, 1).appendNull().build();
            blocks[c] = new RunLengthEncodedBlock(nullValue, n);                    // the probe side of an unmatched build row is NULL (:196-203)
        }
        for (int c = 0; c < buildOutputChannels.length; c++) {
            blocks[probeOutputTypes.size() + c] = unvisited.getBlock(c).getRegion(emitted, n);
        }
        emitted += n;
        return new Page(n, blocks);
    }

    @Override
    public void finish()
    {
        // nothing to flush: the operator ends when the unvisited rows are out
    }

    @Override
    public boolean isFinished()
    {
        return closed;
    }

    @Override
    public void close()
    {
        if (!closed) {
            closed = true;
            unvisited = null;
            onClose.run();
        }
    }
}
