/*
 * The device-resident lookup source a GpuHashBuilderOperator publishes.  It is NOT a drop-in io.trino.operator.LookupSource: that
 * interface is position-at-a-time (getJoinPosition / getNextJoinPosition / appendTo per row, OP/LookupSource.java:27-48), which a
 * device table cannot serve without a round trip per row.  GpuLookupJoinOperator consumes it a page at a time instead — the same
 * JoinProbe + LookupJoinPageBuilder result (OP/LookupJoinOperator.java:333-361), produced by one native call.
 * Not compiled in this repository (no JDK); presto-1_b200/join.py (LookupSource) is the mirror the parity tests drive.
 */
package io.trino.gpu;

import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.type.Type;

import java.util.List;
import java.util.Optional;

import static io.trino.gpu.PrestoGpu.check;

public final class GpuLookupSource
        implements AutoCloseable
{
    private final long ctx;
    private final long handle;
    private final List<Type> buildTypes;

    GpuLookupSource(long ctx, long handle, List<Type> buildTypes)
    {
        this.ctx = ctx;
        this.handle = handle;
        this.buildTypes = List.copyOf(buildTypes);
    }

    private long[] stats()
    {
        long[] s = new long[4];
        check(ctx, PrestoGpu.lookupStats(handle, s));
        return s;
    }

    public int getChannelCount()
    {
        return buildTypes.size();
    }

    public long getJoinPositionCount()
    {
        return stats()[0];
    }

    public long getInMemorySizeInBytes()
    {
        return stats()[2];
    }

    public boolean isEmpty()
    {
        return stats()[0] == 0;
    }

    /** Result of one probe page: probe positions (to select the probe-side output channels) and the build-side blocks. */
    public static final class JoinedPage
    {
        public final int[] probePositions;
        public final int positionCount;
        public final Block[] buildBlocks;

        JoinedPage(int[] probePositions, int positionCount, Block[] buildBlocks)
        {
            this.probePositions = probePositions;
            this.positionCount = positionCount;
            this.buildBlocks = buildBlocks;
        }
    }

    /**
     * JoinProbe over the page + appendTo of `buildOutputChannels` for every match.  Fixed width build channels go through
     * lookupJoinPage (one call); variable width ones through lookupProbe + lookupGatherVarBytes + lookupGather.
     */
    public JoinedPage joinPage(Page probe, int[] probeJoinChannels, List<Type> probeJoinTypes, Optional<Integer> probeHashChannel, int joinType, int[] buildOutputChannels)
    {
        PageArrays keys = probeHashChannel.isPresent()
                ? PageArrays.of(probe, probeJoinChannels, probeJoinTypes, new Block[] {probe.getBlock(probeHashChannel.get())}, new int[] {PrestoGpu.BIGINT})
                : PageArrays.of(probe, probeJoinChannels, probeJoinTypes);
        int[] outTags = new int[buildOutputChannels.length];
        boolean variableWidth = false;
        for (int i = 0; i < outTags.length; i++) {
            outTags[i] = PageArrays.typeTag(buildTypes.get(buildOutputChannels[i]));
            variableWidth |= outTags[i] == PrestoGpu.VARCHAR || outTags[i] == PrestoGpu.VARBINARY;
        }
        int capacity = Math.max(1, probe.getPositionCount());
        long[] produced = new long[1];
        while (true) {
            int[] probeIdx = new int[capacity];
            int status;
            if (!variableWidth) {
                OutputArrays out = OutputArrays.allocate(outTags, capacity);
                status = PrestoGpu.lookupJoinPage(handle, keys.types, keys.values, keys.nulls, keys.offsets, keys.arrayOffsets, keys.positionCount,
                        probeHashChannel.isPresent(), joinType, buildOutputChannels, probeIdx, outTags, out.values, out.nulls, out.offsets, out.valueBytes, produced);
                if (status == PrestoGpu.PG_OK) {
                    int n = (int) produced[0];
                    return new JoinedPage(probeIdx, n, PageArrays.toBlocks(outTags, out, n));
                }
            }
            else {
                long[] buildPos = new long[capacity];
                status = PrestoGpu.lookupProbe(handle, keys.types, keys.values, keys.nulls, keys.offsets, keys.arrayOffsets, keys.positionCount,
                        probeHashChannel.isPresent(), joinType, probeIdx, buildPos, produced);
                if (status == PrestoGpu.PG_OK) {
                    int n = (int) produced[0];
                    long[] varBytes = new long[outTags.length];
                    for (int i = 0; i < outTags.length; i++) {
                        if (outTags[i] == PrestoGpu.VARCHAR || outTags[i] == PrestoGpu.VARBINARY) {
                            long[] b = new long[1];
                            check(ctx, PrestoGpu.lookupGatherVarBytes(handle, buildPos, n, buildOutputChannels[i], b));
                            varBytes[i] = b[0];
                        }
                    }
                    OutputArrays out = OutputArrays.allocate(outTags, n, varBytes);
                    check(ctx, PrestoGpu.lookupGather(handle, buildPos, n, buildOutputChannels, outTags, out.values, out.nulls, out.offsets, out.valueBytes));
                    return new JoinedPage(probeIdx, n, PageArrays.toBlocks(outTags, out, n));
                }
            }
            if (status != PrestoGpu.PG_NEED_MEMORY) {
                check(ctx, status);
            }
            capacity = (int) produced[0];       // more matches than probe rows: the call consumed nothing and said how many rows it needs
        }
    }

    /** OuterLookupSource (OP/OuterLookupSource.java:176-205): from now on every emitted build position is recorded.  Call before the first probe. */
    public void trackVisitedPositions()
    {
        check(ctx, PrestoGpu.lookupTrackVisited(handle));
    }

    /**
     * OuterPositionIterator (OP/OuterLookupSource.java:118-160) in one piece: the build rows no probe has matched, ascending, with
     * the requested build channels materialised (LookupSource.appendTo for each of them).
     */
    public Page unvisitedRows(int[] buildOutputChannels)
    {
        long[] count = new long[1];
        long[] positions = new long[0];
        int status = PrestoGpu.lookupOuterPositions(handle, positions, count);
        if (status == PrestoGpu.PG_NEED_MEMORY) {
            positions = new long[(int) count[0]];
            status = PrestoGpu.lookupOuterPositions(handle, positions, count);
        }
        check(ctx, status);
        int n = (int) count[0];
        if (n == 0) {
            return null;
        }
        int[] outTags = new int[buildOutputChannels.length];
        long[] varBytes = new long[outTags.length];
        for (int i = 0; i < outTags.length; i++) {
            outTags[i] = PageArrays.typeTag(buildTypes.get(buildOutputChannels[i]));
            if (outTags[i] == PrestoGpu.VARCHAR || outTags[i] == PrestoGpu.VARBINARY) {
                long[] b = new long[1];
                check(ctx, PrestoGpu.lookupGatherVarBytes(handle, positions, n, buildOutputChannels[i], b));
                varBytes[i] = b[0];
            }
        }
        OutputArrays out = OutputArrays.allocate(outTags, n, varBytes);
        check(ctx, PrestoGpu.lookupGather(handle, positions, n, buildOutputChannels, outTags, out.values, out.nulls, out.offsets, out.valueBytes));
        return new Page(n, PageArrays.toBlocks(outTags, out, n));
    }

    /** The probe-side channels of the output: every output row repeats probe row probePositions[i]. */
    static Block[] selectProbeRows(Page probe, int[] probeOutputChannels, JoinedPage joined)
    {
        Block[] blocks = new Block[probeOutputChannels.length];
        for (int c = 0; c < blocks.length; c++) {
            blocks[c] = probe.getBlock(probeOutputChannels[c]).getPositions(joined.probePositions, 0, joined.positionCount);   // a DictionaryBlock view, as LookupJoinPageBuilder builds
        }
        return blocks;
    }

    static Block positionsBlock(JoinedPage joined)
    {
        return new IntArrayBlock(joined.positionCount, Optional.empty(), joined.probePositions);
    }

    @Override
    public void close()
    {
        PrestoGpu.lookupDestroy(handle);
    }
}
