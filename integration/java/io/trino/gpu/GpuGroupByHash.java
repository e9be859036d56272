/*
 * GroupByHash on the GPU: the class GroupByHash.createGroupByHash (OP/GroupByHash.java:45-59) returns when GPU execution is on.
 * Work / UpdateMemory: the library refuses a page that could grow the table past the granted bytes (PG_NEED_MEMORY, nothing
 * consumed); process() then asks UpdateMemory for more, exactly where tryRehash would have (OP/BigintGroupByHash.java:295-334).
 * Not compiled in this repository (no JDK); presto-1_b200/groupby.py is the mirror the parity tests drive.
 */
package io.trino.gpu;

import io.trino.operator.GroupByHash;
import io.trino.operator.GroupByIdBlock;
import io.trino.operator.UpdateMemory;
import io.trino.operator.Work;
import io.trino.spi.Page;
import io.trino.spi.PageBuilder;
import io.trino.spi.block.Block;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.type.Type;

import java.util.List;
import java.util.Optional;

import static io.trino.gpu.PrestoGpu.check;

public class GpuGroupByHash
        implements GroupByHash, AutoCloseable
{
    private final long ctx;
    private final long handle;
    private final List<Type> types;
    private final int[] hashChannels;
    private final Optional<Integer> inputHashChannel;
    private final UpdateMemory updateMemory;
    private long grantedBytes;
    // appendValuesTo / getRawHash read through a window of exported keys (one native call per OUTPUT_WINDOW groups, not per group)
    private static final int OUTPUT_WINDOW = 8192;
    private int windowFirst = -1;
    private Block[] windowKeys;
    private long[] windowHash;

    public GpuGroupByHash(long ctx, List<? extends Type> hashTypes, int[] hashChannels, Optional<Integer> inputHashChannel, int expectedSize, UpdateMemory updateMemory)
    {
        this.ctx = ctx;
        this.types = List.copyOf(hashTypes);
        this.hashChannels = hashChannels.clone();
        this.inputHashChannel = inputHashChannel;
        this.updateMemory = updateMemory;
        long[] out = new long[1];
        check(ctx, PrestoGpu.groupbyCreate(ctx, PageArrays.typeTags(types), inputHashChannel.isPresent(), expectedSize, out));
        this.handle = out[0];
        this.grantedBytes = getEstimatedSize();
        if (updateMemory != UpdateMemory.NOOP) {
            check(ctx, PrestoGpu.groupbySetMemoryLimit(handle, grantedBytes));
        }
    }

    private PageArrays keyArrays(Page page)
    {
        if (inputHashChannel.isPresent()) {
            return PageArrays.of(page, hashChannels, types, new Block[] {page.getBlock(inputHashChannel.get())}, new int[] {PrestoGpu.BIGINT});
        }
        return PageArrays.of(page, hashChannels, types);
    }

    /** One attempt; false = the memory limit blocked the page (nothing consumed) and UpdateMemory did not grant more. */
    private boolean attempt(java.util.function.IntSupplier call)
    {
        while (true) {
            int status = call.getAsInt();
            if (status != PrestoGpu.PG_NEED_MEMORY) {
                check(ctx, status);
                windowFirst = -1;
                return true;
            }
            grantedBytes *= 2;                                     // what the next table size needs at most
            check(ctx, PrestoGpu.groupbySetMemoryLimit(handle, grantedBytes));
            if (!updateMemory.update()) {
                grantedBytes /= 2;
                check(ctx, PrestoGpu.groupbySetMemoryLimit(handle, grantedBytes));
                return false;                                      // Work.process() == false: the driver yields and calls again
            }
        }
    }

    @Override
    public Work<?> addPage(Page page)
    {
        PageArrays a = keyArrays(page);
        return new Work<Void>()
        {
            private boolean done;

            @Override
            public boolean process()
            {
                done = attempt(() -> PrestoGpu.groupbyAddPage(handle, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount, inputHashChannel.isPresent()));
                return done;
            }

            @Override
            public Void getResult()
            {
                throw new UnsupportedOperationException();
            }
        };
    }

    @Override
    public Work<GroupByIdBlock> getGroupIds(Page page)
    {
        PageArrays a = keyArrays(page);
        int[] ids = new int[page.getPositionCount()];
        long[] groupCount = new long[1];
        return new Work<GroupByIdBlock>()
        {
            private boolean done;

            @Override
            public boolean process()
            {
                done = attempt(() -> PrestoGpu.groupbyGetGroupIds(handle, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount, inputHashChannel.isPresent(), ids, groupCount));
                return done;
            }

            @Override
            public GroupByIdBlock getResult()
            {
                if (!done) {
                    throw new IllegalStateException("getResult() before process() returned true");
                }
                // GroupByIdBlock wraps a BIGINT block in the reference (GroupByIdBlock.java:30-49); ids fit an int
                long[] wide = new long[ids.length];
                for (int i = 0; i < ids.length; i++) {
                    wide[i] = ids[i];
                }
                return new GroupByIdBlock(groupCount[0], new io.trino.spi.block.LongArrayBlock(ids.length, Optional.empty(), wide));
            }
        };
    }

    @Override
    public boolean contains(int position, Page page, int[] hashChannels)
    {
        Page row = page.getRegion(position, 1);
        PageArrays a = PageArrays.of(row, hashChannels, types);
        boolean[] found = new boolean[1];
        check(ctx, PrestoGpu.groupbyContains(handle, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, 1, false, found));
        return found[0];
    }

    private void loadWindow(int groupId)
    {
        int first = groupId - groupId % OUTPUT_WINDOW;
        int n = Math.min(OUTPUT_WINDOW, getGroupCount() - first);
        int[] tags = PageArrays.typeTags(types);
        long[] varBytes = new long[tags.length];
        for (int k = 0; k < tags.length; k++) {
            if (tags[k] == PrestoGpu.VARCHAR || tags[k] == PrestoGpu.VARBINARY) {
                long[] b = new long[1];
                check(ctx, PrestoGpu.groupbyVarwidthBytes(handle, k, b));     // all groups: an upper bound for the window
                varBytes[k] = b[0];
            }
        }
        OutputArrays out = OutputArrays.allocate(tags, n, varBytes);
        windowHash = new long[n];
        check(ctx, PrestoGpu.groupbyExportKeys(handle, first, n, tags, out.values, out.nulls, out.offsets, out.valueBytes, windowHash));
        windowKeys = PageArrays.toBlocks(tags, out, n);
        windowFirst = first;
    }

    @Override
    public void appendValuesTo(int groupId, PageBuilder pageBuilder, int outputChannelOffset)
    {
        if (windowFirst < 0 || groupId < windowFirst || groupId >= windowFirst + windowHash.length) {
            loadWindow(groupId);
        }
        int at = groupId - windowFirst;
        for (int k = 0; k < types.size(); k++) {
            types.get(k).appendTo(windowKeys[k], at, pageBuilder.getBlockBuilder(outputChannelOffset + k));
        }
        if (inputHashChannel.isPresent()) {
            io.trino.spi.type.BigintType.BIGINT.writeLong(pageBuilder.getBlockBuilder(outputChannelOffset + types.size()), windowHash[at]);
        }
    }

    @Override
    public long getRawHash(int groupId)
    {
        if (windowFirst < 0 || groupId < windowFirst || groupId >= windowFirst + windowHash.length) {
            loadWindow(groupId);
        }
        return windowHash[groupId - windowFirst];
    }

    /**
     * MarkDistinctHash.markDistinctRows (OP/MarkDistinctHash.java:52-69): the page is added, and marks[i] is true for the first row of
     * every group the page introduced.  Used by GpuMarkDistinctOperator and GpuDistinctLimitOperator.
     */
    boolean[] markDistinct(Page page)
    {
        PageArrays a = keyArrays(page);
        boolean[] marks = new boolean[page.getPositionCount()];
        check(ctx, PrestoGpu.groupbyMarkDistinct(handle, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount, inputHashChannel.isPresent(), marks));
        windowFirst = -1;
        return marks;
    }

    /**
     * The BOOLEAN channel HashSemiJoinOperator appends (OP/HashSemiJoinOperator.java:190-218) for the probe channel of `page`:
     * values[i] / isNull[i] under SQL's three-valued IN.  `this` is the ChannelSet's hash (built by GpuSetBuilderOperator).
     */
    void semiJoin(Page page, boolean setContainsNull, boolean[] values, boolean[] isNull)
    {
        PageArrays a = keyArrays(page);
        check(ctx, PrestoGpu.groupbySemiJoin(handle, a.types, a.values, a.nulls, a.offsets, a.arrayOffsets, a.positionCount, inputHashChannel.isPresent(), setContainsNull, values, isNull));
    }

    private long[] stats()
    {
        long[] s = new long[4];
        check(ctx, PrestoGpu.groupbyStats(handle, s));
        return s;
    }

    @Override
    public long getEstimatedSize()
    {
        return stats()[3];          // device bytes (not JVM heap; reported to the operator's system memory context)
    }

    @Override
    public long getHashCollisions()
    {
        return stats()[2];
    }

    @Override
    public double getExpectedHashCollisions()
    {
        return io.trino.util.HashCollisionsEstimator.estimateNumberOfHashCollisions(getGroupCount(), getCapacity());
    }

    @Override
    public List<Type> getTypes()
    {
        return types;
    }

    @Override
    public int getGroupCount()
    {
        return (int) stats()[0];
    }

    @Override
    public int getCapacity()
    {
        return (int) stats()[1];    // the reference's logical capacity schedule, not the device table's slot count
    }

    @Override
    public void close()
    {
        PrestoGpu.groupbyDestroy(handle);
    }
}
