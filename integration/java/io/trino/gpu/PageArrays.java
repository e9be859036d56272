/*
 * A Page as the parallel arrays the natives take (types, values, nulls, offsets, arrayOffsets, positionCount), and the way back:
 * output arrays wrapped into blocks.  Not compiled in this repository (no JDK).
 */
package io.trino.gpu;

import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.block.ByteArrayBlock;
import io.trino.spi.block.Int128ArrayBlock;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.block.LongArrayBlock;
import io.trino.spi.block.VariableWidthBlock;
import io.trino.spi.type.DecimalType;
import io.trino.spi.type.Type;

import java.util.List;
import java.util.Optional;

import static io.airlift.slice.Slices.wrappedBuffer;
import static io.trino.spi.type.BigintType.BIGINT;
import static io.trino.spi.type.BooleanType.BOOLEAN;
import static io.trino.spi.type.DateType.DATE;
import static io.trino.spi.type.DoubleType.DOUBLE;
import static io.trino.spi.type.IntegerType.INTEGER;
import static io.trino.spi.type.VarbinaryType.VARBINARY;

final class PageArrays
{
    final int[] types;
    final Object[] values;
    final Object[] nulls;
    final Object[] offsets;
    final int[] arrayOffsets;
    final long positionCount;

    private PageArrays(int channels, long positionCount)
    {
        this.types = new int[channels];
        this.values = new Object[channels];
        this.nulls = new Object[channels];
        this.offsets = new Object[channels];
        this.arrayOffsets = new int[channels];
        this.positionCount = positionCount;
    }

    /** Channels `channels` of the page, typed by `channelTypes` (the block class alone cannot tell BIGINT from DOUBLE or a short decimal). */
    static PageArrays of(Page page, int[] channels, List<? extends Type> channelTypes)
    {
        PageArrays a = new PageArrays(channels.length, page.getPositionCount());
        for (int i = 0; i < channels.length; i++) {
            a.set(i, page.getBlock(channels[i]), typeTag(channelTypes.get(i)));
        }
        return a;
    }

    /** The same, followed by extra blocks (a precomputed hash channel, a mask, the group ids of a GroupByIdBlock). */
    static PageArrays of(Page page, int[] channels, List<? extends Type> channelTypes, Block[] extra, int[] extraTags)
    {
        PageArrays a = new PageArrays(channels.length + extra.length, page.getPositionCount());
        for (int i = 0; i < channels.length; i++) {
            a.set(i, page.getBlock(channels[i]), typeTag(channelTypes.get(i)));
        }
        for (int i = 0; i < extra.length; i++) {
            a.set(channels.length + i, extra[i], extraTags[i]);
        }
        return a;
    }

    private void set(int i, Block block, int tag)
    {
        BlockArrays arrays = BlockArrays.of(block);      // flattens Dictionary / RLE / Lazy blocks
        types[i] = tag;
        values[i] = arrays.values;
        nulls[i] = arrays.valueIsNull;
        offsets[i] = arrays.offsets;
        arrayOffsets[i] = arrays.arrayOffset;
    }

    static int typeTag(Type type)
    {
        if (type.equals(BIGINT)) {
            return PrestoGpu.BIGINT;
        }
        if (type.equals(INTEGER)) {
            return PrestoGpu.INTEGER;
        }
        if (type.equals(DATE)) {
            return PrestoGpu.DATE;
        }
        if (type.equals(DOUBLE)) {
            return PrestoGpu.DOUBLE;
        }
        if (type.equals(BOOLEAN)) {
            return PrestoGpu.BOOLEAN;
        }
        if (type instanceof DecimalType) {
            return ((DecimalType) type).isShort() ? PrestoGpu.SHORT_DECIMAL : PrestoGpu.LONG_DECIMAL;
        }
        if (type.equals(VARBINARY)) {
            return PrestoGpu.VARBINARY;
        }
        if (type.getJavaType() == io.airlift.slice.Slice.class) {
            return PrestoGpu.VARCHAR;       // varchar(n), char(n) as their raw bytes
        }
        throw new IllegalArgumentException("type not supported by the GPU operators: " + type);
    }

    static int[] typeTags(List<? extends Type> types)
    {
        int[] tags = new int[types.size()];
        for (int i = 0; i < tags.length; i++) {
            tags[i] = typeTag(types.get(i));
        }
        return tags;
    }

    /** Output arrays (OutputArrays) -> blocks of `rows` positions. */
    static Block[] toBlocks(int[] types, OutputArrays out, int rows)
    {
        Block[] blocks = new Block[types.length];
        for (int c = 0; c < blocks.length; c++) {
            Optional<boolean[]> isNull = Optional.ofNullable((boolean[]) out.nulls[c]);
            switch (types[c]) {
                case PrestoGpu.INTEGER:
                case PrestoGpu.DATE:
                    blocks[c] = new IntArrayBlock(rows, isNull, (int[]) out.values[c]);
                    break;
                case PrestoGpu.LONG_DECIMAL:
                    blocks[c] = new Int128ArrayBlock(rows, isNull, (long[]) out.values[c]);
                    break;
                case PrestoGpu.VARCHAR:
                case PrestoGpu.VARBINARY:
                    blocks[c] = new VariableWidthBlock(rows, wrappedBuffer((byte[]) out.values[c]), (int[]) out.offsets[c], isNull);
                    break;
                case PrestoGpu.BOOLEAN:
                    blocks[c] = new ByteArrayBlock(rows, isNull, (byte[]) out.values[c]);
                    break;
                default:
                    blocks[c] = new LongArrayBlock(rows, isNull, (long[]) out.values[c]);
            }
        }
        return blocks;
    }
}
