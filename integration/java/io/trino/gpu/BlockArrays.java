/*
 * Backing arrays of a Block, as the JNI shim wants them.  Dictionary / RLE / lazy blocks are flattened first.
 * The accessors used here are the ones the SPI exposes to encodings (LongArrayBlockEncoding etc. read the same fields
 * through package-private getters; a maintainer places this class in io.trino.spi.block or adds public accessors).
 * Not compiled in this repository (no JDK).
 */
package io.trino.gpu;

import io.trino.spi.block.Block;
import io.trino.spi.block.ByteArrayBlock;
import io.trino.spi.block.DictionaryBlock;
import io.trino.spi.block.Int128ArrayBlock;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.block.LongArrayBlock;
import io.trino.spi.block.RunLengthEncodedBlock;
import io.trino.spi.block.VariableWidthBlock;

final class BlockArrays
{
    int typeTag;
    Object values;          // long[] | int[] | byte[]
    boolean[] valueIsNull;  // null when the block has no nulls
    int[] offsets;          // variable width only
    int arrayOffset;

    static BlockArrays of(Block block)
    {
        block = block.getLoadedBlock();
        if (block instanceof DictionaryBlock || block instanceof RunLengthEncodedBlock) {
            int[] positions = new int[block.getPositionCount()];
            for (int i = 0; i < positions.length; i++) {
                positions[i] = i;
            }
            block = block.copyPositions(positions, 0, positions.length);   // flat copy of the same values
        }
        BlockArrays a = new BlockArrays();
        if (block instanceof LongArrayBlock) {
            a.typeTag = PrestoGpu.BIGINT;     // caller overrides with DOUBLE / SHORT_DECIMAL from the channel's Type
            a.values = BlockFields.longValues((LongArrayBlock) block);
        }
        else if (block instanceof IntArrayBlock) {
            a.typeTag = PrestoGpu.INTEGER;
            a.values = BlockFields.intValues((IntArrayBlock) block);
        }
        else if (block instanceof Int128ArrayBlock) {
            a.typeTag = PrestoGpu.LONG_DECIMAL;
            a.values = BlockFields.int128Values((Int128ArrayBlock) block);
        }
        else if (block instanceof ByteArrayBlock) {
            a.typeTag = PrestoGpu.BOOLEAN;
            a.values = BlockFields.byteValues((ByteArrayBlock) block);
        }
        else if (block instanceof VariableWidthBlock) {
            a.typeTag = PrestoGpu.VARCHAR;
            a.values = BlockFields.sliceBytes((VariableWidthBlock) block);
            a.offsets = BlockFields.offsets((VariableWidthBlock) block);
        }
        else {
            throw new IllegalArgumentException("unsupported block " + block.getClass().getSimpleName());
        }
        a.valueIsNull = BlockFields.valueIsNull(block);
        a.arrayOffset = BlockFields.arrayOffset(block);
        return a;
    }
}
