/*
 * The backing arrays of the SPI's flat block classes.  They are private final fields without public getters
 * (SPI/block/LongArrayBlock.java:37-41 arrayOffset / valueIsNull / values; IntArrayBlock, ByteArrayBlock, Int128ArrayBlock the same;
 * VariableWidthBlock.java:38-48 arrayOffset / slice / offsets / valueIsNull), so this class reads them through VarHandles it
 * resolves once.  A maintainer who can touch the SPI replaces it by package-private accessors in io.trino.spi.block
 * (the block encodings in that package read the same fields directly).  Not compiled in this repository (no JDK).
 */
package io.trino.gpu;

import io.airlift.slice.Slice;
import io.trino.spi.block.Block;
import io.trino.spi.block.ByteArrayBlock;
import io.trino.spi.block.Int128ArrayBlock;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.block.LongArrayBlock;
import io.trino.spi.block.VariableWidthBlock;

import java.lang.invoke.MethodHandles;
import java.lang.invoke.VarHandle;

final class BlockFields
{
    private static final VarHandle LONG_VALUES = field(LongArrayBlock.class, "values", long[].class);
    private static final VarHandle INT_VALUES = field(IntArrayBlock.class, "values", int[].class);
    private static final VarHandle BYTE_VALUES = field(ByteArrayBlock.class, "values", byte[].class);
    private static final VarHandle INT128_VALUES = field(Int128ArrayBlock.class, "values", long[].class);
    private static final VarHandle VAR_SLICE = field(VariableWidthBlock.class, "slice", Slice.class);
    private static final VarHandle VAR_OFFSETS = field(VariableWidthBlock.class, "offsets", int[].class);

    private BlockFields() {}

    private static VarHandle field(Class<?> owner, String name, Class<?> type)
    {
        try {
            return MethodHandles.privateLookupIn(owner, MethodHandles.lookup()).findVarHandle(owner, name, type);
        }
        catch (ReflectiveOperationException e) {
            throw new ExceptionInInitializerError(e);
        }
    }

    private static VarHandle fieldOf(Block block, String name, Class<?> type)
    {
        return field(block.getClass(), name, type);
    }

    static long[] longValues(LongArrayBlock block)
    {
        return (long[]) LONG_VALUES.get(block);
    }

    static int[] intValues(IntArrayBlock block)
    {
        return (int[]) INT_VALUES.get(block);
    }

    static byte[] byteValues(ByteArrayBlock block)
    {
        return (byte[]) BYTE_VALUES.get(block);
    }

    /** long[2 * positionCount]: per position the two words of UnscaledDecimal128Arithmetic's Slice (low word, then sign + high word:
     *  UnscaledDecimal128Arithmetic.java:165-176) — the layout the library's LONG_DECIMAL columns use, no conversion. */
    static long[] int128Values(Int128ArrayBlock block)
    {
        return (long[]) INT128_VALUES.get(block);
    }

    /** The Slice's backing byte[]; offsets[] are relative to the Slice, so a Slice that does not start at byte 0 of its array is copied. */
    static byte[] sliceBytes(VariableWidthBlock block)
    {
        Slice slice = (Slice) VAR_SLICE.get(block);
        if (slice.hasByteArray() && slice.byteArrayOffset() == 0) {
            return slice.byteArray();
        }
        return slice.getBytes();
    }

    static int[] offsets(VariableWidthBlock block)
    {
        return (int[]) VAR_OFFSETS.get(block);
    }

    /** null when the block has no null mask (mayHaveNull() == false). */
    static boolean[] valueIsNull(Block block)
    {
        if (!block.mayHaveNull()) {
            return null;
        }
        return (boolean[]) fieldOf(block, "valueIsNull", boolean[].class).get(block);
    }

    static int arrayOffset(Block block)
    {
        return (int) fieldOf(block, "arrayOffset", int.class).get(block);
    }
}
