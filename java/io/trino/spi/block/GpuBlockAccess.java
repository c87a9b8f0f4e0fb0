/*
 * Bulk access to the backing arrays of the flat block classes, for io.trino.operator.gpu.GpuPageMarshaller.
 * Lives in package io.trino.spi.block because the accessors it needs are package-private there:
 *   LongArrayBlock.getValuesSlice()    (core/trino-spi/src/main/java/io/trino/spi/block/LongArrayBlock.java:269-272)
 *   IntArrayBlock.getValuesSlice()     (IntArrayBlock.java:222-225)
 *   ByteArrayBlock.getValuesSlice()    (ByteArrayBlock.java:222-225)
 *   Int128ArrayBlock.getValuesSlice()  (Int128ArrayBlock.java:230)
 *   AbstractVariableWidthBlock.getRawSlice / getPositionOffset (protected; VariableWidthBlock.java:79)
 * Each returns a Slice that WRAPS the block's array at its arrayOffset (Slices.wrappedLongArray(values, arrayOffset, n)):
 * Slice.getBase() is the array itself and Slice.getAddress() its byte position, so neither the private `values` nor the
 * private `arrayOffset` field has to be opened up. valueIsNull has no bulk accessor in the reference at all; it is gathered
 * with isNull(position), as the reference's own block encoders do (EncoderUtil.encodeNullsAsBits, EncoderUtil.java:33-58).
 * NOT COMPILED IN THIS IMAGE (no JDK, no Maven repository).
 */
package io.trino.spi.block;

import io.airlift.slice.Slice;

import static sun.misc.Unsafe.ARRAY_BYTE_BASE_OFFSET;
import static sun.misc.Unsafe.ARRAY_INT_BASE_OFFSET;
import static sun.misc.Unsafe.ARRAY_LONG_BASE_OFFSET;

public final class GpuBlockAccess
{
    private GpuBlockAccess() {}

    /** Unwraps LazyBlock (LazyBlock.java:262-274) and returns the loaded block. */
    public static Block loaded(Block block)
    {
        return block.getLoadedBlock();
    }

    private static Slice valuesSlice(Block block)
    {
        if (block instanceof LongArrayBlock) {
            return ((LongArrayBlock) block).getValuesSlice();
        }
        if (block instanceof IntArrayBlock) {
            return ((IntArrayBlock) block).getValuesSlice();
        }
        if (block instanceof ByteArrayBlock) {
            return ((ByteArrayBlock) block).getValuesSlice();
        }
        if (block instanceof Int128ArrayBlock) {
            return ((Int128ArrayBlock) block).getValuesSlice();
        }
        if (block instanceof VariableWidthBlock) {
            return ((VariableWidthBlock) block).getRawSlice(0);
        }
        throw new IllegalArgumentException("not a flat block: " + block.getClass().getName());
    }

    /** The backing array of a flat block: long[] (LongArrayBlock, Int128ArrayBlock), int[], byte[] (ByteArrayBlock, var-width payload). */
    public static Object rawValues(Block block)
    {
        return valuesSlice(block).getBase();
    }

    /**
     * Index of position 0 inside rawValues, in ELEMENTS of that array (for Int128ArrayBlock: in longs, i.e. 2 x arrayOffset;
     * for var-width blocks: the byte at which the slice starts — per-position offsets come from variableWidthOffsets).
     */
    public static int rawOffset(Block block)
    {
        Slice slice = valuesSlice(block);
        if (block instanceof LongArrayBlock || block instanceof Int128ArrayBlock) {
            return (int) ((slice.getAddress() - ARRAY_LONG_BASE_OFFSET) / Long.BYTES);
        }
        if (block instanceof IntArrayBlock) {
            return (int) ((slice.getAddress() - ARRAY_INT_BASE_OFFSET) / Integer.BYTES);
        }
        return (int) (slice.getAddress() - ARRAY_BYTE_BASE_OFFSET);
    }

    /** boolean[positionCount] of the block's null flags, or null when it has none (Block.mayHaveNull()). */
    public static boolean[] nullFlags(Block block)
    {
        if (!block.mayHaveNull()) {
            return null;
        }
        int n = block.getPositionCount();
        boolean[] flags = new boolean[n];
        boolean any = false;
        for (int i = 0; i < n; i++) {
            flags[i] = block.isNull(i);
            any |= flags[i];
        }
        return any ? flags : null;
    }

    /**
     * int[positionCount + 1] offsets of a var-width block into rawValues, absolute (rawOffset already added), as the native
     * side expects them (pgpu_column.offsets).
     */
    public static int[] variableWidthOffsets(VariableWidthBlock block)
    {
        int n = block.getPositionCount();
        int base = rawOffset(block);
        int[] offsets = new int[n + 1];
        for (int i = 0; i < n; i++) {
            offsets[i] = base + block.getPositionOffset(i);
        }
        offsets[n] = n == 0 ? base : offsets[n - 1] + block.getSliceLength(n - 1);
        return offsets;
    }

    public static int[] dictionaryIds(DictionaryBlock block)
    {
        return block.getRawIds();          // DictionaryBlock.java:429-441 (ids + idsOffset)
    }

    public static int dictionaryIdsOffset(DictionaryBlock block)
    {
        return block.getRawIdsOffset();
    }
}
