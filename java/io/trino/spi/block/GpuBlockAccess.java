/*
 * Lives in package io.trino.spi.block to reach the package-private bulk accessors of the block classes
 * (LongArrayBlock.java:269 getValuesSlice(), the private long[] values / arrayOffset fields are exposed through
 * getRawValues-style helpers added here). NOT COMPILED IN THIS IMAGE (no JDK).
 */
package io.trino.spi.block;

public final class GpuBlockAccess
{
    private GpuBlockAccess() {}

    /** Unwraps LazyBlock (LazyBlock.java:262-274) and returns the loaded block. */
    public static Block loaded(Block block)
    {
        return block.getLoadedBlock();
    }

    /** Backing array + offset of a flat block without per-position calls. */
    public static Object rawValues(Block block)
    {
        if (block instanceof LongArrayBlock) {
            return ((LongArrayBlock) block).getValuesSlice().getBase();      // long[]
        }
        if (block instanceof IntArrayBlock) {
            return ((IntArrayBlock) block).getValuesSlice().getBase();       // int[]
        }
        if (block instanceof ByteArrayBlock) {
            return ((ByteArrayBlock) block).getValuesSlice().getBase();      // byte[]
        }
        if (block instanceof Int128ArrayBlock) {
            return ((Int128ArrayBlock) block).getValuesSlice().getBase();    // long[2n], sign-magnitude
        }
        if (block instanceof VariableWidthBlock) {
            return ((VariableWidthBlock) block).getRawSlice(0).getBase();    // byte[]
        }
        throw new IllegalArgumentException("not a flat block: " + block.getClass().getName());
    }

    /** Element offset of position 0 inside rawValues (arrayOffset / slice address). */
    public static int rawOffset(Block block)
    {
        if (block instanceof VariableWidthBlock) {
            VariableWidthBlock v = (VariableWidthBlock) block;
            return (int) (v.getRawSlice(0).getAddress() - sun.misc.Unsafe.ARRAY_BYTE_BASE_OFFSET) + v.getPositionOffset(0);
        }
        return AbstractArrayBlockAccess.arrayOffset(block); // one-line accessor per block class (private field arrayOffset)
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
