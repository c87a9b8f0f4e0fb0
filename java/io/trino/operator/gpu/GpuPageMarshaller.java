/*
 * Page <-> raw arrays, the Java half of the boundary between io.trino.spi.Page and libprestogpu (include/prestogpu.h).
 *   addPage:  Page -> per channel (encoding, type, backing arrays) -> PrestoGpuNative.addPage. LazyBlocks are loaded
 *             (LazyBlock.getLoadedBlock, spi/block/LazyBlock.java:262-274), DictionaryBlock / RunLengthEncodedBlock are passed
 *             as they are (ids + dictionary / one-position value block: spi/block/DictionaryBlock.java:40-48,
 *             RunLengthEncodedBlock.java:45-48) so the native side can use its dictionary-aware kernels; nested
 *             dictionaries are flattened first.
 *   toBlocks: the arrays PrestoGpuNative.getOutput returns -> fresh LongArrayBlock / IntArrayBlock / ByteArrayBlock /
 *             Int128ArrayBlock / VariableWidthBlock objects owned by the consumer.
 * NOT COMPILED IN THIS IMAGE (no JDK, no Maven repository); the Python mirror of the same marshalling is presto_b200/spi.py.
 */
package io.trino.operator.gpu;

import io.airlift.slice.Slices;
import io.trino.spi.Page;
import io.trino.spi.TrinoException;
import io.trino.spi.block.Block;
import io.trino.spi.block.ByteArrayBlock;
import io.trino.spi.block.DictionaryBlock;
import io.trino.spi.block.GpuBlockAccess;
import io.trino.spi.block.Int128ArrayBlock;
import io.trino.spi.block.IntArrayBlock;
import io.trino.spi.block.LongArrayBlock;
import io.trino.spi.block.RunLengthEncodedBlock;
import io.trino.spi.block.VariableWidthBlock;
import io.trino.spi.type.BigintType;
import io.trino.spi.type.BooleanType;
import io.trino.spi.type.DateType;
import io.trino.spi.type.DecimalType;
import io.trino.spi.type.DoubleType;
import io.trino.spi.type.IntegerType;
import io.trino.spi.type.Type;
import io.trino.spi.type.VarbinaryType;
import io.trino.spi.type.VarcharType;

import java.util.List;
import java.util.Optional;

import static io.trino.spi.StandardErrorCode.NOT_SUPPORTED;

final class GpuPageMarshaller
{
    // pgpu_encoding / pgpu_type_kind (include/prestogpu.h)
    static final int ENC_PLAIN = 0;
    static final int ENC_DICTIONARY = 1;
    static final int ENC_RLE = 2;
    static final int BOOLEAN = 0;
    static final int BIGINT = 1;
    static final int INTEGER = 2;
    static final int DATE = 3;
    static final int DOUBLE = 4;
    static final int DECIMAL = 5;
    static final int VARCHAR = 6;
    static final int VARBINARY = 7;

    private GpuPageMarshaller() {}

    /** (kind, precision, scale) of a Trino type in the native type system; NOT_SUPPORTED outside the GPU subset. */
    static int[] nativeType(Type type)
    {
        if (type == BooleanType.BOOLEAN) {
            return new int[] {BOOLEAN, 0, 0};
        }
        if (type == BigintType.BIGINT) {
            return new int[] {BIGINT, 0, 0};
        }
        if (type == IntegerType.INTEGER) {
            return new int[] {INTEGER, 0, 0};
        }
        if (type == DateType.DATE) {
            return new int[] {DATE, 0, 0};
        }
        if (type == DoubleType.DOUBLE) {
            return new int[] {DOUBLE, 0, 0};
        }
        if (type instanceof DecimalType) {
            DecimalType decimal = (DecimalType) type;
            return new int[] {DECIMAL, decimal.getPrecision(), decimal.getScale()};
        }
        if (type instanceof VarcharType) {
            VarcharType varchar = (VarcharType) type;
            return new int[] {VARCHAR, varchar.isUnbounded() ? 0 : varchar.getBoundedLength(), 0};
        }
        if (type == VarbinaryType.VARBINARY) {
            return new int[] {VARBINARY, 0, 0};
        }
        throw new TrinoException(NOT_SUPPORTED, "not supported on the GPU path: type " + type);
    }

    static void addPage(long handle, Page page, List<Type> inputTypes)
    {
        int channels = page.getChannelCount();
        int[] encoding = new int[channels];
        int[] kind = new int[channels];
        int[] precision = new int[channels];
        int[] scale = new int[channels];
        Object[] values = new Object[channels];
        int[] valuesOffset = new int[channels];
        boolean[][] valueIsNull = new boolean[channels][];
        int[][] offsets = new int[channels][];
        int[][] ids = new int[channels][];
        int[] idsOffset = new int[channels];
        Object[] dictionaryValues = new Object[channels];
        boolean[][] dictionaryIsNull = new boolean[channels][];
        int[][] dictionaryOffsets = new int[channels][];
        int[] dictionaryPositionCount = new int[channels];

        for (int channel = 0; channel < channels; channel++) {
            int[] nativeType = nativeType(inputTypes.get(channel));
            kind[channel] = nativeType[0];
            precision[channel] = nativeType[1];
            scale[channel] = nativeType[2];
            Block block = GpuBlockAccess.loaded(page.getBlock(channel));
            if (block instanceof DictionaryBlock) {
                DictionaryBlock dictionaryBlock = (DictionaryBlock) block;
                Block dictionary = GpuBlockAccess.loaded(dictionaryBlock.getDictionary());
                if (dictionary instanceof DictionaryBlock || dictionary instanceof RunLengthEncodedBlock) {
                    block = flatten(block);      // nested encodings: materialise (rare: the readers produce one level)
                }
                else {
                    encoding[channel] = ENC_DICTIONARY;
                    ids[channel] = GpuBlockAccess.dictionaryIds(dictionaryBlock);
                    idsOffset[channel] = GpuBlockAccess.dictionaryIdsOffset(dictionaryBlock);
                    fillFlat(dictionary, channel, dictionaryValues, dictionaryIsNull, dictionaryOffsets, null);
                    dictionaryPositionCount[channel] = dictionary.getPositionCount();
                    continue;
                }
            }
            if (block instanceof RunLengthEncodedBlock) {
                Block value = GpuBlockAccess.loaded(((RunLengthEncodedBlock) block).getValue());
                if (value instanceof DictionaryBlock || value instanceof RunLengthEncodedBlock) {
                    block = flatten(block);
                }
                else {
                    encoding[channel] = ENC_RLE;
                    fillFlat(value, channel, dictionaryValues, dictionaryIsNull, dictionaryOffsets, null);
                    dictionaryPositionCount[channel] = 1;
                    continue;
                }
            }
            encoding[channel] = ENC_PLAIN;
            fillFlat(block, channel, values, valueIsNull, offsets, valuesOffset);
        }
        PrestoGpuNative.addPage(handle, page.getPositionCount(), encoding, kind, precision, scale, values, valuesOffset, valueIsNull,
                offsets, ids, idsOffset, dictionaryValues, dictionaryIsNull, dictionaryOffsets, dictionaryPositionCount);
    }

    /**
     * Backing arrays of a flat block. Fixed-width blocks: the array + the element offset of position 0 (valuesOffset; null
     * flags are gathered from position 0, so they carry no offset of their own — the shim applies valuesOffset to both, hence
     * the flags array is padded in front). Var-width blocks: the payload byte[] and ABSOLUTE offsets into it.
     */
    private static void fillFlat(Block block, int channel, Object[] values, boolean[][] isNull, int[][] offsets, int[] valuesOffset)
    {
        values[channel] = GpuBlockAccess.rawValues(block);
        boolean[] flags = GpuBlockAccess.nullFlags(block);
        if (block instanceof VariableWidthBlock) {
            offsets[channel] = GpuBlockAccess.variableWidthOffsets((VariableWidthBlock) block);
            isNull[channel] = flags;
            return;
        }
        int elementOffset = GpuBlockAccess.rawOffset(block);
        if (valuesOffset == null) {
            // dictionary / RLE value blocks have no offset slot in the native call: hand over a trimmed copy when they are views
            if (elementOffset != 0) {
                values[channel] = copyRange(values[channel], elementOffset, block);
            }
            isNull[channel] = flags;
            return;
        }
        // Int128ArrayBlock: rawOffset is in longs (2 per position); the native offset is in POSITIONS
        int positionOffset = block instanceof Int128ArrayBlock ? elementOffset / 2 : elementOffset;
        valuesOffset[channel] = positionOffset;
        if (flags != null && positionOffset != 0) {
            boolean[] padded = new boolean[positionOffset + flags.length];
            System.arraycopy(flags, 0, padded, positionOffset, flags.length);
            flags = padded;
        }
        isNull[channel] = flags;
    }

    private static Object copyRange(Object array, int elementOffset, Block block)
    {
        int n = block.getPositionCount();
        if (array instanceof long[]) {
            int words = block instanceof Int128ArrayBlock ? 2 * n : n;
            long[] out = new long[words];
            System.arraycopy(array, elementOffset, out, 0, words);
            return out;
        }
        if (array instanceof int[]) {
            int[] out = new int[n];
            System.arraycopy(array, elementOffset, out, 0, n);
            return out;
        }
        byte[] out = new byte[n];
        System.arraycopy(array, elementOffset, out, 0, n);
        return out;
    }

    /** Materialises a (nested) dictionary / RLE block into a flat block of the same type. */
    private static Block flatten(Block block)
    {
        int[] positions = new int[block.getPositionCount()];
        for (int i = 0; i < positions.length; i++) {
            positions[i] = i;
        }
        Block copy = block.copyPositions(positions, 0, positions.length);
        while (copy instanceof DictionaryBlock) {
            copy = ((DictionaryBlock) copy).getDictionary().copyPositions(GpuBlockAccess.dictionaryIds((DictionaryBlock) copy),
                    GpuBlockAccess.dictionaryIdsOffset((DictionaryBlock) copy), copy.getPositionCount());
        }
        return copy;
    }

    /** raw = PrestoGpuNative.getOutput(handle): [positionCount, then (values, valueIsNull, offsets) per channel]. */
    static Block[] toBlocks(Object[] raw, List<Type> outputTypes)
    {
        int positionCount = (Integer) raw[0];
        Block[] blocks = new Block[outputTypes.size()];
        for (int channel = 0; channel < blocks.length; channel++) {
            Object values = raw[1 + 3 * channel];
            Optional<boolean[]> isNull = Optional.ofNullable((boolean[]) raw[2 + 3 * channel]);
            int[] offsets = (int[]) raw[3 + 3 * channel];
            int kind = nativeType(outputTypes.get(channel))[0];
            int precision = nativeType(outputTypes.get(channel))[1];
            if (kind == VARCHAR || kind == VARBINARY) {
                blocks[channel] = new VariableWidthBlock(positionCount, Slices.wrappedBuffer((byte[]) values), offsets, isNull);
            }
            else if (kind == BOOLEAN) {
                blocks[channel] = new ByteArrayBlock(positionCount, isNull, (byte[]) values);
            }
            else if (kind == INTEGER || kind == DATE) {
                blocks[channel] = new IntArrayBlock(positionCount, isNull, (int[]) values);
            }
            else if (kind == DECIMAL && precision > 18) {
                blocks[channel] = new Int128ArrayBlock(positionCount, isNull, (long[]) values);
            }
            else {
                blocks[channel] = new LongArrayBlock(positionCount, isNull, (long[]) values);
            }
        }
        return blocks;
    }
}
