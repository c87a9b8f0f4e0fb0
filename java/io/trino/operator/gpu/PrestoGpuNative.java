/*
 * JNI surface of libprestogpu (include/prestogpu.h). NOT COMPILED IN THIS REPOSITORY'S IMAGE (no JDK); this is the
 * binding a Trino 359 maintainer adds next to io.trino.operator. One long handle = one pgpu_pipeline = one Operator.
 */
package io.trino.operator.gpu;

final class PrestoGpuNative
{
    static {
        System.loadLibrary("prestogpu_jni"); // jni/prestogpu_jni.c, links libprestogpu.so
    }

    private PrestoGpuNative() {}

    /** pgpu_pipeline_create: plan text (DESIGN.md "Plan language"); throws TrinoException(NOT_SUPPORTED | COMPILER_ERROR ...). */
    static native long create(byte[] plan, int device);

    /** pgpu_pipeline_destroy — Operator.close(). */
    static native void destroy(long handle);

    /** pgpu_needs_input — Operator.needsInput(). */
    static native boolean needsInput(long handle);

    /**
     * pgpu_add_page — Operator.addInput(Page). One entry per channel:
     * encoding[c] (0 plain, 1 dictionary, 2 RLE), typeKind/precision/scale[c], and the block's backing arrays
     * (long[] / int[] / byte[] / boolean[]; null when absent). The shim pins each array with
     * GetPrimitiveArrayCritical, copies it into the pipeline's pinned staging ring and releases it before returning,
     * so Trino's immutable Blocks are never retained.
     */
    static native void addPage(long handle, int positionCount, int[] encoding, int[] typeKind, int[] precision, int[] scale,
            Object[] values, int[] valuesOffset, boolean[][] valueIsNull, int[][] offsets, int[][] ids, int[] idsOffset,
            Object[] dictionaryValues, boolean[][] dictionaryIsNull, int[][] dictionaryOffsets, int[] dictionaryPositionCount);

    /**
     * pgpu_add_serialized_page — Operator.addInput of a Page in the wire format of PagesSerdeUtil.writeRawPage: address / length of
     * the (uncompressed) payload Slice, off-heap or pinned by the shim for the duration of the call.
     */
    static native void addSerializedPage(long handle, long address, int length, int positionCount);

    /**
     * pgpu_add_serialized_page_marked — a SerializedPage as it arrives from an output buffer: its slice, position count,
     * PageCodecMarker byte (SerializedPage.getPageCodecMarkers()) and getUncompressedSizeInBytes(). LZ4-compressed slices are
     * inflated natively; encrypted pages raise NOT_SUPPORTED.
     */
    static native void addSerializedPageMarked(long handle, long address, int length, int positionCount, int markers, int uncompressedSize);

    /**
     * pgpu_add_parquet_row_group — Operator.addInput of the Page ParquetReader.nextPage would build from one row group. Per channel:
     * the column chunk's bytes as they lie in the file (ColumnChunkMetaData.getStartingPos() .. + getTotalSize(), still
     * compressed: `codec` = the chunk's parquet.thrift CompressionCodec, 0 UNCOMPRESSED, 1 SNAPPY, 2 GZIP, 5 LZ4, 6 ZSTD, 7 LZ4_RAW) as address / length of a direct buffer, the parquet physical type ordinal
     * (PrimitiveTypeName: BOOLEAN 0, INT32 1, INT64 2, DOUBLE 5, BINARY 6, FIXED_LEN_BYTE_ARRAY 7), the column's max definition
     * level (0 or 1), its value count, for FIXED_LEN_BYTE_ARRAY the type length, and the codec.
     */
    static native void addParquetRowGroup(long handle, long rows, long[] address, long[] length, int[] physicalType, int[] maxDefinitionLevel,
            long[] valueCount, int[] typeLength, int[] codec);

    /** pgpu_finish — Operator.finish(). */
    static native void finish(long handle);

    /** pgpu_is_finished — Operator.isFinished(). */
    static native boolean isFinished(long handle);

    /**
     * pgpu_get_output — Operator.getOutput(). Returns null when there is no page now; otherwise
     * [positionCount (Integer), then per channel: values array (long[] / int[] / byte[]; long[2n] for long decimals; the byte[]
     * payload of a var-width channel), boolean[] valueIsNull or null, int[] offsets (var-width) or null]. Arrays are fresh Java
     * arrays (the library-owned pinned buffers are copied out).
     */
    static native Object[] getOutput(long handle);

    /** pgpu_output_channel_count / pgpu_output_channel_type. */
    static native int[] outputTypes(long handle); // kind, precision, scale triples

    /** pgpu_last_error. */
    static native String lastError(long handle);

    /** pgpu_is_blocked — true while device work queued by this operator is still running (Operator.isBlocked() is not done). */
    static native boolean isBlocked(long handle);

    /**
     * pgpu_on_unblocked — runs {@code whenDone} (from a CUDA host-callback thread attached to the JVM for the call) once everything
     * queued so far has completed: the completion of the ListenableFuture the operator hands to the Driver.
     */
    static native void onUnblocked(long handle, Runnable whenDone);
}
