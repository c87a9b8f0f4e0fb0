/*
 * prestogpu.h — the drop-in boundary: a plain C ABI for the B200-native
 * scan -> filter -> project -> (hash-)aggregate pipeline.
 *
 * Nothing like this exists in the reference (Trino 359 is 100 % Java); every entry point
 * below names the reference interface it stands in for (paths under /root/reference/):
 *   main/ = core/trino-main/src/main/java/io/trino/
 *   spi/  = core/trino-spi/src/main/java/io/trino/spi/
 *
 * A `pgpu_pipeline` is ONE operator instance (single owner, called by one thread at a time,
 * possibly a different thread per call — exactly the Driver contract, main/operator/Driver.java:60-62,
 * 284-305).  It fuses what the reference runs as
 *     FilterAndProjectOperator / ScanFilterAndProjectOperator  (main/operator/FilterAndProjectOperator.java:42-88)
 *  -> AggregationOperator                                      (main/operator/AggregationOperator.java:93-186)
 *   | HashAggregationOperator                                  (main/operator/HashAggregationOperator.java:367-518)
 * Any of the stages may be absent from the plan, so each reference operator is also available alone.
 *
 * There is no CPU fallback: a plan outside the supported subset fails in pgpu_pipeline_create
 * with PGPU_NOT_SUPPORTED (the Java shim maps it to TrinoException(NOT_SUPPORTED)).
 */
#ifndef PRESTOGPU_H
#define PRESTOGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Status codes. The non-zero ones map 1:1 onto io.trino.spi.StandardErrorCode values raised on this
 * path (SURVEY.md §8(b) "Errors"). */
typedef enum pgpu_status {
    PGPU_OK = 0,
    PGPU_NOT_SUPPORTED = 1,               /* NOT_SUPPORTED: plan-time rejection                         */
    PGPU_INVALID_ARGUMENT = 2,            /* GENERIC_INTERNAL_ERROR: malformed plan / page              */
    PGPU_NUMERIC_VALUE_OUT_OF_RANGE = 3,  /* main/type/BigintOperators.java:60-110, Decimal overflow    */
    PGPU_DIVISION_BY_ZERO = 4,            /* main/type/BigintOperators.java:86-110                      */
    PGPU_INSUFFICIENT_RESOURCES = 5,      /* main/operator/BigintGroupByHash.java:265-267               */
    PGPU_COMPILER_ERROR = 6,              /* main/sql/gen/ExpressionCompiler.java:137                   */
    PGPU_CUDA_ERROR = 7,                  /* GENERIC_INTERNAL_ERROR: device failure                     */
    PGPU_INVALID_STATE = 8,               /* checkState(...) failures, e.g. addInput after finish       */
    PGPU_INVALID_CAST_ARGUMENT = 9        /* main/type/DoubleOperators.java:155-165                     */
} pgpu_status;

/* Block encodings (spi/block/...): flat arrays, DictionaryBlock, RunLengthEncodedBlock. */
typedef enum pgpu_encoding {
    PGPU_ENC_PLAIN = 0,       /* LongArrayBlock / IntArrayBlock / ByteArrayBlock / Int128ArrayBlock / VariableWidthBlock */
    PGPU_ENC_DICTIONARY = 1,  /* spi/block/DictionaryBlock.java:40-48 : ids[] into `dictionary`                          */
    PGPU_ENC_RLE = 2,         /* spi/block/RunLengthEncodedBlock.java:45-48 : single-position `dictionary` repeated       */
    PGPU_ENC_VARCELL = 3      /* device-resident VARCHAR(<=15) column as fixed 16-byte cells (byte 0 = length, bytes 1..15 =
                                 data): what pgpu_get_output_device returns for VARCHAR channels and pgpu_add_device_page
                                 accepts — the device-to-device form of a short VariableWidthBlock (no reference counterpart) */
} pgpu_encoding;

/* Logical types of the supported subset (spi/type/...). Physical layout in a block:
 *   BOOLEAN  -> ByteArrayBlock   (int8 per position)
 *   INTEGER, DATE -> IntArrayBlock (int32)
 *   BIGINT, DOUBLE (raw IEEE bits, spi/type/DoubleType.java:98-106), DECIMAL(p<=18) -> LongArrayBlock (int64)
 *   DECIMAL(p>18) -> Int128ArrayBlock: 2 x int64 per position, sign-magnitude, low word first,
 *                    sign = top bit of word 1 (spi/type/UnscaledDecimal128Arithmetic.java:30-56)
 *   VARCHAR, VARBINARY -> VariableWidthBlock: bytes + int32 offsets[n+1]                              */
typedef enum pgpu_type_kind {
    PGPU_BOOLEAN = 0,
    PGPU_BIGINT = 1,
    PGPU_INTEGER = 2,
    PGPU_DATE = 3,
    PGPU_DOUBLE = 4,
    PGPU_DECIMAL = 5,
    PGPU_VARCHAR = 6,
    PGPU_VARBINARY = 7
} pgpu_type_kind;

/* One Block (spi/block/Block.java). Pointers are HOST pointers for pgpu_add_page / pgpu_get_output
 * and DEVICE pointers for pgpu_add_device_page / pgpu_get_output_device. */
typedef struct pgpu_column {
    int32_t encoding;        /* pgpu_encoding                                                           */
    int32_t type_kind;       /* pgpu_type_kind                                                          */
    int32_t precision;       /* DECIMAL precision; VARCHAR bound (0 = unbounded); else 0                */
    int32_t scale;           /* DECIMAL scale; else 0                                                   */
    int64_t position_count;  /* positions in THIS block                                                 */
    const void* values;      /* PLAIN: element array (see layout above); var-width: the byte slice      */
    const uint8_t* nulls;    /* PLAIN: Java boolean[] valueIsNull, 1 byte/position, or NULL = no nulls  */
    const int32_t* offsets;  /* PLAIN var-width only: position_count+1 offsets into `values`            */
    const int32_t* ids;      /* DICTIONARY: position_count ids into `dictionary`                        */
    const struct pgpu_column* dictionary; /* DICTIONARY: the dictionary block (PLAIN); RLE: 1-position value block (PLAIN) */
} pgpu_column;

typedef struct pgpu_pipeline pgpu_pipeline;

/* Flags for pgpu_add_page. */
#define PGPU_PAGE_PINNED 1u /* caller guarantees all buffers are page-locked and stay unchanged until
                               pgpu_sync(); the library then skips its staging copy and DMAs directly. */

/* Library / environment ------------------------------------------------------------------------- */
const char* pgpu_version(void);
/* Number of visible CUDA devices (0 when there is no GPU; never an error). */
int32_t pgpu_device_count(void);
/* Last error text of calls that have no pipeline handle (create failures). Thread-local. */
const char* pgpu_last_global_error(void);

/* Plan compilation — replaces ExpressionCompiler.compilePageProcessor (main/sql/gen/ExpressionCompiler.java:
 * 98-137) + AccumulatorCompiler (main/operator/aggregation/AccumulatorCompiler.java:80-140) + the factory
 * constructors LocalExecutionPlanner calls (main/sql/planner/LocalExecutionPlanner.java:1680,1699,3389,3492).
 * `plan` is the textual plan described in DESIGN.md §"Plan language". `device` < 0 compiles only
 * (no GPU needed: used by the CPU test-suite and by build() to warm the kernel cache). */
pgpu_status pgpu_pipeline_create(const uint8_t* plan, size_t plan_len, int32_t device, pgpu_pipeline** out);
/* Pre-compile the kernel variant for one page signature (per channel: pgpu_encoding, and whether the block /
 * its dictionary carries a null vector). Variants are otherwise compiled lazily by the first page that needs
 * them; this entry point lets plan-time code (and the CPU-only test-suite) pay for it up front. */
pgpu_status pgpu_compile_signature(pgpu_pipeline* p, const int32_t* encodings, const uint8_t* nullable, int32_t ncols);
/* Operator.close() (main/operator/Operator.java:97-102). */
void pgpu_pipeline_destroy(pgpu_pipeline* p);
/* Error text for the last failing call on `p` (TrinoException message). */
const char* pgpu_last_error(const pgpu_pipeline* p);

/* Operator protocol (main/operator/Operator.java:21-103) ------------------------------------------- */
/* Operator.needsInput(): 1 / 0. HashAggregationOperator.needsInput (main/operator/HashAggregationOperator.java:367-378). */
int32_t pgpu_needs_input(const pgpu_pipeline* p);
/* Operator.addInput(Page): columns are copied (or DMA'd) before the call returns control of unpinned buffers. */
pgpu_status pgpu_add_page(pgpu_pipeline* p, const pgpu_column* cols, int32_t ncols, int64_t rows, uint32_t flags);
/* Everything pgpu_add_page for a page of this signature and size could WAIT for, done ahead of it: the kernel variant of the
 * signature is compiled, and a staging slot with room for `rows` more positions is ready (the pending super-page is sent and
 * the next slot's previous kernel awaited when it is not). pgpu_add_page then only copies — which is what a JNI caller holding
 * GetPrimitiveArrayCritical needs: no compilation and no device wait inside the critical region (pages larger than a whole
 * staging slot excepted). Optional; pgpu_add_page does the same work itself when it was not called. */
pgpu_status pgpu_reserve_input(pgpu_pipeline* p, const int32_t* encodings, const uint8_t* nullable, int32_t ncols, int64_t rows);
/* Operator.isBlocked() (main/operator/Operator.java:24-35): 1 while device work queued by `p` is still running, i.e. while a
 * pgpu_get_output* / pgpu_sync call would have to wait; never an error. pgpu_on_unblocked registers `fn(user)` to be called
 * (from a CUDA host-callback thread; it must not call back into the library) once everything queued so far has completed —
 * the completion of the ListenableFuture the Java operator returns instead of blocking the Driver thread. */
typedef void (*pgpu_callback)(void* user);
int32_t pgpu_is_blocked(const pgpu_pipeline* p);
pgpu_status pgpu_on_unblocked(pgpu_pipeline* p, pgpu_callback fn, void* user);
/* Operator.addInput of a Page in the reference's wire format: `data` (host) is the uncompressed payload of a SerializedPage —
 * what PagesSerdeUtil.writeRawPage produces (main/execution/buffer/PagesSerdeUtil.java:45-62): int32 channel count, then per
 * block the length-prefixed encoding name and body (LONG_ARRAY / INT_ARRAY / BYTE_ARRAY / INT128_ARRAY / VARIABLE_WIDTH, and
 * DICTIONARY / RLE over those; spi/block/{Long,Int,...}ArrayBlockEncoding.java etc.). Only the block headers are read on the host: the payload goes
 * to the device in ONE copy and is decoded there (null bit stream -> flags, non-null-only values -> positions). Pages with codec
 * markers go through pgpu_add_serialized_page_marked. `data` may be reused when the call returns. */
pgpu_status pgpu_add_serialized_page(pgpu_pipeline* p, const uint8_t* data, size_t len, int64_t position_count);
/* The same for a SerializedPage as it travels (main/execution/buffer/SerializedPage.java): `markers` = its PageCodecMarker set
 * (PageCodecMarker.java:28-29: COMPRESSED = bit 0, ENCRYPTED = bit 1), `uncompressed_size` = getUncompressedSizeInBytes().
 * COMPRESSED slices are LZ4 blocks (PagesSerde.java:156-165, airlift Lz4Decompressor) and are inflated on the host, as
 * PagesSerde.deserialize does, before the payload goes to the device; ENCRYPTED pages are PGPU_NOT_SUPPORTED (the spill cipher
 * is the caller's). */
#define PGPU_PAGE_MARKER_COMPRESSED 1
#define PGPU_PAGE_MARKER_ENCRYPTED 2
pgpu_status pgpu_add_serialized_page_marked(pgpu_pipeline* p, const uint8_t* data, size_t len, int64_t position_count, int32_t markers,
                                            int64_t uncompressed_size);
/* Operator.getOutput() in the same wire format: *data / *len = the payload PagesSerdeUtil.writeRawPage would write for the
 * output page (library-owned pinned memory, valid until the next call on this pipeline), *rows = its position count, or
 * -1 when there is no output now. Null flags are bit-packed, nullable fixed-width blocks carry only their non-null values,
 * var-width blocks their end offsets and bytes; all of it is packed on the device. A PARTIAL step's payload is what the
 * reference's exchange would carry to the FINAL step (feed it to pgpu_add_serialized_page of a FINAL pipeline). */
pgpu_status pgpu_get_serialized_output(pgpu_pipeline* p, const uint8_t** data, size_t* len, int64_t* rows);
/* One column chunk of a Parquet row group, as it lies in the file: from the first page header of the chunk (dictionary page if
 * there is one) to its end — ColumnChunkMetaData.getStartingPos() / getTotalSize() as ParquetReader.readColumnChunk slices it
 * (lib/trino-parquet/src/main/java/io/trino/parquet/reader/ParquetReader.java:244-268). */
typedef struct pgpu_parquet_chunk {
    const uint8_t* data;          /* host pointer                                                            */
    size_t len;
    int32_t physical_type;        /* parquet.thrift Type: 0 BOOLEAN, 1 INT32, 2 INT64, 5 DOUBLE, 6 BYTE_ARRAY, 7 FIXED_LEN_BYTE_ARRAY */
    int32_t max_definition_level; /* 0 = required, 1 = optional flat column (nested columns are rejected)     */
    int64_t num_values;           /* ColumnMetaData.num_values = rows of the row group for a flat column     */
    int32_t type_length;          /* FIXED_LEN_BYTE_ARRAY: bytes per value (1..16, a DECIMAL's unscaled value); else 0 */
    int32_t codec;                /* parquet.thrift CompressionCodec of the chunk's pages: 0 UNCOMPRESSED, 1 SNAPPY, 2 GZIP, 5 LZ4 (Hadoop-framed), 6 ZSTD, 7 LZ4_RAW (others:
                                   * PGPU_NOT_SUPPORTED). Compressed pages are inflated on the host, as the reference does
                                   * (ParquetCompressionUtils.java:43-70), then take the same path                          */
} pgpu_parquet_chunk;
/* Operator.addInput of the Page that ParquetReader.nextPage / readBlock -> PrimitiveColumnReader.readPrimitive would build from
 * one row group (.../reader/ParquetReader.java:211-268, PrimitiveColumnReader.java:148-317): chunk c feeds input channel c
 * (INT32 -> INTEGER/DATE, or widened to BIGINT/short DECIMAL as IntColumnReader / ShortDecimalColumnReader.java:52-62 do; INT64 ->
 * BIGINT/short DECIMAL; DOUBLE; BOOLEAN; BYTE_ARRAY -> VARCHAR/VARBINARY with the bounded-VARCHAR truncation of
 * BinaryColumnReader.java:44-58; FIXED_LEN_BYTE_ARRAY(<= 8) -> short DECIMAL, (9..16) -> long DECIMAL: big-endian two's complement
 * unscaled value, ParquetTypeUtils.getShortDecimalValue :226-240, LongDecimalColumnReader.java:58. The file's decimal must have the
 * channel's scale: the rescaling casts of those readers are not built). Only page headers and the run headers of the RLE / bit-packed hybrid streams are
 * read on the host; each chunk goes to the device in ONE copy and definition levels, dictionary ids and booleans are expanded
 * there. A chunk whose data pages are all dictionary encoded is consumed as a DictionaryBlock (ids + dictionary page), never
 * materialised. Data pages v1 and v2; PLAIN, PLAIN_DICTIONARY, RLE_DICTIONARY, RLE (levels, booleans). Rejected with
 * PGPU_NOT_SUPPORTED: BROTLI / LZO pages, nested columns, DELTA_* / BYTE_STREAM_SPLIT encodings, INT96 / FLOAT,
 * var-width chunks that are not dictionary encoded throughout. */
pgpu_status pgpu_add_parquet_row_group(pgpu_pipeline* p, const pgpu_parquet_chunk* chunks, int32_t nchunks, int64_t rows);
/* What pgpu_add_parquet_row_group reads on the host, without a GPU: info[0] data pages, [1] dictionary entries (-1: none), [2] every
 * data page dictionary encoded, [3] non-null values, [4] definition-level runs kept (0: no NULLs), [5] value runs, [6] some page
 * dictionary encoded, [7] rows. Same status / message as the ingest would give for a chunk it rejects. */
pgpu_status pgpu_parquet_inspect_chunk(const pgpu_parquet_chunk* chunk, int64_t* info, char* err, size_t err_len);
/* The run tables the ingest hands to its device kernels, for checking them without a GPU: which = 0 definition levels, 1 values
 * (dictionary ids / booleans), 2 data pages {start = first non-null value, src = byte offset of the values, count = non-null
 * values, width = 1 if dictionary encoded}. Entry = {int64 start; uint64 src (bit offset of a bit-packed run | value of an RLE
 * run); uint32 count; uint32 bit width | packed << 31}. *n = entries available, at most `cap` are written to `out`. */
pgpu_status pgpu_parquet_chunk_runs(const pgpu_parquet_chunk* chunk, int32_t which, void* out, int64_t cap, int64_t* n);
/* Same, with device-resident blocks (no copy). Buffers must stay valid until pgpu_sync(). */
pgpu_status pgpu_add_device_page(pgpu_pipeline* p, const pgpu_column* cols, int32_t ncols, int64_t rows);
/* Operator.finish(): no more input; may be called repeatedly (main/operator/Driver.java:368-372). */
pgpu_status pgpu_finish(pgpu_pipeline* p);
/* Operator.isFinished(). */
int32_t pgpu_is_finished(const pgpu_pipeline* p);
/* Operator.getOutput(): *rows = -1 means "no page now" (Java null). Output blocks live in library-owned
 * pinned host memory and stay valid until the next pgpu_get_output / pgpu_pipeline_destroy on `p`.
 * Device-side arithmetic faults (overflow, division by zero) surface here or in pgpu_finish. */
pgpu_status pgpu_get_output(pgpu_pipeline* p, pgpu_column* out_cols, int32_t max_cols, int32_t* ncols, int64_t* rows);
/* As pgpu_get_output but the blocks stay in HBM (device pointers). */
pgpu_status pgpu_get_output_device(pgpu_pipeline* p, pgpu_column* out_cols, int32_t max_cols, int32_t* ncols, int64_t* rows);
/* Block until all device work queued by `p` is complete; reports sticky device-side errors. */
pgpu_status pgpu_sync(pgpu_pipeline* p);
/* Drop all aggregation state and return to the just-created state (new TaskContext per benchmark
 * iteration, testing/trino-benchmark/.../AbstractOperatorBenchmark.java:279-335). */
pgpu_status pgpu_reset(pgpu_pipeline* p);

/* Introspection --------------------------------------------------------------------------------- */
/* Output schema: number of output channels and their types. */
int32_t pgpu_output_channel_count(const pgpu_pipeline* p);
pgpu_status pgpu_output_channel_type(const pgpu_pipeline* p, int32_t channel, int32_t* type_kind, int32_t* precision, int32_t* scale);
/* Generated CUDA source of the most recently compiled kernel variant (for reviews / profiles). */
const char* pgpu_generated_source(const pgpu_pipeline* p);
/* Counters: kernels launched / bytes copied H2D and D2H / device-time (ms, CUDA events, only when
 * timing was enabled with pgpu_set_timing) since creation or the last pgpu_reset. */
typedef struct pgpu_stats {
    int64_t kernel_launches;
    int64_t h2d_bytes;
    int64_t d2h_bytes;
    int64_t input_rows;
    int64_t input_pages;
    int64_t jit_compiles;
    int64_t jit_cache_hits;
    double  device_ms;      /* sum over timed page kernels */
    int64_t table_capacity; /* slots of the HBM hash table (0 = none) */
    int64_t table_rehashes;
    int64_t super_pages;    /* launches over coalesced host pages (pgpu_add_page / pgpu_add_serialized_page appends) */
    int64_t table_bucketed; /* 1: the HBM table keeps 8 neighbouring integer keys per slot (clustered keys), 0: one key per slot */
    int64_t table_slots_claimed; /* slots in use at the last look (output time / table growth); buckets for a bucketed table */
    int64_t bucket_offset;  /* bucketed table: buckets are cut at 8n - bucket_offset (chosen by the key-clustering probe) */
    int64_t pages_merged;   /* filter + project with (merge-min-rows / merge-min-bytes): small output pages appended to the
                             * buffer instead of going out on their own (MergePages.java:147 appendPage) */
    int64_t partition_sweeps; /* long device-resident pages whose rows were reordered by table slice before the hash-aggregation
                               * kernel ran (tables far larger than L2; DESIGN.md "partition sweep") */
    int64_t streamed_runs;    /* rows of the run buffer: device pages whose keys arrive in key order are aggregated run by run,
                               * without a hash table (DESIGN.md "streaming runs"); 0 = the table path */
} pgpu_stats;
pgpu_status pgpu_get_stats(const pgpu_pipeline* p, pgpu_stats* out);
pgpu_status pgpu_set_timing(pgpu_pipeline* p, int32_t enabled);
/* The CUDA stream (cudaStream_t) the pipeline launches on, so callers can bracket it with events. */
void* pgpu_stream(const pgpu_pipeline* p);

/* Partial -> final exchange (multi-GPU) -----------------------------------------------------------
 * Mirrors PartitionedOutputOperator (main/operator/PartitionedOutputOperator.java:56-133) between a
 * Step.PARTIAL and a Step.FINAL HashAggregationOperator: pgpu_comm_init / pgpu_exchange_partial below run the whole exchange
 * (DESIGN.md "Multi-GPU"); the building blocks they use are exported too. */
/* Hash-partition the complete output of a finishing PARTIAL-step pipeline into `nparts` (<= 64) contiguous
 * row ranges, device resident. After the call out_cols describe ONE device page sorted by partition and
 * partition i occupies rows [part_row_offsets[i], part_row_offsets[i] + part_row_counts[i]) of every column (host arrays of
 * nparts+1 / nparts entries). Offsets are multiples of 16 rows, so each partition of each column is a 16-byte aligned
 * device block that pgpu_add_device_page accepts; part_row_offsets[nparts] is the padded end. All columns live in ONE
 * device allocation (see pgpu_ipc_export). The pipeline is finished afterwards.
 * The partition function is the reference's: (rawHash & 0x7fff_ffff_ffff_ffff) % nparts with
 * rawHash = 31 * previous + hash(key channel) (main/operator/HashGenerator.java:24-35,
 * InterpretedHashGenerator.java:66-74, spi/type/AbstractLongType.java:126-130). Key types: BIGINT, INTEGER,
 * DATE, DOUBLE; other key types are rejected with PGPU_NOT_SUPPORTED (VARCHAR keys hash with airlift's
 * XxHash64, which is not in the reference tree). */
pgpu_status pgpu_partition_output_device(pgpu_pipeline* p, int32_t nparts, pgpu_column* out_cols, int32_t max_cols,
                                         int32_t* ncols, int64_t* part_row_offsets, int64_t* part_row_counts);
/* The same partition function for ONE row on the host (kinds: 0 = int64, 1 = int32, 2 = double bits). */
int32_t pgpu_partition_of(const int64_t* key_bits, const uint8_t* is_null, const int32_t* kinds, int32_t nkeys, int32_t nparts);

/* The exchange itself, for one-process-per-GPU deployments on one box. A communicator is the rank's membership of a job:
 * ranks rendezvous through the POSIX shared-memory segment `name` (unique per job; created by rank 0, unlinked once everybody
 * has joined) and map each other's HBM through CUDA IPC over NVLink / NVSwitch. No collective library is involved.
 * pgpu_comm_init blocks until all `nranks` ranks have called it. device < 0 joins the rendezvous only (CPU tests). */
typedef struct pgpu_comm pgpu_comm;
pgpu_status pgpu_comm_init(int32_t nranks, int32_t rank, int32_t device, const char* name, pgpu_comm** out);
/* Waits (at most 30 s) until every peer has finished reading this rank's partition arena of the last exchange: destroy the
 * communicator BEFORE the PARTIAL pipeline whose output it exchanged (the arena is that pipeline's memory). */
void pgpu_comm_destroy(pgpu_comm* c);
const char* pgpu_comm_last_error(const pgpu_comm* c);
/* All ranks have arrived (host-side, through the segment). */
pgpu_status pgpu_comm_barrier(pgpu_comm* c);
/* PartitionedOutputOperator -> exchange -> HashAggregationOperator(Step.FINAL).addInput for one step, called by every rank:
 * `partial` is a PARTIAL (or INTERMEDIATE) step pipeline that has received pgpu_finish; `final_pipeline` is a FINAL-step
 * pipeline in its just-created / reset state whose input channels are the partial plan's output channels. On return the
 * partial pipeline is finished and the FINAL pipeline is finishing: drain it with pgpu_get_output*.
 * The shape is chosen from the partial plan alone (so that every rank takes the same one):
 *   gather — global aggregations and plans whose group count fits an inbox slot (<= 4096 rows; Q1, Q6): every rank's partial
 *            rows are pushed into every rank's inbox by a kernel and merged by the FINAL plan's gather kernel, with no host
 *            round trip; every rank ends with the complete result. More groups than a slot holds fail with
 *            PGPU_INSUFFICIENT_RESOURCES at pgpu_get_output (raise expected-groups to get the repartition shape).
 *   repartition — the partial rows are hash-partitioned with the reference's partition function (see
 *            pgpu_partition_output_device); each rank's FINAL page kernel reads its partition of every peer's rows straight
 *            out of that peer's HBM. The result stays partitioned by key.
 * *rows_received = rows this rank's FINAL step consumed (-1: known on the device only, gather shape); *bytes_sent = bytes that
 * crossed NVLink on this rank's behalf. Either may be NULL. */
pgpu_status pgpu_exchange_partial(pgpu_comm* c, pgpu_pipeline* partial, pgpu_pipeline* final_pipeline, int64_t* rows_received, int64_t* bytes_sent);
/* Range-aware split of the repartition shape (plans with ONE integer grouping key; PGPU_TUNE_RANGE_SPLIT=0 turns it off): every
 * rank publishes the [min, max] of its partial keys; a group whose key lies in no other rank's range exists on this rank alone, so it
 * skips the exchange and the second hash build and is finalised straight out of the partial table (by the same plan compiled as a
 * SINGLE step), appended behind the FINAL pipeline's own groups. With row-range shards of a table clustered on the key (l_orderkey)
 * only the boundary groups move; with keys spread over all ranks everything is exchanged as before. Every group still comes out
 * exactly once across the ranks, but the output is then not partitioned by key hash (no consumer may rely on an order:
 * InMemoryHashAggregationBuilder.java:309-312 / SURVEY.md 8(b)). Returns the groups the last exchange finalised in place. */
int64_t pgpu_comm_private_groups(const pgpu_comm* c);

/* Peer access over NVLink / NVSwitch for one-process-per-GPU deployments (cudaIpc*). The owner exports the allocation
 * that contains dev_ptr (`handle`, PGPU_IPC_HANDLE_BYTES bytes, plus dev_ptr's byte offset inside it); a peer process maps it
 * and may hand `mapped_base + offset (+ rows * width)` to pgpu_add_device_page of its own pipeline: the kernel then reads the
 * rows straight out of the owner's HBM, no staging copy (this is the exchange between PARTIAL and FINAL, standing in for
 * PartitionedOutputOperator -> ExchangeClient, main/operator/PartitionedOutputOperator.java:56-133). The owner must keep
 * the buffer unchanged until the peers are done (a barrier in the caller); a handle must not be opened by its exporter. */
#define PGPU_IPC_HANDLE_BYTES 64
pgpu_status pgpu_ipc_export(int32_t device, const void* dev_ptr, uint8_t* handle, int64_t* offset);
pgpu_status pgpu_ipc_open(int32_t device, const uint8_t* handle, void** mapped_base);
pgpu_status pgpu_ipc_close(int32_t device, void* mapped_base);

/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it). Decodes ONE nullable fixed-width block of the wire format
 * (what pgpu_add_serialized_page does per such block: spi/block/LongArrayBlockEncoding.java:45-60 + EncoderUtil.java:33-70) with
 * scratch memory that lives in a file-scope device array of the library instead of an allocation: `null_bits` = the (positions + 7) / 8
 * bytes of the null bit stream (most significant bit first, 1 = NULL), `compact_values` = the non-null values only, `width` = 1 / 4 /
 * 8 / 16 bytes per value -> values_out[positions] (zero at NULL positions), nulls_out[positions] (1 = NULL). All four are DEVICE
 * pointers supplied by the caller (compact_values and values_out aligned to `width`); `stream` is a cudaStream_t (NULL = the default
 * stream); the three kernels are queued on it and the call returns without waiting. Calls no cudaMalloc / cudaFree / allocating
 * library routine. Limits: positions <= PGPU_EXPAND_STATIC_MAX_POSITIONS (larger counts are rejected with PGPU_INVALID_ARGUMENT and a
 * text in pgpu_last_global_error()); every GPU that loads the library holds the 66 560-byte scratch array until exit; ALL calls on a
 * device share it, so calls that overlap on the GPU (different streams without an ordering between them) race — serialise them. */
#define PGPU_EXPAND_STATIC_MAX_POSITIONS 16777216 /* 2^24 positions = 16 384 blocks of 1024 */
pgpu_status pgpu_expand_nulls_static(int32_t device, void* stream, const uint8_t* null_bits, int64_t positions, const void* compact_values,
                                     int32_t width, void* values_out, uint8_t* nulls_out);

/* Pinned host memory helpers (what the JNI shim stages heap arrays into). */
void* pgpu_host_alloc(size_t bytes);
void pgpu_host_free(void* ptr);
/* Device memory helpers for callers without torch. */
void* pgpu_device_alloc(int32_t device, size_t bytes);
void pgpu_device_free(int32_t device, void* ptr);
#ifdef __cplusplus
}
#endif
#endif /* PRESTOGPU_H */
