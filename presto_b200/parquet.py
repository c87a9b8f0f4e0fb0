"""Row groups of a Parquet file as the column chunks pgpu_add_parquet_row_group takes.

The footer (FileMetaData: schema, row groups, chunk offsets) is read with pyarrow — the reference reads it with
MetadataReader.readFooter (lib/trino-parquet/src/main/java/io/trino/parquet/reader/MetadataReader.java:95-140) and a Trino
integration would hand the same numbers over from its own ParquetMetadata. The column chunk itself is sliced out of the file
untouched, [starting position, + total size), exactly as ParquetReader.readColumnChunk does (.../reader/ParquetReader.java:
244-268); nothing in it is decoded here."""
import io
from typing import Dict, Iterator, List, Tuple, Union

PHYSICAL = {"BOOLEAN": 0, "INT32": 1, "INT64": 2, "INT96": 3, "FLOAT": 4, "DOUBLE": 5, "BYTE_ARRAY": 6, "FIXED_LEN_BYTE_ARRAY": 7}  # parquet.thrift Type
# parquet.thrift CompressionCodec, keyed by pyarrow's names: Arrow calls the raw LZ4 block codec (thrift LZ4_RAW = 7) "LZ4" and the
# deprecated Hadoop-framed one (thrift LZ4 = 5) "LZ4_HADOOP"
CODEC = {"UNCOMPRESSED": 0, "SNAPPY": 1, "GZIP": 2, "LZO": 3, "BROTLI": 4, "LZ4_HADOOP": 5, "ZSTD": 6, "LZ4": 7, "LZ4_RAW": 7}
# (chunk bytes, physical type, max definition level, num_values, type_length of FIXED_LEN_BYTE_ARRAY[, compression codec if not UNCOMPRESSED])
Chunk = Tuple


def row_groups(source: Union[str, bytes]) -> Iterator[Tuple[int, Dict[str, Chunk]]]:
    """-> (row count, {column path: chunk}) per row group of the file (a path or the file's bytes)"""
    import pyarrow.parquet as pq
    data = open(source, "rb").read() if isinstance(source, str) else bytes(source)
    md = pq.ParquetFile(io.BytesIO(data)).metadata
    for g in range(md.num_row_groups):
        rg = md.row_group(g)
        cols: Dict[str, Chunk] = {}
        for j in range(rg.num_columns):
            c = rg.column(j)
            start = c.data_page_offset
            if c.has_dictionary_page and c.dictionary_page_offset:
                start = min(start, c.dictionary_page_offset)
            chunk = (data[start:start + c.total_compressed_size], PHYSICAL[c.physical_type], md.schema.column(j).max_definition_level, c.num_values,
                     md.schema.column(j).length if c.physical_type == "FIXED_LEN_BYTE_ARRAY" else 0)
            codec = CODEC.get(c.compression, 99)
            cols[c.path_in_schema] = chunk + (codec,) if codec else chunk   # the codec is only spelled out for compressed chunks
        yield rg.num_rows, cols


def add_file(operator, source: Union[str, bytes], columns: List[str]) -> int:
    """addInput of every row group of a file (columns = column paths in the plan's channel order); -> rows added.
    The caller drains getOutput() between calls if the plan can fill up (partial aggregation limits)."""
    total = 0
    for rows, cols in row_groups(source):
        operator.addParquetRowGroup([cols[c] for c in columns], rows)
        total += rows
    return total
