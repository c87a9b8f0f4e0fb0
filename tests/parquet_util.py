"""Test helpers for Parquet column chunks: the chunk bytes and their metadata out of a file (footer read by pyarrow — the
reference reads it with MetadataReader.readFooter, lib/trino-parquet/.../reader/MetadataReader.java — the column chunk
itself is sliced out of the file untouched, as ParquetReader.readColumnChunk does, .../reader/ParquetReader.java:244-268)."""
import os

import numpy as np
import pyarrow.parquet as pq

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PHYSICAL = {"BOOLEAN": 0, "INT32": 1, "INT64": 2, "INT96": 3, "FLOAT": 4, "DOUBLE": 5, "BYTE_ARRAY": 6, "FIXED_LEN_BYTE_ARRAY": 7}


def chunks_of(path):
    """-> [(row count, {column name: (chunk bytes, physical type, max definition level, num_values)})] per row group"""
    data = open(path, "rb").read() if isinstance(path, str) else path
    import io
    md = pq.ParquetFile(io.BytesIO(data)).metadata
    out = []
    for g in range(md.num_row_groups):
        rg = md.row_group(g)
        cols = {}
        for j in range(rg.num_columns):
            c = rg.column(j)
            start = c.dictionary_page_offset if c.has_dictionary_page and c.dictionary_page_offset else c.data_page_offset
            start = min(start, c.data_page_offset)
            sc = md.schema.column(j)
            cols[c.path_in_schema] = (data[start:start + c.total_compressed_size], PHYSICAL[c.physical_type], sc.max_definition_level, c.num_values)
        out.append((rg.num_rows, cols))
    return out


def expected_of(name):
    z = np.load(os.path.join(GOLDEN, name + ".expected.npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}
