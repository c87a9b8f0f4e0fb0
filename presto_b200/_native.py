"""ctypes binding of include/prestogpu.h. There is no fallback: a missing library is a hard error."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libprestogpu.so")
BENCH_LIB_PATH = os.path.join(_HERE, "libprestogpu_bench.so")   # include/prestogpu_bench.h: generators, copies, L2 flush (bench / tests)

# pgpu_status
OK, NOT_SUPPORTED, INVALID_ARGUMENT, NUMERIC_VALUE_OUT_OF_RANGE, DIVISION_BY_ZERO = 0, 1, 2, 3, 4
INSUFFICIENT_RESOURCES, COMPILER_ERROR, CUDA_ERROR, INVALID_STATE, INVALID_CAST_ARGUMENT = 5, 6, 7, 8, 9
STATUS_NAMES = {0: "OK", 1: "NOT_SUPPORTED", 2: "INVALID_ARGUMENT", 3: "NUMERIC_VALUE_OUT_OF_RANGE", 4: "DIVISION_BY_ZERO",
                5: "GENERIC_INSUFFICIENT_RESOURCES", 6: "COMPILER_ERROR", 7: "CUDA_ERROR", 8: "INVALID_STATE",
                9: "INVALID_CAST_ARGUMENT"}
ENC_PLAIN, ENC_DICTIONARY, ENC_RLE, ENC_VARCELL = 0, 1, 2, 3
PAGE_PINNED = 1

# PGPU_LI_* column ids
LI = dict(orderkey=0, quantity=1, extendedprice=2, discount=3, tax=4, shipdate=5, returnflag=6, linestatus=7,
          iota_offsets=8, discount_id=9, tax_id=10, quantity_id=11, returnflag_id=12, linestatus_id=13,
          orderkey_shuffled=14)


class Column(ctypes.Structure):
    pass


Column._fields_ = [
    ("encoding", ctypes.c_int32), ("type_kind", ctypes.c_int32), ("precision", ctypes.c_int32), ("scale", ctypes.c_int32),
    ("position_count", ctypes.c_int64), ("values", ctypes.c_void_p), ("nulls", ctypes.c_void_p),
    ("offsets", ctypes.c_void_p), ("ids", ctypes.c_void_p), ("dictionary", ctypes.POINTER(Column)),
]


class ParquetChunk(ctypes.Structure):
    """pgpu_parquet_chunk"""
    _fields_ = [("data", ctypes.c_char_p), ("len", ctypes.c_size_t), ("physical_type", ctypes.c_int32), ("max_definition_level", ctypes.c_int32),
                ("num_values", ctypes.c_int64), ("type_length", ctypes.c_int32), ("codec", ctypes.c_int32)]


class Stats(ctypes.Structure):
    _fields_ = [("kernel_launches", ctypes.c_int64), ("h2d_bytes", ctypes.c_int64), ("d2h_bytes", ctypes.c_int64),
                ("input_rows", ctypes.c_int64), ("input_pages", ctypes.c_int64), ("jit_compiles", ctypes.c_int64),
                ("jit_cache_hits", ctypes.c_int64), ("device_ms", ctypes.c_double), ("table_capacity", ctypes.c_int64),
                ("table_rehashes", ctypes.c_int64), ("super_pages", ctypes.c_int64), ("table_bucketed", ctypes.c_int64), ("table_slots_claimed", ctypes.c_int64), ("bucket_offset", ctypes.c_int64), ("pages_merged", ctypes.c_int64), ("partition_sweeps", ctypes.c_int64), ("streamed_runs", ctypes.c_int64)]


# every symbol include/prestogpu.h declares: (name, restype, argtypes)
_P = ctypes.c_void_p
_COLS = ctypes.POINTER(Column)
SYMBOLS = [
    ("pgpu_version", ctypes.c_char_p, []),
    ("pgpu_device_count", ctypes.c_int32, []),
    ("pgpu_last_global_error", ctypes.c_char_p, []),
    ("pgpu_pipeline_create", ctypes.c_int32, [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int32, ctypes.POINTER(_P)]),
    ("pgpu_compile_signature", ctypes.c_int32, [_P, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_uint8), ctypes.c_int32]),
    ("pgpu_pipeline_destroy", None, [_P]),
    ("pgpu_last_error", ctypes.c_char_p, [_P]),
    ("pgpu_needs_input", ctypes.c_int32, [_P]),
    ("pgpu_add_page", ctypes.c_int32, [_P, _COLS, ctypes.c_int32, ctypes.c_int64, ctypes.c_uint32]),
    ("pgpu_reserve_input", ctypes.c_int32, [_P, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_uint8), ctypes.c_int32, ctypes.c_int64]),
    ("pgpu_is_blocked", ctypes.c_int32, [_P]),
    ("pgpu_on_unblocked", ctypes.c_int32, [_P, ctypes.c_void_p, ctypes.c_void_p]),
    ("pgpu_add_device_page", ctypes.c_int32, [_P, _COLS, ctypes.c_int32, ctypes.c_int64]),
    ("pgpu_finish", ctypes.c_int32, [_P]),
    ("pgpu_is_finished", ctypes.c_int32, [_P]),
    ("pgpu_get_output", ctypes.c_int32, [_P, _COLS, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_get_output_device", ctypes.c_int32, [_P, _COLS, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_sync", ctypes.c_int32, [_P]),
    ("pgpu_reset", ctypes.c_int32, [_P]),
    ("pgpu_output_channel_count", ctypes.c_int32, [_P]),
    ("pgpu_output_channel_type", ctypes.c_int32, [_P, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    ("pgpu_generated_source", ctypes.c_char_p, [_P]),
    ("pgpu_get_stats", ctypes.c_int32, [_P, ctypes.POINTER(Stats)]),
    ("pgpu_set_timing", ctypes.c_int32, [_P, ctypes.c_int32]),
    ("pgpu_stream", ctypes.c_void_p, [_P]),
    ("pgpu_partition_output_device", ctypes.c_int32, [_P, ctypes.c_int32, _COLS, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_add_serialized_page", ctypes.c_int32, [_P, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int64]),
    ("pgpu_add_serialized_page_marked", ctypes.c_int32, [_P, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int64, ctypes.c_int32, ctypes.c_int64]),
    ("pgpu_get_serialized_output", ctypes.c_int32, [_P, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_add_parquet_row_group", ctypes.c_int32, [_P, ctypes.POINTER(ParquetChunk), ctypes.c_int32, ctypes.c_int64]),
    ("pgpu_parquet_inspect_chunk", ctypes.c_int32, [ctypes.POINTER(ParquetChunk), ctypes.POINTER(ctypes.c_int64), ctypes.c_char_p, ctypes.c_size_t]),
    ("pgpu_parquet_chunk_runs", ctypes.c_int32, [ctypes.POINTER(ParquetChunk), ctypes.c_int32, ctypes.c_void_p, ctypes.c_int64, ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_comm_init", ctypes.c_int32, [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_char_p, ctypes.POINTER(_P)]),
    ("pgpu_comm_destroy", None, [_P]),
    ("pgpu_comm_last_error", ctypes.c_char_p, [_P]),
    ("pgpu_comm_barrier", ctypes.c_int32, [_P]),
    ("pgpu_exchange_partial", ctypes.c_int32, [_P, _P, _P, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_comm_private_groups", ctypes.c_int64, [_P]),
    ("pgpu_ipc_export", ctypes.c_int32, [ctypes.c_int32, ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_ipc_open", ctypes.c_int32, [ctypes.c_int32, ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]),
    ("pgpu_ipc_close", ctypes.c_int32, [ctypes.c_int32, ctypes.c_void_p]),
    ("pgpu_partition_of", ctypes.c_int32, [ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_uint8), ctypes.POINTER(ctypes.c_int32), ctypes.c_int32, ctypes.c_int32]),
    ("pgpu_expand_nulls_static", ctypes.c_int32, [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]),
    ("pgpu_host_alloc", ctypes.c_void_p, [ctypes.c_size_t]),
    ("pgpu_host_free", None, [ctypes.c_void_p]),
    ("pgpu_device_alloc", ctypes.c_void_p, [ctypes.c_int32, ctypes.c_size_t]),
    ("pgpu_device_free", None, [ctypes.c_int32, ctypes.c_void_p]),
]

# every symbol include/prestogpu_bench.h declares (libprestogpu_bench.so)
BENCH_SYMBOLS = [
    ("pgpu_bench_host_copy", ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]),
    ("pgpu_tpch_order_count", ctypes.c_int64, [ctypes.c_double]),
    ("pgpu_tpch_lineitem_rows", ctypes.c_int32, [ctypes.c_double, ctypes.c_int64, ctypes.c_int64, ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_tpch_lineitem_device", ctypes.c_int32, [ctypes.c_int32, ctypes.c_double, ctypes.c_int64, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_tpch_lineitem_host", ctypes.c_int32, [ctypes.c_double, ctypes.c_int64, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]),
    ("pgpu_memcpy_h2d", ctypes.c_int32, [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]),
    ("pgpu_memcpy_d2h", ctypes.c_int32, [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]),
    ("pgpu_gen_lineitem_device", ctypes.c_int32, [ctypes.c_int32, ctypes.c_int32, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]),
    ("pgpu_gen_lineitem_host", ctypes.c_int32, [ctypes.c_int32, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]),
    ("pgpu_flush_l2", ctypes.c_int32, [ctypes.c_int32, ctypes.c_void_p, ctypes.c_size_t]),
]

_lib = None


class _Lib:
    """The product library, with the bench / test plumbing of libprestogpu_bench.so reachable under the same handle (loaded on
    first use of one of its symbols, so that product code paths never map it)."""

    def __init__(self, handle):
        self._handle = handle
        self._bench = None

    def __getattr__(self, name):
        h = self.__dict__["_handle"]
        if any(name == b[0] for b in BENCH_SYMBOLS):
            if self.__dict__["_bench"] is None:
                self.__dict__["_bench"] = bench_lib()
            return getattr(self.__dict__["_bench"], name)
        return getattr(h, name)


_bench = None


def bench_lib():
    """libprestogpu_bench.so alone (generators, synchronous copies, L2 flush): what the CPU reference arm of bench.py uses to make
    its input, so that it never maps the product library."""
    global _bench
    if _bench is None:
        if not os.path.exists(BENCH_LIB_PATH):
            raise RuntimeError("presto_b200: %s is missing. Build it with `python -m presto_b200.build`." % BENCH_LIB_PATH)
        b = ctypes.CDLL(BENCH_LIB_PATH)
        for n, res, args in BENCH_SYMBOLS:
            fn = getattr(b, n)
            fn.restype = res
            fn.argtypes = args
        _bench = b
    return _bench


def lib():
    """Load libprestogpu.so (built in-tree by presto_b200.build). Raises if it is missing: no fallback exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "presto_b200: native library %s is missing. Build it with `python -m presto_b200.build` "
                "(or __graft_entry__.build()); there is no CPU fallback." % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = _Lib(handle)
    return _lib
