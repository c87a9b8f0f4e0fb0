"""ctypes wrapper of oracle/_build/liboracle.so — TEST INFRASTRUCTURE ONLY (the CPU restatement of the reference)."""
import ctypes
import os
import subprocess

from presto_b200 import _native as N
from presto_b200.operators import TrinoException
from presto_b200.spi import Page, block_from_column

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
_lib = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        h = ctypes.CDLL(LIB)
        P = ctypes.c_void_p
        C = ctypes.POINTER(N.Column)
        h.oracle_create.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(P), ctypes.c_char_p, ctypes.c_size_t]
        h.oracle_destroy.argtypes = [P]
        h.oracle_last_error.argtypes = [P]
        h.oracle_last_error.restype = ctypes.c_char_p
        h.oracle_needs_input.argtypes = [P]
        h.oracle_is_finished.argtypes = [P]
        h.oracle_finish.argtypes = [P]
        h.oracle_add_page.argtypes = [P, C, ctypes.c_int, ctypes.c_int64]
        h.oracle_get_output.argtypes = [P, C, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int64)]
        h.oracle_group_ids.argtypes = [ctypes.c_char_p, ctypes.c_size_t, C, ctypes.c_int, ctypes.c_int64, ctypes.c_int64,
                                       ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]
        U = ctypes.POINTER(ctypes.c_uint64)
        h.oracle_dec_add_with_overflow.argtypes = [U, U, U, ctypes.POINTER(ctypes.c_int64)]
        h.oracle_dec_binary.argtypes = [ctypes.c_int, U, U, U]
        h.oracle_dec_rescale.argtypes = [U, ctypes.c_int, U]
        h.oracle_dec_compare.argtypes = [U, U]
        h.oracle_dec_overflows.argtypes = [U]
        _lib = h
    return _lib


class OracleOperator:
    """Same protocol as presto_b200.operators.GpuPipelineOperator, executed by the CPU oracle."""

    def __init__(self, plan: str):
        self._lib = lib()
        h = ctypes.c_void_p()
        err = ctypes.create_string_buffer(4096)
        raw = plan.encode()
        st = self._lib.oracle_create(raw, len(raw), ctypes.byref(h), err, 4096)
        if st:
            raise TrinoException(st, err.value.decode())
        self._h = h

    def _check(self, st):
        if st:
            raise TrinoException(st, self._lib.oracle_last_error(self._h).decode())

    def needsInput(self):
        return bool(self._lib.oracle_needs_input(self._h))

    def addInput(self, page: Page):
        keep = []
        cols = (N.Column * len(page.blocks))()
        for i, b in enumerate(page.blocks):
            b._fill(cols[i], keep)
        self._check(self._lib.oracle_add_page(self._h, cols, len(page.blocks), page.position_count))

    def addRaw(self, cols, ncols, rows):
        self._check(self._lib.oracle_add_page(self._h, cols, ncols, rows))

    def getOutput(self):
        cols = (N.Column * 64)()
        nc = ctypes.c_int()
        rows = ctypes.c_int64()
        self._check(self._lib.oracle_get_output(self._h, cols, 64, ctypes.byref(nc), ctypes.byref(rows)))
        if rows.value < 0:
            return None
        return Page(*[block_from_column(cols[i]) for i in range(nc.value)], positionCount=rows.value)

    def finish(self):
        self._lib.oracle_finish(self._h)

    def isFinished(self):
        return bool(self._lib.oracle_is_finished(self._h))

    def close(self):
        if self._h:
            self._lib.oracle_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
