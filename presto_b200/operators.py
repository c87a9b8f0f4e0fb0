"""GPU operators behind the reference's Operator / OperatorFactory contract
(core/trino-main/src/main/java/io/trino/operator/Operator.java:21-103, OperatorFactory.java:18-50).

  GpuFilterAndProjectOperator   <- FilterAndProjectOperator.java:42-88 (PageProcessor: PageFilter + PageProjections)
  GpuAggregationOperator        <- AggregationOperator.java:93-186
  GpuHashAggregationOperator    <- HashAggregationOperator.java:57-135 (factory arguments), :367-518 (protocol)
  GpuPipelineOperator           the fused filter -> project -> aggregate chain in one native operator

All of them are thin shells over one native pgpu_pipeline (include/prestogpu.h). Method names are the
reference's (needsInput/addInput/getOutput/finish/isFinished/close) so tests read like the reference's own.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence

from . import _native as N
from .relational import RowExpression
from .spi import Block, Page, Type, block_from_column


class TrinoException(RuntimeError):
    """io.trino.spi.TrinoException with the StandardErrorCode name in `.code`."""

    def __init__(self, status: int, message: str):
        self.status = status
        self.code = N.STATUS_NAMES.get(status, str(status))
        super().__init__("%s: %s" % (self.code, message))


@dataclass
class Aggregation:
    """One accumulator binding: function(arg channel) [mask channel] -> output type.
    Stands for AccumulatorFactory built by InternalAggregationFunction.bind(channels, maskChannel)."""
    function: str                 # sum count avg min max
    outputType: Type
    inputChannel: Optional[int] = None   # None: count(*)
    inputType: Optional[Type] = None
    maskChannel: Optional[int] = None

    def sexpr(self) -> str:
        s = "(agg %s %s" % (self.function, self.outputType)
        if self.inputChannel is not None:
            s += " (arg %d %s)" % (self.inputChannel, self.inputType)
        if self.maskChannel is not None:
            s += " (mask %d)" % self.maskChannel
        return s + ")"


def plan_text(inputTypes: Sequence[Type], filter: Optional[RowExpression] = None,
              projections: Optional[Sequence[RowExpression]] = None, step: Optional[str] = None,
              groupByChannels: Sequence[int] = (), aggregations: Sequence[Aggregation] = (),
              expectedGroups: int = 10000, maxPartialGroups: int = 0, maxOutputRows: int = 0,
              mergeMinRows: int = 0, mergeMinBytes: int = 0, mergeMaxBytes: int = 0) -> str:
    """mergeMinRows / mergeMinBytes: the session properties filter_and_project_min_output_page_row_count / _size
    (SystemSessionProperties.java) that ScanFilterAndProjectOperator and FilterAndProjectOperator hand to MergePages; 0/0 = off."""
    parts = ["(plan", " (input %s)" % " ".join(str(t) for t in inputTypes)]
    if filter is not None:
        parts.append(" (filter %s)" % filter.sexpr())
    if projections is not None:
        parts.append(" (project %s)" % " ".join(p.sexpr() for p in projections))
    if step is not None:
        parts.append(" (aggregate (step %s)" % step.lower())
        if groupByChannels:
            parts.append(" (group %s)" % " ".join(str(c) for c in groupByChannels))
        parts.append(" (expected-groups %d)" % expectedGroups)
        for a in aggregations:
            parts.append(" " + a.sexpr())
        parts.append(")")
    opts = []
    if maxPartialGroups:
        opts.append("(max-partial-groups %d)" % maxPartialGroups)
    if maxOutputRows:
        opts.append("(max-output-rows %d)" % maxOutputRows)
    if mergeMinRows:
        opts.append("(merge-min-rows %d)" % mergeMinRows)
    if mergeMinBytes:
        opts.append("(merge-min-bytes %d)" % mergeMinBytes)
    if mergeMaxBytes:
        opts.append("(merge-max-bytes %d)" % mergeMaxBytes)
    if opts:
        parts.append(" (options %s)" % " ".join(opts))
    parts.append(")")
    return "\n".join(parts)


class GpuPipelineOperator:
    """One native pipeline. device < 0 compiles the plan only (plan-time validation without a GPU)."""

    def __init__(self, plan: str, device: int = 0):
        self._lib = N.lib()
        self.plan = plan
        self.device = device
        h = ctypes.c_void_p()
        raw = plan.encode()
        st = self._lib.pgpu_pipeline_create(raw, len(raw), device, ctypes.byref(h))
        if st != N.OK:
            raise TrinoException(st, self._lib.pgpu_last_global_error().decode())
        self._h = h
        self._keep: list = []

    # -- Operator ------------------------------------------------------------------------------
    def needsInput(self) -> bool:
        return bool(self._lib.pgpu_needs_input(self._h))

    def _check(self, st: int):
        if st != N.OK:
            raise TrinoException(st, self._lib.pgpu_last_error(self._h).decode())

    @staticmethod
    def _marshal(blocks: Sequence[Block]):
        keep: list = []
        cols = (N.Column * len(blocks))()
        for i, b in enumerate(blocks):
            b._fill(cols[i], keep)
        return cols, keep

    def addInput(self, page: Page, pinned: bool = False):
        cols, keep = self._marshal(page.blocks)
        self._keep = keep if pinned else []
        self._check(self._lib.pgpu_add_page(self._h, cols, len(page.blocks), page.position_count,
                                            N.PAGE_PINNED if pinned else 0))

    def addDeviceInput(self, cols, ncols: int, rows: int):
        """cols: ctypes array of pgpu_column holding DEVICE pointers."""
        self._check(self._lib.pgpu_add_device_page(self._h, cols, ncols, rows))

    def addSerializedInput(self, payload: bytes, positionCount: int):
        """addInput of a page in the reference's wire format (the uncompressed payload of a SerializedPage,
        main/execution/buffer/PagesSerdeUtil.java:45-62); decoded on the device."""
        self._check(self._lib.pgpu_add_serialized_page(self._h, payload, len(payload), positionCount))

    def addSerializedPage(self, slice_: bytes, positionCount: int, markers: int = 0, uncompressedSizeInBytes: int = 0):
        """addInput of a SerializedPage as it travels (main/execution/buffer/SerializedPage.java): markers = PageCodecMarker set
        (1 = COMPRESSED: the slice is an LZ4 block, inflated on the host as PagesSerde.deserialize does; 2 = ENCRYPTED: rejected)."""
        self._check(self._lib.pgpu_add_serialized_page_marked(self._h, slice_, len(slice_), positionCount, markers, uncompressedSizeInBytes))

    def addParquetRowGroup(self, chunks, rows: int):
        """addInput of the Page the reference's ParquetReader.nextPage would build from one row group
        (lib/trino-parquet/.../reader/ParquetReader.java:211-268): chunks[c] = (column chunk bytes as they lie in the file,
        parquet physical type; max definition level; num_values[; FIXED_LEN_BYTE_ARRAY length[; compression codec]]) feeds input channel c. Page and run headers are
        read on the host, everything else is decoded on the device (pgpu_add_parquet_row_group)."""
        arr = (N.ParquetChunk * len(chunks))()
        keep = []
        for i, chunk in enumerate(chunks):
            data, ptype, max_def, num_values = chunk[:4]
            data = bytes(data)
            keep.append(data)
            arr[i].data, arr[i].len, arr[i].physical_type, arr[i].max_definition_level, arr[i].num_values = data, len(data), ptype, max_def, num_values
            arr[i].type_length = chunk[4] if len(chunk) > 4 else 0
            arr[i].codec = chunk[5] if len(chunk) > 5 else 0
        self._check(self._lib.pgpu_add_parquet_row_group(self._h, arr, len(chunks), rows))

    def getSerializedOutput(self):
        """getOutput() in the reference's wire format: (payload bytes, positionCount) or None."""
        data, n, rows = ctypes.c_void_p(), ctypes.c_size_t(), ctypes.c_int64()
        self._check(self._lib.pgpu_get_serialized_output(self._h, ctypes.byref(data), ctypes.byref(n), ctypes.byref(rows)))
        if rows.value < 0:
            return None
        return ctypes.string_at(data.value, n.value), rows.value

    def getOutput(self) -> Optional[Page]:
        n = max(self._lib.pgpu_output_channel_count(self._h), 1)
        cols = (N.Column * n)()
        ncols = ctypes.c_int32()
        rows = ctypes.c_int64()
        self._check(self._lib.pgpu_get_output(self._h, cols, n, ctypes.byref(ncols), ctypes.byref(rows)))
        if rows.value < 0:
            return None
        blocks = [block_from_column(cols[i]) for i in range(ncols.value)]
        return Page(*blocks, positionCount=rows.value)

    def finish(self):
        self._check(self._lib.pgpu_finish(self._h))

    def isFinished(self) -> bool:
        return bool(self._lib.pgpu_is_finished(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.pgpu_pipeline_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- extras --------------------------------------------------------------------------------
    def sync(self):
        self._check(self._lib.pgpu_sync(self._h))

    def reset(self):
        self._check(self._lib.pgpu_reset(self._h))

    def stats(self) -> N.Stats:
        s = N.Stats()
        self._check(self._lib.pgpu_get_stats(self._h, ctypes.byref(s)))
        return s

    def setTiming(self, enabled: bool):
        self._check(self._lib.pgpu_set_timing(self._h, 1 if enabled else 0))

    def generatedSource(self) -> str:
        return self._lib.pgpu_generated_source(self._h).decode()

    def compileSignature(self, encodings: Sequence[int], nullable: Sequence[bool]):
        n = len(encodings)
        e = (ctypes.c_int32 * n)(*encodings)
        u = (ctypes.c_uint8 * n)(*[1 if x else 0 for x in nullable])
        self._check(self._lib.pgpu_compile_signature(self._h, e, u, n))

    def outputTypes(self) -> List[Type]:
        out = []
        for c in range(self._lib.pgpu_output_channel_count(self._h)):
            k, p, s = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
            self._lib.pgpu_output_channel_type(self._h, c, ctypes.byref(k), ctypes.byref(p), ctypes.byref(s))
            out.append(Type(k.value, p.value, s.value))
        return out


class GpuFilterAndProjectOperator(GpuPipelineOperator):
    def __init__(self, inputTypes: Sequence[Type], filter: Optional[RowExpression], projections: Sequence[RowExpression],
                 device: int = 0):
        super().__init__(plan_text(inputTypes, filter, projections), device)


class GpuAggregationOperator(GpuPipelineOperator):
    """Global aggregation: exactly one output row at finish (AggregationOperator.java:164-186)."""

    def __init__(self, inputTypes: Sequence[Type], step: str, aggregations: Sequence[Aggregation], device: int = 0):
        super().__init__(plan_text(inputTypes, step=step, aggregations=aggregations), device)


class GpuHashAggregationOperator(GpuPipelineOperator):
    def __init__(self, inputTypes: Sequence[Type], groupByChannels: Sequence[int], step: str,
                 aggregations: Sequence[Aggregation], expectedGroups: int = 10000, maxPartialGroups: int = 0,
                 maxOutputRows: int = 0, device: int = 0):
        super().__init__(plan_text(inputTypes, step=step, groupByChannels=groupByChannels, aggregations=aggregations,
                                   expectedGroups=expectedGroups, maxPartialGroups=maxPartialGroups,
                                   maxOutputRows=maxOutputRows), device)


class GpuOperatorFactory:
    """OperatorFactory.java:18-50: createOperator / noMoreOperators / duplicate."""

    def __init__(self, operatorId: int, plan: str, device: int = 0):
        self.operatorId = operatorId
        self.plan = plan
        self.device = device
        self.closed = False
        GpuPipelineOperator(plan, -1).close()  # plan-time rejection happens here, as in the planner

    def createOperator(self, driverContext=None) -> GpuPipelineOperator:
        if self.closed:
            raise TrinoException(N.INVALID_STATE, "Factory is already closed")
        return GpuPipelineOperator(self.plan, self.device)

    def noMoreOperators(self):
        self.closed = True

    def duplicate(self) -> "GpuOperatorFactory":
        return GpuOperatorFactory(self.operatorId, self.plan, self.device)


def toPages(operator: GpuPipelineOperator, pages: Sequence[Page]) -> List[Page]:
    """The hand-driving loop of the reference's operator tests
    (core/trino-main/src/test/java/io/trino/operator/OperatorAssertion.java:62-138)."""
    out: List[Page] = []
    it = iter(pages)
    pending = next(it, None)
    loops = 0
    while pending is not None:
        if operator.needsInput():
            operator.addInput(pending)
            pending = next(it, None)
        p = operator.getOutput()
        if p is not None and p.position_count:
            out.append(p)
        loops += 1
        assert loops < 1_000_000, "operator makes no progress"
    operator.finish()
    loops = 0
    while not operator.isFinished():
        p = operator.getOutput()
        if p is not None and p.position_count:
            out.append(p)
        loops += 1
        assert loops < 1_000_000, "operator does not finish"
    return out
