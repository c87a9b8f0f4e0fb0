"""Partial -> final aggregation exchange across the GPUs of one box (one process per GPU, torch.distributed).

Mirrors what the reference runs between a Step.PARTIAL and a Step.FINAL HashAggregationOperator:
PartitionedOutputOperator -> exchange -> final operator (core/trino-main/src/main/java/io/trino/operator/
PartitionedOutputOperator.java:56-133, HashAggregationOperator.java:367-518 with Step.FINAL, whose addIntermediate is
deserialize + combine, operator/aggregation/AccumulatorCompiler.java:674-767).

The exchange itself lives behind the C ABI (pgpu_comm_init / pgpu_exchange_partial, csrc/comm.inc): `Comm` below is a thin
caller. Two shapes, chosen by the library from the partial plan:
  * gather  — small group count (Q1, Q6): every rank's push kernel stores its partial rows into every rank's inbox over
              NVLink and every rank's FINAL gather kernel merges them: no host round trip, no collective; every rank ends
              with the final result (the all-reduce shape, without assuming keys line up by slot).
  * repartition — large group count: rows of the partial output are hash-partitioned on the device by the reference's
              own partition function and each rank's FINAL page kernel reads its partition of every peer's rows straight
              out of that peer's HBM (CUDA IPC); the result stays partitioned by key, as the output of Trino's final stage.
What remains in Python: the NCCL baselines the bench can switch to (`--exchange nccl`: packed all-gather / all_to_all_single
per column) and the host-side restatements the CPU test-suite drives over gloo.
"""
from __future__ import annotations

import ctypes
import pickle
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N
from .spi import Block, Page, Type, block_of

_FRAME = 32 * 1024
PART_KIND = {1: 0, 2: 1, 3: 1, 4: 2}  # type kind -> pgpu_partition_of kind (BIGINT, INTEGER, DATE, DOUBLE)


# ------------------------------------------------------------------------------------------------ the native exchange
class Comm:
    """One rank's membership of a job on one box (pgpu_comm): rendezvous through the shared-memory segment `name`, peers'
    HBM mapped through CUDA IPC. exchange(partial, final) = PartitionedOutputOperator -> exchange -> FINAL.addInput for one
    step; afterwards `final` is finishing and is drained with getOutput()."""

    def __init__(self, world: int, rank: int, device: int, name: str):
        self._lib = N.lib()
        h = ctypes.c_void_p()
        st = self._lib.pgpu_comm_init(world, rank, device, name.encode(), ctypes.byref(h))
        if st != N.OK:
            from .operators import TrinoException
            raise TrinoException(st, self._lib.pgpu_last_global_error().decode())
        self._h = h
        self.world, self.rank, self.device = world, rank, device
        self.recv_rows = 0
        self.sent_bytes = 0
        self.private_groups = 0

    def _check(self, st: int):
        if st != N.OK:
            from .operators import TrinoException
            raise TrinoException(st, self._lib.pgpu_comm_last_error(self._h).decode())

    def exchange(self, partial, final) -> int:
        rows, sent = ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.pgpu_exchange_partial(self._h, partial._h, final._h, ctypes.byref(rows), ctypes.byref(sent)))
        self.recv_rows, self.sent_bytes = rows.value, sent.value
        self.private_groups = int(self._lib.pgpu_comm_private_groups(self._h))   # finalised in place by the range-aware split
        return rows.value

    run = exchange   # the name the older Python transports used

    def barrier(self):
        self._check(self._lib.pgpu_comm_barrier(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.pgpu_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def gather_signature(final_plan: str):
    """(encodings, nullable) of the pages the gather shape feeds a FINAL plan: every channel carries a null vector, VARCHAR
    channels arrive as 16-byte cells. For pre-compiling the variant (GpuPipelineOperator.compileSignature) at plan time."""
    at = final_plan.index("(input ") + len("(input ")
    depth, tok, types = 0, "", []
    for ch in final_plan[at:]:
        if ch == "(":
            depth += 1
        if ch == ")":
            if depth == 0:
                break
            depth -= 1
        if ch.isspace() and depth == 0:
            if tok:
                types.append(tok)
            tok = ""
        else:
            tok += ch
    if tok:
        types.append(tok)
    enc = [N.ENC_VARCELL if t.startswith("(varchar") or t == "varchar" else N.ENC_PLAIN for t in types]
    return enc, [True] * len(enc)


# ------------------------------------------------------------------------------------------------ helpers
class _DeviceBytes:
    """Zero-copy view of library-owned device memory for torch (torch.as_tensor understands __cuda_array_interface__)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def device_bytes_tensor(ptr: int, nbytes: int, device):
    import torch
    if nbytes == 0 or not ptr:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_DeviceBytes(ptr, nbytes), device=device)


def _pages_to_blob(pages: Sequence[Page]) -> bytes:
    payload = []
    for p in pages:
        cols = []
        for b in p.blocks:
            cols.append((b.type.kind, b.type.precision, b.type.scale, b.to_pylist()))
        payload.append((p.position_count, cols))
    return pickle.dumps(payload, protocol=4)


def _blob_to_pages(blob: bytes) -> List[Page]:
    out = []
    for rows, cols in pickle.loads(blob):
        blocks = [block_of(Type(k, pr, sc), vals) for k, pr, sc, vals in cols]
        out.append(Page(*blocks, positionCount=rows))
    return out


# ------------------------------------------------------------------------------------------------ gather (small)
def gather_partial_pages(pages: Sequence[Page], device=None, group=None) -> List[Page]:
    """All ranks receive the partial pages of all ranks, in rank order (small group counts: the pages are a few rows)."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return list(pages)
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device("cpu")
    blob = np.frombuffer(_pages_to_blob(pages), dtype=np.uint8)
    # one collective in the common case: [8-byte length | payload] in a fixed 32 KB frame; every rank sees every length,
    # so all ranks agree on whether a second, larger round is needed
    cap = _FRAME
    rounds = 0
    while True:
        frame = np.zeros(cap, dtype=np.uint8)
        frame[:8] = np.frombuffer(np.int64(blob.size).tobytes(), dtype=np.uint8)
        n = min(blob.size, cap - 8)
        frame[8:8 + n] = blob[:n]
        send = torch.from_numpy(frame).to(dev)
        recv = torch.empty(world * cap, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(recv, send, group=group)
        recv_h = recv.cpu().numpy()
        sizes = [int(np.frombuffer(recv_h[r * cap:r * cap + 8].tobytes(), dtype=np.int64)[0]) for r in range(world)]
        rounds += 1
        if max(sizes) <= cap - 8:
            break
        cap = (max(sizes) + 8 + 4095) // 4096 * 4096
        assert rounds < 3
    out: List[Page] = []
    for r in range(world):
        out.extend(_blob_to_pages(recv_h[r * cap + 8:r * cap + 8 + sizes[r]].tobytes()))
    return out


# Raw (no Python pages) variant of the gather shape: output blocks are packed straight from the library-owned host
# buffers into one frame and handed to the FINAL pipeline as host columns pointing into the received frame.
_WIDTH = {0: 1, 1: 8, 2: 4, 3: 4, 4: 8, 5: 8}


def _col_width(col) -> int:
    return 16 if (col.type_kind == 5 and col.precision > 18) else _WIDTH[col.type_kind]


def pack_output(op) -> bytes:
    """Drain a finishing operator through pgpu_get_output and pack its pages:
    [npages:i64] then per page [ncols:i64][rows:i64] and per column [kind, precision, scale, flags(1 nulls | 2 offsets), nbytes]
    (5 x i64) followed by values, nulls, offsets — each padded to 16 bytes."""
    L = N.lib()
    n = max(L.pgpu_output_channel_count(op._h), 1)
    cols = (N.Column * n)()
    nc, rows = ctypes.c_int32(), ctypes.c_int64()
    parts: List[bytes] = []
    npages = 0

    def pad(b: bytes) -> bytes:
        return b + b"\0" * (-len(b) % 16)

    while not L.pgpu_is_finished(op._h):
        op._check(L.pgpu_get_output(op._h, cols, n, ctypes.byref(nc), ctypes.byref(rows)))
        r = rows.value
        if r <= 0:
            continue
        npages += 1
        parts.append(np.array([nc.value, r], dtype=np.int64).tobytes())
        for c in range(nc.value):
            col = cols[c]
            var = col.type_kind in (6, 7)
            offs = ctypes.string_at(col.offsets, (r + 1) * 4) if var else b""
            nbytes = int(np.frombuffer(offs[-4:], dtype=np.int32)[0]) if var else r * _col_width(col)
            flags = (1 if col.nulls else 0) | (2 if var else 0)
            parts.append(np.array([col.type_kind, col.precision, col.scale, flags, nbytes], dtype=np.int64).tobytes())
            parts.append(pad(ctypes.string_at(col.values, nbytes) if nbytes else b""))
            if col.nulls:
                parts.append(pad(ctypes.string_at(col.nulls, r)))
            if var:
                parts.append(pad(offs))
    return np.array([npages], dtype=np.int64).tobytes() + b"".join(parts)


def add_packed(final_op, frame: np.ndarray) -> int:
    """Feed the pages of one packed frame (uint8 array, kept alive by the caller) to `final_op` as host pages."""
    L = N.lib()
    base = frame.ctypes.data
    hdr = frame[:8].view(np.int64)
    pos, total = 8, 0
    for _ in range(int(hdr[0])):
        ncols, rows = (int(x) for x in frame[pos:pos + 16].view(np.int64))
        pos += 16
        cols = (N.Column * ncols)()
        for c in range(ncols):
            kind, prec, scale, flags, nbytes = (int(x) for x in frame[pos:pos + 40].view(np.int64))
            pos += 40
            col = cols[c]
            col.encoding, col.type_kind, col.precision, col.scale, col.position_count = N.ENC_PLAIN, kind, prec, scale, rows
            col.values = base + pos
            pos += nbytes + (-nbytes % 16)
            if flags & 1:
                col.nulls = base + pos
                pos += rows + (-rows % 16)
            if flags & 2:
                col.offsets = base + pos
                pos += (rows + 1) * 4 + (-((rows + 1) * 4) % 16)
        final_op._check(L.pgpu_add_page(final_op._h, cols, ncols, rows, 0))
        total += rows
    return total


def all_gather_frames(frame: bytes, device=None, group=None):
    """One padded all-gather of every rank's frame; returns a list of uint8 arrays (one per rank, rank order)."""
    import torch
    import torch.distributed as dist
    blob = np.frombuffer(frame, dtype=np.uint8)
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [blob.copy()]
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device("cpu")
    cap, rounds = _FRAME, 0
    while True:
        buf = np.zeros(cap, dtype=np.uint8)
        buf[:8] = np.frombuffer(np.int64(blob.size).tobytes(), dtype=np.uint8)
        n = min(blob.size, cap - 16)
        buf[16:16 + n] = blob[:n]
        recv = torch.empty(world * cap, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(recv, torch.from_numpy(buf).to(dev), group=group)
        recv_h = recv.cpu().numpy()
        sizes = [int(recv_h[r * cap:r * cap + 8].view(np.int64)[0]) for r in range(world)]
        rounds += 1
        if max(sizes) <= cap - 16:
            return [recv_h[r * cap + 16:r * cap + 16 + sizes[r]] for r in range(world)]
        cap = (max(sizes) + 16 + 4095) // 4096 * 4096
        assert rounds < 3


def final_over_pages(final_operator, pages: Sequence[Page]) -> List[Page]:
    from .operators import toPages
    return toPages(final_operator, [p for p in pages if p.position_count])


# ------------------------------------------------------------------------------------------------ repartition (host logic)
def partition_of_row(key_values: Sequence, key_types: Sequence[Type], nparts: int) -> int:
    """The reference's partition function for one row of key values (None = NULL)."""
    n = len(key_values)
    bits = (ctypes.c_int64 * n)()
    nulls = (ctypes.c_uint8 * n)()
    kinds = (ctypes.c_int32 * n)()
    for i, (v, t) in enumerate(zip(key_values, key_types)):
        if t.kind not in PART_KIND:
            raise ValueError("hash partitioning on %s keys is not supported" % t)
        kinds[i] = PART_KIND[t.kind]
        if v is None:
            nulls[i] = 1
        elif t.kind == 4:
            bits[i] = int(np.array([v], dtype=np.float64).view(np.int64)[0])
        else:
            bits[i] = int(v)
    return N.lib().pgpu_partition_of(bits, nulls, kinds, n, nparts)


def partition_pages_host(pages: Sequence[Page], nkeys: int, nparts: int) -> List[List[Page]]:
    """Host restatement of PartitionedOutputOperator.partitionPage for small pages: one page list per partition."""
    buckets: List[List[list]] = [[] for _ in range(nparts)]
    types: Optional[List[Type]] = None
    for p in pages:
        types = [b.type for b in p.blocks]
        for row in p.to_rows():
            buckets[partition_of_row(row[:nkeys], types[:nkeys], nparts)].append(row)
    out: List[List[Page]] = []
    for rows in buckets:
        if not rows or types is None:
            out.append([])
            continue
        cols = list(zip(*rows))
        out.append([Page(*[block_of(t, list(c)) for t, c in zip(types, cols)], positionCount=len(rows))])
    return out


def exchange_pages_host(per_dest: Sequence[Sequence[Page]], device=None, group=None) -> List[Page]:
    """All-to-all of small host pages (per_dest[r] goes to rank r); returns what this rank received, in source-rank order."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [p for pages in per_dest for p in pages]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    blobs = [_pages_to_blob(list(pages)) for pages in per_dest]
    gathered: List[Optional[list]] = [None] * world
    dist.all_gather_object(gathered, blobs, group=group)
    out: List[Page] = []
    for src in range(world):
        out.extend(_blob_to_pages(gathered[src][rank]))
    return out


# ------------------------------------------------------------------------------------------------ repartition (device)
class DeviceRepartition:
    """PARTIAL pipeline -> hash partition on the device -> FINAL pipeline of the owning rank.

    `partial` must be finishing (finish() called); `final` is a FINAL-step GpuPipelineOperator whose input channels are the
    partial pipeline's output channels. After run() the final operator holds this rank's partition of the groups.

    This is the NCCL BASELINE of the repartition shape (one all_to_all_single per column into receive buffers, then the FINAL
    pipeline over those); the product path is Comm.exchange (peer memory, no collective, no staging copy)."""

    def __init__(self, device: int, world: int, rank: int, group=None, transport: str = "nccl"):
        self.device, self.world, self.rank, self.group = device, world, rank, group
        self.transport = transport if world > 1 else "local"
        self.recv_keepalive: list = []
        self.sent_bytes = 0       # bytes this rank's partitions put on the wire (nccl) / peers read from it is symmetric (peer)
        self.recv_rows = 0

    def close(self):
        self.recv_keepalive = []

    @staticmethod
    def _width(t) -> int:
        w = 16 if t.is_long_decimal else {0: 1, 1: 8, 2: 4, 3: 4, 4: 8, 5: 8}.get(t.kind)
        if w is None:
            raise RuntimeError("device repartition of %s channels is not supported" % t)
        return w

    def _feed(self, final, types, bases, start, count):
        """add one partition (rows [start, start+count) of buffers `bases` = [(values, nulls)] per channel) to FINAL"""
        nc = len(types)
        fcols = (N.Column * nc)()
        for c, t in enumerate(types):
            fc = fcols[c]
            fc.encoding = N.ENC_PLAIN
            fc.type_kind, fc.precision, fc.scale = t.kind, t.precision, t.scale
            fc.position_count = count
            fc.values = bases[c][0] + start * self._width(t)
            fc.nulls = bases[c][1] + start
        final.addDeviceInput(fcols, nc, count)

    def run(self, partial, final) -> int:
        L = N.lib()
        self.sent_bytes = 0
        ncap = max(L.pgpu_output_channel_count(partial._h), 1)
        cols = (N.Column * ncap)()
        nc = ctypes.c_int32()
        offs = (ctypes.c_int64 * (self.world + 1))()
        cnts = (ctypes.c_int64 * self.world)()
        partial._check(L.pgpu_partition_output_device(partial._h, self.world, cols, ncap, ctypes.byref(nc), offs, cnts))
        types = partial.outputTypes()[:nc.value]
        own = [(cols[c].values or 0, cols[c].nulls or 0) for c in range(nc.value)]
        if self.transport == "local":
            total = 0
            for part in range(self.world):
                if cnts[part]:
                    self._feed(final, types, own, offs[part], cnts[part])
                    total += cnts[part]
        else:
            total = self._run_nccl(final, types, own, offs, cnts)
        self.recv_rows = total
        final.finish()
        return total

    # -- NCCL transport (baseline) -----------------------------------------------------------------------------------------
    def _run_nccl(self, final, types, own, offs, cnts) -> int:
        import torch
        import torch.distributed as dist
        dev = torch.device("cuda", self.device)
        pad_send = [offs[i + 1] - offs[i] for i in range(self.world)]           # padded partitions, multiples of 16 rows
        send_t = torch.tensor([v for i in range(self.world) for v in (pad_send[i], cnts[i])], dtype=torch.int64, device=dev)
        recv_t = torch.empty(2 * self.world, dtype=torch.int64, device=dev)
        dist.all_to_all_single(recv_t, send_t, group=self.group)
        rt = recv_t.cpu().tolist()
        pad_recv, real_recv = rt[0::2], rt[1::2]
        tot_send, tot_recv = int(offs[self.world]), int(sum(pad_recv))
        self.recv_keepalive = []
        bases = []
        for c, t in enumerate(types):
            pair = []
            for ptr, w in ((own[c][0], self._width(t)), (own[c][1], 1)):
                src = device_bytes_tensor(ptr, tot_send * w, dev)
                dst = torch.empty(max(tot_recv * w, 16), dtype=torch.uint8, device=dev)
                dist.all_to_all_single(dst[:tot_recv * w], src, [r * w for r in pad_recv], [r * w for r in pad_send], group=self.group)
                self.sent_bytes += (tot_send - pad_send[self.rank]) * w
                self.recv_keepalive.append(dst)
                pair.append(dst.data_ptr())
            bases.append(tuple(pair))
        torch.cuda.current_stream(dev).synchronize()  # the FINAL pipeline reads the received buffers on its own stream
        total, start = 0, 0
        for src in range(self.world):
            if real_recv[src]:
                self._feed(final, types, bases, start, real_recv[src])
                total += real_recv[src]
            start += pad_recv[src]
        return total
