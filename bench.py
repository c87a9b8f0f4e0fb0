#!/usr/bin/env python
"""bench.py — TPC-H Q1 (BASELINE.json configs[1]) lineitem rows/s on N B200s, with the kernel roofline, a CPU baseline, an
end-to-end leg at the reference's page shape and the other north_star configs as sub-records.

One "step" = one full pass of the fused scan->filter->project->hash-aggregate pipeline over the rank's lineitem shard
(device-resident for `value`; 1 M-row host pages in ordinary pageable memory through pgpu_add_page for `e2e`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--workload q1|q6|highcard|highcard_shuffled|q1dict] [--rows R]

Without --workload the line is the Q1 headline plus `configs` = {q6, highcard, highcard_shuffled, q1dict} sub-records
(N > 1: {highcard}), every record checked against the CPU oracle on a prefix of the very rows it timed.
"""
import argparse
import ctypes
import json
import os
import struct
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SF100_ROWS = 600_037_902  # TPC-H lineitem at sf100
SUB_CONFIGS = ("q6", "highcard", "highcard_shuffled", "q1dict")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["q1", "q6", "highcard", "highcard_shuffled", "q1dict"],
                    help="one workload only (default: the Q1 headline + the other configs as sub-records)")
    ap.add_argument("--rows", type=int, default=SF100_ROWS, help="lineitem rows per GPU")
    ap.add_argument("--data", default="synthetic", choices=["synthetic", "dbgen"],
                    help="synthetic: counter-based columns with the TPC-H value distributions (--rows per GPU); dbgen: the tpch connector's real rows "
                         "(spec-conformant generator, --sf per GPU; rows per GPU follow from the order range)")
    ap.add_argument("--sf", type=float, default=100.0, help="--data dbgen: TPC-H scale factor per GPU")
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows of the end-to-end (host pages) leg; 0 = same as --rows if RAM allows")
    ap.add_argument("--page-rows", type=int, default=1_000_000, help="rows per host page of the end-to-end leg")
    ap.add_argument("--verify-rows", type=int, default=64_000_000, help="prefix of the timed rows checked against the CPU oracle (0 = skip)")
    ap.add_argument("--cpu-rows-per-core", type=int, default=4_000_000, help="--impl reference: rows per host thread and step")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1, large group counts: FINAL reads the peers' partitions over NVLink (peer) or NCCL all-to-all first (nccl)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="headline only")
    return ap.parse_args()


def workload_def(name, rows, dbgen_sf=None):
    wl = _workload_def(name, rows)
    wl["name"] = name
    wl["rows"] = rows
    if dbgen_sf is not None:
        from presto_b200 import tpch
        if name == "highcard_shuffled":
            raise SystemExit("--data dbgen has no shuffled-key variant")
        if name == "q1dict":
            raise SystemExit("--data dbgen host pages are plain blocks: use --workload q1 (the device leg of q1dict is available through DevicePage)")
        make = {"q1": tpch.tpch_q1_page, "q6": tpch.tpch_q6_page, "highcard": tpch.tpch_highcard_page}[name]
        # CPU legs address pages by row range: map it onto the order range of the same position (4.0008 rows per order)
        norders = tpch.tpch_order_count(dbgen_sf)
        wl["page"] = lambda first_row, n, total: make(dbgen_sf, (first_row // 4) % max(norders - max(n // 4, 1), 1), max(n // 4, 1))
        wl["label"] = wl["label"].replace("synthetic lineitem", "the tpch connector's lineitem rows (dbgen)")
    return wl


def _workload_def(name, rows):
    from presto_b200 import tpch
    from presto_b200.spi import DOUBLE, LongArrayBlock, VariableWidthBlock
    import numpy as np
    if name == "q1":
        return dict(plan=tpch.q1_plan, spec=tpch.Q1_DEVICE_SPEC, bytes_per_row=tpch.Q1_BYTES_PER_ROW, page=tpch.q1_page,
                    label="TPC-H Q1 (GROUP BY returnflag, linestatus; 8 aggregates) over synthetic lineitem", nkeys=2)
    if name == "q1dict":
        rf = VariableWidthBlock.from_values(tpch.VARCHAR1, ["A", "N", "R"])
        ls = VariableWidthBlock.from_values(tpch.VARCHAR1, ["F", "O"])
        qty = LongArrayBlock(DOUBLE, np.arange(1, 51, dtype=np.float64))
        disc = LongArrayBlock(DOUBLE, np.arange(0, 11, dtype=np.float64) / 100.0)
        tax = LongArrayBlock(DOUBLE, np.arange(0, 9, dtype=np.float64) / 100.0)
        from presto_b200.spi import DATE
        spec = [(tpch.VARCHAR1, "dict", "returnflag_id", rf), (tpch.VARCHAR1, "dict", "linestatus_id", ls),
                (DOUBLE, "dict", "quantity_id", qty), (DOUBLE, "plain", "extendedprice"), (DOUBLE, "dict", "discount_id", disc),
                (DOUBLE, "dict", "tax_id", tax), (DATE, "plain", "shipdate")]
        return dict(plan=tpch.q1_plan, spec=spec, bytes_per_row=32, page=tpch.q1_page_encoded,
                    label="TPC-H Q1 over dictionary-encoded lineitem pages (config 4)", nkeys=2)
    if name == "q6":
        return dict(plan=tpch.q6_plan, spec=tpch.Q6_DEVICE_SPEC, bytes_per_row=tpch.Q6_BYTES_PER_ROW, page=tpch.q6_page,
                    label="TPC-H Q6 (filter + project + global SUM) over synthetic lineitem", nkeys=None)
    shuffled = name == "highcard_shuffled"
    return dict(plan=lambda step="single": tpch.highcard_plan(max(rows // 4, 1000), step),
                spec=tpch.HC_SHUFFLED_DEVICE_SPEC if shuffled else tpch.HC_DEVICE_SPEC, bytes_per_row=tpch.HC_BYTES_PER_ROW,
                page=lambda a, b, c: tpch.highcard_page(a, b, c, shuffled),
                label="GROUP BY l_orderkey SUM(l_quantity), %s keys (config 3)" % ("shuffled" if shuffled else "clustered"), nkeys=1)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm, mx, reasons = [], 0.0, set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx = max(mx, float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s; MEASURED_PEAKS.json absent)"


def measured_traffic(name):
    """DRAM bytes per row of the workload's dominant kernel from the last `ncu --set full` capture (profiles/traffic.json,
    written by tools/collect_profiles.py from dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        return json.load(f).get(name)


# ------------------------------------------------------------------------------------------------ host pages
def h2d_ceiling(device, mb=512, reps=6):
    """Plain pinned host -> device copy bandwidth of this process's GPU (GB/s): what the link gives when nothing else is in the way."""
    import torch
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(n, dtype=torch.uint8, device="cuda:%d" % device)
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    e1.record()
    e1.synchronize()
    return reps * n / (e0.elapsed_time(e1) / 1e3) / 1e9


def host_copy_ceiling(mb=256, seconds=0.4):
    """pageable -> pinned through the library's own staging copy (worker pool, non-temporal stores), GB/s: what the host memory
    system gives the ingest path before any byte crosses PCIe"""
    import numpy as np
    import torch
    from presto_b200 import _native as N
    n = mb << 20
    src = np.ones(n, dtype=np.uint8)
    dst = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    L = N.bench_lib()
    L.pgpu_bench_host_copy(dst.data_ptr(), src.ctypes.data, n)
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < seconds:
        L.pgpu_bench_host_copy(dst.data_ptr(), src.ctypes.data, n)
        reps += 1
    return reps * n / (time.perf_counter() - t0) / 1e9


class HostTable:
    """The rank's shard as a Trino worker holds it: one ordinary (pageable) numpy array per block array, cut into pages of
    page_rows positions. The bytes are the device columns copied back, so host pages and device page hold identical rows."""
    WIDTH = {0: 1, 1: 8, 2: 4, 3: 4, 4: 8}

    def __init__(self, device, wl, dp, rows):
        import numpy as np
        from presto_b200 import _native as N
        self.N, self.np = N, np
        L = N.lib()
        self.rows = rows
        self.spec = wl["spec"]
        self.ncol = len(self.spec)
        self.arrays = []       # per channel: dict(values=, offsets=, ids=)
        self.base = (N.Column * self.ncol)()
        self._keep = []
        self.bytes_per_row = 0

        def back(dev_ptr, count, dtype):
            a = np.empty(count, dtype=dtype)
            if L.pgpu_memcpy_d2h(device, a.ctypes.data, dev_ptr, a.nbytes):
                raise RuntimeError("d2h copy failed")
            return a

        for i, s in enumerate(self.spec):
            src = dp.cols[i]
            ctypes.memmove(ctypes.byref(self.base[i]), ctypes.byref(src), ctypes.sizeof(N.Column))
            self.base[i].values = None
            self.base[i].offsets = None
            self.base[i].ids = None
            self.base[i].nulls = None
            self.base[i].dictionary = None
            t, enc = s[0], s[1]
            a = {}
            if enc == "plain":
                w = self.WIDTH[t.kind]
                a["values"] = back(src.values, rows, {1: np.uint8, 4: np.int32, 8: np.int64}[w])
                self.bytes_per_row += w
            elif enc == "varchar1":
                a["values"] = back(src.values, rows, np.uint8)
                a["offsets"] = back(src.offsets, rows + 1, np.int32)
                self.bytes_per_row += 5
            else:  # dict: ids from the device, the dictionary block straight from the host block
                a["ids"] = back(src.ids, rows, np.int32)
                d = N.Column()
                keep = []
                s[3]._fill(d, keep)
                self._keep.append((d, keep))
                a["dictionary"] = d
                a["dict_block"] = s[3]
                self.bytes_per_row += 4
            self.arrays.append(a)

    def pages(self, page_rows, rows=None):
        """[(pgpu_column array of HOST pointers, positions)] — built up front, like Block objects a caller already holds"""
        N = self.N
        rows = self.rows if rows is None else min(rows, self.rows)
        out = []
        for r0 in range(0, rows, page_rows):
            n = min(page_rows, rows - r0)
            cols = (N.Column * self.ncol)()
            ctypes.memmove(cols, self.base, ctypes.sizeof(cols))
            for i, a in enumerate(self.arrays):
                c = cols[i]
                c.position_count = n
                if "ids" in a:
                    c.ids = a["ids"].ctypes.data + 4 * r0
                    c.dictionary = ctypes.pointer(a["dictionary"])
                elif "offsets" in a:   # a slice of a VariableWidthBlock: absolute offsets into the block's bytes (arrayOffset != 0)
                    c.values = a["values"].ctypes.data
                    c.offsets = a["offsets"].ctypes.data + 4 * r0
                else:
                    c.values = a["values"].ctypes.data + a["values"].itemsize * r0
            out.append((cols, n))
        return out

    def serialized_pages(self, page_rows, rows=None):
        """[(payload bytes, positions)]: PagesSerdeUtil.writeRawPage of the same pages (main/execution/buffer/PagesSerdeUtil.java:45-62;
        blocks without nulls: spi/block/LongArrayBlockEncoding.java:39-60, VariableWidthBlockEncoding.java:37-58,
        DictionaryBlockEncoding.java:36-58) — what an upstream worker's output buffer would deliver."""
        np = self.np
        rows = self.rows if rows is None else min(rows, self.rows)

        def name(b):
            return struct.pack("<i", len(b)) + b

        def flat(arr, n):
            return [name({1: b"BYTE_ARRAY", 4: b"INT_ARRAY", 8: b"LONG_ARRAY"}[arr.itemsize]), struct.pack("<i", n), b"\x00", arr.tobytes()]

        def varwidth(data, offs, n):
            ends = (offs[1:n + 1].astype(np.int64) - int(offs[0])).astype("<i4")
            total = int(ends[-1]) if n else 0
            return [name(b"VARIABLE_WIDTH"), struct.pack("<i", n), ends.tobytes(), b"\x00", struct.pack("<i", total),
                    data[int(offs[0]):int(offs[0]) + total].tobytes()]

        out = []
        for r0 in range(0, rows, page_rows):
            n = min(page_rows, rows - r0)
            parts = [struct.pack("<i", self.ncol)]
            for a in self.arrays:
                if "ids" in a:
                    d = a["dict_block"]
                    parts += [name(b"DICTIONARY"), struct.pack("<i", n)]
                    parts += varwidth(d.data, d.offsets, d.position_count) if hasattr(d, "offsets") else flat(d.values, d.position_count)
                    parts += [a["ids"][r0:r0 + n].tobytes(), struct.pack("<qqq", 0x1234, 0x5678, 1)]
                elif "offsets" in a:
                    parts += varwidth(a["values"], a["offsets"][r0:r0 + n + 1], n)
                else:
                    parts += flat(a["values"][r0:r0 + n], n)
            out.append((b"".join(parts), n))
        return out

    def exact_view(self, name, rows):
        """numpy view of a plain column's first `rows` values (float64 for DOUBLE channels)"""
        np = self.np
        for s, a in zip(self.spec, self.arrays):
            if s[2] == name:
                v = a["values"][:rows]
                return v.view(np.float64) if s[0].kind == 4 else v
        raise KeyError(name)


def drain_output(op, to_host):
    """Operator.getOutput() until isFinished(), through the raw C ABI (no Python-side page copies). Returns rows."""
    from presto_b200 import _native as N
    L = N.lib()
    n = max(L.pgpu_output_channel_count(op._h), 1)
    cols = (N.Column * n)()
    nc, rows = ctypes.c_int32(), ctypes.c_int64()
    total = 0
    fn = L.pgpu_get_output if to_host else L.pgpu_get_output_device
    while not L.pgpu_is_finished(op._h):
        op._check(fn(op._h, cols, n, ctypes.byref(nc), ctypes.byref(rows)))
        if rows.value > 0:
            total += rows.value
    return total


def time_host_pages(op, pages, steps, to_host=True, warmup=1, flags=0):
    """One step = Operator protocol over the page stream: needsInput / addInput(page) per page, finish, getOutput until finished.
    Wall clock from the first addInput to the last output page on the host."""
    from presto_b200 import _native as N
    L = N.lib()
    h = op._h
    ncol = len(pages[0][0])
    add, need = L.pgpu_add_page, L.pgpu_needs_input
    total = sum(n for _, n in pages)
    t0 = 0.0
    for it in range(warmup + steps):
        if it == warmup:
            t0 = time.perf_counter()
        op.reset()
        for cols, n in pages:
            if not need(h):
                raise RuntimeError("operator refused input")
            st = add(h, cols, ncol, n, flags)
            if st:
                op._check(st)
        op.finish()
        out_rows = drain_output(op, to_host=to_host)
        if not to_host:
            op.sync()
    dt = time.perf_counter() - t0
    st = op.stats()
    return dict(rows_per_s=total * steps / dt, seconds_per_step=dt / steps, rows=total, pages=len(pages), steps=steps,
                h2d_bytes_per_step=int(st.h2d_bytes), d2h_bytes_per_step=int(st.d2h_bytes), super_pages_per_step=int(st.super_pages),
                launches_per_step=int(st.kernel_launches), result_rows=out_rows)


def time_serialized_pages(op, payloads, steps, to_host=True, warmup=1):
    from presto_b200 import _native as N
    L = N.lib()
    h = op._h
    add = L.pgpu_add_serialized_page
    total = sum(n for _, n in payloads)
    t0 = 0.0
    for it in range(warmup + steps):
        if it == warmup:
            t0 = time.perf_counter()
        op.reset()
        for data, n in payloads:
            st = add(h, data, len(data), n)
            if st:
                op._check(st)
        op.finish()
        out_rows = drain_output(op, to_host=to_host)
    dt = time.perf_counter() - t0
    st = op.stats()
    return dict(rows_per_s=total * steps / dt, seconds_per_step=dt / steps, rows=total, pages=len(payloads), steps=steps,
                h2d_bytes_per_step=int(st.h2d_bytes), d2h_bytes_per_step=int(st.d2h_bytes), super_pages_per_step=int(st.super_pages),
                launches_per_step=int(st.kernel_launches), result_rows=out_rows)


# ------------------------------------------------------------------------------------------------ CPU oracle legs
def oracle_run(wl, pages, cores):
    """The CPU oracle (oracle/, the C++ restatement of the reference's loops — the reference itself is Java and there is no JVM
    on this box) over `pages` = [(pgpu_column array, positions)] on `cores` threads: one PARTIAL pipeline per thread over its
    share of the pages, then one FINAL pipeline over the partial outputs (the analogue of task_concurrency drivers + the
    partial -> final exchange). Returns (seconds, result pages)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_op import OracleOperator
    from presto_b200.operators import toPages
    partial_plan = wl["plan"]("partial")
    ncol = len(pages[0][0])
    shards = [pages[t::cores] for t in range(cores)]
    outs = [None] * cores

    def work(t):
        op = OracleOperator(partial_plan)
        res = []
        for cols, n in shards[t]:
            op.addRaw(cols, ncol, n)
            p = op.getOutput()
            if p is not None and p.position_count:
                res.append(p)
        op.finish()
        while not op.isFinished():
            p = op.getOutput()
            if p is not None and p.position_count:
                res.append(p)
        op.close()
        outs[t] = res

    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(t,)) for t in range(cores)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    fop = OracleOperator(final_plan_for(wl, partial_plan))
    result = toPages(fop, [p for o in outs for p in o])
    fop.close()
    return time.perf_counter() - t0, result


def cpu_run(wl, rows_per_core, cores, repeats=1):
    """--impl reference: the oracle over freshly generated host pages (1 M-row pages, rows_per_core per thread)."""
    total = rows_per_core * cores
    page_rows = 1_000_000
    from presto_b200.operators import GpuPipelineOperator
    pages = []
    keep = []
    for t in range(cores):
        r0, end = t * rows_per_core, (t + 1) * rows_per_core
        while r0 < end:
            n = min(page_rows, end - r0)
            pg = wl["page"](r0, n, total)
            cols, k = GpuPipelineOperator._marshal(pg.blocks)
            keep.append((pg, k))
            pages.append((cols, pg.position_count))
            r0 += n
    total = sum(n for _, n in pages)
    best, result = None, None
    for _ in range(repeats):
        dt, result = oracle_run(wl, pages, cores)
        best = dt if best is None else min(best, dt)
    return total / best, total, result


def final_plan_for(wl, partial_plan):
    """FINAL-step plan whose input channels are the PARTIAL plan's output channels."""
    from presto_b200.operators import Aggregation, plan_text
    from presto_b200 import tpch
    from presto_b200.spi import BIGINT, DOUBLE
    name = wl["name"]
    if name in ("q1", "q1dict"):
        # partial output: returnflag, linestatus, then per aggregate its intermediate channels
        inp = [tpch.VARCHAR1, tpch.VARCHAR1] + [BIGINT, DOUBLE] * 4 + [DOUBLE, BIGINT] * 3 + [BIGINT]
        aggs = [Aggregation("sum", DOUBLE, 2, DOUBLE), Aggregation("sum", DOUBLE, 4, DOUBLE), Aggregation("sum", DOUBLE, 6, DOUBLE),
                Aggregation("sum", DOUBLE, 8, DOUBLE), Aggregation("avg", DOUBLE, 10, DOUBLE), Aggregation("avg", DOUBLE, 12, DOUBLE),
                Aggregation("avg", DOUBLE, 14, DOUBLE), Aggregation("count", BIGINT, 16, BIGINT)]
        return plan_text(inp, step="final", groupByChannels=[0, 1], aggregations=aggs)
    if name == "q6":
        return plan_text([BIGINT, DOUBLE], step="final", aggregations=[Aggregation("sum", DOUBLE, 0, DOUBLE)])
    return plan_text([BIGINT, BIGINT, DOUBLE], step="final", groupByChannels=[0], aggregations=[Aggregation("sum", DOUBLE, 1, DOUBLE)],
                     expectedGroups=max(wl.get("rows", 40_000_000) // 4, 1000))


# ------------------------------------------------------------------------------------------------ correctness of the timed data
def _channels(pages):
    """result pages -> per channel (values ndarray, nulls ndarray | None), rows concatenated"""
    import numpy as np
    if not pages:
        return []
    out = []
    for c in range(len(pages[0].blocks)):
        bl = [p.blocks[c] for p in pages]
        if hasattr(bl[0], "offsets"):
            vals = np.array([v if v is not None else "" for b in bl for v in b.to_pylist()])
        else:
            vals = np.concatenate([b.values.view(np.float64) if b.type.kind == 4 else b.values for b in bl])
        nulls = None
        if any(b.valueIsNull is not None for b in bl):
            nulls = np.concatenate([b.valueIsNull if b.valueIsNull is not None else np.zeros(b.position_count, np.uint8) for b in bl])
        out.append((vals, nulls))
    return out


def _close(a, b, rtol):
    import numpy as np
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return bool(np.all(np.abs(a - b) <= rtol * np.maximum(np.abs(a), np.abs(b))))


def gpu_result_pages(op, dp, ncol, rows):
    """SINGLE-step result of the device page's first `rows` rows, as host pages"""
    from presto_b200 import _native as N
    cols = (N.Column * ncol)()
    ctypes.memmove(cols, dp.cols, ctypes.sizeof(cols))
    for i in range(ncol):
        cols[i].position_count = rows
    op.reset()
    op.addDeviceInput(cols, ncol, rows)
    op.finish()
    out = []
    while not op.isFinished():
        p = op.getOutput()
        if p is not None and p.position_count:
            out.append(p)
    return out


def verify_prefix(wl, op, dp, host, rows, cores):
    """Check the pipeline on the rows it was timed on: the first `rows` rows of the device page through the GPU, the same rows
    (the host copy of the same bytes) through the CPU oracle on `cores` threads, and — Q1 / Q6 — through numpy with 80-bit
    accumulation as the correctly rounded reference. Keys and counts bit-exact; DOUBLE sums within 1e-12 relative of the exact
    sum (GPU) and within the sequential-summation bound n * 2^-53 (oracle). Returns (record, oracle seconds)."""
    import numpy as np
    name = wl["name"]
    ncol = len(wl["spec"])
    gpu = _channels(gpu_result_pages(op, dp, ncol, rows))
    secs, ora_pages = oracle_run(wl, host.pages(1_000_000, rows), cores)
    ora = _channels(ora_pages)
    rec = {"verified": False, "verified_rows": rows, "oracle_seconds": secs, "tolerance": "keys / counts bit-exact; DOUBLE sums 1e-12 relative"}
    seq = max(1e-12, rows * 2.0 ** -53)
    ok = len(gpu) == len(ora) and len(gpu) > 0 and len(gpu[0][0]) == len(ora[0][0])
    if not ok:
        rec["why"] = "result shapes differ"
        return rec, secs
    if name in ("q1", "q1dict", "q6"):
        ld = np.longdouble
        if name == "q6":
            e, d, q, ship = (host.exact_view(k, rows) for k in ("extendedprice", "discount", "quantity", "shipdate"))
            sel = (ship >= 8766) & (ship < 9131) & (d >= 0.05) & (d <= 0.07) & (q < 24.0)
            exact = float(np.sum((e * d)[sel], dtype=ld))
            ok = _close(gpu[0][0], [exact], 1e-12) and _close(ora[0][0], [exact], seq)
            rec["groups"] = 1
        else:
            if name == "q1":
                rf, ls = host.arrays[0]["values"][:rows], host.arrays[1]["values"][:rows]
                q, e, d, t, ship = (host.exact_view(k, rows) for k in ("quantity", "extendedprice", "discount", "tax", "shipdate"))
            else:
                rf = np.frombuffer(b"ANR", dtype=np.uint8)[host.arrays[0]["ids"][:rows]]
                ls = np.frombuffer(b"FO", dtype=np.uint8)[host.arrays[1]["ids"][:rows]]
                q = np.arange(1, 51, dtype=np.float64)[host.arrays[2]["ids"][:rows]]
                d = (np.arange(0, 11, dtype=np.float64) / 100.0)[host.arrays[4]["ids"][:rows]]
                t = (np.arange(0, 9, dtype=np.float64) / 100.0)[host.arrays[5]["ids"][:rows]]
                e, ship = host.exact_view("extendedprice", rows), host.exact_view("shipdate", rows)
            dp_ = e * (1.0 - d)
            ch = dp_ * (1.0 + t)
            sel = ship <= 10471
            code = rf.astype(np.int32) * 256 + ls
            exact = {}
            for k in np.unique(code[sel]):
                m = sel & (code == k)
                n = int(m.sum())
                s = [float(np.sum(v[m], dtype=ld)) for v in (q, e, dp_, ch, d)]
                exact[(chr(k >> 8), chr(k & 255))] = s[:4] + [s[0] / n, s[1] / n, s[4] / n, n]

            def rows_of(ch_):
                return {(str(ch_[0][0][i]), str(ch_[1][0][i])): [ch_[c][0][i] for c in range(2, 10)] for i in range(len(ch_[0][0]))}
            g, o = rows_of(gpu), rows_of(ora)
            ok = set(g) == set(o) == set(exact)
            for k in exact if ok else ():
                ok &= int(g[k][7]) == int(o[k][7]) == exact[k][7]                       # count(*): bit-exact
                ok &= _close(g[k][:7], exact[k][:7], 1e-12) and _close(o[k][:7], exact[k][:7], seq)
            rec["groups"] = len(exact)
    else:
        gk, ok_ = np.argsort(gpu[0][0], kind="stable"), np.argsort(ora[0][0], kind="stable")
        ok = bool(np.array_equal(gpu[0][0][gk], ora[0][0][ok_])) and _close(gpu[1][0][gk], ora[1][0][ok_], 1e-12)
        q = host.exact_view("quantity", rows)
        ok &= _close([float(np.sum(gpu[1][0], dtype=np.longdouble))], [float(np.sum(q, dtype=np.longdouble))], 1e-12)  # sum of sums
        rec["groups"] = int(len(gk))
    rec["verified"] = bool(ok)
    return rec, secs


# ------------------------------------------------------------------------------------------------ GPU legs
_COMM = {}


def get_comm(device, world, rank):
    """the rank's pgpu_comm (one per process; every pipeline pair of the run exchanges through it)"""
    from presto_b200 import exchange as X
    if world == 1:
        return None
    if "c" not in _COMM:
        import torch.distributed as td
        token = [("%s_%d_%d" % (os.environ.get("MASTER_PORT", "0"), os.getpid(), time.time_ns())) if rank == 0 else None]
        td.broadcast_object_list(token, src=0)      # a segment name no earlier (crashed) job can have left behind
        _COMM["c"] = X.Comm(world, rank, device, "pgpu_bench_" + token[0])
    return _COMM["c"]


def make_exchange(args, device, world, rank, small, leg):
    """--exchange peer (default): the native exchange (pgpu_exchange_partial); nccl: the Python NCCL baselines"""
    from presto_b200 import exchange as X
    if world == 1:
        return None
    if args.exchange == "peer":
        return get_comm(device, world, rank)
    return None if small else X.DeviceRepartition(device, world, rank, transport="nccl")


def exchange_and_final(op, final, small, repart, tdev, state):
    """N > 1: the partial -> final exchange of one step (op has finished its shard); leaves the result drained.
    `repart` is the exchange object: Comm (native: gather or repartition, chosen by the library from the plan), None (NCCL
    all-gather baseline, small results) or DeviceRepartition (NCCL all-to-all baseline)."""
    from presto_b200 import exchange as X
    final.reset()
    if isinstance(repart, X.Comm):
        repart.exchange(op, final)                  # device-side exchange; FINAL is left finishing
        state["result_rows"] = drain_output(final, to_host=small)
        state["exchange_bytes"] = repart.sent_bytes
        state["private_groups"] = repart.private_groups
    elif small:
        frame = X.pack_output(op)                                       # waits for the shard's kernels, drains the partial page
        frames = X.all_gather_frames(frame, device=tdev)                # one NCCL all-gather of a 32 KB frame per rank
        for f in frames:
            X.add_packed(final, f)                                      # FINAL step: addIntermediate over every rank's partial page
        final.finish()
        state["result_rows"] = drain_output(final, to_host=True)
    else:
        repart.run(op, final)
        state["result_rows"] = drain_output(final, to_host=False)
        state["exchange_bytes"] = repart.sent_bytes


def run_device_steps(args, wl, device, dist, steps, warmup):
    """`value`: inputs resident in HBM before the timed region. One step = reset, one pass of the fused pipeline over the
    rank's shard, and (N > 1) the partial -> final exchange. Timed with CUDA events recorded on an idle device on both
    sides (barrier + synchronize), max over ranks taken by the caller."""
    import torch
    from presto_b200 import tpch
    from presto_b200.operators import GpuPipelineOperator
    rows = wl["rows"]
    world, rank = dist["world"], dist["rank"]
    first = rank * rows
    total = rows * world
    dp = tpch.DevicePage(device, wl["spec"], rows, first_row=first, total_rows=total, dbgen=wl.get("dbgen"))
    assert dp.rows == rows
    single = world == 1
    op = GpuPipelineOperator(wl["plan"]("single" if single else "partial"), device)
    final = None if single else GpuPipelineOperator(final_plan_for(wl, None), device)
    ncol = len(wl["spec"])
    small = wl["name"] in ("q1", "q1dict", "q6")   # result is a handful of rows: host output / gather exchange
    tdev = torch.device("cuda", device)
    repart = make_exchange(args, device, world, rank, small, "value_" + wl["name"])
    state = {"result_rows": 0, "exchange_bytes": 0}

    def one_step():
        op.reset()
        op.addDeviceInput(dp.cols, ncol, rows)
        op.finish()
        if single:
            state["result_rows"] = drain_output(op, to_host=small)
            return
        exchange_and_final(op, final, small, repart, tdev, state)

    def barrier():
        if world > 1:
            import torch.distributed as td
            td.barrier()
        torch.cuda.synchronize(device)

    def launches():
        n = op.stats().kernel_launches
        return n + (final.stats().kernel_launches if final is not None else 0)

    for _ in range(warmup):
        one_step()
    sampler = ClockSampler(device)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    nl = 0
    for _ in range(steps):
        one_step()
        nl += launches()
    barrier()
    e1.record()
    e1.synchronize()
    wall = time.perf_counter() - t0
    dev_ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    # dominant kernel alone: per-launch CUDA-event durations on the pipeline's stream (pgpu_set_timing brackets every page-kernel launch)
    op.setTiming(True)
    kern_ms, kern_launches = 0.0, 0
    ksteps = max(3, min(steps, 5))
    for _ in range(ksteps):
        op.reset()
        op.addDeviceInput(dp.cols, ncol, rows)
        op.sync()
        st = op.stats()
        kern_ms += st.device_ms
        kern_launches += st.kernel_launches
    op.setTiming(False)
    info = dict(dev_ms=dev_ms, wall_s=wall, launches=nl, clocks=clocks, kern_ms=kern_ms, kern_launches=kern_launches,
                kern_steps=ksteps, result_rows=state["result_rows"], table_capacity=op.stats().table_capacity,
                exchange_bytes=state["exchange_bytes"], private_groups=state.get("private_groups", 0),
                partition_sweeps=op.stats().partition_sweeps, streamed_runs=op.stats().streamed_runs)
    if repart is not None and not hasattr(repart, "exchange"):
        repart.close()
    if final is not None:
        final.close()
    return info, dp, op


def e2e_rows_for(args, wl, world, want):
    """rows of the host-page leg: bounded by host memory (every rank of the box keeps its shard in pageable memory)"""
    rows = want
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 64 << 30
    need = rows * wl["bytes_per_row"] * 1.1
    if need * world > 0.5 * avail:
        rows = int(0.5 * avail / world / (wl["bytes_per_row"] * 1.1))
    return rows


def run_e2e(args, wl, device, dp, dist, steps, full):
    """`e2e`: the same pass fed the way a Trino driver feeds an operator: pages of --page-rows positions in ordinary pageable
    memory (Java heap arrays cannot be pinned by the caller) through pgpu_add_page(flags = 0); the library coalesces them into
    super-pages, stages them through its pinned ring with worker threads and overlaps that with the DMA and the kernels. The
    H2D copy of every input byte, the exchange (N > 1) and the D2H read of the result are inside the timed region (wall clock
    around the public API call sequence, bracketed by barriers; the caller takes the max over ranks).
    `full`: also time 64 K-row pages, the wire format and the caller-pinned single-page ceiling."""
    import torch
    from presto_b200 import _native as N
    from presto_b200.operators import GpuPipelineOperator
    L = N.lib()
    world, rank = dist["world"], dist["rank"]
    rows = min(e2e_rows_for(args, wl, world, args.e2e_rows or wl["rows"]), dp.rows)
    if world > 1:                            # every rank runs the same number of rows
        import torch.distributed as td
        t = torch.tensor([rows], device="cuda", dtype=torch.int64)
        td.all_reduce(t, op=td.ReduceOp.MIN)
        rows = int(t.item())
    host = HostTable(device, wl, dp, rows)
    single = world == 1
    small = wl["name"] in ("q1", "q1dict", "q6")
    def sync_ranks():
        if world > 1:
            import torch.distributed as td
            td.barrier()

    sync_ranks()                                  # every rank measures its ceilings while all the others measure theirs
    ceiling = h2d_ceiling(device)
    sync_ranks()
    copy_ceiling = host_copy_ceiling()
    sync_ranks()
    rec = {"unit": "rows/s", "config": {"page_rows": args.page_rows, "pinned": False, "host_memory": "pageable numpy arrays, one per block array; pages are consecutive slices",
                                        "entry_point": "pgpu_add_page(flags=0)", "copy_threads": int(os.environ.get("PGPU_COPY_THREADS", "0")) or "library default",
                                        "rows_per_gpu": rows},
           "h2d_ceiling_gbs": ceiling, "host_copy_ceiling_gbs": copy_ceiling,
           "ceilings": "measured concurrently on all ranks: pinned cudaMemcpy H2D, and the staging copy alone (pageable -> pinned, no GPU); the leg needs both for every byte"}
    if single:
        op = GpuPipelineOperator(wl["plan"]("single"), device)
        r = time_host_pages(op, host.pages(args.page_rows), steps, to_host=True)
        rec.update(value=r["rows_per_s"], h2d_bytes_per_step=r["h2d_bytes_per_step"], d2h_bytes_per_step=r["d2h_bytes_per_step"],
                   rows_per_step=r["rows"], steps=steps, seconds_per_step=r["seconds_per_step"], super_pages_per_step=r["super_pages_per_step"],
                   launches_per_step=r["launches_per_step"], gb_per_s=r["rows_per_s"] * host.bytes_per_row / 1e9)
        rec["frac_of_h2d_ceiling"] = rec["gb_per_s"] / ceiling
        rec["frac_of_host_copy_ceiling"] = rec["gb_per_s"] / copy_ceiling
        if full:
            variants = {}
            r = time_host_pages(op, host.pages(65536), max(1, steps - 1), to_host=True)
            variants["page_rows_65536"] = {"value": r["rows_per_s"], "gb_per_s": r["rows_per_s"] * host.bytes_per_row / 1e9, "pages_per_step": r["pages"],
                                           "super_pages_per_step": r["super_pages_per_step"]}
            srows = min(rows, 96_000_000)
            payloads = host.serialized_pages(args.page_rows, srows)
            r = time_serialized_pages(op, payloads, max(1, steps - 1), to_host=True)
            variants["serialized_pages"] = {"value": r["rows_per_s"], "gb_per_s": r["rows_per_s"] * host.bytes_per_row / 1e9, "rows_per_step": srows,
                                            "page_rows": args.page_rows, "entry_point": "pgpu_add_serialized_page", "super_pages_per_step": r["super_pages_per_step"]}
            del payloads
            rec["variants"] = variants
            # ceiling of the path: ONE caller-pinned page (a shape no Trino caller produces): DMA straight from the caller's buffers
            prow = min(rows, 128_000_000)
            pinned, bufs = (N.Column * host.ncol)(), []
            ctypes.memmove(pinned, host.pages(prow, prow)[0][0], ctypes.sizeof(pinned))
            for i, a in enumerate(host.arrays):
                for field, key, cnt in (("values", "values", prow), ("offsets", "offsets", prow + 1), ("ids", "ids", prow)):
                    if key in a:
                        src = a[key][:cnt]
                        hp = L.pgpu_host_alloc(max(src.nbytes, 16))
                        if not hp:
                            raise MemoryError("pinned allocation failed")
                        ctypes.memmove(hp, src.ctypes.data, src.nbytes)
                        bufs.append(hp)
                        setattr(pinned[i], field, hp)
            r = time_host_pages(op, [(pinned, prow)], max(1, steps - 1), to_host=True, flags=N.PAGE_PINNED)
            rec["ceiling_pinned_single_page"] = {"value": r["rows_per_s"], "gb_per_s": r["rows_per_s"] * host.bytes_per_row / 1e9, "rows_per_step": prow}
            op.sync()
            for hp in bufs:
                L.pgpu_host_free(hp)
        op.close()
        return rec, host
    # N > 1: PARTIAL over the host pages, then the exchange and FINAL
    op = GpuPipelineOperator(wl["plan"]("partial"), device)
    final = GpuPipelineOperator(final_plan_for(wl, None), device)
    tdev = torch.device("cuda", device)
    repart = make_exchange(args, device, world, rank, small, "e2e_" + wl["name"])
    state = {"result_rows": 0, "exchange_bytes": 0}
    pages = host.pages(args.page_rows)
    ncol = host.ncol

    def barrier():
        import torch.distributed as td
        td.barrier()
        torch.cuda.synchronize(device)

    def one_step():
        op.reset()
        for cols, n in pages:
            op._check(L.pgpu_add_page(op._h, cols, ncol, n, 0))
        op.finish()
        exchange_and_final(op, final, small, repart, tdev, state)
        if not small:                         # the device-resident result is read back: that is the step's D2H
            final.sync()

    one_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_step()
    barrier()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device="cuda", dtype=torch.float64)
    import torch.distributed as td
    td.all_reduce(t, op=td.ReduceOp.MAX)
    dt = float(t.item())
    t = torch.tensor([ceiling, copy_ceiling], device="cuda", dtype=torch.float64)
    td.all_reduce(t, op=td.ReduceOp.SUM)
    st = op.stats()
    d2h = st.d2h_bytes + final.stats().d2h_bytes
    rec.update(value=rows * world * steps / dt, h2d_bytes_per_step=int(st.h2d_bytes) * world, d2h_bytes_per_step=int(d2h) * world,
               rows_per_step=rows * world, steps=steps, seconds_per_step=dt / steps, gb_per_s=rows * world * steps * host.bytes_per_row / dt / 1e9,
               h2d_ceiling_gbs=float(t[0].item()), host_copy_ceiling_gbs=float(t[1].item()))
    rec["frac_of_h2d_ceiling"] = rec["gb_per_s"] / rec["h2d_ceiling_gbs"]
    rec["frac_of_host_copy_ceiling"] = rec["gb_per_s"] / rec["host_copy_ceiling_gbs"]
    op.close()
    if repart is not None and not hasattr(repart, "exchange"):
        repart.close()
    final.close()
    return rec, host


def run_workload(args, name, device, dist, headline):
    """One record: device-resident value + roofline, e2e over host pages, verification against the CPU oracle."""
    import torch
    world, rank = dist["world"], dist["rank"]
    rows = args.rows
    wl = workload_def(name, rows, args.sf * world if args.data == "dbgen" else None)
    if args.data == "dbgen":
        wl["dbgen"] = args.dbgen
    steps = args.steps if headline else max(3, min(args.steps, 10))
    warmup = args.warmup if headline else 3
    info, dp, op = run_device_steps(args, wl, device, dist, steps, warmup)
    ms_per_step = info["dev_ms"] / steps
    total_rows = rows * world
    if world > 1:
        import torch.distributed as td
        t = torch.tensor([ms_per_step], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        ms_per_step = float(t.item())
        if args.data == "dbgen":        # order-range shards differ by a few rows
            t = torch.tensor([rows], device="cuda", dtype=torch.int64)
            td.all_reduce(t)
            total_rows = int(t.item())
    value = total_rows / (ms_per_step / 1000.0)
    peak, peak_src = measured_peak()
    kern_ms_step = info["kern_ms"] / info["kern_steps"]
    launches_step = info["kern_launches"] / info["kern_steps"]
    alg_bytes_step = rows * wl["bytes_per_row"]
    achieved = alg_bytes_step / (kern_ms_step / 1e3) / 1e9
    kname = "pgpu_k_page_staged (cp.async.bulk ring + register-resident groups)" if name in ("q1", "q1dict") else "pgpu_k_page (fused scan+filter+project+aggregate, direct loads)"
    if info.get("partition_sweeps"):
        kname = "pgpu_k_part_count + pgpu_k_part_scatter + pgpu_k_page_sweep (rows reordered by table slice, then the fused kernel with L2-resident table slices; kernel_ms_per_step is the sum of the three)"
    if info.get("streamed_runs"):
        kname = ("pgpu_k_runs_count + pgpu_k_runs_write (streaming runs: the keys arrive in key order, one buffer row per run, no hash table; "
                 "kernel_ms_per_step is the sum of the two and of the offsets scan between them; they read the rows twice, the algorithmic bytes count them once)")
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "bytes_per_launch": alg_bytes_step / max(launches_step, 1), "avg_launch_ms": kern_ms_step / max(launches_step, 1),
                "kernel_ms_per_step": kern_ms_step, "launches_per_step": launches_step, "kernel": kname, "peak_source": peak_src,
                "algorithmic_bytes_per_row": wl["bytes_per_row"]}
    tr = measured_traffic(name)
    if tr:
        dram_step = tr["dram_bytes_per_row"] * rows
        roofline["traffic"] = dram_step / max(launches_step, 1)
        roofline["traffic_source"] = tr.get("source")
        roofline["achieved_dram"] = dram_step / (kern_ms_step / 1e3) / 1e9
        roofline["frac_dram"] = roofline["achieved_dram"] / peak
        if name == "q6":   # the kernel skips extendedprice sectors of unselected rows: its roofline position is the DRAM bytes it really moves
            roofline["frac_algorithmic"] = roofline["frac"]
            roofline["frac"] = roofline["frac_dram"]
    config = {"workload": wl["label"], "rows_per_gpu": rows, "scale_factor_equivalent": round(rows / 6_001_215, 2),
              "bytes_per_row": wl["bytes_per_row"], "partitioning": "row-range shard per GPU, partial -> final aggregation",
              "l2_policy": "inputs (%.1f GB per GPU) are far larger than the 126 MB L2: no flush needed" % (rows * wl["bytes_per_row"] / 1e9)}
    rec = {"metric": "lineitem_rows_per_second", "value": value, "unit": "rows/s", "ms_per_step": ms_per_step, "steps": steps, "warmup": warmup,
           "gb_per_s": value * wl["bytes_per_row"] / 1e9, "roofline": roofline, "config": config, "gpu_launches": info["launches"],
           "clocks": info["clocks"], "table_capacity_slots": info["table_capacity"], "result_rows": info["result_rows"],
           "partition_sweeps_per_step": info.get("partition_sweeps", 0), "streamed_runs_per_step": info.get("streamed_runs", 0)}
    if world > 1:
        small = name in ("q1", "q1dict", "q6")
        rec["exchange"] = {"shape": ("gather: push kernel stores the partial rows into every rank's inbox over NVLink, FINAL gather kernel merges them; no host round trip (pgpu_exchange_partial)" if args.exchange == "peer" else "gather: one NCCL all-gather of packed partial pages, FINAL on every rank") if small else ("hash repartition; FINAL kernel reads each peer's partition over NVLink (CUDA IPC peer memory, pgpu_exchange_partial)" if args.exchange == "peer" else "hash repartition + all_to_all_single per column + FINAL per rank"),
                           "bytes_sent_per_rank_per_step": info["exchange_bytes"],
                           "groups_finalised_in_place_per_rank": info.get("private_groups", 0)}
    host = None
    if not args.no_e2e:
        if not headline and not args.e2e_rows:
            args_e2e_rows, args.e2e_rows = args.e2e_rows, min(rows, 64_000_000)   # sub-records: a 64 M-row sample of the shard
            rec["e2e"], host = run_e2e(args, wl, device, dp, dist, steps=2, full=False)
            args.e2e_rows = args_e2e_rows
        else:
            rec["e2e"], host = run_e2e(args, wl, device, dp, dist, steps=max(2, min(steps, 3)), full=headline)
    # correctness of the timed rows + the CPU baseline (rank 0, N = 1)
    if rank == 0 and world == 1 and not args.no_cpu and args.verify_rows > 0:
        vrows = min(args.verify_rows if headline else min(args.verify_rows, 16_000_000), rows)
        if host is None or host.rows < vrows:
            host = HostTable(device, wl, dp, vrows)
        cores = os.cpu_count() or 1
        v, secs = verify_prefix(wl, op, dp, host, vrows, cores)
        rec.update(verified=v["verified"], verified_rows=v["verified_rows"], verification=v)
        cpu = {"value": vrows / secs, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": "the first %d rows of the timed shard (host copy of the same bytes), 1 M-row pages, one PARTIAL pipeline per thread + FINAL merge" % vrows}
        if headline:
            srows = min(vrows, 4_000_000)
            s1, _ = oracle_run(wl, host.pages(1_000_000, srows), 1)
            cpu["single_thread_value"] = srows / s1
            cpu["single_thread_sample"] = "%d rows, one thread (the harness's one split / one driver)" % srows
        rec["cpu_baseline"] = cpu
    op.close()
    dp.free()
    del host
    return rec


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    if "PGPU_COPY_THREADS" not in os.environ:   # staging-copy workers of this process: the box's cores shared by its ranks
        os.environ["PGPU_COPY_THREADS"] = str(max(2, min(16, cores // max(world, 1))))
    if args.data == "dbgen":
        from presto_b200 import tpch
        per_gpu = int(1_500_000 * args.sf)
        total_sf = args.sf * world
        first_order = rank * per_gpu
        n_orders = min(per_gpu, tpch.tpch_order_count(total_sf) - first_order)
        if args.impl == "ours":
            args.dbgen = (total_sf, first_order, n_orders)
            args.rows = tpch.tpch_rows(*args.dbgen)
    data = "synthetic" if args.data == "synthetic" else "synthetic: TPC-H dbgen rows (the tpch connector's generator, restated; sf %g per GPU)%s" % (args.sf, ", generated on the device" if args.impl == "ours" else "")
    headline_name = args.workload or "q1"

    if args.impl == "reference":
        # the reference's CPU implementation of the path: Java, no JVM here -> the oracle port on all host cores
        if rank != 0:
            return
        # build what this arm needs without mapping the product library: the oracle, and libprestogpu_bench.so for the input rows
        from presto_b200 import build as B
        B.build()
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
        wl = workload_def(headline_name, args.cpu_rows_per_core, args.sf * world if args.data == "dbgen" else None)
        per_core = args.cpu_rows_per_core
        vals = []
        for i in range(args.warmup + args.steps):
            v, total, _ = cpu_run(wl, per_core, cores)
            if i >= args.warmup:
                vals.append(v)
        value = sum(vals) / len(vals)
        sample = "%d rows (%d per core x %d cores) of the same synthetic lineitem per step, 1 M-row pages, partial per thread + final merge" % (per_core * cores, per_core, cores)
        config = {"workload": wl["label"], "rows_per_gpu": args.rows, "rows_per_step_timed": per_core * cores, "bytes_per_row": wl["bytes_per_row"],
                  "partitioning": "row-range shard per host thread, partial -> final aggregation"}
        line = {"impl": "reference", "metric": "lineitem_rows_per_second", "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1000.0 * per_core * cores / value, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": data, "config": config,
                "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "reference Java benchmarks (HandTpchQuery1/6) cannot run: no JVM in this image; this is the C++ oracle port"}
        print(json.dumps(line))
        return

    import torch
    device = local_rank
    torch.cuda.set_device(device)
    dist = {"rank": rank, "world": world}
    if world > 1:
        import torch.distributed as td
        td.init_process_group("nccl", device_id=torch.device("cuda", device))
    rec = run_workload(args, headline_name, device, dist, headline=True)
    configs = {}
    if args.workload is None and not args.no_configs and args.data == "synthetic":
        for name in (SUB_CONFIGS if world == 1 else ("highcard",)):
            try:
                configs[name] = run_workload(args, name, device, dist, headline=False)
            except Exception as e:  # a failing sub-record must not take the headline with it
                configs[name] = {"error": "%s: %s" % (type(e).__name__, e)}
                if world > 1:
                    raise
    if rank == 0:
        line = {"metric": rec["metric"], "value": rec["value"], "unit": "rows/s", "n_gpus": world, "steps": rec["steps"], "warmup": rec["warmup"],
                "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": data}
        for k in ("config", "gb_per_s", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks", "table_capacity_slots", "result_rows",
                  "verified", "verified_rows", "verification", "exchange"):
            line[k] = rec.get(k)
        if configs:
            line["configs"] = configs
        print(json.dumps(line))
    if "c" in _COMM:
        _COMM.pop("c").close()
    if world > 1:
        import torch.distributed as td
        td.destroy_process_group()


if __name__ == "__main__":
    main()
