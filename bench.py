#!/usr/bin/env python
"""bench.py — TPC-H Q1 (BASELINE.json configs[1]) lineitem rows/s on N B200s, with the kernel roofline and a CPU
baseline. One "step" = one full pass of the fused scan->filter->project->hash-aggregate pipeline over the
synthetic lineitem table (device-resident for `value`, pinned host pages for `e2e`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload q1|q6|highcard|q1dict]
                  [--rows R]
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SF100_ROWS = 600_037_902  # TPC-H lineitem at sf100
# dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the ncu --set full captures under
# profiles/ (bytes; None = not captured for that workload)
# stored as measured DRAM bytes / algorithmic bytes of the captured launch (profiles/r01e_q1_staged_ncu_raw.txt)
TRAFFIC_RATIO = {"q1": 1.0014}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="q1", choices=["q1", "q6", "highcard", "highcard_shuffled", "q1dict"])
    ap.add_argument("--rows", type=int, default=SF100_ROWS, help="lineitem rows per GPU")
    ap.add_argument("--data", default="synthetic", choices=["synthetic", "dbgen"],
                    help="synthetic: counter-based columns with the TPC-H value distributions (--rows per GPU); dbgen: the tpch connector's real rows "
                         "(spec-conformant generator, --sf per GPU; rows per GPU follow from the order range)")
    ap.add_argument("--sf", type=float, default=100.0, help="--data dbgen: TPC-H scale factor per GPU")
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows of the end-to-end (host pages) leg; 0 = same as --rows if RAM allows")
    ap.add_argument("--cpu-rows-per-core", type=int, default=4_000_000)
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1, large group counts: FINAL reads the peers' partitions over NVLink (peer) or NCCL all-to-all first (nccl)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def workload_def(name, rows, dbgen_sf=None):
    wl = _workload_def(name, rows)
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


# ------------------------------------------------------------------------------------------------ CPU baseline
def cpu_run(wl, rows_per_core, cores, repeats=1):
    """The CPU oracle (oracle/, the C++ restatement of the reference's loops — the reference itself is Java and there
    is no JVM on this box) on all host cores: one PARTIAL pipeline per thread over its own row range, page-at-a-time
    (1 M-row pages), then one FINAL pipeline over the partial outputs (the analogue of task_concurrency drivers +
    the partial->final exchange). Data generation is outside the timed region. Returns rows/s."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_op import OracleOperator
    from presto_b200.operators import toPages
    from presto_b200.spi import Page
    total = rows_per_core * cores
    page_rows = 1_000_000
    shards = []
    for t in range(cores):
        pages, r0 = [], t * rows_per_core
        end = r0 + rows_per_core
        while r0 < end:
            n = min(page_rows, end - r0)
            pages.append(wl["page"](r0, n, total))
            r0 += n
        shards.append(pages)
    total = sum(p.position_count for pages in shards for p in pages)
    partial_plan = wl["plan"]("partial")
    best = None
    result = None
    for _ in range(repeats):
        outs = [None] * cores

        def work(t):
            op = OracleOperator(partial_plan)
            outs[t] = toPages(op, shards[t])
            op.close()

        t0 = time.perf_counter()
        th = [threading.Thread(target=work, args=(t,)) for t in range(cores)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        # FINAL step over the partial pages
        from presto_b200.tpch import plan_text  # noqa: F401
        final_plan = final_plan_for(wl, partial_plan)
        fop = OracleOperator(final_plan)
        pages = [p for o in outs for p in o]
        result = toPages(fop, pages)
        fop.close()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return total / best, total, result


def final_plan_for(wl, partial_plan):
    """FINAL-step plan whose input channels are the PARTIAL plan's output channels."""
    from presto_b200.operators import Aggregation, GpuPipelineOperator, plan_text
    import re
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


# ------------------------------------------------------------------------------------------------ GPU legs
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


def make_exchange(args, device, world, rank, small, leg):
    from presto_b200 import exchange as X
    if world == 1:
        return None
    if not small:
        return X.DeviceRepartition(device, world, rank, transport=args.exchange)
    if args.exchange == "peer":
        return X.PeerGather(device, world, rank, "pgpu_mail_%s_%s" % (os.environ.get("MASTER_PORT", "0"), leg))
    return None


def exchange_and_final(op, final, small, repart, tdev, state):
    """N > 1: the partial -> final exchange of one step (op has finished its shard); leaves the result drained.
    `repart` is the exchange object: PeerGather / None (small results), DeviceRepartition (large group counts)."""
    from presto_b200 import exchange as X
    trace = state.get("trace")
    t0 = time.perf_counter()
    final.reset()
    t1 = time.perf_counter()
    if small and repart is not None:
        if trace is not None:
            op.sync()
            trace["kernels"] = trace.get("kernels", 0.0) + time.perf_counter() - t1
            t1 = time.perf_counter()
        repart.run(op, final)                       # FINAL kernel reads every rank's partial rows over NVLink
        t2 = time.perf_counter()
        state["result_rows"] = drain_output(final, to_host=True)
        if trace is not None:
            trace["peer_gather+final"] = trace.get("peer_gather+final", 0.0) + t2 - t1
            trace["final_output"] = trace.get("final_output", 0.0) + time.perf_counter() - t2
    elif small:
        if trace is not None:
            op.sync()
            trace["kernels"] = trace.get("kernels", 0.0) + time.perf_counter() - t1
            t1 = time.perf_counter()
        frame = X.pack_output(op)                                       # waits for the shard's kernels, drains the partial page
        t2 = time.perf_counter()
        frames = X.all_gather_frames(frame, device=tdev)                # one NCCL all-gather of a 32 KB frame per rank
        t3 = time.perf_counter()
        for f in frames:
            X.add_packed(final, f)                                      # FINAL step: addIntermediate over every rank's partial page
        final.finish()
        t4 = time.perf_counter()
        state["result_rows"] = drain_output(final, to_host=True)
        t5 = time.perf_counter()
        if trace is not None:
            for k, v in (("final_reset", t1 - t0), ("pack_output", t2 - t1), ("all_gather", t3 - t2), ("final_add", t4 - t3), ("final_output", t5 - t4)):
                trace[k] = trace.get(k, 0.0) + v
    else:
        repart.run(op, final)
        state["result_rows"] = drain_output(final, to_host=False)
        state["exchange_bytes"] = repart.sent_bytes


def run_device_steps(args, wl, device, dist):
    """`value`: inputs resident in HBM before the timed region. One step = reset, one pass of the fused pipeline over the
    rank's shard, and (N > 1) the partial -> final exchange. Timed with CUDA events recorded on an idle device on both
    sides (barrier + synchronize), max over ranks taken by the caller."""
    import torch
    from presto_b200 import _native as N
    from presto_b200 import exchange as X
    from presto_b200 import tpch
    from presto_b200.operators import GpuPipelineOperator, toPages
    L = N.lib()
    rows = args.rows
    world, rank = dist["world"], dist["rank"]
    first = rank * rows
    total = rows * world
    dp = tpch.DevicePage(device, wl["spec"], rows, first_row=first, total_rows=total, dbgen=getattr(args, "dbgen", None))
    assert dp.rows == rows
    single = world == 1
    op = GpuPipelineOperator(wl["plan"]("single" if single else "partial"), device)
    final = None if single else GpuPipelineOperator(final_plan_for(wl, None), device)
    ncol = len(wl["spec"])
    small = wl["name"] in ("q1", "q1dict", "q6")   # result is a handful of rows: host output / gather exchange
    tdev = torch.device("cuda", device)
    repart = make_exchange(args, device, world, rank, small, "value")
    state = {"result_rows": 0, "exchange_bytes": 0}
    if os.environ.get("PGPU_BENCH_TRACE"):
        state["trace"] = {}

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

    for _ in range(args.warmup):
        one_step()
    if state.get("trace") is not None:
        state["trace"].clear()
    sampler = ClockSampler(device)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    nl = 0
    for _ in range(args.steps):
        one_step()
        nl += launches()
    barrier()
    e1.record()
    e1.synchronize()
    wall = time.perf_counter() - t0
    dev_ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    if state.get("trace") and rank == 0:
        print("exchange trace (ms/step): " + ", ".join("%s %.3f" % (k, v * 1e3 / args.steps) for k, v in state["trace"].items()), file=sys.stderr)
    # dominant kernel alone: per-launch CUDA-event durations (pgpu_set_timing brackets every page-kernel launch)
    op.setTiming(True)
    kern_ms, kern_launches = 0.0, 0
    ksteps = max(3, min(args.steps, 5))
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
                exchange_bytes=state["exchange_bytes"])
    if repart is not None:
        repart.close()
    if final is not None:
        final.close()
    return info, dp, op


def run_e2e_steps(args, wl, device, dp_device, steps, warmup, dist):
    """`e2e`: the same pass through pgpu_add_page with HOST pages in pinned memory; H2D copies, the exchange (N > 1) and the
    D2H read of the result are inside the timed region (wall clock around the public API call sequence, bracketed by
    barriers; the caller takes the max over ranks)."""
    import numpy as np
    import torch
    from presto_b200 import _native as N
    from presto_b200 import exchange as X
    from presto_b200 import tpch
    from presto_b200.operators import GpuPipelineOperator
    L = N.lib()
    world, rank = dist["world"], dist["rank"]
    rows = args.e2e_rows or args.rows
    # bound by available host memory (pinned): keep below 40 % of MemAvailable
    need = rows * wl["bytes_per_row"] * 1.05
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 64 << 30
    if need * world > 0.4 * avail:           # all ranks pin host memory of the same box
        rows = int(0.4 * avail / world / (wl["bytes_per_row"] * 1.05))
    rows = min(rows, dp_device.rows)
    if world > 1:                            # every rank runs the same number of rows
        import torch.distributed as td
        t = torch.tensor([rows], device="cuda", dtype=torch.int64)
        td.all_reduce(t, op=td.ReduceOp.MIN)
        rows = int(t.item())
    # host pages = the device columns copied back into pinned host buffers (identical bytes, no slow host generation)
    cols = (N.Column * len(wl["spec"]))()
    hostbufs = []
    h2d = 0

    def pinned_copy(dev_ptr, nbytes):
        nonlocal h2d
        hp = L.pgpu_host_alloc(max(nbytes, 16))
        if not hp:
            raise MemoryError("pinned allocation of %d bytes failed" % nbytes)
        hostbufs.append(hp)
        if L.pgpu_memcpy_d2h(device, hp, dev_ptr, nbytes):
            raise RuntimeError("d2h copy failed")
        h2d += nbytes
        return hp

    width = {0: 1, 1: 8, 2: 4, 3: 4, 4: 8}
    dict_cols = []
    for i, s in enumerate(wl["spec"]):
        src, dst = dp_device.cols[i], cols[i]
        ctypes.memmove(ctypes.byref(dst), ctypes.byref(src), ctypes.sizeof(N.Column))
        dst.position_count = rows
        t = s[0]
        if s[1] == "plain":
            dst.values = pinned_copy(src.values, rows * width[t.kind])
        elif s[1] == "varchar1":
            dst.values = pinned_copy(src.values, rows)
            dst.offsets = pinned_copy(src.offsets, (rows + 1) * 4)
        else:  # dict: ids from device, dictionary block straight from the host block
            dst.ids = pinned_copy(src.ids, rows * 4)
            d = N.Column()
            keep = []
            s[3]._fill(d, keep)
            dict_cols.append((d, keep))
            dst.dictionary = ctypes.pointer(d)
    single = world == 1
    op = GpuPipelineOperator(wl["plan"]("single" if single else "partial"), device)
    final = None if single else GpuPipelineOperator(final_plan_for(wl, None), device)
    small = wl["name"] in ("q1", "q1dict", "q6")
    tdev = torch.device("cuda", device)
    repart = make_exchange(args, device, world, rank, small, "e2e")
    state = {"result_rows": 0, "exchange_bytes": 0}

    def one_step():
        op.reset()
        op._check(L.pgpu_add_page(op._h, cols, len(wl["spec"]), rows, N.PAGE_PINNED))
        op.finish()
        if single:
            return drain_output(op, to_host=True)
        exchange_and_final(op, final, small, repart, tdev, state)
        if not small:                         # the device-resident result is read back: that is the step's D2H
            final.sync()
        return state["result_rows"]

    def barrier():
        if world > 1:
            import torch.distributed as td
            td.barrier()
        torch.cuda.synchronize(device)

    for _ in range(warmup):
        one_step()
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(steps):
        one_step()
        d2h = op.stats().d2h_bytes + (final.stats().d2h_bytes if final is not None else 0)
    barrier()
    dt = time.perf_counter() - t0
    st = op.stats()
    op.close()
    if repart is not None:
        repart.close()
    if final is not None:
        final.close()
    for hp in hostbufs:
        L.pgpu_host_free(hp)
    return dict(rows=rows, seconds=dt, steps=steps, h2d_bytes_per_step=int(st.h2d_bytes), d2h_bytes_per_step=int(d2h))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.data == "dbgen":
        from presto_b200 import tpch
        per_gpu = int(1_500_000 * args.sf)
        total_sf = args.sf * world
        first_order = rank * per_gpu
        n_orders = min(per_gpu, tpch.tpch_order_count(total_sf) - first_order)
        if args.impl == "ours":
            args.dbgen = (total_sf, first_order, n_orders)
            args.rows = tpch.tpch_rows(*args.dbgen)
    wl = workload_def(args.workload, args.rows, args.sf * world if args.data == "dbgen" else None)
    wl["dbgen_sf"] = args.sf * world if args.data == "dbgen" else None
    data = "synthetic" if args.data == "synthetic" else "synthetic: TPC-H dbgen rows (the tpch connector's generator, restated; sf %g per GPU)%s" % (args.sf, ", generated on the device" if args.impl == "ours" else "")
    wl["name"] = args.workload
    wl["rows"] = args.rows if args.impl == "ours" else args.cpu_rows_per_core
    cores = os.cpu_count() or 1
    metric = "lineitem_rows_per_second"
    config = {"workload": wl["label"], "rows_per_gpu": args.rows, "scale_factor_equivalent": round(args.rows / 6_001_215, 2),
              "bytes_per_row": wl["bytes_per_row"], "partitioning": "row-range shard per GPU, partial -> final aggregation",
              "l2_policy": "inputs (%.1f GB per GPU) are far larger than the 126 MB L2: no flush needed" % (args.rows * wl["bytes_per_row"] / 1e9)}

    if args.impl == "reference":
        # the reference's CPU implementation of the path: Java, no JVM here -> the oracle port on all host cores
        if rank != 0:
            return
        from __graft_entry__ import build
        build()
        per_core = args.cpu_rows_per_core
        vals = []
        for i in range(args.warmup + args.steps):
            v, total, _ = cpu_run(wl, per_core, cores)
            if i >= args.warmup:
                vals.append(v)
        value = sum(vals) / len(vals)
        sample = "%d rows (%d per core x %d cores) of the same synthetic lineitem per step, 1 M-row pages, partial per thread + final merge" % (per_core * cores, per_core, cores)
        line = {"impl": "reference", "metric": metric, "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps,
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
    info, dp, op = run_device_steps(args, wl, device, dist)
    ms_per_step = info["dev_ms"] / args.steps
    if world > 1:
        import torch.distributed as td
        t = torch.tensor([ms_per_step], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        ms_per_step = float(t.item())
    total_rows = args.rows * world
    if world > 1 and args.data == "dbgen":        # order-range shards differ by a few rows
        t = torch.tensor([args.rows], device="cuda", dtype=torch.int64)
        td.all_reduce(t)
        total_rows = int(t.item())
    value = total_rows / (ms_per_step / 1000.0)
    peak, peak_src = measured_peak()
    kern_rows = args.rows * info["kern_steps"]
    achieved = kern_rows * wl["bytes_per_row"] / (info["kern_ms"] / 1000.0) / 1e9
    kname = "pgpu_k_page_staged (cp.async.bulk ring + register-resident groups)" if args.workload in ("q1", "q1dict") else "pgpu_k_page (fused scan+filter+project+aggregate, direct loads)"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": (TRAFFIC_RATIO[args.workload] * kern_rows * wl["bytes_per_row"] / max(info["kern_launches"], 1)) if args.workload in TRAFFIC_RATIO else None,
                "bytes_per_launch": kern_rows * wl["bytes_per_row"] / max(info["kern_launches"], 1),
                "avg_launch_ms": info["kern_ms"] / max(info["kern_launches"], 1),
                "kernel": kname, "peak_source": peak_src,
                "launches_timed": info["kern_launches"], "algorithmic_bytes_per_row": wl["bytes_per_row"]}
    e2e = None
    if not args.no_e2e:
        e = run_e2e_steps(args, wl, device, dp, steps=max(2, min(args.steps, 5)), warmup=1, dist=dist)
        if world > 1:
            t = torch.tensor([e["seconds"]], device="cuda", dtype=torch.float64)
            td.all_reduce(t, op=td.ReduceOp.MAX)
            e["seconds"] = float(t.item())
        e2e = {"value": e["rows"] * world * e["steps"] / e["seconds"], "unit": "rows/s", "h2d_bytes_per_step": e["h2d_bytes_per_step"] * world,
               "d2h_bytes_per_step": e["d2h_bytes_per_step"] * world, "rows_per_step": e["rows"] * world, "steps": e["steps"],
               "gb_per_s": e["h2d_bytes_per_step"] * world * e["steps"] / e["seconds"] / 1e9}
    op.close()
    dp.free()
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, total, _ = cpu_run(wl, args.cpu_rows_per_core, cores)
        cpu = {"value": v, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": "%d rows (%d per core) of the same synthetic lineitem, 1 M-row pages, partial per thread + final merge" % (total, args.cpu_rows_per_core)}
    if rank == 0:
        line = {"metric": metric, "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": data, "config": config, "gb_per_s": value * wl["bytes_per_row"] / 1e9,
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": info["launches"], "clocks": info["clocks"],
                "table_capacity_slots": info["table_capacity"], "result_rows": info["result_rows"],
                "exchange": None if world == 1 else {"shape": ("gather: every rank's FINAL kernel reads all partial pages over NVLink (CUDA IPC peer memory, shared-memory mailbox)" if args.exchange == "peer" else "gather: one NCCL all-gather of packed partial pages, FINAL on every rank") if args.workload in ("q1", "q1dict", "q6") else ("hash repartition; FINAL kernel reads each peer's partition over NVLink (CUDA IPC peer memory)" if args.exchange == "peer" else "hash repartition + all_to_all_single per column + FINAL per rank"),
                                                     "bytes_sent_per_rank_per_step": info["exchange_bytes"]}}
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as td
        td.destroy_process_group()


if __name__ == "__main__":
    main()



