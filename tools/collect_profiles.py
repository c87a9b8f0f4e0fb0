#!/usr/bin/env python
"""Copy the judged summaries of the last evidence visit (tools/gpu_trip_r2i.sh) from gpurun_out/ (scratch) into profiles/ (tracked):
bench JSON lines, ncu launch lists, per capture the details page, selected raw metrics and the SASS hot-spot summary, and
profiles/traffic.json — measured DRAM bytes per row of every workload's dominant kernel, which bench.py reads for `roofline.traffic`.
Usage: python tools/collect_profiles.py <tag>   (e.g. r02)"""
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1]
os.makedirs(P, exist_ok=True)

# rows of the captured (steady-state, full-chunk) launch: hash-aggregation pages are launched in chunks of kHashChunkRows = 64 Mi rows
# (csrc/pipeline.cpp), rounded down to whole staged tiles (Q1: 1920 rows, dictionary Q1: 960); a global aggregation is one launch
CHUNK = 64 << 20
LAUNCH_ROWS = {"q1": CHUNK // 1920 * 1920, "q1dict": CHUNK // 960 * 960, "q6": 600_037_902, "highcard_table": CHUNK, "highcard_shuffled": CHUNK,
               "highcard_buckets": CHUNK, "highcard_shuffled_scatter": 600_037_902, "highcard_shuffled_plain": CHUNK,
               "highcard_runs_count": 600_037_902, "highcard_runs_write": 600_037_902}
ALG_BYTES = {"q1": 46, "q1dict": 32, "q6": 28, "highcard_table": 16, "highcard_shuffled": 16, "highcard_buckets": 16, "highcard_shuffled_scatter": 32,
             "highcard_shuffled_plain": 16, "highcard_runs_count": 16, "highcard_runs_write": 16}
KEEP = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_atom.sum", "lts__t_sectors_srcunit_tex_op_red.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio")


def copy(name, dst=None):
    src = os.path.join(G, name)
    if os.path.exists(src) and os.path.getsize(src):
        shutil.copy(src, os.path.join(P, "%s_%s" % (tag, dst or name)))
        return True
    return False


for f in ("bench_default.json", "bench_ref.json", "bench_highcard_buckets.json", "bench_highcard_buckets_casfirst.json", "randtable.txt",
          "bench_highcard_shuffled_nosweep.json", "parquet_ingest_NONE.json", "parquet_ingest_SNAPPY.json", "parquet_ingest_LZ4.json", "parquet_ingest_GZIP.json",
          "launches_q1.csv", "launches_highcard.csv", "ingest_serialized.json"):
    copy(f)
for t in (16, 8, 4, 2, 1):
    copy("ingest_t%d.json" % t)

traffic = {}
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
lines = []
for name in ("q1", "q6", "q1dict", "highcard_table", "highcard_shuffled", "highcard_shuffled_scatter", "highcard_shuffled_plain", "highcard_buckets", "highcard_emit",
             "highcard_runs_count", "highcard_runs_write", "highcard_runs_emit"):
    rep = os.path.join(G, name + "_full.ncu-rep")
    if not os.path.exists(rep):
        continue
    det = subprocess.run(["ncu", "-i", rep, "--page", "details"], stdout=subprocess.PIPE, text=True).stdout
    open(os.path.join(P, "%s_%s_ncu_details.txt" % (tag, name)), "w").write(det)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    vals = {}
    with open(os.path.join(P, "%s_%s_ncu_raw.txt" % (tag, name)), "w") as out:
        for i, h in enumerate(r[0]):
            if h in KEEP:
                out.write("%s [%s] = %s\n" % (h, r[1][i], r[2][i]))
                vals[h] = (r[1][i], r[2][i])
    hot = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_hot.py"), rep], stdout=subprocess.PIPE, text=True).stdout
    open(os.path.join(P, "%s_%s_ncu_hot.txt" % (tag, name)), "w").write(hot)

    def to_bytes(key):
        unit, v = vals.get(key, ("byte", "0"))
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)
        return float(v.replace(",", "")) * scale

    def to_ms(key):
        unit, v = vals.get(key, ("ms", "0"))
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1.0)
        return float(v.replace(",", "")) * scale

    if name in LAUNCH_ROWS:
        dram = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
        ms = to_ms("gpu__time_duration.sum")
        rows = LAUNCH_ROWS[name]
        traffic[name] = {"dram_bytes_per_row": dram / rows, "dram_bytes_per_launch": dram, "launch_rows": rows, "launch_ms_under_ncu": ms,
                         "algorithmic_bytes_per_row": ALG_BYTES[name], "ratio_to_algorithmic": dram / rows / ALG_BYTES[name],
                         "dram_gbs_under_ncu": dram / ms / 1e6 if ms else None,
                         "source": "profiles/%s_%s_ncu_raw.txt (ncu --set full --clock-control none, one steady-state launch)" % (tag, name)}
        lines.append("| %s | %.0f M | %.3f | %.2f | %.2f | %.2f | %.0f | %.2f |" % (name, rows / 1e6, ms, dram / 1e9, dram / rows, dram / rows / ALG_BYTES[name],
                                                                                 dram / ms / 1e6, dram / ms / 1e6 / peak))
# Workloads whose step is several kernels: DRAM bytes of one whole step from the launch list
# (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum over bench.py --workload W --steps 1):
# GROUP BY over keys in no order = partition sweep (count + scan + scatter + sweeps), GROUP BY over keys in key order = streaming
# runs (count + 2 scan kernels + write). The emit kernel of finish() is not part of either.
def step_traffic(name, first, last, must, what):
    global lines
    ll = os.path.join(G, "launches_%s.csv" % name)
    if not os.path.exists(ll):
        return
    shutil.copy(ll, os.path.join(P, "%s_launches_%s.csv" % (tag, name)))
    rr = [r for r in csv.reader(open(ll)) if len(r) > 10]
    if not rr or "Metric Name" not in rr[0]:
        return
    hdr = rr[0]
    ik, im, iv, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
    launches = {}
    for r_ in rr[1:]:
        launches.setdefault(int(r_[iid]), {"kernel": r_[ik]})[r_[im]] = float(r_[iv].replace(",", ""))
    ids = sorted(launches)
    starts = [i for i in ids if launches[i]["kernel"].startswith(first)]
    step = []
    for st_ in reversed(starts):  # the last COMPLETE step (the capture may end inside one)
        step, complete = [], False
        for i in ids[ids.index(st_):]:
            if launches[i]["kernel"].startswith(last):
                complete = True
                break
            step.append(launches[i])
        if complete:
            break
        step = []
    if not step or not any(x["kernel"].startswith(must) for x in step) or "dram__bytes_read.sum" not in step[0]:
        return
    dram = sum(x.get("dram__bytes_read.sum", 0) + x.get("dram__bytes_write.sum", 0) for x in step)
    ms = sum(x.get("gpu__time_duration.sum", 0) for x in step) / 1e6
    rows = LAUNCH_ROWS["q6"]
    kernels = {}
    for x in step:
        kernels[x["kernel"]] = kernels.get(x["kernel"], 0) + 1
    traffic[name] = {"dram_bytes_per_row": dram / rows, "dram_bytes_per_launch": dram / len(step), "launch_rows": rows, "launch_ms_under_ncu": ms,
                     "algorithmic_bytes_per_row": 16, "ratio_to_algorithmic": dram / rows / 16, "dram_gbs_under_ncu": dram / ms / 1e6,
                     "launches": len(step),
                     "source": "profiles/%s_launches_%s.csv (ncu launch list of one step: %s)" % (tag, name, " + ".join("%d x %s" % (n, k) for k, n in kernels.items()))}
    lines = [l for l in lines if not l.startswith("| %s |" % name)]
    lines.append("| %s (whole step: %s) | %.0f M | %.3f | %.2f | %.2f | %.2f | %.0f | %.2f |" % (name, what, rows / 1e6, ms, dram / 1e9, dram / rows, dram / rows / 16, dram / ms / 1e6, dram / ms / 1e6 / peak))


step_traffic("highcard_shuffled", "pgpu_k_part_count", "pgpu_k_emit", "pgpu_k_page_sweep", "count + scatter + sweeps")
step_traffic("highcard", "pgpu_k_runs_count", "pgpu_k_runs_emit", "pgpu_k_runs_write", "streaming runs: count + scan + write")
with open(os.path.join(P, "traffic.json"), "w") as f:
    json.dump(traffic, f, indent=1, sort_keys=True)
    f.write("\n")

with open(os.path.join(P, "%s_summary.md" % tag), "w") as out:
    src = os.path.join(G, "bench_default.json")
    if os.path.exists(src) and os.path.getsize(src):
        d = json.loads(open(src).read().strip().splitlines()[-1])
        out.write("## bench.py default line (1 x B200; every record verified against the CPU oracle on a prefix of the timed rows)\n\n")
        out.write("| workload | ms/step | rows/s | kernel ms/step | roofline frac | e2e rows/s (1 M-row pageable pages) | e2e / H2D ceiling | verified rows | CPU port rows/s (cores) |\n|---|---|---|---|---|---|---|---|---|\n")
        recs = [("q1 (headline)", d)] + list((d.get("configs") or {}).items())
        for name, v in recs:
            if "error" in v:
                out.write("| %s | error: %s |\n" % (name, v["error"]))
                continue
            e, c = v.get("e2e") or {}, v.get("cpu_baseline") or {}
            out.write("| %s | %.3f | %.4g | %.3f | %.3f | %s | %s | %s %s | %s |\n" % (
                name, v["ms_per_step"], v["value"], v["roofline"]["kernel_ms_per_step"], v["roofline"]["frac"],
                ("%.4g" % e["value"]) if e else "", ("%.2f" % e["frac_of_h2d_ceiling"]) if e.get("frac_of_h2d_ceiling") else "",
                v.get("verified"), v.get("verified_rows"), ("%.4g (%d)" % (c["value"], c["cores"])) if c else ""))
        e = d.get("e2e") or {}
        out.write("\nQ1 end to end: H2D ceiling %.1f GB/s; 64 K-row pages %.4g rows/s; wire format %.4g rows/s; caller-pinned single page (ceiling of the path) %.4g rows/s; CPU port single thread %.4g rows/s.\n" % (
            e.get("h2d_ceiling_gbs", 0), e.get("variants", {}).get("page_rows_65536", {}).get("value", 0), e.get("variants", {}).get("serialized_pages", {}).get("value", 0),
            e.get("ceiling_pinned_single_page", {}).get("value", 0), (d.get("cpu_baseline") or {}).get("single_thread_value", 0)))
    out.write("\n## ncu, one steady-state launch per workload (DRAM bytes are cold-cache and serialised: ratios matter, not absolute times)\n\n")
    out.write("| kernel of | launch rows | ms under ncu | DRAM GB | DRAM B/row | x algorithmic | DRAM GB/s | of %.0f GB/s |\n|---|---|---|---|---|---|---|---|\n" % peak)
    out.write("\n".join(lines) + "\n")
    for extra, title in (("bench_highcard_buckets.json", "bucketed table, load-then-CAS leaders"), ("bench_highcard_buckets_casfirst.json", "bucketed table, CAS-first leaders")):
        src = os.path.join(G, extra)
        if os.path.exists(src) and os.path.getsize(src):
            v = json.loads(open(src).read().strip().splitlines()[-1])
            out.write("\nGROUP BY l_orderkey, %s (PGPU_TUNE_BUCKETS=1): %.3f ms/step, kernel %.3f ms/step.\n" % (title, v["ms_per_step"], v["roofline"]["kernel_ms_per_step"]))
    src = os.path.join(G, "bench_highcard_shuffled_nosweep.json")
    if os.path.exists(src) and os.path.getsize(src):
        v = json.loads(open(src).read().strip().splitlines()[-1])
        out.write("\nGROUP BY shuffled l_orderkey without the partition sweep (PGPU_TUNE_SWEEP=0): %.3f ms/step, kernel %.3f ms/step.\n" % (v["ms_per_step"], v["roofline"]["kernel_ms_per_step"]))
    pqs = [os.path.join(G, "parquet_ingest_%s.json" % c) for c in ("NONE", "SNAPPY", "LZ4", "GZIP")]
    if any(os.path.exists(x) and os.path.getsize(x) for x in pqs):
        out.write("\n## Parquet ingest (tools/parquet_ingest_bench.py: Q1 over the same rows as plain blocks and as Parquet row groups, COMPRESSION = NONE / SNAPPY / LZ4 / GZIP)\n\n```\n")
        for x in pqs:
            if os.path.exists(x) and os.path.getsize(x):
                out.write(open(x).read().strip() + "\n")
        out.write("```\n")
    rt = os.path.join(G, "randtable.txt")
    if os.path.exists(rt):
        out.write("\n## random 32-byte slot updates into an 8.6 GB table (tools/microbench/randtable.cu): the floor under shuffled keys\n\n```\n" + open(rt).read() + "```\n")
    out.write("\n## ingest thread sweep (tools/ingest_bench.py, Q1, 128 M rows, pageable pages through pgpu_add_page)\n\n| copy threads | page rows | rows/s | GB/s | H2D ceiling GB/s |\n|---|---|---|---|---|\n")
    for t in (16, 8, 4, 2, 1):
        src = os.path.join(G, "ingest_t%d.json" % t)
        if os.path.exists(src) and os.path.getsize(src):
            v = json.loads(open(src).read().strip().splitlines()[-1])
            for r_ in v.get("add_page", []):
                out.write("| %d | %d | %.4g | %.1f | %.1f |\n" % (t, r_["page_rows"], r_["rows_per_s"], r_["gb_per_s"], v["h2d_ceiling_gbs"]))
print(open(os.path.join(P, "%s_summary.md" % tag)).read())
