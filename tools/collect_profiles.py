#!/usr/bin/env python
"""Copy the judged summaries of the last tools/gpu_check.sh visit from gpurun_out/ (scratch) into profiles/ (tracked):
bench JSON lines, the ncu launch list, ncu details pages, raw DRAM-traffic metrics and the SASS hot-spot summary.
Usage: python tools/collect_profiles.py <tag>   (e.g. r01c)"""
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
rows = []
for w in ("q1", "q6", "highcard", "highcard_shuffled", "q1dict", "ref"):
    src = os.path.join(G, "bench_%s.json" % w)
    if os.path.exists(src) and os.path.getsize(src):
        shutil.copy(src, os.path.join(P, "%s_bench_%s.json" % (tag, w)))
        d = json.loads(open(src).read().strip().splitlines()[-1])
        rows.append((w, d))
if os.path.exists(os.path.join(G, "launches_q1.csv")):
    shutil.copy(os.path.join(G, "launches_q1.csv"), os.path.join(P, "%s_q1_launches.csv" % tag))
if os.path.exists(os.path.join(G, "launches_highcard.csv")):
    shutil.copy(os.path.join(G, "launches_highcard.csv"), os.path.join(P, "%s_highcard_launches.csv" % tag))
for rep, name in (("q1_full", "q1_staged"), ("hc_full", "highcard"), ("q6_full", "q6"), ("q1dict_full", "q1dict_staged")):
    f = os.path.join(G, rep + ".ncu-rep")
    if not os.path.exists(f):
        continue
    det = subprocess.run(["ncu", "-i", f, "--page", "details"], stdout=subprocess.PIPE, text=True).stdout
    open(os.path.join(P, "%s_%s_ncu_details.txt" % (tag, name)), "w").write(det)
    raw = subprocess.run(["ncu", "-i", f, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    keep = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
            "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__throughput.avg.pct_of_peak_sustained_elapsed")
    with open(os.path.join(P, "%s_%s_ncu_raw.txt" % (tag, name)), "w") as out:
        for i, h in enumerate(r[0]):
            if h in keep:
                out.write("%s [%s] = %s\n" % (h, r[1][i], r[2][i]))
    hot = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_hot.py"), f], stdout=subprocess.PIPE, text=True).stdout
    open(os.path.join(P, "%s_%s_ncu_hot.txt" % (tag, name)), "w").write(hot)
with open(os.path.join(P, "%s_summary.md" % tag), "w") as out:
    out.write("| workload | ms/step | rows/s | GB/s (algorithmic) | kernel roofline frac | e2e rows/s | launches/step |\n|---|---|---|---|---|---|---|\n")
    for w, d in rows:
        if d.get("impl") == "reference":
            out.write("| CPU oracle port, %d cores (Q1) | %.1f | %.3g | | | | |\n" % (d["cpu_baseline"]["cores"], d["ms_per_step"], d["value"]))
            continue
        e = d.get("e2e") or {}
        out.write("| %s | %.2f | %.3g | %.0f | %.3f | %s | %d |\n" % (w, d["ms_per_step"], d["value"], d["gb_per_s"], d["roofline"]["frac"],
                                                                    ("%.3g" % e["value"]) if e else "", d["gpu_launches"] // d["steps"]))
print(open(os.path.join(P, "%s_summary.md" % tag)).read())
