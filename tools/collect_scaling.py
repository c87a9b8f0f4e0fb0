#!/usr/bin/env python
"""Copy the multi-GPU bench lines of tools/gpu_trip2.sh visits (gpurun_out/scale<N>_<tag>.json) into profiles/ and write a
scaling table. Usage: python tools/collect_scaling.py <tag> [single-GPU bench tag]"""
import glob
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1]
base_tag = sys.argv[2] if len(sys.argv) > 2 else tag
single = {}
for w, name in (("q1", "q1"), ("q1", "q1_nccl"), ("q6", "q6"), ("q1dict", "q1dict"), ("highcard", "hc_peer"), ("highcard", "hc_nccl"), ("highcard_shuffled", "hcs_peer")):
    f = os.path.join(P, "%s_bench_%s.json" % (base_tag, w))
    if os.path.exists(f):
        single[name] = json.loads(open(f).read().strip().splitlines()[-1])
rows = []
for f in sorted(glob.glob(os.path.join(G, "scale*_*.json"))):
    b = os.path.basename(f)
    n, name = b[5:].split("_", 1)
    name = name[:-5]
    txt = open(f).read().strip()
    if not txt:
        continue
    d = json.loads(txt.splitlines()[-1])
    shutil.copy(f, os.path.join(P, "%s_%s" % (tag, b)))
    rows.append((name, int(n), d))
with open(os.path.join(P, "%s_scaling.md" % tag), "w") as out:
    out.write("Weak scaling, %s rows per GPU (sf100 shard each), one process per GPU under torchrun; `value` = all ranks' rows / max-over-ranks device time.\n\n" % "600 037 902")
    out.write("| workload / exchange | GPUs | ms/step | rows/s | vs 1 GPU | efficiency | e2e rows/s | exchange bytes sent per rank |\n|---|---|---|---|---|---|---|---|\n")
    for name in ("q1", "q1_nccl", "q6", "q1dict", "hc_peer", "hc_nccl", "hcs_peer"):
        s = single.get(name)
        if s:
            out.write("| %s | 1 | %.2f | %.3g | 1.00 | 1.00 | %s | |\n" % (name, s["ms_per_step"], s["value"], ("%.3g" % s["e2e"]["value"]) if s.get("e2e") else ""))
        for nm, n, d in sorted(rows, key=lambda r: r[1]):
            if nm != name:
                continue
            sp = d["value"] / s["value"] if s else float("nan")
            out.write("| %s | %d | %.2f | %.3g | %.2f | %.2f | %s | %s |\n" % (name, n, d["ms_per_step"], d["value"], sp, sp / n, ("%.3g" % d["e2e"]["value"]) if d.get("e2e") else "",
                                                                     (d.get("exchange") or {}).get("bytes_sent_per_rank_per_step", "")))
    big = [(nm, n, d) for nm, n, d in rows if nm.endswith("_sf1000")]
    if big:
        out.write("\nConfig 5 — sf1000 (5 999 989 709 lineitem rows in total, device resident, split over the GPUs; 1 and 2 GPUs cannot hold it: "
                  "276 GB of Q1 columns, and one page of VARCHAR(1) bytes is limited to 2^31 rows by the int32 offsets):\n\n")
        out.write("| workload | GPUs | rows per GPU | ms/step | rows/s | GB/s (algorithmic) |\n|---|---|---|---|---|---|\n")
        for nm, n, d in sorted(big, key=lambda r: (r[0], r[1])):
            out.write("| %s | %d | %d | %.2f | %.3g | %.0f |\n" % (nm, n, d["config"]["rows_per_gpu"], d["ms_per_step"], d["value"], d["gb_per_s"]))
print(open(os.path.join(P, "%s_scaling.md" % tag)).read())
