#!/usr/bin/env python
"""Summarise the SASS source page of an .ncu-rep: executed-instruction totals by opcode class and the hottest
instructions.  Usage: python tools/ncu_hot.py report.ncu-rep"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rd = csv.DictReader(io.StringIO("\n".join(lines[start:])))
rows = [r for r in rd if r.get("Address")]
tot = 0
byop = defaultdict(int)
samples = defaultdict(int)
stall_cols = [c for c in rows[0].keys() if c.startswith("stall_") and "Not Issued" not in c]
stalls = defaultdict(int)
for r in rows:
    n = int(r["Instructions Executed"] or 0)
    tot += n
    src = r["Source"].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    op = op.split(".")[0]
    byop[op] += n
    samples[op] += int(r["# Samples"] or 0)
    for c in stall_cols:
        stalls[c] += int(r[c] or 0)
print("total warp instructions executed:", tot)
print("%-10s %14s %6s %9s" % ("opcode", "executed", "%", "samples"))
for op, n in sorted(byop.items(), key=lambda x: -x[1])[:28]:
    print("%-10s %14d %6.2f %9d" % (op, n, 100.0 * n / max(tot, 1), samples[op]))
ts = sum(stalls.values())
print("stall samples:", ", ".join("%s=%.1f%%" % (k[6:], 100.0 * v / max(ts, 1)) for k, v in sorted(stalls.items(), key=lambda x: -x[1])[:8]))
print("hottest instructions (samples, executed, sass):")
for r in sorted(rows, key=lambda r: -int(r["# Samples"] or 0))[:25]:
    print("  %6s %12s  %s  %s" % (r["# Samples"], r["Instructions Executed"], r["Address"][-6:], r["Source"].strip()[:80]))
