#!/usr/bin/env python
"""Compile the canonical plans into a scratch cache and print register / shared-memory usage of every kernel
(cuobjdump -res-usage). Usage: python tools/kernel_info.py [q1|q6|highcard|q1dict ...]"""
import glob
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
d = tempfile.mkdtemp(prefix="pgpu_ki_")
os.environ["PGPU_CACHE_DIR"] = d
os.environ["PGPU_KEEP_SOURCE"] = "1"
from presto_b200 import build as B  # noqa: E402
B.build()
from presto_b200 import tpch  # noqa: E402
from presto_b200.operators import GpuPipelineOperator  # noqa: E402

which = sys.argv[1:] or ["q1", "q6", "highcard", "q1dict"]
for w in which:
    before = set(glob.glob(d + "/*.cubin"))
    if w == "q1":
        GpuPipelineOperator(tpch.q1_plan(), -1).close()
    elif w == "q6":
        GpuPipelineOperator(tpch.q6_plan(), -1).close()
    elif w == "highcard":
        GpuPipelineOperator(tpch.highcard_plan(150_000_000), -1).close()
    elif w == "q1dict":
        op = GpuPipelineOperator(tpch.q1_plan(), -1)
        before = set(glob.glob(d + "/*.cubin"))
        op.compileSignature([1, 1, 1, 0, 1, 1, 0], [False] * 7)
        op.close()
    for c in sorted(set(glob.glob(d + "/*.cubin")) - before):
        print("==", w, c)
        out = subprocess.run(["cuobjdump", "-res-usage", c], stdout=subprocess.PIPE, text=True).stdout
        for l in out.splitlines():
            if "Function" in l or "REG" in l:
                print("  ", l.strip())
