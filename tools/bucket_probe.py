#!/usr/bin/env python
"""How well the bucketed table packs clustered keys: slots claimed vs groups, at a few sizes (diagnostic)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from presto_b200 import tpch
from presto_b200.operators import GpuPipelineOperator
import bench as B
for rows in (1_000_000, 32_000_000, 200_000_000):
    dp = tpch.DevicePage(0, tpch.HC_DEVICE_SPEC, rows)
    op = GpuPipelineOperator(tpch.highcard_plan(rows // 4), 0)
    op.addDeviceInput(dp.cols, 2, rows)
    op.finish()
    n = B.drain_output(op, to_host=False)
    st = op.stats()
    print("rows", rows, "groups", n, "bucketed", st.table_bucketed, "offset", st.bucket_offset, "slots claimed", st.table_slots_claimed,
          "capacity", st.table_capacity, "rehashes", st.table_rehashes, "groups/slot", round(n / max(st.table_slots_claimed, 1), 2))
    op.close(); dp.free()
