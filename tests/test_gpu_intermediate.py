"""INTERMEDIATE aggregation step on the GPU: PARTIAL -> INTERMEDIATE -> FINAL equals SINGLE, and the states the INTERMEDIATE
pipeline emits equal the oracle's for the same partial pages (counts and keys bit-exact, DOUBLE sums within the stated bound)."""
import pytest

from presto_b200 import tpch
from presto_b200.operators import GpuPipelineOperator
from test_oracle_protocol import intermediate_chain, q1_stage_plan
from util import assert_rows_equal_ignore_order, rows_of, run_gpu, run_oracle, sequential_bound

pytestmark = pytest.mark.gpu


def test_gpu_partial_intermediate_final_equals_single():
    total, pages, inter, final = intermediate_chain(lambda plan: GpuPipelineOperator(plan, 0))
    single = run_gpu(tpch.q1_plan(), pages)
    assert all(len(pg.blocks) == 17 for pg in inter)
    got = rows_of(final)
    assert_rows_equal_ignore_order(got, single, nkeys=2, rtol=sequential_bound(total))
    assert sorted(r[-1] for r in got) == sorted(r[-1] for r in single)
    assert_rows_equal_ignore_order(got, run_oracle(tpch.q1_plan(), pages), nkeys=2, rtol=sequential_bound(total))


def test_gpu_intermediate_states_equal_oracle():
    total = 50_000
    partial = []
    for a, n in ((0, 20_000), (20_000, 30_000)):
        op = GpuPipelineOperator(tpch.q1_plan("partial"), 0)
        from presto_b200.operators import toPages
        partial += toPages(op, [tpch.q1_page(a, n, total)])
        op.close()
    plan = q1_stage_plan("intermediate")
    assert_rows_equal_ignore_order(run_gpu(plan, partial), run_oracle(plan, partial), nkeys=2, rtol=sequential_bound(total))
