"""CUDA path (through the C ABI) against the reference's known-answer tests (tests/kat_cases.py) and, on the same
inputs, against the CPU oracle."""
import pytest

import kat_cases
from util import assert_rows_equal_ignore_order, assert_rows_equal_ordered, run_gpu, run_oracle

pytestmark = pytest.mark.gpu
CASES = kat_cases.all_cases()


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_known_answer(case):
    name, plan, pages, expected, nkeys = case
    actual = run_gpu(plan, pages)
    oracle = run_oracle(plan, pages)
    if nkeys == "ordered":
        assert_rows_equal_ordered(actual, expected)      # SelectedPositions are ascending: output keeps input order
        assert_rows_equal_ordered(actual, oracle)
    else:
        assert_rows_equal_ignore_order(actual, expected, nkeys=nkeys)
        assert_rows_equal_ignore_order(actual, oracle, nkeys=nkeys)
