"""CUDA path against the reference's scalar operator tests (tests/operator_cases.py) and, on the same inputs, the CPU oracle:
BIGINT / INTEGER / DOUBLE / DECIMAL arithmetic, comparisons and casts evaluated by the generated kernels, bit-exact, and the
faults they must raise (overflow, division by zero, invalid cast)."""
import pytest

import operator_cases
from presto_b200.operators import TrinoException
from util import assert_rows_equal_ordered, run_gpu, run_oracle

pytestmark = pytest.mark.gpu
VALUES, ERRORS = operator_cases.value_cases(), operator_cases.error_cases()


@pytest.mark.parametrize("case", VALUES, ids=[c[0] for c in VALUES])
def test_gpu_operator_values(case):
    name, plan, pages, expected, _ = case
    actual = run_gpu(plan, pages)
    assert_rows_equal_ordered(actual, expected)
    assert_rows_equal_ordered(actual, run_oracle(plan, pages))


@pytest.mark.parametrize("case", ERRORS, ids=[c[0] for c in ERRORS])
def test_gpu_operator_errors(case):
    name, plan, pages, code = case
    with pytest.raises(TrinoException) as e:
        run_gpu(plan, pages)
    assert e.value.code == code
