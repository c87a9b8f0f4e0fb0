"""Pins the oracle's scalar operators against the reference's operator tests (tests/operator_cases.py)."""
import pytest

import operator_cases
from presto_b200.operators import TrinoException
from util import assert_rows_equal_ordered, run_oracle

VALUES, ERRORS = operator_cases.value_cases(), operator_cases.error_cases()


@pytest.mark.parametrize("case", VALUES, ids=[c[0] for c in VALUES])
def test_oracle_operator_values(case):
    name, plan, pages, expected, _ = case
    assert_rows_equal_ordered(run_oracle(plan, pages), expected)


@pytest.mark.parametrize("case", ERRORS, ids=[c[0] for c in ERRORS])
def test_oracle_operator_errors(case):
    name, plan, pages, code = case
    with pytest.raises(TrinoException) as e:
        run_oracle(plan, pages)
    assert e.value.code == code
