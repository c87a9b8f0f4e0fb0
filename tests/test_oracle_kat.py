"""Pins the CPU oracle (oracle/oracle.cpp) against the reference's own known-answer tests (tests/kat_cases.py)."""
import pytest

import kat_cases
from util import assert_rows_equal_ignore_order, assert_rows_equal_ordered, run_oracle

CASES = kat_cases.all_cases()


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_known_answer(case):
    name, plan, pages, expected, nkeys = case
    actual = run_oracle(plan, pages)
    if nkeys == "ordered":
        assert_rows_equal_ordered(actual, expected)
    else:
        assert_rows_equal_ignore_order(actual, expected, nkeys=nkeys)
