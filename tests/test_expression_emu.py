"""The expression code the plan compiler generates for the GPU, executed on the CPU from the same source (tests/emu/expr_emu.py):
every filter + project known-answer case over plain fixed-width blocks — scalar operators (tests/operator_cases.py), NULLIF / CASE
grids and TestPageProcessor-style cases (tests/kat_cases.py) — must give the literal expectations of the reference's tests, and
the error cases must raise the device error bit the host maps to the expected StandardErrorCode. This checks the generated
straight-line CUDA and the checked-arithmetic / decimal helpers without a GPU; the -m gpu tests check the real execution."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import expr_emu  # noqa: E402
import kat_cases  # noqa: E402
import operator_cases  # noqa: E402
from presto_b200.operators import GpuPipelineOperator  # noqa: E402
from util import assert_rows_equal_ordered  # noqa: E402

ERR_BITS = {"NUMERIC_VALUE_OUT_OF_RANGE": 1, "DIVISION_BY_ZERO": 2, "INVALID_CAST_ARGUMENT": 32}


def emulatable(plan, pages):
    return "(aggregate" not in plan and all(type(b).__name__ in ("LongArrayBlock", "IntArrayBlock", "ByteArrayBlock", "Int128ArrayBlock") for p in pages for b in p.blocks)


def execute(plan, pages):
    op = GpuPipelineOperator(plan, device=-1)
    try:
        rows, err = [], 0
        out_types = op.outputTypes()
        for page in pages:
            op.compileSignature([0] * len(page.blocks), [bool(b.mayHaveNull()) for b in page.blocks])
            lib = expr_emu.build(op.generatedSource())
            r, e = expr_emu.run(lib, page, out_types)
            rows += r
            err |= e
        return rows, err
    finally:
        op.close()


VALUES = [c for c in operator_cases.value_cases() + kat_cases.conditional_cases() + kat_cases.filter_project_cases() if emulatable(c[1], c[2])]
ERRORS = [c for c in operator_cases.error_cases() if emulatable(c[1], c[2])]


def test_coverage_of_the_emulated_cases():
    assert len(VALUES) >= 48 and len(ERRORS) >= 28


@pytest.mark.parametrize("case", VALUES, ids=[c[0] for c in VALUES])
def test_generated_expression_code_values(case):
    name, plan, pages, expected, _ = case
    rows, err = execute(plan, pages)
    assert err == 0
    assert_rows_equal_ordered(rows, expected)


@pytest.mark.parametrize("case", ERRORS, ids=[c[0] for c in ERRORS])
def test_generated_expression_code_errors(case):
    name, plan, pages, code = case
    rows, err = execute(plan, pages)
    assert err & ERR_BITS[code], (err, code)
