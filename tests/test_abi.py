"""The C ABI boundary without a GPU: the library loads, exports every symbol include/prestogpu.h declares, compiles
plans (device = -1) and rejects everything outside the supported subset at plan time (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from presto_b200 import _native as N
from presto_b200 import tpch
from presto_b200.operators import (Aggregation, GpuFilterAndProjectOperator, GpuHashAggregationOperator, GpuOperatorFactory,
                                   GpuPipelineOperator, TrinoException, plan_text)
from presto_b200.relational import CallExpression, SpecialForm, call, compare, constant, field
from presto_b200.spi import BIGINT, BOOLEAN, DOUBLE, VARCHAR, Page, block_of, createDecimalType, createVarcharType

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "prestogpu.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(pgpu_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 30
    handle = ctypes.CDLL(N.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(handle, name), "libprestogpu.so does not export %s" % name
    bound = {s[0] for s in N.SYMBOLS}
    assert declared == bound, "ctypes binding and header disagree: %s" % sorted(declared ^ bound)


def test_stats_struct_mirrors_the_header_field_by_field():
    """pgpu_get_stats fills the caller's struct: the ctypes mirror must name the header's fields in the header's order"""
    h = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "prestogpu.h")).read()
    body = h[h.index("typedef struct pgpu_stats"):h.index("} pgpu_stats;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    declared = re.findall(r"\b(int64_t|double)\s+(\w+);", body)
    mirror = [(name, ct) for name, ct in N.Stats._fields_]
    assert [n for _, n in declared] == [n for n, _ in mirror]
    for (ctype, name), (_, ct) in zip(declared, mirror):
        assert ct is (ctypes.c_double if ctype == "double" else ctypes.c_int64), name
    assert "streamed_runs" in [n for n, _ in mirror] and "partition_sweeps" in [n for n, _ in mirror]


def test_bench_plumbing_lives_in_its_own_library():
    """generators, synchronous copies and the L2 flush are not part of the drop-in boundary (include/prestogpu_bench.h,
    libprestogpu_bench.so): the product library exports none of them"""
    header = open(os.path.join(ROOT, "include", "prestogpu_bench.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(pgpu_[a-z0-9_]+)\s*\(", header))
    assert declared == {s[0] for s in N.BENCH_SYMBOLS}
    product, bench = ctypes.CDLL(N.LIB_PATH), ctypes.CDLL(N.BENCH_LIB_PATH)
    for name in sorted(declared):
        assert hasattr(bench, name), "libprestogpu_bench.so does not export %s" % name
        assert not hasattr(product, name), "libprestogpu.so exports bench symbol %s" % name


def test_no_device_means_zero_devices_not_an_error():
    L = N.lib()
    assert L.pgpu_device_count() >= 0
    assert b"sm_100a" in L.pgpu_version()


def test_product_never_touches_the_oracle():
    # the product package must not import, link or execute anything under oracle/
    pkg = os.path.join(ROOT, "presto_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle/" not in src and "liboracle" not in src and "oracle_op" not in src, f


def test_compile_only_pipeline_refuses_pages():
    op = GpuPipelineOperator(tpch.q6_plan(), device=-1)
    assert "pgpu_k_page" in op.generatedSource()
    with pytest.raises(TrinoException) as e:
        op.addInput(tpch.q6_page(0, 10, 10))
    assert e.value.code == "INVALID_STATE"
    op.close()


def test_canonical_plans_compile_for_every_page_signature():
    for plan in (tpch.q1_plan(), tpch.q1_plan("partial"), tpch.q6_plan(), tpch.highcard_plan(1000), tpch.highcard_plan(1000, "partial")):
        op = GpuPipelineOperator(plan, device=-1)
        n = len(re.search(r"\(input ([^\n]*)\)\n", plan).group(1).replace("(varchar 1)", "v").split())
        op.compileSignature([1] * n, [True] * n)      # all dictionary blocks with null vectors
        op.compileSignature([2] * n, [False] * n)     # all RLE
        assert op.stats().jit_compiles + op.stats().jit_cache_hits >= 3
        op.close()


def test_conditional_forms_compile_for_sm_100a():
    """NULL_IF / SWITCH plans of the known-answer suite: generated CUDA compiles (NVRTC, sm_100a) for plain and all-nullable pages"""
    import kat_cases
    for name, plan, pages, _, _ in kat_cases.conditional_cases():
        op = GpuPipelineOperator(plan, device=-1)
        n = len(pages[0].blocks)
        op.compileSignature([0] * n, [b.mayHaveNull() for b in pages[0].blocks])
        op.compileSignature([1] * n, [True] * n)
        assert "null-if" in plan or "switch" in plan, name
        op.close()


def test_intermediate_step_compiles_and_keeps_the_state_schema():
    from test_oracle_protocol import q1_stage_plan
    op = GpuPipelineOperator(q1_stage_plan("intermediate"), device=-1)
    assert [t.kind for t in op.outputTypes()] == [6, 6] + [1, 4] * 4 + [4, 1] * 3 + [1]     # same as the PARTIAL step's output
    op.compileSignature([0] * 17, [True] * 17)
    op.close()


def test_output_schema():
    op = GpuPipelineOperator(tpch.q1_plan(), device=-1)
    kinds = [t.kind for t in op.outputTypes()]
    assert kinds == [6, 6, 4, 4, 4, 4, 4, 4, 4, 1]
    op.close()
    op = GpuPipelineOperator(tpch.q1_plan("partial"), device=-1)
    # PARTIAL: keys, then per aggregate its intermediate channels (StateCompiler ROW order): sum -> (count, sum), avg -> (sum, count)
    assert [t.kind for t in op.outputTypes()] == [6, 6] + [1, 4] * 4 + [4, 1] * 3 + [1]
    op.close()
    d = createDecimalType(12, 2)
    op = GpuHashAggregationOperator([BIGINT, d], [0], "single", [Aggregation("sum", createDecimalType(38, 2), 1, d), Aggregation("avg", d, 1, d)], device=-1)
    t = op.outputTypes()
    assert (t[1].kind, t[1].precision, t[1].scale) == (5, 38, 2) and (t[2].precision, t[2].scale) == (12, 2)
    op.close()


REJECTED = [
    ("unbounded varchar key", lambda: plan_text([VARCHAR, BIGINT], step="single", groupByChannels=[0], aggregations=[Aggregation("count", BIGINT)])),
    ("varchar(64) key", lambda: plan_text([createVarcharType(64), BIGINT], step="single", groupByChannels=[0], aggregations=[Aggregation("count", BIGINT)])),
    ("unsupported special form", lambda: plan_text([BIGINT], None, [SpecialForm("DEREFERENCE", BIGINT, [field(0, BIGINT), constant(1, BIGINT)])])),
    ("when outside switch", lambda: plan_text([BIGINT], None, [SpecialForm("WHEN", BIGINT, [field(0, BIGINT), constant(1, BIGINT)])])),
    ("null-if of differing types", lambda: plan_text([BIGINT, DOUBLE], None, [SpecialForm("NULL_IF", BIGINT, [field(0, BIGINT), field(1, DOUBLE)])])),
    ("switch without when", lambda: plan_text([BIGINT], None, [SpecialForm("SWITCH", BIGINT, [field(0, BIGINT), constant(1, BIGINT)])])),
    ("switch operand of another type", lambda: plan_text([BIGINT, DOUBLE], None, [SpecialForm("SWITCH", BIGINT, [field(0, BIGINT), SpecialForm("WHEN", BIGINT, [field(1, DOUBLE), constant(1, BIGINT)])])])),
    ("switch result of another type", lambda: plan_text([BIGINT], None, [SpecialForm("SWITCH", BIGINT, [field(0, BIGINT), SpecialForm("WHEN", DOUBLE, [field(0, BIGINT), constant(1.0, DOUBLE)])])])),
    ("switch computing varchar", lambda: plan_text([BIGINT, createVarcharType(3)], None, [SpecialForm("SWITCH", createVarcharType(3), [field(0, BIGINT), SpecialForm("WHEN", createVarcharType(3), [constant(1, BIGINT), field(1, createVarcharType(3))])])])),
    ("unknown function", lambda: plan_text([DOUBLE], None, [CallExpression("sqrt", DOUBLE, [field(0, DOUBLE)])])),
    ("unknown aggregate", lambda: plan_text([DOUBLE], step="single", aggregations=[Aggregation("approx_distinct", BIGINT, 0, DOUBLE)])),
    ("decimal division", lambda: plan_text([createDecimalType(10, 2)], None, [call("divide", createDecimalType(10, 2), field(0, createDecimalType(10, 2)), field(0, createDecimalType(10, 2)))])),
    ("unknown step", lambda: plan_text([BIGINT], step="streaming", aggregations=[Aggregation("count", BIGINT)])),
    ("intermediate step over raw input", lambda: plan_text([DOUBLE], step="intermediate", aggregations=[Aggregation("sum", DOUBLE, 0, DOUBLE)])),
]


@pytest.mark.parametrize("what,make", REJECTED, ids=[r[0] for r in REJECTED])
def test_plan_time_rejection(what, make):
    with pytest.raises(TrinoException) as e:
        GpuPipelineOperator(make(), device=-1)
    assert e.value.code in ("NOT_SUPPORTED", "INVALID_ARGUMENT"), e.value
    assert str(e.value)


def test_malformed_plan_is_invalid_argument():
    for bad in ("", "(plan", "(plan (input bigint) (filter (in 3 bigint)))", "(plan (input bigint) (project (call add bigint (in 0 bigint))))"):
        with pytest.raises(TrinoException) as e:
            GpuPipelineOperator(bad, device=-1)
        assert e.value.code in ("INVALID_ARGUMENT", "NOT_SUPPORTED")


def test_operator_factory_contract():
    # OperatorFactory.java:18-50: createOperator after noMoreOperators is an error; duplicate() gives a fresh factory
    f = GpuOperatorFactory(0, tpch.q6_plan(), device=-1)
    op = f.createOperator()
    assert op.needsInput() and not op.isFinished()
    op.close()
    g = f.duplicate()
    f.noMoreOperators()
    with pytest.raises(TrinoException):
        f.createOperator()
    g.createOperator().close()
    with pytest.raises(TrinoException):
        GpuOperatorFactory(1, plan_text([VARCHAR], step="single", groupByChannels=[0], aggregations=[Aggregation("count", BIGINT)]), device=-1)


def test_null_arguments_do_not_crash():
    L = N.lib()
    assert L.pgpu_pipeline_create(None, 0, -1, None) == N.INVALID_ARGUMENT
    assert L.pgpu_needs_input(None) == 0
    assert L.pgpu_is_finished(None) == 0
    assert L.pgpu_finish(None) == N.INVALID_ARGUMENT
    L.pgpu_pipeline_destroy(None)
    assert L.pgpu_output_channel_count(None) == 0


def test_plan_parser_under_sanitizers(tmp_path):
    """csrc/plan.cpp compiled with AddressSanitizer + UBSan (tests/emu/plan_fuzz.cpp): 400 random edits of each benchmark / known-answer
    plan are accepted or rejected with a PlanError — no crash, no out-of-bounds access, no undefined behaviour"""
    import shutil
    import subprocess
    import kat_cases
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "plan_fuzz")
    r = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-I" + os.path.join(root, "include"), "-o", exe,
                        os.path.join(root, "tests", "emu", "plan_fuzz.cpp"), os.path.join(root, "presto_b200", "csrc", "plan.cpp")], capture_output=True, text=True)
    if r.returncode != 0 and "asan" in r.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert r.returncode == 0, r.stderr
    plans = [tpch.q1_plan(), tpch.q1_plan("partial"), tpch.q6_plan(), tpch.q1_decimal_plan(), tpch.highcard_plan(1000)] + [c[1] for c in kat_cases.conditional_cases()]
    accepted = rejected = 0
    for i, plan in enumerate(plans):
        path = tmp_path / "plan.txt"
        path.write_text(plan)
        r = subprocess.run([exe, str(path), "400", str(i)], capture_output=True, text=True)
        assert r.returncode == 0, (i, r.returncode, r.stderr[-2000:])
        a, j = map(int, r.stdout.split())
        accepted += a
        rejected += j
    assert accepted >= len(plans) and rejected > 1000


def test_code_generator_under_sanitizers(tmp_path):
    """csrc/plan.cpp + codegen.cpp compiled with AddressSanitizer + UBSan (tests/emu/codegen_asan.cpp): the kernel source of every
    benchmark / known-answer / operator plan is generated for every sampled combination of block encodings and null vectors"""
    import shutil
    import subprocess
    import kat_cases
    import operator_cases
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "codegen_asan")
    csrc = os.path.join(root, "presto_b200", "csrc")
    r = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-o", exe,
                        os.path.join(root, "tests", "emu", "codegen_asan.cpp"), os.path.join(csrc, "plan.cpp"), os.path.join(csrc, "codegen.cpp")], capture_output=True, text=True)
    if r.returncode != 0 and "asan" in r.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert r.returncode == 0, r.stderr
    plans = [tpch.q1_plan(), tpch.q1_plan("partial"), tpch.q6_plan(), tpch.q1_decimal_plan(), tpch.highcard_plan(1000), tpch.highcard_plan(1 << 22)]
    plans += [c[1] for c in kat_cases.all_cases()] + [c[1] for c in operator_cases.value_cases()]
    paths = []
    for i, plan in enumerate(plans):
        path = tmp_path / ("plan%d.txt" % i)
        path.write_text(plan)
        paths.append(str(path))
    r = subprocess.run([exe] + paths, capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stderr[-3000:])
    generated, rejected, size = map(int, r.stdout.split())
    assert generated > 10 * len(plans) and size > 1 << 20


def test_intermediate_step_differs_from_final_only_in_the_output_section():
    """the kernel source generated for an INTERMEDIATE step is the FINAL step's except for the output section (which writes the
    accumulator cells out as states, like PARTIAL, instead of finalising them): the input side — state channels into cells — is the
    code that has run on the GPU as the FINAL step since round 1"""
    import difflib
    from test_oracle_protocol import q1_stage_plan

    def source(plan, n):
        op = GpuPipelineOperator(plan, device=-1)
        op.compileSignature([0] * n, [False] * n)
        s = [line.strip() for line in op.generatedSource().splitlines() if not line.startswith("//")]
        op.close()
        return s
    final, inter = source(q1_stage_plan("final"), 17), source(q1_stage_plan("intermediate"), 17)
    diff = [d for d in difflib.unified_diff(final, inter, lineterm="", n=0) if d[:1] in "+-" and d[:3] not in ("+++", "---")]
    assert len(diff) > 20
    assert all("out.c[" in d or d[1:].strip() in ("{", "}") or "NO =" in d for d in diff), [d for d in diff if "out.c[" not in d][:5]
    states = [d for d in diff if d.startswith("+") and "cells[" in d]
    assert len(states) == 8 and all("/" not in d for d in states)     # 8 aggregates, none divided (avg stays sum + count)
