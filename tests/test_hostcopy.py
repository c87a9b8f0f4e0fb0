"""The staging-copy pool (presto_b200/csrc/hostcopy.cpp) on the CPU: tests/emu/hostcopy_test.cpp copies random batches (bytes and
rebased int32 offsets, unaligned sources) through hostCopy from one and from several caller threads and compares every byte; it
also plants single faults in offset / id arrays and expects the pool's CHECK_ASCENDING32 / CHECK_IDS32 tasks (the wire-format
validation) to find exactly those."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def binary(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("hostcopy") / "hostcopy_test")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "presto_b200", "csrc"), os.path.join(ROOT, "tests", "emu", "hostcopy_test.cpp"),
                    os.path.join(ROOT, "presto_b200", "csrc", "hostcopy.cpp"), "-o", out], check=True)
    return out


@pytest.mark.parametrize("threads,callers", [(1, 1), (4, 1), (4, 3), (8, 2)])
def test_host_copy_pool(binary, threads, callers):
    env = dict(os.environ, PGPU_COPY_THREADS=str(threads))
    r = subprocess.run([binary, str(callers), "30"], env=env, stdout=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    assert "mismatches=0" in r.stdout
