"""The JNI shim (jni/prestogpu_jni.c) and the Java natives it implements (java/io/trino/operator/gpu/PrestoGpuNative.java).
No JDK exists in this image, so the shim is compiled against a stand-in jni.h (tests/emu/jni_stub: the JNI types and the entry
points the shim uses, with the specification's signatures) — enough to prove that it is complete C, that every native the Java
class declares is defined with the JNI-mangled name, and that it only calls C ABI functions include/prestogpu.h declares."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "tests", "emu", "_build", "prestogpu_jni.o")


def _compile():
    os.makedirs(os.path.dirname(OBJ), exist_ok=True)
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Wno-unused-parameter", "-Werror", "-fPIC", "-c", "-I", os.path.join(ROOT, "tests", "emu", "jni_stub"),
           "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "jni", "prestogpu_jni.c"), "-o", OBJ]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return subprocess.run(["nm", OBJ], stdout=subprocess.PIPE, text=True, check=True).stdout


def test_every_declared_native_is_defined():
    java = open(os.path.join(ROOT, "java", "io", "trino", "operator", "gpu", "PrestoGpuNative.java")).read()
    java = re.sub(r"/\*.*?\*/", "", java, flags=re.S)
    declared = set(re.findall(r"static\s+native\s+[\w\[\]\.]+\s+(\w+)\s*\(", java))
    assert len(declared) >= 13
    nm = _compile()
    defined = set(re.findall(r" T Java_io_trino_operator_gpu_PrestoGpuNative_(\w+)", nm))
    assert declared == defined, "declared but not defined: %s; defined but not declared: %s" % (sorted(declared - defined), sorted(defined - declared))


def test_shim_calls_only_the_product_abi():
    nm = _compile()
    used = set(re.findall(r" U (pgpu_\w+)", nm))
    header = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "prestogpu.h")).read(), flags=re.S)
    declared = set(re.findall(r"\b(pgpu_[a-z0-9_]+)\s*\(", header))
    assert used and used <= declared, sorted(used - declared)
    for must in ("pgpu_reserve_input", "pgpu_add_page", "pgpu_get_output", "pgpu_finish", "pgpu_on_unblocked"):
        assert must in used


def test_no_device_wait_inside_the_critical_region():
    """between GetPrimitiveArrayCritical and ReleasePrimitiveArrayCritical the shim may only call pgpu_add_page (memory copies
    after pgpu_reserve_input) — no JNI call, no other library call"""
    src = open(os.path.join(ROOT, "jni", "prestogpu_jni.c")).read()
    body = src[src.index("Java_io_trino_operator_gpu_PrestoGpuNative_addPage"):]
    body = body[:body.index("\n}\n")]
    first_pin = body.index("PIN(aValues")
    release = body.index("for (int i = npinned")
    assert body.index("pgpu_reserve_input") < first_pin
    region = body[first_pin:release]
    assert re.findall(r"\bpgpu_\w+\(", region) == ["pgpu_add_page("]
    assert "(*env)->" not in region.replace("(*env)->GetPrimitiveArrayCritical", "")
