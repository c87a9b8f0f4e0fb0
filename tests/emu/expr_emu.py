"""TEST INFRASTRUCTURE: runs the expression code the plan compiler generates — P::eval_project / P::store_row of a filter + project
plan, together with the scalar part of presto_b200/csrc/device/pgpu_device.cuh (checked arithmetic, decimal helpers, casts) — on
the CPU, compiled as plain C++ from the very source NVRTC compiles for the GPU. What is restated here is only the frame around it:
body_compact's "every thread evaluates 4 consecutive rows, selected rows are stored in position order" becomes a loop.
Plain fixed-width input blocks only (what the operator / conditional known-answer cases use)."""
import ctypes
import hashlib
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
BUILD = os.path.join(HERE, "_build")

PRELUDE = r'''
#define PGPU_HOST_EMULATION 1
#include <cuda_runtime.h>   // int4 and friends; host side only
#include <cmath>
#include <cstring>
#undef __device__
#undef __global__
#undef __forceinline__
#undef __launch_bounds__
#define __device__
#define __global__
#define __forceinline__ inline
#define __launch_bounds__(...)
// device intrinsics of the scalar section, as the PTX ISA defines them
static inline long long __double_as_longlong(double d) { long long r; memcpy(&r, &d, 8); return r; }
static inline double __longlong_as_double(long long v) { double r; memcpy(&r, &v, 8); return r; }
static inline long long __mul64hi(long long a, long long b) { return (long long)(((__int128)a * (__int128)b) >> 64); }
static inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) { return (unsigned)((((unsigned long long)hi << 32) | lo) >> (sh & 31)); }
template <typename T> static inline T __ldg(const T* p) { return *p; }
#include "pgpu_device.cuh"
'''

DRIVER = r'''
extern "C" long long emu_filter_project(const ColArg* cols, long long rows, const OutCol* outs, unsigned* err) {
    PageArgs<P::NC> a;
    for (int c = 0; c < P::NC; c++) a.c[c] = cols[c];
    a.rows = rows;
    OutArgs<P::NO> out;
    for (int c = 0; c < P::NO; c++) out.c[c] = outs[c];
    out.capacity = rows;
    long long pos = 0;
    for (long long base = 0; base < rows; base += 4) {   // one iteration = one GPU thread of body_compact
        bool sel[4];
        P::Proj pv[4];
        unsigned e = 0;
        P::eval_project(a, base, rows, sel, pv, e);
        *err |= e;
        for (int r = 0; r < 4; r++)
            if (sel[r]) P::store_row(out, pos++, pv[r]);
    }
    return pos;
}
'''


class ColArg(ctypes.Structure):
    _fields_ = [("values", ctypes.c_void_p), ("nulls", ctypes.c_void_p), ("offsets", ctypes.c_void_p), ("ids", ctypes.c_void_p), ("dvalues", ctypes.c_void_p),
                ("dnulls", ctypes.c_void_p), ("doffsets", ctypes.c_void_p), ("dcount", ctypes.c_int32), ("pad", ctypes.c_int32)]


class OutCol(ctypes.Structure):
    _fields_ = [("values", ctypes.c_void_p), ("nulls", ctypes.c_void_p)]


def build(generated_source):
    """generated kernel source (GpuPipelineOperator.generatedSource()) -> ctypes library exporting emu_filter_project"""
    cut = generated_source.find('extern "C" __global__')
    assert cut > 0 and "body_compact" in generated_source[cut:], "not a filter + project plan"
    body = generated_source[:cut].replace('#include "pgpu_device.cuh"', "")
    text = PRELUDE + body + DRIVER
    dev = open(os.path.join(ROOT, "presto_b200", "csrc", "device", "pgpu_device.cuh"), "rb").read()
    key = hashlib.sha1(text.encode() + dev).hexdigest()[:20]
    os.makedirs(BUILD, exist_ok=True)
    so = os.path.join(BUILD, "expr_%s.so" % key)
    if not os.path.exists(so):
        src = os.path.join(BUILD, "expr_%s.cpp" % key)
        open(src, "w").write(text)
        cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
        r = subprocess.run(["g++", "-O0", "-std=c++17", "-shared", "-fPIC", "-w", "-ffp-contract=off", "-x", "c++", "-I" + os.path.join(cuda, "include"),
                            "-I" + os.path.join(ROOT, "presto_b200", "csrc", "device"), "-o", so, src], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]
    return ctypes.CDLL(so)


_WIDTH = {0: 1, 1: 8, 2: 4, 3: 4, 4: 8}


def width_of(t):
    return _WIDTH[t.kind] if t.kind != 5 else (8 if t.precision <= 18 else 16)


def run(lib, page, output_types):
    """-> (rows as tuples of Python values, error bits)"""
    n = page.position_count
    pad = (n + 7) // 4 * 4 + 16
    keep, cols = [], (ColArg * len(page.blocks))()
    for i, b in enumerate(page.blocks):
        w = width_of(b.type)
        raw = np.zeros(pad * w, np.uint8)
        raw[:n * w] = np.ascontiguousarray(b.values).view(np.uint8).reshape(-1)[:n * w]
        keep.append(raw)
        cols[i].values = raw.ctypes.data
        if b.mayHaveNull():
            nl = np.zeros(pad, np.uint8)
            nl[:n] = np.asarray(b.valueIsNull, np.uint8)[:n]
            keep.append(nl)
            cols[i].nulls = nl.ctypes.data
    outs = (OutCol * len(output_types))()
    bufs = []
    for i, t in enumerate(output_types):
        v, z = np.zeros(max(n, 1) * width_of(t), np.uint8), np.zeros(max(n, 1), np.uint8)
        bufs.append((v, z))
        outs[i].values, outs[i].nulls = v.ctypes.data, z.ctypes.data
    err = ctypes.c_uint32(0)
    lib.emu_filter_project.restype = ctypes.c_longlong
    m = lib.emu_filter_project(cols, ctypes.c_longlong(n), outs, ctypes.byref(err))
    columns = []
    for (v, z), t in zip(bufs, output_types):
        w = width_of(t)
        if t.kind == 0:
            vals = [bool(x) for x in v[:m]]
        elif t.kind in (2, 3):
            vals = [int(x) for x in v[:m * 4].view("<i4")]
        elif t.kind == 4:
            vals = [float(x) for x in v[:m * 8].view("<f8")]
        elif w == 8:
            vals = [int(x) for x in v[:m * 8].view("<i8")]
        else:
            words = v[:m * 16].view("<u8").reshape(-1, 2)
            vals = [(-1 if int(hi) >> 63 else 1) * (int(lo) | ((int(hi) & (2 ** 63 - 1)) << 64)) for lo, hi in words]
        columns.append([None if zz else x for x, zz in zip(vals, z[:m])])
    return list(zip(*columns)) if columns else [], int(err.value)
