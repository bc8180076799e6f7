"""ctypes binding of the raster-path oracle.  Test infrastructure only.

  oracle()     oracle/libref_draw3d.so       plain-C restatement (always buildable from this repo)
  reference()  oracle/_ref/libref_draw3d.so  the reference's own code compiled natively; built by
               oracle/Makefile when /root/reference is present, otherwise the prebuilt file that
               travelled with the snapshot, otherwise None.
Both export the same two entry points (oracle/ref_draw3d.h).
"""
import ctypes as C
import glob
import gzip
import os
import re
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden", "raster")

_libs = {}


def _bind(path):
    L = C.CDLL(path)
    L.ref_draw3d_render.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                    C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    L.ref_draw3d_binning.restype = C.c_int64
    L.ref_draw3d_binning.argtypes = [C.c_char_p, C.c_uint64] + [C.c_uint32] * 4 + [
        C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p, C.POINTER(C.c_uint64)]
    L.ref_draw3d_kernel_arg_size.restype = C.c_uint32
    L.ref_draw3d_prim_size.restype = C.c_uint32
    return L


def _make():
    if "made" not in _libs:
        subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)
        _libs["made"] = True


def oracle():
    if "oracle" not in _libs:
        _make()
        _libs["oracle"] = _bind(os.path.join(ORACLE_DIR, "libref_draw3d.so"))
    return _libs["oracle"]


def reference():
    if "ref" not in _libs:
        _make()
        p = os.path.join(ORACLE_DIR, "_ref", "libref_draw3d.so")
        _libs["ref"] = _bind(p) if os.path.exists(p) else None
    return _libs["ref"]


def render(L, blob, width, height, threads=4, depth=True):
    """-> (colour [H, W] u32, depth [H, W] u32, kernel seconds); row 0 = framebuffer row 0."""
    fb = np.zeros((height, width), np.uint32)
    zb = np.zeros((height, width), np.uint32)
    ks = C.c_double()
    rc = L.ref_draw3d_render(blob, len(blob), width, height, threads, 5, fb.ctypes.data,
                             zb.ctypes.data if depth else None, C.byref(ks))
    assert rc == 0, f"ref_draw3d_render -> {rc}"
    return fb, zb, ks.value


def binning(L, blob, draw, width, height, cap=64 << 20):
    tb = np.zeros(cap, np.uint8)
    pb = np.zeros(cap, np.uint8)
    ts, ps = C.c_uint64(cap), C.c_uint64(cap)
    n = L.ref_draw3d_binning(blob, len(blob), draw, width, height, 5, tb.ctypes.data, C.byref(ts),
                             pb.ctypes.data, C.byref(ps))
    assert n >= 0, f"ref_draw3d_binning -> {n}"
    return int(n), tb[:ts.value].tobytes(), pb[:ps.value].tobytes()


def trace_names():
    return sorted(os.path.basename(p)[:-7] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.trc.gz")))


def trace_blob(name):
    with gzip.open(os.path.join(GOLDEN_DIR, name + ".trc.gz"), "rb") as f:
        return f.read()


def goldens():
    """[(trace name, size, path)] for every golden image of the reference that has a trace."""
    out = []
    for p in sorted(glob.glob(os.path.join(GOLDEN_DIR, "*_ref_*.npy"))):
        m = re.match(r"(\w+)_ref_(\d+)\.npy$", os.path.basename(p))
        out.append((m.group(1), int(m.group(2)), p))
    return out


def golden_framebuffer(path):
    """The golden PNG as framebuffer rows (the reference saves bottom-up, draw3d/main.cpp:383-386)."""
    return np.load(path)[::-1]


def num_draws(blob):
    return int(np.frombuffer(blob[8:12], np.uint32)[0])
