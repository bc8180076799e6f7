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
    L.ref_draw3d_set_om_override.argtypes = [C.POINTER(C.c_uint32)]
    L.ref_draw3d_set_om_override.restype = None
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


def render(L, blob, width, height, threads=4, depth=True, om_override=None):
    """-> (colour [H, W] u32, depth [H, W] u32, kernel seconds); row 0 = framebuffer row 0.
    om_override = (blend mode, blend func, blend constant, logic op) DCR values for blending draws, or None."""
    fb = np.zeros((height, width), np.uint32)
    zb = np.zeros((height, width), np.uint32)
    ks = C.c_double()
    L.ref_draw3d_set_om_override((C.c_uint32 * 4)(*om_override) if om_override is not None else None)
    try:
        rc = L.ref_draw3d_render(blob, len(blob), width, height, threads, 5, fb.ctypes.data,
                                 zb.ctypes.data if depth else None, C.byref(ks))
    finally:
        L.ref_draw3d_set_om_override(None)
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


# ---------------------------------------------------------------- state mutation (widens what the traces cover)
STATE_FIELDS = ["color_enabled", "color_format", "color_writemask", "depth_test", "depth_writemask", "depth_format",
                "depth_func", "stencil_test", "stencil_func", "stencil_zpass", "stencil_zfail", "stencil_fail",
                "stencil_ref", "stencil_mask", "stencil_writemask", "texture_enabled", "texture_envmode",
                "texture_minfilter", "texture_magfilter", "texture_addressU", "texture_addressV", "blend_enabled",
                "blend_src", "blend_dst"]
FORMAT_BPP = {1: 1, 2: 1, 3: 2, 4: 2, 5: 4, 6: 2, 8: 2}  # cocogfx ePixelFormat -> bytes per texel


def _walk(blob):
    """-> ([(texture header offset, format, width, height, size)], [offset of each draw's 24 state words])"""
    import struct
    nd, nt = struct.unpack_from("<II", blob, 8)
    p, texs, draws = 16, [], []
    for _ in range(nt):
        tid, fmt, w, h, sz = struct.unpack_from("<5I", blob, p)
        texs.append((p, fmt, w, h, sz))
        p += 20 + ((sz + 3) & ~3)
    for _ in range(nd):
        draws.append(p)
        nv, npr = struct.unpack_from("<II", blob, p + 112 + 4 + 24)
        p += 112 + 4 + 24 + 8 + nv * 40 + npr * 12
    return texs, draws


def mutate(blob, texture_format=None, only_draws=None, **states):
    """A copy of the trace with the given state fields overwritten in every draw (or `only_draws`), and
    optionally every texture re-labelled as `texture_format` (skipped where the texels would not fit)."""
    import struct
    out = bytearray(blob)
    texs, draws = _walk(blob)
    for i, off in enumerate(draws):
        if only_draws is not None and i not in only_draws:
            continue
        for k, v in states.items():
            struct.pack_into("<I", out, off + 4 * STATE_FIELDS.index(k), v)
    if texture_format is not None:
        for off, fmt, w, h, sz in texs:
            if w * h * FORMAT_BPP[texture_format] <= sz:
                struct.pack_into("<I", out, off + 4, texture_format)
    return bytes(out)


def state_variants():
    """(label, kwargs for mutate) -- every value the app can program into the OM / TEX units."""
    v = []
    for f in range(8):
        v.append((f"depth_func={f}", dict(depth_test=1, depth_writemask=1, depth_func=f)))
    v.append(("depth_nowrite", dict(depth_test=1, depth_writemask=0, depth_func=1)))
    for s in range(11):
        v.append((f"blend_src={s}", dict(blend_enabled=1, blend_src=s, blend_dst=5)))
    for d in range(11):
        v.append((f"blend_dst={d}", dict(blend_enabled=1, blend_src=4, blend_dst=d)))
    for func, zpass, zfail, fail, ref, mask, wm in [(7, 1, 0, 0, 0x5a, 0xff, 0xff), (1, 2, 3, 4, 0x80, 0xf0, 0x0f),
                                                    (5, 5, 1, 2, 0xff, 0xff, 0xff), (2, 3, 3, 3, 0xff, 0xff, 0xf0),
                                                    (0, 4, 5, 1, 7, 0x3c, 0xff), (3, 2, 2, 5, 1, 0xff, 0xaa)]:
        v.append((f"stencil={func},{zpass},{zfail},{fail}", dict(stencil_test=1, stencil_func=func, stencil_zpass=zpass,
                                                               stencil_zfail=zfail, stencil_fail=fail, stencil_ref=ref,
                                                               stencil_mask=mask, stencil_writemask=wm)))
    for fmt in (1, 2, 3, 4, 5, 6, 8):
        for mag in (1, 2):
            for addr in (0, 1):
                v.append((f"tex fmt={fmt} mag={mag} addr={addr}",
                          dict(texture_format=fmt, texture_magfilter=mag, texture_addressU=addr)))
    for wm in (0x0, 0x5, 0xA, 0x7):
        v.append((f"color_writemask={wm:#x}", dict(color_writemask=wm)))
    for ce, env in [(1, 3), (1, 2), (0, 3), (0, 2)]:
        v.append((f"color_enabled={ce} envmode={env}", dict(color_enabled=ce, texture_envmode=env)))
    v.append(("no texture", dict(texture_enabled=0, color_enabled=1)))
    return v


def om_override_variants():
    """(label, (blend mode, blend func, blend constant, logic op)): the output-merger state the reference app never
    programs (it is MODE_ADD only, main.cpp:271-279) but its units implement -- sim/common/graphics.cpp:366-404
    (DoLogicOp, 16 ops), :455-472 (CONST_* factors), :477-524 (SUB / REV_SUB / MIN / MAX / LOGICOP), :611 (constant)."""
    ADD, SUB, REV_SUB, MIN, MAX, LOGICOP = range(6)
    ONE, SRC_A, ONE_MINUS_SRC_A, DST_RGB = 1, 6, 7, 4
    CONST_RGB, ONE_MINUS_CONST_RGB, CONST_A, ONE_MINUS_CONST_A = 10, 11, 12, 13

    def func(src_rgb, src_a, dst_rgb, dst_a):
        return dst_a << 24 | dst_rgb << 16 | src_a << 8 | src_rgb
    v = []
    cst = 0x40C08020
    for m in (SUB, REV_SUB, MIN, MAX):
        v.append((f"mode={m}", (m << 16 | m, func(SRC_A, SRC_A, ONE_MINUS_SRC_A, ONE_MINUS_SRC_A), 0, 0)))
        v.append((f"mode rgb={m} a=ADD one/one", (ADD << 16 | m, func(ONE, ONE, ONE, ONE), 0, 0)))
    v.append(("mode rgb=ADD a=SUB", (SUB << 16 | ADD, func(SRC_A, ONE, DST_RGB, ONE), 0, 0)))
    for op in range(16):
        v.append((f"logic_op={op}", (LOGICOP << 16 | LOGICOP, func(ONE, ONE, ONE, ONE), 0, op)))
    v.append(("logic_op rgb only", (ADD << 16 | LOGICOP, func(ONE, SRC_A, ONE, ONE_MINUS_SRC_A), 0, 6)))
    for f in (CONST_RGB, ONE_MINUS_CONST_RGB, CONST_A, ONE_MINUS_CONST_A):
        v.append((f"src factor={f}", (ADD << 16 | ADD, func(f, f, ONE_MINUS_SRC_A, ONE_MINUS_SRC_A), cst, 0)))
        v.append((f"dst factor={f}", (ADD << 16 | ADD, func(SRC_A, SRC_A, f, f), cst, 0)))
        v.append((f"factor={f} REV_SUB", (REV_SUB << 16 | REV_SUB, func(f, ONE, ONE, f), 0xFF102030, 0)))
    return v
