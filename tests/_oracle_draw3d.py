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


# ---- the reference's sub-unit regression apps (tests/regression/{tex,raster,om}); fixtures from
# tests/golden/make_unit_fixtures.py ------------------------------------------------------------------------------
UNITS_DIR = os.path.join(ROOT, "tests", "golden", "units")
TEX_DCRS_OFFSET = 24  # draw3d_kernel_arg_t.tex_dcrs (include/skybox_draw3d.h), stage 0 first
TEX_ADDR, TEX_LOGDIM, TEX_FORMAT, TEX_FILTER, TEX_WRAP, TEX_MIPOFF0 = 0, 1, 2, 3, 4, 5  # index = DCR - TEX_STATE_BEGIN - 1


def unit_fixtures(unit):
    return np.load(os.path.join(UNITS_DIR, unit + ".npz"))


def tex_kernel_arg(fmt, width, height, wrap, filt, mips):
    """A draw3d kernel_arg_t image whose stage-0 texture state is what tests/regression/tex/main.cpp:243-252 programs."""
    arg = np.zeros(oracle().ref_draw3d_kernel_arg_size() // 4, np.uint32)
    t = arg[TEX_DCRS_OFFSET // 4:]
    t[TEX_LOGDIM] = ((int(height).bit_length() - 1) << 16) | (int(width).bit_length() - 1)
    t[TEX_FORMAT] = fmt
    t[TEX_WRAP] = (wrap << 16) | wrap
    t[TEX_FILTER] = 1 if filt else 0
    for i, off in enumerate(mips):
        t[TEX_MIPOFF0 + i] = off
    return arg


def tex_image(arg, chain, dst_width, dst_height, tasks, trilinear=0):
    """oracle/ref_draw3d.c: ref_draw3d_tex_image, the restatement of tests/regression/tex/kernel.cpp."""
    L = oracle()
    L.ref_draw3d_tex_image.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
    L.ref_draw3d_tex_image.restype = None
    chain = np.ascontiguousarray(chain, np.uint8)
    out = np.zeros((dst_height, dst_width), np.uint32)
    L.ref_draw3d_tex_image(arg.ctypes.data, chain.ctypes.data, dst_width, dst_height, tasks, trilinear, out.ctypes.data)
    return out


def reference_tex():
    """oracle/_ref/libref_tex.so (the reference's own texture app, oracle/ref_tex_harness.cpp), or None."""
    if "ref_tex" not in _libs:
        _make()
        p = os.path.join(ORACLE_DIR, "_ref", "libref_tex.so")
        L = None
        if os.path.exists(p):
            L = C.CDLL(p)
            L.ref_tex_render.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p] + [C.c_uint32] * 9 + [C.c_void_p]
        _libs["ref_tex"] = L
    return _libs["ref_tex"]

VX_OM_DEPTH_FUNC_ALWAYS = 0  # include/skybox_draw3d.h:73
OM_DCRS_OFFSET = 196  # draw3d_kernel_arg_t.om_dcrs; index = DCR - VX_DCR_OM_STATE_BEGIN
(OM_CBUF_ADDR, OM_CBUF_PITCH, OM_CBUF_WRITEMASK, OM_ZBUF_ADDR, OM_ZBUF_PITCH, OM_DEPTH_FUNC, OM_DEPTH_WRITEMASK,
 OM_STENCIL_FUNC, OM_STENCIL_ZPASS, OM_STENCIL_ZFAIL, OM_STENCIL_FAIL, OM_STENCIL_REF, OM_STENCIL_MASK,
 OM_STENCIL_WRITEMASK, OM_BLEND_MODE, OM_BLEND_FUNC, OM_BLEND_CONST, OM_LOGIC_OP) = range(18)

TEX_TO_COCOGFX = {0: 5, 1: 4, 2: 6, 3: 8, 4: 3, 5: 2, 6: 1}  # VX_TEX_FORMAT_* -> cocogfx ePixelFormat (format.hpp:24-37)
TEX_BPP = {0: 4, 1: 2, 2: 2, 3: 2, 4: 2, 5: 1, 6: 1}


def quad_trace(width, height, texture=None, bilinear=False, repeat=False, color=(1.0, 1.0, 1.0, 1.0)):
    """One draw call (SKBXTRC1, tests/golden/make_raster_fixtures.py) of two triangles that cover the whole
    width x height viewport, texture coordinates (0, 0) at framebuffer pixel (0, 0): with a texture
    (= (VX_TEX_FORMAT, texel bytes of level 0, w, h)) and texture environment REPLACE every fragment (x, y) samples at
    ((x + 0.5) / width, (y + 0.5) / height) -- the image tests/regression/tex renders, here through binning, raster,
    attribute interpolation, sampler, texenv and output merger; without one the fragments carry the vertex colour
    (tests/regression/om's constant-colour box)."""
    import struct
    out = b"SKBXTRC1" + struct.pack("<II", 1, 1 if texture else 0)
    if texture:
        fmt, pix, w, h = texture
        pix = bytes(pix)
        out += struct.pack("<5I", 1, TEX_TO_COCOGFX[fmt], w, h, len(pix)) + pix + b"\0" * (-len(pix) % 4)
    filt = 2 if bilinear else 1   # SRT_FILTER_LINEAR / NEAREST
    addr = 0 if repeat else 1     # SRT_ADDRESS_WRAP / CLAMP
    states = [0 if texture else 1, 5, 0xF] + [0] * 12 + [1 if texture else 0, 2, filt, filt, addr, addr, 0, 0, 0]
    out += struct.pack("<24I", *states) + struct.pack("<4f", 0, 0, 0, 0) + struct.pack("<I", 1 if texture else 0)
    out += struct.pack("<4i2f", 0, width, 0, height, 0.0, 1.0)
    out += struct.pack("<II", 4, 2)
    for x, y, u, v in ((-1, -1, 0, 0), (1, -1, 1, 0), (-1, 1, 0, 1), (1, 1, 1, 1)):
        out += struct.pack("<10f", x, y, 0, 1, *color, u, v)
    return out + struct.pack("<6I", 0, 1, 2, 2, 1, 3)


def tex_unit_cases():
    """[(id, trace blob, w, h, expected image or None)] for the texture fixtures: the 19 golden PNGs and every
    reference-rendered vector of level 0 whose coordinates the raster path reproduces exactly (power-of-two sizes;
    clamp / repeat -- a trace cannot ask for mirror); the other level-0 vectors come with expected = None (the path
    computes their coordinates differently: compare with the oracle's frame instead)."""
    T = unit_fixtures("tex")

    def tex_of(name, fmt):
        w, h = (int(x) for x in T[f"tex/{name}/f{fmt}/dim"])
        return fmt, T[f"tex/{name}/f{fmt}"][:w * h * TEX_BPP[fmt]].tobytes(), w, h

    cases = []
    for f in range(7):
        cases.append((f"toad_ref_f{f}", quad_trace(64, 64, tex_of("toad", f)), 64, 64, T[f"golden/toad_ref_f{f}"]))
    for name in ("soccer", "palette4", "palette16", "palette64"):
        for g in range(3):
            t = tex_of(name, 0)
            cases.append((f"{name}_ref_g{g}", quad_trace(t[2], t[3], t, bilinear=g > 0), t[2], t[3], T[f"golden/{name}_ref_g{g}"]))
    for key in sorted(k for k in T.files if k.startswith("run/")):
        m = re.match(r"run/(\w+)/f(\d)/w(\d)/g(\d)/(\d+)x(\d+)/t(\d+)$", key)
        name, (fmt, wrap, filt, dw, dh, _) = m.group(1), (int(x) for x in m.groups()[1:])
        t = tex_of(name, fmt)
        if wrap == 2 or dw < t[2] or dh < t[3]:
            continue
        exact = (dw & (dw - 1)) == 0 and (dh & (dh - 1)) == 0
        cases.append((key[4:].replace("/", "-"), quad_trace(dw, dh, t, bilinear=bool(filt), repeat=bool(wrap)), dw, dh,
                      T[key] if exact else None))
    return cases
