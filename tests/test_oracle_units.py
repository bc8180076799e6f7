"""The reference's three sub-unit regression apps (tests/regression/{tex,raster,om}) and their golden PNGs
 against the oracle.  CPU only.

Fixtures: tests/golden/units/*.npz, made by tests/golden/make_unit_fixtures.py in the build container from the
reference tree: the golden PNGs decoded, the texture images the reference's own loader / converter / mip generator
produce for its input PNGs, and vectors rendered by the reference's own kernel (oracle/_ref/libref_tex.so) for states
the goldens do not reach.  That script asserts that the compiled reference kernel reproduces all 19 texture goldens."""
import re

import numpy as np
import pytest

import _oracle_draw3d as od

TEX = od.unit_fixtures("tex")
GOLDEN_CASES = [("toad", f, 0, f"toad_ref_f{f}") for f in range(7)] + [
    (name, 0, g, f"{name}_ref_g{g}") for name in ("soccer", "palette4", "palette16", "palette64") for g in range(3)]
RUN_KEYS = sorted(k for k in TEX.files if k.startswith("run/"))


def texture(name, fmt):
    w, h = (int(x) for x in TEX[f"tex/{name}/f{fmt}/dim"])
    return TEX[f"tex/{name}/f{fmt}"], TEX[f"tex/{name}/f{fmt}/mips"], w, h


@pytest.mark.parametrize("name,fmt,filt,golden", GOLDEN_CASES, ids=[c[3] for c in GOLDEN_CASES])
def test_tex_app_restatement_reproduces_the_reference_golden(name, fmt, filt, golden):
    """ci/regression.sh.in:131-156: `tex -i<name>.png -f<fmt> -g<filt> -r<golden>.png`, CompareImages tolerance 0
    (tex/main.cpp:268).  -g2 only ever reaches the bilinear unit state (tex/main.cpp:247; kernel_arg.filter is never
    set), and at scale 1 every tap sits on a texel centre, which is why the g0 / g1 / g2 goldens are equal."""
    chain, mips, w, h = texture(name, fmt)
    arg = od.tex_kernel_arg(fmt, w, h, 0, filt, mips)
    got = od.tex_image(arg, chain, w, h, 4)
    want = TEX["golden/" + golden]
    assert got.shape == want.shape and (got != want).sum() == 0


@pytest.mark.parametrize("key", RUN_KEYS)
def test_tex_app_restatement_equals_the_reference_kernel_run_in_the_container(key):
    """States the goldens do not reach -- repeat / mirror wrap, magnification, off-centre bilinear taps, mip levels
    1 and 2, ragged task splits -- against images the reference's kernel itself rendered (fixture key = state)."""
    m = re.match(r"run/(\w+)/f(\d)/w(\d)/g(\d)/(\d+)x(\d+)/t(\d+)$", key)
    name, fmt, wrap, filt, dw, dh, tasks = m.group(1), *[int(x) for x in m.groups()[1:]]
    chain, mips, w, h = texture(name, fmt)
    got = od.tex_image(od.tex_kernel_arg(fmt, w, h, wrap, filt, mips), chain, dw, dh, tasks)
    assert (got != TEX[key]).sum() == 0


def test_tex_fixtures_cover_more_than_one_mip_level_and_every_state():
    states = {tuple(int(x) for x in re.match(r"run/\w+/f(\d)/w(\d)/g(\d)/", k).groups()) for k in RUN_KEYS}
    assert {s[0] for s in states} == set(range(7)) and {s[1] for s in states} == {0, 1, 2} and {s[2] for s in states} == {0, 1}
    assert any(k.startswith("run/soccer/") and "/32x32/" in k for k in RUN_KEYS)   # level 1
    assert any(k.startswith("run/soccer/") and "/16x16/" in k for k in RUN_KEYS)   # level 2


def test_trilinear_branch_blends_two_levels():
    """kernel.cpp:67-84 (`filter == 2`, unreachable from the reference app): frac = 0 returns the lower level, and
    between two levels every channel stays within the two levels' values."""
    chain, mips, w, h = texture("soccer", 0)
    arg = od.tex_kernel_arg(0, w, h, 0, 1, mips)
    assert np.array_equal(od.tex_image(arg, chain, 32, 32, 4, trilinear=1), od.tex_image(arg, chain, 32, 32, 4))  # frac = 0
    lo, mid = od.tex_image(arg, chain, 24, 24, 4), od.tex_image(arg, chain, 24, 24, 4, trilinear=1)  # level 1, frac > 0
    assert (lo != mid).any()


@pytest.mark.skipif(od.reference_tex() is None, reason="oracle/_ref/libref_tex.so not built (no /root/reference here)")
def test_committed_vectors_still_equal_the_compiled_reference():
    """Drift check where the compiled reference is present: three stored vectors are rendered again."""
    import ctypes as C
    L = od.reference_tex()
    for key in RUN_KEYS[::max(1, len(RUN_KEYS) // 3)]:
        m = re.match(r"run/(\w+)/f(\d)/w(\d)/g(\d)/(\d+)x(\d+)/t(\d+)$", key)
        name, fmt, wrap, filt, dw, dh, tasks = m.group(1), *[int(x) for x in m.groups()[1:]]
        chain, mips, w, h = texture(name, fmt)
        chain, mips = np.ascontiguousarray(chain), np.ascontiguousarray(mips)
        out = np.zeros((dh, dw), np.uint32)
        assert L.ref_tex_render(chain.ctypes.data, chain.size, mips.ctypes.data, mips.size, w, h, fmt, wrap, filt, dw, dh,
                                tasks, out.ctypes.data) == 0
        assert np.array_equal(out, TEX[key]), key


# ---- raster unit: tests/regression/raster writes white where vx_rast reports coverage -------------------------------
@pytest.mark.parametrize("size", [8, 16, 32, 64, 128])
@pytest.mark.parametrize("which", ["oracle", "reference"])
def test_raster_unit_golden_is_the_coverage_of_the_triangle_frame(which, size):
    """raster/main.cpp renders raster/triangle.cgltrace (the same file as draw3d/triangle.cgltrace) with a kernel that
    stores 0xffffffff for every covered pixel over a 0xff000000 clear (raster/kernel.cpp:8-38, main.cpp:40) and saves
    bottom-up (:196-197): its golden is exactly the set of pixels the draw3d frame touches."""
    L = od.oracle() if which == "oracle" else od.reference()
    if L is None:
        pytest.skip("oracle/_ref/libref_draw3d.so not built (no /root/reference here)")
    fb, _, _ = od.render(L, od.trace_blob("triangle"), size, size, threads=2)
    gold = od.unit_fixtures("raster")[f"triangle_ref_{size}"][::-1]  # -> framebuffer rows
    assert set(np.unique(gold)) == {0xFF000000, 0xFFFFFFFF}
    assert np.array_equal(fb != 0xFF000000, gold == 0xFFFFFFFF)


# ---- output merger: tests/regression/om --------------------------------------------------------------------------------
@pytest.mark.parametrize("size", [8, 16, 32, 64, 128])
def test_om_unit_golden_default_state_writes_the_source_colour(size):
    """om/main.cpp default run (ci/regression.sh.in:165): depth func ALWAYS without depth write over a checkerboard
    depth buffer, stencil KEEP, blend ADD with factors ONE / ZERO, write mask 0xf (:130-171); the kernel sends colour
    0xffffffff at depth 0.5 for every pixel (om/kernel.cpp:15-32, main.cpp:36-37) over a 0 clear -> whitebox_<n>.png."""
    import ctypes as C
    L = od.oracle()
    arg = np.zeros(L.ref_draw3d_kernel_arg_size() // 4, np.uint32)
    om = arg[od.OM_DCRS_OFFSET // 4:]
    om[od.OM_CBUF_WRITEMASK] = 0xF
    om[od.OM_DEPTH_FUNC] = od.VX_OM_DEPTH_FUNC_ALWAYS
    om[od.OM_STENCIL_FUNC] = od.VX_OM_DEPTH_FUNC_ALWAYS
    om[od.OM_STENCIL_MASK] = 0xFF
    om[od.OM_BLEND_MODE] = 0  # ADD / ADD
    om[od.OM_BLEND_FUNC] = (0 << 24) | (0 << 16) | (1 << 8) | 1  # dst ZERO, src ONE
    L.ref_draw3d_blend.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    L.ref_draw3d_blend.restype = C.c_uint32
    L.ref_draw3d_depth_stencil.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
    half, out = 1 << 23, np.zeros((size, size), np.uint32)
    for y in range(size):
        for x in range(size):
            dst_depth = 0 if (x & 1) == (y & 1) else int(0.99 * (1 << 24))  # main.cpp:254-262
            ds = C.c_uint32()
            if L.ref_draw3d_depth_stencil(arg.ctypes.data, half, dst_depth, C.byref(ds)):  # ALWAYS: every pixel passes
                out[y, x] = L.ref_draw3d_blend(arg.ctypes.data, 0xFFFFFFFF, 0x00000000)
    assert np.array_equal(out, od.unit_fixtures("om")[f"whitebox_{size}"])


# ---- the same fixtures through the whole raster path ----------------------------------------------------------------------
UNIT_CASES = od.tex_unit_cases()


@pytest.mark.parametrize("case", UNIT_CASES, ids=[c[0] for c in UNIT_CASES])
def test_textured_quad_through_the_raster_path_reproduces_the_tex_unit_images(case):
    """A viewport-filling textured quad (od.quad_trace) drawn by the draw3d path samples at the tex app's coordinates:
    binning, coverage, attribute interpolation, sampler, texenv REPLACE and output merger together must give the tex
    unit's golden PNG / reference-rendered vector, in the restatement and in the compiled reference."""
    name, blob, w, h, want = case
    fo, _, _ = od.render(od.oracle(), blob, w, h, threads=2)
    if want is not None:
        assert (fo != want).sum() == 0
    if od.reference() is not None:
        fr, _, _ = od.render(od.reference(), blob, w, h, threads=2)
        assert np.array_equal(fo, fr)


@pytest.mark.parametrize("size", [8, 16, 32, 64, 128])
def test_white_quad_through_the_raster_path_is_the_om_unit_golden(size):
    fo, _, _ = od.render(od.oracle(), od.quad_trace(size, size), size, size, threads=2)
    assert np.array_equal(fo, od.unit_fixtures("om")[f"whitebox_{size}"])
    if od.reference() is not None:
        assert np.array_equal(od.render(od.reference(), od.quad_trace(size, size), size, size, threads=2)[0], fo)
