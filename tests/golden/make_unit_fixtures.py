#!/usr/bin/env python3
"""Make the fixtures of the reference's three sub-unit regression apps (tests/regression/{tex,raster,om}) under
tests/golden/units/ from the reference tree.

Run in the build container only (reads /root/reference and loads oracle/_ref/libref_tex.so, the reference's own
texture app compiled by oracle/Makefile; neither exists on the GPU box):

    python tests/golden/make_unit_fixtures.py [/root/reference]

What it writes (numpy .npz, compressed):
  tex.npz     per input image of tests/regression/tex and texture format 0..6 the texture image the reference app
              uploads -- cocogfx LoadImage's conversion + GenerateMipmaps, run from the reference's own code
              (key  tex/<image>/f<format>  = the mip chain as bytes,  .../mips = its offsets,  .../dim = (w, h));
              every golden PNG of that directory decoded to A8R8G8B8 words, row 0 = top of the PNG, which is row 0
              of the app's destination image (it saves without a flip, tex/main.cpp:147)
              (key  golden/<file name without .png>);
              and vectors the golden PNGs do not cover, rendered HERE by the reference's own kernel through
              oracle/ref_tex_harness.cpp: other scales (mip levels 1 and 2, magnification), wrap modes, bilinear
              off the texel centres (key  run/<image>/f<format>/w<wrap>/g<filter>/<dst w>x<dst h>/t<tasks>).
  raster.npz  tests/regression/raster/triangle_ref_{8..128}.png decoded the same way (white = covered)
  om.npz      tests/regression/om/whitebox_{8..128}.png
"""
import ctypes as C
import os
import sys

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "units")
ROOT = os.path.dirname(os.path.dirname(HERE))

INPUTS = ["toad", "soccer", "palette4", "palette16", "palette64"]
# vectors beyond the goldens: (image, format, wrap, filter, dst w, dst h, tasks)
EXTRA = (
    [("toad", f, w, g, 64, 64, 4) for f in range(7) for w in (0, 1, 2) for g in (0, 1)] +          # all states at 1:1
    [("toad", f, 0, g, 128, 128, 16) for f in (0, 1, 3) for g in (0, 1)] +                          # magnification
    [("toad", f, w, 1, 100, 90, 7) for f in (0, 2, 4) for w in (0, 1, 2)] +                        # off-centre taps
    [("soccer", 0, w, g, 32, 32, 8) for w in (0, 1) for g in (0, 1)] +                             # mip level 1
    [("soccer", 1, 0, 1, 16, 16, 3), ("soccer", 0, 2, 1, 24, 40, 5), ("palette16", 0, 1, 1, 64, 64, 64),
     ("palette4", 0, 2, 1, 32, 32, 1), ("palette64", 5, 1, 1, 48, 48, 6), ("palette64", 6, 0, 0, 20, 20, 20)])


def png_pixels(path):
    """-> (uint8 [H, W, channels] in the byte order cocogfx's LoadPNG produces (BGR / BGRA, png.cpp:78), channels)."""
    from PIL import Image
    im = Image.open(path)
    if im.mode == "RGB":
        a = np.asarray(im)[..., ::-1]
    else:
        a = np.asarray(im.convert("RGBA"))[..., [2, 1, 0, 3]]
    return np.ascontiguousarray(a, np.uint8), a.shape[2]


def png_to_argb(path):
    from PIL import Image
    im = np.asarray(Image.open(path).convert("RGBA")).astype(np.uint32)
    return (im[..., 3] << 24) | (im[..., 0] << 16) | (im[..., 1] << 8) | im[..., 2]


def reference_tex():
    import subprocess
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), f"REFERENCE={REF}"], check=True)
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_tex.so"))
    L.ref_tex_make_texture.restype = C.c_int64
    L.ref_tex_make_texture.argtypes = [C.c_void_p] + [C.c_uint32] * 4 + [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32,
                                                                         C.POINTER(C.c_uint32)]
    L.ref_tex_render.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p] + [C.c_uint32] * 9 + [C.c_void_p]
    return L


def make_texture(L, pix, channels, fmt):
    h, w = pix.shape[:2]
    out = np.zeros(w * h * 4 * 2, np.uint8)
    mips = np.zeros(16, np.uint32)
    n = C.c_uint32()
    size = L.ref_tex_make_texture(pix.ctypes.data, channels, w, h, fmt, out.ctypes.data, out.size, mips.ctypes.data, 16,
                                  C.byref(n))
    assert size > 0, (size, fmt)
    return out[:size].copy(), mips[:n.value].copy()


def reference_render(L, chain, mips, w, h, fmt, wrap, filt, dw, dh, tasks):
    out = np.zeros((dh, dw), np.uint32)
    rc = L.ref_tex_render(chain.ctypes.data, chain.size, mips.ctypes.data, mips.size, w, h, fmt, wrap, filt, dw, dh, tasks,
                          out.ctypes.data)
    assert rc == 0
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    L = reference_tex()
    tdir = os.path.join(REF, "tests/regression/tex")
    tex = {}
    textures = {}
    for name in INPUTS:
        pix, ch = png_pixels(os.path.join(tdir, name + ".png"))
        h, w = pix.shape[:2]
        for fmt in range(7):
            chain, mips = make_texture(L, pix, ch, fmt)
            textures[name, fmt] = (chain, mips, w, h)
            tex[f"tex/{name}/f{fmt}"] = chain
            tex[f"tex/{name}/f{fmt}/mips"] = mips
            tex[f"tex/{name}/f{fmt}/dim"] = np.array([w, h], np.uint32)
    for fn in sorted(os.listdir(tdir)):
        if fn.endswith(".png") and "_ref_" in fn:
            tex["golden/" + fn[:-4]] = png_to_argb(os.path.join(tdir, fn))
    for name, fmt, wrap, filt, dw, dh, tasks in EXTRA:
        chain, mips, w, h = textures[name, fmt]
        tex[f"run/{name}/f{fmt}/w{wrap}/g{filt}/{dw}x{dh}/t{tasks}"] = reference_render(L, chain, mips, w, h, fmt, wrap, filt,
                                                                                      dw, dh, tasks)
    # the reference app's own command lines (ci/regression.sh.in:131-156) through the same harness must give its goldens
    for fmt in range(7):
        chain, mips, w, h = textures["toad", fmt]
        got = reference_render(L, chain, mips, w, h, fmt, 0, 0, w, h, 4)
        assert np.array_equal(got, tex[f"golden/toad_ref_f{fmt}"]), f"harness != toad_ref_f{fmt}.png"
    for name in ("soccer", "palette4", "palette16", "palette64"):
        chain, mips, w, h = textures[name, 0]
        for g in range(3):
            got = reference_render(L, chain, mips, w, h, 0, 0, g, w, h, 4)
            assert np.array_equal(got, tex[f"golden/{name}_ref_g{g}"]), f"harness != {name}_ref_g{g}.png"
    np.savez_compressed(os.path.join(OUT, "tex.npz"), **tex)
    print(f"tex.npz: {len(tex)} arrays, {os.path.getsize(os.path.join(OUT, 'tex.npz'))} B; the harness reproduces all "
          f"{sum(k.startswith('golden/') for k in tex)} golden PNGs")
    for unit, pattern in (("raster", "triangle_ref_"), ("om", "whitebox_")):
        d = os.path.join(REF, "tests/regression", unit)
        arrays = {fn[:-4]: png_to_argb(os.path.join(d, fn)) for fn in sorted(os.listdir(d))
                  if fn.endswith(".png") and fn.startswith(pattern)}
        np.savez_compressed(os.path.join(OUT, unit + ".npz"), **arrays)
        print(f"{unit}.npz: {sorted(arrays)}")


if __name__ == "__main__":
    main()
