#!/usr/bin/env python3
"""Make the raster-path (draw3d) fixtures under tests/golden/raster/ from the reference tree.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):

    python tests/golden/make_raster_fixtures.py [/root/reference]

What it writes, per trace of tests/regression/draw3d/*.cgltrace (the reference's own test inputs):
  <name>.trc.gz   the draw-call trace re-encoded as the flat little-endian "SKBXTRC1" layout that
                  skybox_rt_b200/csrc/host/draw3d_trace.c and oracle/ref_draw3d.c read.  Field order is
                  the reference's serialize() order (third_party/cocogfx/src/cgltrace.cpp:24-142).
                  The .cgltrace files are Boost XML archives; nothing but their field values is kept.
and, per golden image of the reference that has a trace (tests/regression/draw3d/*_ref_*.png):
  <name>_ref_<size>.npy   decoded A8R8G8B8 pixels, uint32 [H, W], row 0 = TOP of the PNG as stored
                          (the reference saves its framebuffer bottom-up, draw3d/main.cpp:383-386).

Vertices: the reference keeps them in an unordered_map keyed by id (cgltrace.hpp:148); here they are
written densely in ascending key order and the primitives' indices are remapped to positions.
"""
import base64
import gzip
import os
import struct
import sys
import xml.etree.ElementTree as ET

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "raster")

STATE_FIELDS = [
    "color_enabled", "color_format", "color_writemask", "depth_test", "depth_writemask", "depth_format",
    "depth_func", "stencil_test", "stencil_func", "stencil_zpass", "stencil_zfail", "stencil_fail",
    "stencil_ref", "stencil_mask", "stencil_writemask", "texture_enabled", "texture_envmode",
    "texture_minfilter", "texture_magfilter", "texture_addressU", "texture_addressV", "blend_enabled",
    "blend_src", "blend_dst",
]


def f32(text):
    return np.float32(float(text))


def convert(path):
    root = ET.parse(path).getroot().find("cgltrace")
    draws = [it for it in root.find("drawcalls") if it.tag == "item"]
    texs = [it for it in root.find("textures") if it.tag == "item"]
    out = bytearray()
    out += b"SKBXTRC1" + struct.pack("<II", len(draws), len(texs))
    for it in texs:
        tid = int(it.find("first").text)
        t = it.find("second")
        size = int(t.find("size").text)
        pix = base64.b64decode("".join(t.find("pixels").text.split()))[:size]
        assert len(pix) == size, (len(pix), size)
        out += struct.pack("<5I", tid, int(t.find("format").text), int(t.find("width").text),
                           int(t.find("height").text), size)
        out += pix + b"\0" * (-size % 4)
    for d in draws:
        st = d.find("states")
        out += struct.pack("<24I", *[int(st.find(f).text) & 0xFFFFFFFF for f in STATE_FIELDS])
        ec = st.find("texture_envcolor")
        out += struct.pack("<4f", *[f32(ec.find(c).text) for c in "rgba"])
        out += struct.pack("<I", int(d.find("texture_id").text))
        vp = d.find("viewport")
        out += struct.pack("<4i2f", *[int(vp.find(k).text) for k in ("left", "right", "top", "bottom")],
                           f32(vp.find("near").text), f32(vp.find("far").text))
        verts = {}
        for it in d.find("vertices"):
            if it.tag != "item":
                continue
            v = it.find("second")
            rec = [f32(v.find("pos").find(c).text) for c in "xyzw"]
            rec += [f32(v.find("color").find(c).text) for c in "rgba"]
            rec += [f32(v.find("texcoord").find(c).text) for c in "uv"]
            verts[int(it.find("first").text)] = rec
        keys = sorted(verts)
        pos = {k: i for i, k in enumerate(keys)}
        prims = [it for it in d.find("primitives") if it.tag == "item"]
        out += struct.pack("<II", len(keys), len(prims))
        for k in keys:
            out += struct.pack("<10f", *verts[k])
        for p in prims:
            out += struct.pack("<3I", *[pos[int(p.find(c).text)] for c in ("i0", "i1", "i2")])
    return bytes(out), len(draws), len(texs)


def png_to_argb(path):
    from PIL import Image
    im = np.asarray(Image.open(path).convert("RGBA")).astype(np.uint32)
    return (im[..., 3] << 24) | (im[..., 0] << 16) | (im[..., 1] << 8) | im[..., 2]


def main():
    os.makedirs(OUT, exist_ok=True)
    d3 = os.path.join(REF, "tests/regression/draw3d")
    for fn in sorted(os.listdir(d3)):
        if fn.endswith(".cgltrace"):
            blob, nd, nt = convert(os.path.join(d3, fn))
            dst = os.path.join(OUT, fn[:-9] + ".trc.gz")
            with gzip.GzipFile(dst, "wb", mtime=0) as f:
                f.write(blob)
            print(f"{fn}: {nd} draws, {nt} textures, {len(blob)} B -> {os.path.getsize(dst)} B")
        elif fn.endswith(".png") and "_ref_" in fn and os.path.exists(os.path.join(d3, fn.split("_ref_")[0] + ".cgltrace")):
            np.save(os.path.join(OUT, fn[:-4] + ".npy"), png_to_argb(os.path.join(d3, fn)))


def copy_raw_traces():
    """Two of the reference's trace files verbatim (test DATA, 4 KB and 52 KB), so that the C reader of the
    Boost-XML form (csrc/host/draw3d_cgltrace.c) is exercised where /root/reference does not exist."""
    import shutil
    for fn in ("triangle.cgltrace", "box.cgltrace"):
        shutil.copyfile(os.path.join(REF, "tests/regression/draw3d", fn), os.path.join(OUT, fn))


if __name__ == "__main__":
    copy_raw_traces()
    main()
