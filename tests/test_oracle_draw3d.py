"""CPU tests of the raster path: the oracle against the reference's golden images and (when it is
available) against the reference's own code compiled natively; the product's host pre-pass (trace
loader, binning) against both; the C-ABI layouts."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

import _oracle_draw3d as od

GOLDENS = od.goldens()
NAMES = od.trace_names()


def test_fixture_inventory():
    # the 9 traces of tests/regression/draw3d and their 17 golden images
    assert NAMES == ["box", "carnival", "evilskull", "mouse", "polybump", "scene", "tekkaman", "triangle", "vase"]
    assert len(GOLDENS) == 17


@pytest.mark.parametrize("name,size,path", GOLDENS, ids=[f"{n}@{s}" for n, s, _ in GOLDENS])
def test_oracle_reproduces_reference_golden(name, size, path):
    fb, _, _ = od.render(od.oracle(), od.trace_blob(name), size, size, threads=2)
    gold = od.golden_framebuffer(path)
    assert gold.shape == fb.shape
    assert int((fb != gold).sum()) == 0  # the reference's own compare allows 1/255 (draw3d/main.cpp:507)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_compiled_reference(name):
    ref = od.reference()
    if ref is None:
        pytest.skip("oracle/_ref/libref_draw3d.so not built (needs /root/reference)")
    blob = od.trace_blob(name)
    for w, h in [(96, 96), (200, 120), (333, 517)]:
        fo, zo, _ = od.render(od.oracle(), blob, w, h)
        fr, zr, _ = od.render(ref, blob, w, h)
        assert np.array_equal(fo, fr), (name, w, h, int((fo != fr).sum()))
        assert np.array_equal(zo, zr), (name, w, h)
    for d in range(od.num_draws(blob)):
        assert od.binning(od.oracle(), blob, d, 300, 200) == od.binning(ref, blob, d, 300, 200)


def test_oracle_is_thread_count_independent():
    blob = od.trace_blob("vase")
    a = od.render(od.oracle(), blob, 160, 160, threads=1)
    b = od.render(od.oracle(), blob, 160, 160, threads=7)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_layout_sizes_match_reference():
    # tests/regression/draw3d/common.h kernel_arg_t and sim/common/graphics.h rast_prim_t
    assert od.oracle().ref_draw3d_kernel_arg_size() == 288
    assert od.oracle().ref_draw3d_prim_size() == 120
    ref = od.reference()
    if ref is not None:
        assert ref.ref_draw3d_kernel_arg_size() == 288
        assert ref.ref_draw3d_prim_size() == 120


# ---------------------------------------------------------------- product host pre-pass (no GPU)
@pytest.fixture(scope="module")
def d3(native):
    from skybox_rt_b200 import draw3d
    return draw3d


@pytest.mark.parametrize("name", NAMES)
def test_product_binning_is_byte_identical(d3, name):
    blob = od.trace_blob(name)
    t = d3.Trace(blob=blob)
    checkers = [od.oracle()] + ([od.reference()] if od.reference() is not None else [])
    for w, h in [(128, 128), (300, 200), (37, 61)]:
        for d in range(t.num_drawcalls):
            n, tb, pb, pairs = t.binning(d, w, h)
            for L in checkers:
                assert (n, tb, pb) == od.binning(L, blob, d, w, h), (name, w, h, d)
            assert pairs * 4 + n * 8 == len(tb)


def test_wide_tile_headers_carry_the_same_lists(d3):
    t = d3.Trace(blob=od.trace_blob("mouse"))
    for d in range(t.num_drawcalls):
        n, tb, pb, pairs = t.binning(d, 640, 480)
        nw, tbw, pbw, pairsw = t.binning(d, 640, 480, fmt=d3.TILEBUF_WIDE)
        assert (n, pb, pairs) == (nw, pbw, pairsw)
        hdr = np.frombuffer(tb[:n * 8], np.uint16).reshape(n, 4)
        wide = np.frombuffer(tbw[:n * 16], np.uint32).reshape(n, 4)
        pids = np.frombuffer(tb[n * 8:], np.uint32)
        pidsw = np.frombuffer(tbw[n * 16:], np.uint32)
        assert np.array_equal(pids, pidsw)
        for i in range(n):
            assert (wide[i, 0] & 0xFFFF, wide[i, 0] >> 16, wide[i, 1]) == (hdr[i, 0], hdr[i, 1], hdr[i, 3])
            # reference offset counts u32s from the end of header i; the wide one from the first pid
            assert hdr[i, 2] == (n - 1 - i) * 2 + wide[i, 2]


def test_reference_tile_format_refuses_to_overflow(d3):
    from skybox_rt_b200 import vortex as vx
    t = d3.Trace(blob=od.trace_blob("carnival"))
    with pytest.raises(vx.VortexError, match="-2"):
        t.binning(0, 8192, 8192)  # 65536 tiles: the 16-bit pids_offset of the first header would wrap
    n, tb, pb, pairs = t.binning(0, 8192, 8192, fmt=d3.TILEBUF_WIDE)
    assert n == 65536 and pairs == 131072


def test_rank_split_partitions_the_tiles(d3):
    t = d3.Trace(blob=od.trace_blob("evilskull"))
    for d in range(t.num_drawcalls):
        n, tb, pb, pairs = t.binning(d, 512, 384, fmt=d3.TILEBUF_WIDE)
        full = np.frombuffer(tb[:n * 16], np.uint32).reshape(n, 4)
        seen, tot_pairs = [], 0
        for r in range(3):
            nr, tbr, pbr, pr = t.binning(d, 512, 384, fmt=d3.TILEBUF_WIDE, rank=r, num_ranks=3)
            assert pbr == pb  # primitives are replicated, ids keep their meaning
            h = np.frombuffer(tbr[:nr * 16], np.uint32).reshape(nr, 4)
            assert all(((x >> 16) % 3) == r for x in h[:, 0])
            seen += [int(x) for x in h[:, 0]]
            tot_pairs += pr
        assert sorted(seen) == sorted(int(x) for x in full[:, 0]) and tot_pairs == pairs


def test_trace_loader_rejects_bad_blobs(d3):
    from skybox_rt_b200 import vortex as vx
    blob = od.trace_blob("triangle")
    with pytest.raises(vx.VortexError):
        d3.Trace(blob=b"NOTATRACE" + blob[9:])
    with pytest.raises(vx.VortexError):
        d3.Trace(blob=blob[:len(blob) - 5])  # truncated primitive list
    bad = bytearray(blob)
    struct.pack_into("<I", bad, len(bad) - 4, 99)  # vertex index out of range (vertices.at() throws upstream)
    with pytest.raises(vx.VortexError):
        d3.Trace(blob=bytes(bad))
    assert d3.Trace(blob=blob).num_drawcalls == 1


def test_draw3d_libraries_export_their_api(native):
    host = C.CDLL(native["host"])
    for sym in ("srt_trace_load", "srt_trace_load_cgltrace", "srt_trace_from_blob", "srt_trace_free", "srt_trace_texture", "srt_drawcall_at",
                "srt_draw3d_binning", "srt_draw3d_binning_rank", "srt_binning_free", "srt_format_bpp", "srt_format_to_vx"):
        assert hasattr(host, sym), sym
    C.CDLL(native["vortex"], mode=C.RTLD_GLOBAL)
    app = C.CDLL(native["draw3d"])
    for sym in ("srt_draw3d_open", "srt_draw3d_close", "srt_draw3d_clear", "srt_draw3d_render", "srt_draw3d_prepare",
                "srt_draw3d_submit", "srt_draw3d_read", "srt_draw3d_stats", "srt_draw3d_color_buffer",
                "srt_draw3d_depth_buffer"):
        assert hasattr(app, sym), sym
    assert os.path.exists(os.path.join(os.path.dirname(native["vortex"]), "..", "bin", "draw3d"))


VARIANTS = od.state_variants()


@pytest.mark.parametrize("label,kw", VARIANTS, ids=[v[0] for v in VARIANTS])
def test_oracle_state_space_matches_compiled_reference(label, kw):
    """The traces only use a corner of the fixed-function state (LESS/LEQUAL depth, SRC_ALPHA blending, three
    texture formats, bilinear + repeat).  Every other value the app can program is exercised by overwriting the
    state words of a trace; the restatement must agree with the reference's own code on each."""
    ref = od.reference()
    if ref is None:
        pytest.skip("oracle/_ref/libref_draw3d.so not built (needs /root/reference)")
    for name, size in (("vase", 96), ("evilskull", 80)):
        blob = od.mutate(od.trace_blob(name), **kw)
        fo, zo, _ = od.render(od.oracle(), blob, size, size, threads=2)
        fr, zr, _ = od.render(ref, blob, size, size, threads=2)
        assert np.array_equal(fo, fr) and np.array_equal(zo, zr), (label, name)


OM_VARIANTS = od.om_override_variants()


@pytest.mark.parametrize("label,ov", OM_VARIANTS, ids=[v[0] for v in OM_VARIANTS])
def test_oracle_blend_equations_match_compiled_reference(label, ov):
    """Blend equations other than ADD, the 16 logic ops, CONST_* factors and the blend constant: the restatement
    against the reference's own Blender (sim/common/graphics.cpp:366-524, :603-636) through the OM override hook."""
    ref = od.reference()
    if ref is None:
        pytest.skip("oracle/_ref/libref_draw3d.so not built (needs /root/reference)")
    for name, size in (("vase", 96), ("mouse", 80)):
        blob = od.mutate(od.trace_blob(name), blend_enabled=1, blend_src=4, blend_dst=5)
        fo, zo, _ = od.render(od.oracle(), blob, size, size, threads=2, om_override=ov)
        fr, zr, _ = od.render(ref, blob, size, size, threads=2, om_override=ov)
        assert np.array_equal(fo, fr) and np.array_equal(zo, zr), (label, name)
    # the override really reaches the blender: a non-ADD state must change the frame
    base, _, _ = od.render(ref, blob, size, size, threads=2)
    if label not in ("logic_op=3",):  # COPY: src replaces dst, may coincide with opaque draws
        assert not np.array_equal(base, fr) or label.startswith("src factor") or label.startswith("dst factor"), label


@pytest.mark.parametrize("name", ["triangle", "box"])
def test_c_reader_of_reference_cgltrace_files(d3, name):
    """The reference's own Boost-XML archives, read by the C loader (no Boost), give byte for byte the trace the
    fixture converter made from the same file (floats via strtof, base64 texture payload, vertex-id remap)."""
    t = d3.Trace(path=os.path.join(od.GOLDEN_DIR, name + ".cgltrace"))
    assert t.blob == od.trace_blob(name)
    ref_dir = "/root/reference/tests/regression/draw3d"
    if name == "box" and os.path.isdir(ref_dir):  # in the build container: every trace the reference ships
        for n in od.trace_names():
            assert d3.Trace(path=os.path.join(ref_dir, n + ".cgltrace")).blob == od.trace_blob(n), n


def test_c_reader_rejects_malformed_archives(d3, tmp_path):
    from skybox_rt_b200 import vortex as vx
    xml = open(os.path.join(od.GOLDEN_DIR, "triangle.cgltrace"), "rb").read()
    for i, bad in enumerate([xml[:len(xml) // 2], xml.replace(b"<i2>", b"<j2>"), b"<?xml version='1.0'?><nothing/>"]):
        p = tmp_path / f"bad{i}.cgltrace"
        p.write_bytes(bad)
        with pytest.raises(vx.VortexError):
            d3.Trace(path=str(p))


def test_draw_arrays_expose_the_trace_geometry(d3):
    """Trace.draw_arrays (the input of the device pre-pass) = the vertex / index arrays of the flat trace blob."""
    import struct
    blob = od.trace_blob("box")
    t = d3.Trace(blob=blob)
    _, draws = od._walk(blob)
    for d, off in enumerate(draws):
        near, far = struct.unpack_from("<2f", blob, off + 112 + 4 + 16)
        nv, npr = struct.unpack_from("<II", blob, off + 112 + 4 + 24)
        v, p, n, f = t.draw_arrays(d)
        assert v.shape == (nv, 10) and p.shape == (npr, 3) and (n, f) == (near, far)
        base = off + 112 + 4 + 24 + 8
        assert v.tobytes() == blob[base:base + nv * 40] and p.tobytes() == blob[base + nv * 40:base + nv * 40 + npr * 12]
        assert p.max() < nv
    with pytest.raises(IndexError):
        t.draw_arrays(len(draws))
