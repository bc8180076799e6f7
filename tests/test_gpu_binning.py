"""Triangle setup + tile binning on the GPU (vx_cuda_draw3d_binning, csrc/raster_binning.cu) against the host
pre-pass (csrc/host/draw3d_binning.c), which test_oracle_draw3d.py pins byte-for-byte to the reference's
graphics::Binning (sim/common/gfxutil.cpp:103-276) compiled natively.  Integer / byte work: the bar is identical
bytes for primbuf and tilebuf."""
import numpy as np
import pytest

import _oracle_draw3d as od
from skybox_rt_b200 import vortex as vx

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def d3(native):
    from skybox_rt_b200 import draw3d
    return draw3d


@pytest.fixture(scope="module")
def dev(native):
    with vx.Device() as d:
        yield d


def same(a, b, what):
    assert a[0] == b[0], f"{what}: {a[0]} tiles on the GPU, {b[0]} on the host"
    assert a[3] == b[3], f"{what}: {a[3]} tile-primitive pairs on the GPU, {b[3]} on the host"
    assert a[2] == b[2], f"{what}: primbuf differs"
    assert a[1] == b[1], f"{what}: tilebuf differs"


@pytest.mark.parametrize("name", od.trace_names())
def test_every_trace_small_sizes_reference_format(d3, dev, name):
    blob = od.trace_blob(name)
    t = d3.Trace(blob=blob)
    for size in (8, 32, 128, 200):
        for d in range(od.num_draws(blob)):
            same(t.gpu_binning(dev, d, size, size), t.binning(d, size, size), f"{name}@{size} draw {d}")


@pytest.mark.parametrize("name,w,h", [("mouse", 4096, 4096), ("vase", 2048, 1024), ("carnival", 1000, 3000),
                                      ("evilskull", 4096, 4096), ("scene", 8192, 8192)])
def test_large_frames_wide_format(d3, dev, name, w, h):
    blob = od.trace_blob(name)
    t = d3.Trace(blob=blob)
    for d in range(od.num_draws(blob)):
        same(t.gpu_binning(dev, d, w, h, fmt=d3.TILEBUF_WIDE), t.binning(d, w, h, fmt=d3.TILEBUF_WIDE), f"{name} draw {d}")


@pytest.mark.parametrize("ranks", [2, 3, 8])
def test_rank_dealing(d3, dev, ranks):
    blob = od.trace_blob("tekkaman")
    t = d3.Trace(blob=blob)
    for r in range(ranks):
        for d in range(od.num_draws(blob)):
            same(t.gpu_binning(dev, d, 1024, 1024, fmt=d3.TILEBUF_WIDE, rank=r, num_ranks=ranks),
                 t.binning(d, 1024, 1024, fmt=d3.TILEBUF_WIDE, rank=r, num_ranks=ranks), f"rank {r}/{ranks} draw {d}")


@pytest.mark.parametrize("logsize", [3, 4, 6])
def test_other_tile_sizes(d3, dev, logsize):
    t = d3.Trace(blob=od.trace_blob("polybump"))
    same(t.gpu_binning(dev, 0, 512, 384, fmt=d3.TILEBUF_WIDE, tile_logsize=logsize),
         t.binning(0, 512, 384, fmt=d3.TILEBUF_WIDE, tile_logsize=logsize), f"tile 2^{logsize}")


def test_long_lists_sort_in_global_memory(d3, dev):
    """More than 4096 primitives in one tile: the per-tile sort leaves shared memory."""
    rng = np.random.default_rng(5)
    n = 6000
    v = np.zeros((3 * n, 10), np.float32)
    v[:, 0:2] = rng.uniform(-0.9, 0.9, (3 * n, 2))
    v[:, 2] = rng.uniform(-1, 1, 3 * n)
    v[:, 3] = 1.0
    v[:, 4:8] = rng.uniform(0, 1, (3 * n, 4))
    v[:, 8:10] = rng.uniform(0, 4, (3 * n, 2))
    p = np.arange(3 * n, dtype=np.uint32).reshape(n, 3)
    got = d3.gpu_binning(dev, v, p, 0.0, 1.0, 64, 64, d3.TILEBUF_WIDE)
    # host pre-pass on the same arrays, through a synthetic one-draw trace
    blob = synthetic_trace(v, p)
    same(got, d3.Trace(blob=blob).binning(0, 64, 64, fmt=d3.TILEBUF_WIDE), "6000 random triangles at 64x64")
    assert got[3] > 4096 * 2


def synthetic_trace(v, p, near=0.0, far=1.0):
    """One draw call, no textures, in the SKBXTRC1 layout (tests/golden/make_raster_fixtures.py)."""
    import struct
    states = [1, 5, 0xF] + [0] * 21
    return (b"SKBXTRC1" + struct.pack("<II", 1, 0) + struct.pack("<24I", *states) + struct.pack("<4f", 0, 0, 0, 0) +
            struct.pack("<I", 0) + struct.pack("<4i", 0, 0, 0, 0) + struct.pack("<2f", near, far) +
            struct.pack("<II", len(v), len(p)) + np.ascontiguousarray(v, np.float32).tobytes() +
            np.ascontiguousarray(p, np.uint32).tobytes())


def test_degenerate_offscreen_and_extreme_coordinates(d3, dev):
    """Zero-area, w = 0, behind-the-eye, far off-screen and non-finite vertices convert as on the host CPU."""
    rng = np.random.default_rng(11)
    n = 4000
    v = np.zeros((3 * n, 10), np.float32)
    v[:, 0:3] = rng.uniform(-3, 3, (3 * n, 3))
    v[:, 3] = rng.choice([1.0, 1.0, 0.5, 0.0, -1.0, 1e-30, 1e30], 3 * n)
    v[::17, 0] = 1e38
    v[::29, 1] = -1e38
    v[::53, 0] = np.inf
    v[::97, 1] = np.nan
    v[3:3 * 200:3] = v[4:3 * 200:3]  # repeated vertices: zero determinant
    v[:, 4:10] = rng.uniform(-2, 2, (3 * n, 6))
    p = np.arange(3 * n, dtype=np.uint32).reshape(n, 3)
    for w, h in ((128, 128), (1920, 1080)):
        got = d3.gpu_binning(dev, v, p, 0.1, 0.9, w, h, d3.TILEBUF_WIDE)
        same(got, d3.Trace(blob=synthetic_trace(v, p, 0.1, 0.9)).binning(0, w, h, fmt=d3.TILEBUF_WIDE), f"{w}x{h}")


def test_errors_and_empty_inputs(d3, dev):
    v = np.zeros((3, 10), np.float32)
    v[:, 3] = 1
    v[1, 0] = v[2, 1] = 0.5
    p = np.array([[0, 1, 2]], np.uint32)
    assert d3.gpu_binning(dev, v, np.zeros((0, 3), np.uint32), 0, 1, 64, 64)[0] == 0       # no primitives
    assert d3.gpu_binning(dev, v, np.array([[0, 0, 0]], np.uint32), 0, 1, 64, 64) == (0, b"", b"", 0)  # all dropped
    with pytest.raises(vx.VortexError):
        d3.gpu_binning(dev, v, np.array([[0, 1, 3]], np.uint32), 0, 1, 64, 64)              # index past the vertices
    with pytest.raises(vx.VortexError):
        d3.gpu_binning(dev, v, p, 0, 1, 0, 64)                                              # bad size
    with pytest.raises(vx.VortexError):
        d3.gpu_binning(dev, v, p, 0, 1, 64, 64, rank=2, num_ranks=2)                        # bad rank
    # the reference's 16-bit header fields: refused with -2 exactly where the host pre-pass refuses
    t = d3.Trace(blob=od.trace_blob("carnival"))
    with pytest.raises(vx.VortexError, match="-2"):
        t.gpu_binning(dev, 0, 8192, 8192)
    with pytest.raises(vx.VortexError, match="-2"):
        t.binning(0, 8192, 8192)
    assert dev.mem_info()[1] == 0  # nothing leaked by the error paths


# ---------------------------------------------------------------- the app flow with the pre-pass on the device
ALL_SW = 7


@pytest.mark.parametrize("name,size,path", od.goldens())
def test_reference_goldens_with_gpu_binning(d3, name, size, path):
    t = d3.Trace(blob=od.trace_blob(name))
    with d3.Draw3D(size, size, flags=ALL_SW | d3.GPU_BINNING) as app:
        fb = app.frame(t)
    gold = od.golden_framebuffer(path)
    assert np.array_equal(fb, gold), f"{(fb != gold).sum()} pixels differ from the reference golden"


@pytest.mark.parametrize("name,w,h", [("mouse", 2048, 2048), ("vase", 4096, 4096), ("tekkaman", 1333, 777)])
def test_large_frames_same_as_host_binned(d3, name, w, h):
    t = d3.Trace(blob=od.trace_blob(name))
    with d3.Draw3D(w, h, flags=ALL_SW | d3.WIDE_TILES) as app:
        want_c, want_z = app.frame(t), app.read(depth=True)
        st = app.stats()
        want = (st["prims"], st["tiles"], st["tile_prims"])
    with d3.Draw3D(w, h, flags=ALL_SW | d3.WIDE_TILES | d3.GPU_BINNING) as app:
        got_c, got_z = app.frame(t), app.read(depth=True)
        st = app.stats()
        assert (st["prims"], st["tiles"], st["tile_prims"]) == want
    assert np.array_equal(got_c, want_c) and np.array_equal(got_z, want_z)


def test_gpu_binning_rank_strips_and_prepared_path(d3):
    t = d3.Trace(blob=od.trace_blob("evilskull"))
    w = h = 512
    with d3.Draw3D(w, h, flags=ALL_SW) as app:
        full = app.frame(t)
    out = np.zeros_like(full)
    for r in range(2):
        with d3.Draw3D(w, h, flags=ALL_SW | d3.GPU_BINNING, rank=r, num_ranks=2) as app:
            app.clear()
            app.prepare(t)      # binned once on the device, buffers stay resident
            app.submit()
            app.wait()
            part = app.read()
        rows = [y for y in range(h) if (y // 32) % 2 == r]
        out[rows] = part[rows]
    assert np.array_equal(out, full)


def test_cli_with_gpu_binning(native, tmp_path):
    import os
    import subprocess
    from PIL import Image
    exe = os.path.join(os.path.dirname(native["vortex"]), "..", "bin", "draw3d")
    gold = np.load(os.path.join(od.GOLDEN_DIR, "box_ref_128.npy"))
    rgba = np.stack([(gold >> 16) & 255, (gold >> 8) & 255, gold & 255, gold >> 24], -1).astype(np.uint8)
    ref_png = str(tmp_path / "ref.png")
    Image.fromarray(rgba, "RGBA").save(ref_png)
    p = subprocess.run([exe, "-t", os.path.join(od.GOLDEN_DIR, "box.trc.gz"), "-w", "128", "-h", "128", "-x", "-y", "-u",
                        "-g", "-o", str(tmp_path / "out.png"), "-r", ref_png], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "PASSED!" in p.stdout, p.stdout + p.stderr
