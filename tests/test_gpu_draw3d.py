"""GPU parity tests of the raster path: the sm_100a tile kernel, reached through the vx_* C ABI by the C
host flow of apps/draw3d, against the reference's golden images and the CPU oracle.  Bar: bit-exact
colour AND depth/stencil words (the reference's own compare tolerates 1/255, draw3d/main.cpp:507)."""
import os
import subprocess

import numpy as np
import pytest

import _oracle_draw3d as od

pytestmark = pytest.mark.gpu

GOLDENS = od.goldens()
NAMES = od.trace_names()
ALL_SW = 7  # -x -y -u


@pytest.fixture(scope="module")
def d3(native):
    from skybox_rt_b200 import draw3d
    from skybox_rt_b200 import vortex
    vortex.lib()  # raises if libvortex.so is missing: no fallback
    return draw3d


def checker():
    """The reference's own code compiled natively (oracle/_ref/libref_draw3d.so).  It is git-ignored, built where
    /root/reference exists and shipped to the GPU box as a prebuilt file; when it is absent the parity tests FAIL
    rather than quietly compare the product with this repo's own restatement."""
    ref = od.reference()
    if ref is None:
        pytest.fail("oracle/_ref/libref_draw3d.so is missing: build it with `make -C oracle` where /root/reference is "
                    "mounted (it travels to the GPU box with the snapshot); refusing to fall back to the restatement")
    return ref


@pytest.mark.parametrize("name,size,path", GOLDENS, ids=[f"{n}@{s}" for n, s, _ in GOLDENS])
def test_reference_goldens_bit_exact(d3, name, size, path):
    with d3.Draw3D(size, size, flags=ALL_SW) as app:
        fb = app.frame(d3.Trace(blob=od.trace_blob(name)))
    gold = od.golden_framebuffer(path)
    assert int((fb != gold).sum()) == 0


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("w,h", [(64, 64), (200, 120), (333, 517), (1000, 700)])
def test_colour_and_depth_match_oracle(d3, name, w, h):
    blob = od.trace_blob(name)
    with d3.Draw3D(w, h, flags=ALL_SW) as app:
        fb, zb = app.frame(d3.Trace(blob=blob), depth=True)
    fo, zo, _ = od.render(checker(), blob, w, h, threads=8)
    assert int((fb != fo).sum()) == 0, (name, w, h)
    assert int((zb != zo).sum()) == 0, (name, w, h)


@pytest.mark.parametrize("name", ["vase", "mouse", "tekkaman"])
def test_hardware_unit_path_equals_software_path(d3, name):
    # without -x -y -u the units read the device DCR file instead of the kernel_arg mirrors
    t = d3.Trace(blob=od.trace_blob(name))
    with d3.Draw3D(256, 192, flags=0) as app:
        hw = app.frame(t, depth=True)
    with d3.Draw3D(256, 192, flags=ALL_SW) as app:
        sw = app.frame(t, depth=True)
    assert np.array_equal(hw[0], sw[0]) and np.array_equal(hw[1], sw[1])


@pytest.mark.parametrize("name", ["vase", "polybump", "tekkaman"])
@pytest.mark.parametrize("w,h", [(1280, 1024), (300, 200)])  # above / below the device-clear threshold of 2^20 pixels
def test_auto_flow_and_async_read_back_equal_the_reference_flow(d3, name, w, h):
    """SRT_DRAW3D_AUTO (device clear, device binning for large draws, no wait between draws) and the asynchronous
    read-back change how the frame gets made and fetched, not one word of it: two app objects alternate frames the
    way the pipelined e2e loop of bench_draw3d.py does."""
    import torch
    from skybox_rt_b200 import vortex as vx
    t = d3.Trace(blob=od.trace_blob(name))
    with vx.Device() as dev:
        ref_app = d3.Draw3D(w, h, flags=ALL_SW | d3.WIDE_TILES, device=dev)
        ref, zref = ref_app.frame(t, depth=True)
        ref_app.close()
        apps = [d3.Draw3D(w, h, flags=ALL_SW | d3.WIDE_TILES | d3.AUTO, device=dev) for _ in range(2)]
        host = [torch.empty(h * w, dtype=torch.int32).pin_memory() for _ in range(2)]
        tickets = [None, None]
        for i in range(4):
            a = apps[i & 1]
            if tickets[i & 1] is not None:
                a.read_wait(tickets[i & 1])
                assert np.array_equal(host[i & 1].numpy().view(np.uint32).reshape(h, w), ref)
            a.clear()
            a.render(t)
            tickets[i & 1] = a.read_async(host[i & 1])
        for j in (0, 1):
            apps[j].read_wait(tickets[j])
            assert np.array_equal(host[j].numpy().view(np.uint32).reshape(h, w), ref)
        _, z = apps[1].read(depth=True)
        assert np.array_equal(z, zref)
        for a in apps:
            a.close()


@pytest.mark.parametrize("name", ["vase", "mouse", "scene", "tekkaman"])
def test_auto_flow_bins_on_worker_threads_without_changing_a_word(d3, name):
    """Under SRT_DRAW3D_AUTO the host pre-pass of a frame's draws runs on a thread pool beside the submission
    (apps/draw3d/draw3d_app.c: bin_pool_t).  Ten frames in a row, then draw ranges (render(start, end), which is how
    the per-draw measurements drive it), must equal the serial reference flow in colour and depth."""
    t = d3.Trace(blob=od.trace_blob(name))
    w, h = 640, 400
    n = t.num_drawcalls
    with d3.Draw3D(w, h, flags=ALL_SW) as app:
        ref, zref = app.frame(t, depth=True)
        app.clear()
        app.render(t, 0, n // 2)
        half, zhalf = app.read(depth=True)
    with d3.Draw3D(w, h, flags=ALL_SW | d3.AUTO) as app:
        for i in range(10):
            fb, zb = app.frame(t, depth=True)
            assert np.array_equal(fb, ref) and np.array_equal(zb, zref), i
        app.clear()
        app.render(t, 0, n // 2)
        fb, zb = app.read(depth=True)
        assert np.array_equal(fb, half) and np.array_equal(zb, zhalf)
        if n > 1:
            app.render(t, n // 2 + 1, n - 1)  # the rest of the frame on top
            fb, zb = app.read(depth=True)
            assert np.array_equal(fb, ref) and np.array_equal(zb, zref)


@pytest.mark.parametrize("name", ["evilskull", "polybump"])
def test_prepared_resubmission_equals_per_draw_render(d3, name):
    t = d3.Trace(blob=od.trace_blob(name))
    with d3.Draw3D(400, 300, flags=ALL_SW | d3.WIDE_TILES) as app:
        ref = app.frame(t, depth=True)
        app.prepare(t)
        for _ in range(2):  # the same prepared frame twice: clear, submit all draws back to back, wait
            app.clear()
            app.submit()
            app.wait()
            got = app.read(depth=True)
            assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])


@pytest.mark.parametrize("name,size", [("mouse", 2048), ("vase", 4096), ("carnival", 4096)])
def test_full_size_frames_match_oracle(d3, name, size):
    """BASELINE-sized frames (wide tile headers).  The oracle walks its own tile lists, so it stays a
    valid checker where the reference's 16-bit tile header format ends."""
    blob = od.trace_blob(name)
    with d3.Draw3D(size, size, flags=ALL_SW | d3.WIDE_TILES) as app:
        fb, zb = app.frame(d3.Trace(blob=blob), depth=True)
    fo, zo, _ = od.render(od.oracle(), blob, size, size, threads=os.cpu_count() or 8)
    assert int((fb != fo).sum()) == 0 and int((zb != zo).sum()) == 0


def test_rank_strips_compose_the_single_gpu_frame(d3):
    blob = od.trace_blob("evilskull")
    t = d3.Trace(blob=blob)
    w, h, n = 320, 272, 3
    with d3.Draw3D(w, h, flags=ALL_SW) as app:
        full, zfull = app.frame(t, depth=True)
    out = np.full((h, w), 0xFF000000, np.uint32)
    for r in range(n):
        with d3.Draw3D(w, h, flags=ALL_SW, rank=r, num_ranks=n) as app:
            part = app.frame(t)
        rows = [y for y in range(h) if (y // 32) % n == r]
        out[rows] = part[rows]
        other = [y for y in range(h) if (y // 32) % n != r]
        assert (part[other] == 0xFF000000).all()  # a rank never touches another rank's tile rows
    assert np.array_equal(out, full)


def test_fragment_counters(d3):
    from skybox_rt_b200 import vortex as vx
    t = d3.Trace(blob=od.trace_blob("triangle"))
    with d3.Draw3D(128, 128, flags=ALL_SW) as app:
        app.dev.dcr_write(vx.VX_DCR_RT_STATS, 1)
        fb = app.frame(t)
        frags = app.dev.mpm_query(d3.VX_CSR_RASTER_FRAGMENTS)
        tiles = app.dev.mpm_query(d3.VX_CSR_RASTER_TILES)
    # one opaque triangle, no depth test: every shaded fragment is a written pixel
    assert frags == int((fb != 0xFF000000).sum()) and tiles == t.binning(0, 128, 128)[0]


def test_bad_state_is_refused_not_rendered(d3):
    from skybox_rt_b200 import vortex as vx
    t = d3.Trace(blob=od.trace_blob("triangle"))
    with d3.Draw3D(64, 64, flags=0) as app:
        app.prepare(t)
        app.submit()
        app.wait()
        # an arguments buffer that is too small for kernel_arg_t must fail vx_start (callbacks.inc conventions)
        small = app.dev.upload(np.zeros(16, np.uint8))
        img = np.zeros(4, np.uint64)
        img[0], img[1], img[2], img[3] = 0x80000000, 0x80000040, 0x5452584F42594B53, 2 | (1 << 32)
        k = app.dev.upload_kernel(img.tobytes())
        with pytest.raises(vx.VortexError):
            app.dev.start(k, small)


def test_draw3d_app_compares_against_reference_golden(native, tmp_path):
    """The C program end to end: trace file in, PNG out, exit code = mismatching pixels (main.cpp:505-514)."""
    from PIL import Image
    exe = os.path.join(os.path.dirname(native["vortex"]), "..", "bin", "draw3d")
    gold = np.load(os.path.join(od.GOLDEN_DIR, "vase_ref_128.npy"))
    rgba = np.stack([(gold >> 16) & 255, (gold >> 8) & 255, gold & 255, gold >> 24], -1).astype(np.uint8)
    ref_png, out_png = str(tmp_path / "vase_ref_128.png"), str(tmp_path / "out.png")
    Image.fromarray(rgba, "RGBA").save(ref_png)
    p = subprocess.run([exe, "-t", os.path.join(od.GOLDEN_DIR, "vase.trc.gz"), "-w", "128", "-h", "128", "-x", "-y",
                        "-u", "-o", out_png, "-r", ref_png], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "PASSED!" in p.stdout
    got = np.asarray(Image.open(out_png).convert("RGBA"))
    assert np.array_equal(got, rgba)


VARIANTS = od.state_variants()


@pytest.mark.parametrize("label,kw", VARIANTS, ids=[v[0] for v in VARIANTS])
def test_state_space_matches_oracle(d3, label, kw):
    """Every depth function, blend factor, stencil function/op, texture format x filter x address mode,
    colour write mask and texture-environment switch the app can program (generic kernel paths)."""
    for name, w, h in (("vase", 160, 128), ("evilskull", 100, 100)):
        blob = od.mutate(od.trace_blob(name), **kw)
        with d3.Draw3D(w, h, flags=ALL_SW) as app:
            fb, zb = app.frame(d3.Trace(blob=blob), depth=True)
        fo, zo, _ = od.render(checker(), blob, w, h, threads=4)
        assert int((fb != fo).sum()) == 0 and int((zb != zo).sum()) == 0, (label, name)


OM_VARIANTS = od.om_override_variants()


@pytest.mark.parametrize("label,ov", OM_VARIANTS, ids=[v[0] for v in OM_VARIANTS])
def test_blend_equations_logic_ops_and_constant_factors(d3, label, ov):
    """Row R9's other half: SUB / REV_SUB / MIN / MAX / LOGICOP equations, the 16 logic ops, the CONST_* factors and
    the blend constant (sim/common/graphics.cpp:366-404, :455-472, :477-524, :603-636).  The reference app never
    programs them (main.cpp:271-279), so product and compiled reference both take the four OM DCR values as an
    override for every blending draw; blending is forced on in every draw of the trace."""
    for name, w, h in (("vase", 160, 128), ("mouse", 130, 100)):
        blob = od.mutate(od.trace_blob(name), blend_enabled=1, blend_src=4, blend_dst=5)
        for flags in (ALL_SW, 0):  # software-unit path (kernel_arg_t mirrors) and hardware-unit path (device DCR file)
            with d3.Draw3D(w, h, flags=flags) as app:
                app.set_om_override(ov)
                fb, zb = app.frame(d3.Trace(blob=blob), depth=True)
            fo, zo, _ = od.render(checker(), blob, w, h, threads=4, om_override=ov)
            assert int((fb != fo).sum()) == 0 and int((zb != zo).sum()) == 0, (label, name, flags)


def test_specialised_kernels_equal_the_generic_kernel(d3):
    t = d3.Trace(blob=od.trace_blob("mouse"))
    frames = []
    for generic in (0, 1):
        with d3.Draw3D(512, 512, flags=ALL_SW) as app:
            app.dev.dcr_write(0x805, generic)  # VX_DCR_RASTER_EXT_GENERIC
            frames.append(app.frame(t, depth=True))
    assert np.array_equal(frames[0][0], frames[1][0]) and np.array_equal(frames[0][1], frames[1][1])


def test_kernel_vxbin_file_runs_the_reference_flow(d3, native):
    """An unmodified reference app calls vx_upload_kernel_file("kernel.vxbin"); lib/draw3d.vxbin is that file.
    Drive one draw by hand through the raw vx_* calls, exactly as tests/regression/draw3d/main.cpp does."""
    import ctypes as C

    from skybox_rt_b200 import vortex as vx
    L = vx.lib()
    blob = od.trace_blob("triangle")
    t = d3.Trace(blob=blob)
    W = H = 64
    n, tb, pb, _ = t.binning(0, W, H)
    with vx.Device() as dev:
        k = C.c_void_p()
        path = os.path.join(os.path.dirname(native["vortex"]), "draw3d.vxbin").encode()
        assert L.vx_upload_kernel_file(dev.handle, path, C.byref(k)) == 0
        kernel = vx.Buffer(dev, k, 32)
        cbuf, zbuf = dev.alloc(W * H * 4), dev.alloc(W * H * 4)
        cbuf.write(np.full(W * H, 0xFF000000, np.uint32))
        zbuf.write(np.full(W * H, 0xFFFFFFFF, np.uint32))
        tile, prim = dev.alloc(len(tb), vx.VX_MEM_READ), dev.alloc(len(pb), vx.VX_MEM_READ)
        tile.write(tb)
        prim.write(pb)
        arg = np.zeros(288 // 4, np.uint32)  # kernel_arg_t of draw3d/common.h, mirrors left at zero: hardware-unit path
        arg[268 // 4] = 6                     # log_num_tasks
        arg[272 // 4], arg[276 // 4] = prim.address & 0xFFFFFFFF, prim.address >> 32
        arg8 = arg.view(np.uint8)
        arg8[281] = 1                         # color_enabled
        RAS, OM = 28, 34                      # VX_DCR_RASTER_STATE_BEGIN, VX_DCR_OM_STATE_BEGIN
        for a, v in [(RAS + 0, (tile.address // 64) & 0xFFFFFFFF), (RAS + 1, n), (RAS + 2, (prim.address // 64) & 0xFFFFFFFF),
                     (RAS + 3, 120), (RAS + 4, W << 16), (RAS + 5, H << 16), (OM + 0, (cbuf.address // 64) & 0xFFFFFFFF),
                     (OM + 1, W * 4), (OM + 2, 0xFFFFFFFF), (OM + 5, 0), (OM + 6, 0), (OM + 7, 0), (OM + 8, 0), (OM + 10, 0),
                     (OM + 11, 0), (OM + 12, 0xFF), (OM + 13, 0), (OM + 14, 0), (OM + 15, (1 << 8) | 1)]:
            dev.dcr_write(a, v)
        args = dev.upload(arg)
        dev.start(kernel, args)
        dev.ready_wait()
        fb = cbuf.read(np.empty((H, W), np.uint32))
    gold = od.golden_framebuffer(os.path.join(od.GOLDEN_DIR, "triangle_ref_64.npy"))
    assert np.array_equal(fb, gold)


@pytest.mark.parametrize("name", ["mouse", "evilskull", "scene"])
def test_pixel_per_thread_kernel_equals_quad_per_thread_kernel(d3, name):
    t = d3.Trace(blob=od.trace_blob(name))
    frames = []
    for ppt in (4, 1):
        with d3.Draw3D(700, 500, flags=ALL_SW) as app:
            app.dev.dcr_write(0x806, ppt)  # VX_DCR_RASTER_EXT_PPT
            frames.append(app.frame(t, depth=True))
    assert np.array_equal(frames[0][0], frames[1][0]) and np.array_equal(frames[0][1], frames[1][1])


def test_device_fill_and_device_clear(d3):
    """vx_cuda_fill32 (any 4-byte aligned range) and the app's device-side clear give the host-upload clear's frame."""
    from skybox_rt_b200 import vortex as vx
    with vx.Device() as dev:
        n = 1000
        buf = dev.alloc(n * 4)
        base = np.arange(n, dtype=np.uint32)
        for off, cnt, val in [(0, n, 0xff000000), (1, 1, 7), (3, 10, 0xdeadbeef), (5, 123, 1), (997, 3, 2), (4, 0, 9)]:
            buf.write(base)
            buf.fill32(val, off * 4, cnt * 4)
            want = base.copy()
            want[off:off + cnt] = val
            assert np.array_equal(buf.read(np.empty(n, np.uint32)), want), (off, cnt)
        for bad in [(2, 4), (0, 6), (0, n * 4 + 4)]:
            with pytest.raises(vx.VortexError):
                buf.fill32(0, *bad)
        buf.free()
    t = d3.Trace(blob=od.trace_blob("vase"))
    with d3.Draw3D(200, 120, flags=ALL_SW) as app:
        want_c, want_z = app.frame(t), app.read(depth=True)
    with d3.Draw3D(200, 120, flags=ALL_SW | d3.DEVICE_CLEAR) as app:
        for _ in range(2):  # the second frame starts from a dirty target
            got_c, got_z = app.frame(t), app.read(depth=True)
    assert np.array_equal(got_c, want_c) and np.array_equal(got_z, want_z)


class RawDraw:
    """One draw of the triangle trace driven through the raw vx_* calls (the flow of the test above), with the tile and
    primitive buffers supplied by the caller -- for the checks vx_start and the tile kernel make on them."""

    def __init__(self, d3, native, size=64):
        import ctypes as C

        from skybox_rt_b200 import vortex as vx
        self.vx, self.size = vx, size
        self.trace = d3.Trace(blob=od.trace_blob("triangle"))
        self.dev = vx.Device()
        k = C.c_void_p()
        path = os.path.join(os.path.dirname(native["vortex"]), "draw3d.vxbin").encode()
        assert vx.lib().vx_upload_kernel_file(self.dev.handle, path, C.byref(k)) == 0
        self.kernel = vx.Buffer(self.dev, k, 32)
        self.cbuf = self.dev.alloc(size * size * 4)

    def draw(self, n, tile, prim):
        """tile / prim: device buffers.  Raises VortexError where vx_start or vx_ready_wait report a failure."""
        vx, dev, W = self.vx, self.dev, self.size
        self.cbuf.write(np.full(W * W, 0xFF000000, np.uint32))
        arg = np.zeros(288 // 4, np.uint32)
        arg[268 // 4] = 6
        arg[272 // 4], arg[276 // 4] = prim.address & 0xFFFFFFFF, prim.address >> 32
        arg.view(np.uint8)[281] = 1
        RAS, OM = 28, 34
        for a, v in [(RAS + 0, (tile.address // 64) & 0xFFFFFFFF), (RAS + 1, n), (RAS + 2, (prim.address // 64) & 0xFFFFFFFF),
                     (RAS + 3, 120), (RAS + 4, W << 16), (RAS + 5, W << 16), (OM + 0, (self.cbuf.address // 64) & 0xFFFFFFFF),
                     (OM + 1, W * 4), (OM + 2, 0xFFFFFFFF), (OM + 5, 0), (OM + 6, 0), (OM + 7, 0), (OM + 8, 0), (OM + 10, 0),
                     (OM + 11, 0), (OM + 12, 0xFF), (OM + 13, 0), (OM + 14, 0), (OM + 15, (1 << 8) | 1)]:
            dev.dcr_write(a, v)
        args = dev.upload(arg)
        try:
            dev.start(self.kernel, args)
            dev.ready_wait()
            return self.cbuf.read(np.empty((W, W), np.uint32))
        finally:
            args.free()

    def upload(self, data):
        b = self.dev.alloc(max(len(data), 4), self.vx.VX_MEM_READ)
        b.write(np.frombuffer(bytes(data), np.uint8))
        return b

    def close(self):
        self.dev.close()


def test_tile_buffers_of_another_tile_size_are_refused(d3, native):
    """The kernel is built for 32 x 32 tiles and the buffer formats do not say what they were binned with: a buffer that
    vx_cuda_draw3d_binning made at another size, or one the app declares in VX_DCR_RASTER_EXT_TILE_LOGSIZE, fails
    vx_start instead of rendering garbage."""
    import ctypes as C

    from skybox_rt_b200 import vortex as vx
    r = RawDraw(d3, native)
    try:
        gold = od.golden_framebuffer(os.path.join(od.GOLDEN_DIR, "triangle_ref_64.npy"))
        n5, tb5, pb5, _ = r.trace.binning(0, 64, 64)
        n4, tb4, pb4, _ = r.trace.binning(0, 64, 64, tile_logsize=4)
        t5, p5, t4, p4 = r.upload(tb5), r.upload(pb5), r.upload(tb4), r.upload(pb4)
        assert np.array_equal(r.draw(n5, t5, p5), gold)            # undeclared: accepted as before
        r.dev.dcr_write(d3.VX_DCR_RASTER_EXT_TILE_LOGSIZE, 4)       # host-binned at 16 x 16, declared
        with pytest.raises(vx.VortexError):
            r.draw(n4, t4, p4)
        r.dev.dcr_write(d3.VX_DCR_RASTER_EXT_TILE_LOGSIZE, 5)
        assert np.array_equal(r.draw(n5, t5, p5), gold)
        r.dev.dcr_write(d3.VX_DCR_RASTER_EXT_TILE_LOGSIZE, 0)
        # device-binned: the buffer itself remembers its size
        verts, prims, near, far = r.trace.draw_arrays(0)
        vb, pbuf = r.dev.upload(np.ascontiguousarray(verts, np.float32)), r.dev.upload(np.ascontiguousarray(prims, np.uint32))
        for logsize, ok in ((4, False), (5, True)):
            tb, prb, info = C.c_void_p(), C.c_void_p(), d3.BinningInfo()
            assert vx.lib().vx_cuda_draw3d_binning(r.dev.handle, vb.handle, len(verts), pbuf.handle, len(prims), near, far, 64, 64,
                                                   logsize, d3.TILEBUF_REFERENCE, 0, 1, C.byref(tb), C.byref(prb), C.byref(info)) == 0
            tile, prim = vx.Buffer(r.dev, tb, info.tilebuf_size), vx.Buffer(r.dev, prb, info.primbuf_size)
            if ok:
                assert np.array_equal(r.draw(info.num_tiles, tile, prim), gold)
            else:
                with pytest.raises(vx.VortexError):
                    r.draw(info.num_tiles, tile, prim)
    finally:
        r.close()


def test_malformed_tile_and_primitive_buffers_are_reported_not_faulted(d3, native, capfd):
    """A primitive id past the primitive buffer, a primitive list that leaves the tile buffer and a primitive buffer
    without one whole record end as -1 from vx_ready_wait / vx_start with a message -- not as a sticky CUDA fault:
    the next well-formed draw on the same device gives the golden frame."""
    from skybox_rt_b200 import vortex as vx
    r = RawDraw(d3, native)
    try:
        gold = od.golden_framebuffer(os.path.join(od.GOLDEN_DIR, "triangle_ref_64.npy"))
        n, tb, pb, _ = r.trace.binning(0, 64, 64)
        good_t, good_p = r.upload(tb), r.upload(pb)
        assert np.array_equal(r.draw(n, good_t, good_p), gold)
        words = np.frombuffer(tb, np.uint32).copy()
        bad_pid = words.copy()
        bad_pid[2 * n] = 0x00FFFFFF                                   # first id of the first list (rast_tile_header_t = 8 B)
        with pytest.raises(vx.VortexError):
            r.draw(n, r.upload(bad_pid.tobytes()), good_p)
        assert "primitive id lies outside" in capfd.readouterr().err
        assert np.array_equal(r.draw(n, good_t, good_p), gold)        # the device is alive and the flag was cleared
        bad_list = words.copy()
        bad_list[2 * (n - 1) + 1] |= 0xFFFF0000                        # pids_count of the last tile = 65535
        with pytest.raises(vx.VortexError):
            r.draw(n, r.upload(bad_list.tobytes()), good_p)
        assert "primitive list lies outside" in capfd.readouterr().err
        assert np.array_equal(r.draw(n, good_t, good_p), gold)
        with pytest.raises(vx.VortexError):                            # 64 bytes hold no 120-byte record: refused at vx_start
            r.draw(n, good_t, r.upload(pb[:64]))
        assert np.array_equal(r.draw(n, good_t, good_p), gold)
    finally:
        r.close()
