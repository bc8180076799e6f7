"""Runs in a subprocess of tests/test_gpu_ref_stub.py with
    SKYBOX_RT_LIB=oracle/_ref/libvortex.so   the REFERENCE's own host stub (runtime/stub/vortex.cpp + utils.cpp)
    VORTEX_DRIVER=cuda                       so that its vx_dev_open dlopen()s "libvortex-cuda.so" ...
    LD_LIBRARY_PATH=skybox_rt_b200/lib       ... which is this repo's back-end under the reference's plug-in name.
Every vx_* call below therefore goes  reference stub -> callbacks_t -> CUDA back-end  (runtime/stub/vortex.cpp:58-182),
including the stub's own dcr_initialize at open, its upload helpers (stub/utils.cpp:49-155) and its vx_dump_perf at
close.  Prints one line per check; the parent asserts on them."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from skybox_rt_b200 import raytrace as rt  # noqa: E402
from skybox_rt_b200 import vortex as vx  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
LIBDIR = os.path.join(ROOT, "skybox_rt_b200", "lib")


def main():
    L = vx.lib()
    assert os.path.realpath(vx.LIB_PATH) == os.path.realpath(os.path.join(ROOT, "oracle", "_ref", "libvortex.so"))
    assert not hasattr(L, "vx_dev_init") and not hasattr(L, "vx_cuda_stream"), "this must be the reference stub"

    # ---- ray-tracing frame: the Python twin of apps/raytrace/main.c, kernel image from the .vxbin FILE ----------
    soup = rt.scene_cornell()
    nodes, tris = rt.build_bvh2(soup)
    with vx.Device() as dev:
        assert dev.dcr_read(0x005) == 0  # VX_DCR_BASE_MPM_CLASS written by the stub's dcr_initialize
        k = C.c_void_p()
        assert L.vx_upload_kernel_file(dev.handle, os.path.join(LIBDIR, "raytrace.vxbin").encode(), C.byref(k)) == 0
        tracer = rt.RayTracer(nodes, tris, device=dev)
        tracer.kernel.free()
        tracer.kernel = vx.Buffer(dev, k, 32)
        for variant in sorted(vx.RT_VARIANTS):
            dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
            fb = tracer.render(rt.make_arg(64, 64, 2, spp=1, seed=1234))
            gold = np.load(os.path.join(GOLDEN, "cornell_64_bounce.npy"))
            assert np.array_equal(fb, gold), f"variant {variant}: {(fb != gold).sum()} pixels differ"
        c = tracer.counters()
        assert c["rays_primary"] == 64 * 64 and c["elapsed_ns"] > 0
        tracer.close()
        print("STUB-OK raytrace: cornell 64x64 bounce == golden through the reference stub, all variants")

        # ---- raster draw: the raw call sequence of tests/regression/draw3d/main.cpp:198-357 ----------------------
        import _oracle_draw3d as od
        from skybox_rt_b200 import draw3d as d3
        t = d3.Trace(blob=od.trace_blob("triangle"))
        W = H = 64
        n, tb, pb, _ = t.binning(0, W, H)
        k = C.c_void_p()
        assert L.vx_upload_kernel_file(dev.handle, os.path.join(LIBDIR, "draw3d.vxbin").encode(), C.byref(k)) == 0
        kernel = vx.Buffer(dev, k, 32)
        cbuf, zbuf = dev.alloc(W * H * 4), dev.alloc(W * H * 4)
        cbuf.write(np.full(W * H, 0xFF000000, np.uint32))
        zbuf.write(np.full(W * H, 0xFFFFFFFF, np.uint32))
        tile, prim = dev.alloc(len(tb), vx.VX_MEM_READ), dev.alloc(len(pb), vx.VX_MEM_READ)
        tile.write(tb)
        prim.write(pb)
        arg = np.zeros(288 // 4, np.uint32)
        arg[268 // 4] = 6
        arg[272 // 4], arg[276 // 4] = prim.address & 0xFFFFFFFF, prim.address >> 32
        arg.view(np.uint8)[281] = 1
        RAS, OM = 28, 34
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
        assert np.array_equal(fb, gold), f"{(fb != gold).sum()} pixels differ"
        print("STUB-OK draw3d: triangle 64x64 == reference golden through the reference stub")
        mx = C.c_uint32()
        assert L.vx_check_occupancy(dev.handle, 1, C.byref(mx)) == 0
        sys.stdout.flush()
    # vx_dev_close has run the STUB's vx_dump_perf(stdout): its "PERF: instrs=..." line follows in the output
    print("STUB-OK closed")


if __name__ == "__main__":
    main()
