"""Device BVH build time (vx_cuda_rt_build_lbvh) against the host builder, soup already on the device.
Run on the GPU box:  python profiles/lbvh_build_time.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skybox_rt_b200 import raytrace as rt  # noqa: E402
from skybox_rt_b200 import vortex as vx  # noqa: E402


def main():
    print(f"{'triangles':>10} {'host SAH s':>11} {'device ms':>10} {'Mtri/s':>8}")
    with vx.Device() as dev:
        L = vx.lib()
        for qx, qy in ((250, 200), (1000, 500), (2500, 2000)):
            soup = rt.scene_terrain(qx, qy)
            t0 = time.perf_counter()
            if len(soup) <= 1000000:
                rt.build_bvh2(soup)
                host_s = time.perf_counter() - t0
            else:
                host_s = float("nan")  # minutes on one core; not timed here
            sb = dev.upload(soup)
            best = 1e9
            for it in range(4):
                nb, tb, info = C.c_void_p(), C.c_void_p(), (C.c_uint32 * 2)()
                dev.ready_wait()
                t0 = time.perf_counter()
                assert L.vx_cuda_rt_build_lbvh(dev.handle, sb.handle, len(soup), C.byref(nb), C.byref(tb), info) == 0
                dev.ready_wait()
                ms = (time.perf_counter() - t0) * 1e3
                if it:
                    best = min(best, ms)
                L.vx_mem_free(nb)
                L.vx_mem_free(tb)
            sb.free()
            print(f"{len(soup):>10} {host_s:>11.2f} {best:>10.3f} {len(soup) / best / 1e3:>8.0f}")


def first_frame():
    """Scene change to first pixels: build + BVH2 -> Node4 repack + one 64x64 frame, wall clock."""
    print()
    print(f"{'triangles':>10} {'builder':>8} {'repack':>7} {'build ms':>9} {'first frame ms (incl. repack)':>30}")
    for qx, qy in ((1000, 500), (2500, 2000)):
        soup = rt.scene_terrain(qx, qy)
        for builder in ("host", "device"):
            if builder == "host" and len(soup) > 1000000:
                continue
            for repack in (1, 0):
                for it in range(2):  # second pass: pools and modules warm
                    t0 = time.perf_counter()
                    if builder == "host":
                        nodes, tris = rt.build_bvh2(soup)
                        t = rt.RayTracer(nodes, tris)
                    else:
                        t = rt.RayTracer(soup=soup)
                    t.dev.ready_wait()
                    build_ms = (time.perf_counter() - t0) * 1e3
                    t.dev.dcr_write(vx.VX_DCR_RT_REPACK, repack)
                    t0 = time.perf_counter()
                    t.render(rt.make_arg(64, 64, 0, scene="terrain"))
                    frame_ms = (time.perf_counter() - t0) * 1e3
                    t.close()
                print(f"{len(soup):>10} {builder:>8} {'host' if repack else 'device':>7} {build_ms:>9.1f} {frame_ms:>30.2f}")


if __name__ == "__main__":
    main()
    first_frame()
