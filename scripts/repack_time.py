"""First-frame cost of a scene (= the BVH2 -> Node4 repack) with the three repack paths: VX_DCR_RT_REPACK 0 = one
cooperative launch on the device, 2 = one set of launches per tree level (round 1), 1 = host.  Prints wall ms of the
first vx_start + wait minus a steady-state frame.   python scripts/repack_time.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skybox_rt_b200 import raytrace as rt  # noqa: E402
from skybox_rt_b200 import vortex as vx  # noqa: E402

for name, quads in (("100K", (250, 200)), ("1M", (1000, 500)), ("10M", (2500, 2000))):
    soup = rt.scene_terrain(*quads)
    nodes, tris = rt.build_bvh2(soup)
    for mode in (0, 2, 1):
        if mode == 1 and name == "10M":
            continue
        best = 1e9
        for rep in range(3):
            with rt.RayTracer(nodes, tris) as t:
                t.dev.dcr_write(vx.VX_DCR_RT_REPACK, mode)
                arg = rt.make_arg(64, 64, 0, scene="terrain")
                t0 = time.perf_counter()
                t.render(arg)
                first = (time.perf_counter() - t0) * 1e3
                t0 = time.perf_counter()
                t.render(arg)
                steady = (time.perf_counter() - t0) * 1e3
                best = min(best, first - steady)
        print(f"{name} triangles, repack mode {mode}: {best:.3f} ms")
