"""Diagnostic: throughput of the explicit-ray kernels on coherent (camera) and incoherent (random)
ray streams over the C3 scene.  usage: python scripts/diag_rays.py [variant] [tune]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from skybox_rt_b200 import raytrace as rt, vortex as vx
from skybox_rt_b200.formats import RAY

variant = int(sys.argv[1]) if len(sys.argv) > 1 else 0
tune = int(sys.argv[2]) if len(sys.argv) > 2 else 0
soup = rt.scene_terrain(1000, 500)
nodes, tris = rt.build_bvh2(soup)
W, H = 1920, 1080
a = rt.make_arg(W, H, 0, scene="terrain")
# camera rays in 8x4 tile order
tx, ty = (W + 7) // 8, (H + 3) // 4
j = np.arange(tx * ty * 32, dtype=np.int64)
tile, l = j >> 5, j & 31
x = (tile % tx) * 8 + (l & 7); y = (tile // tx) * 4 + (l >> 3)
ok = (x < W) & (y < H); x, y = x[ok], y[ok]
fx = ((x + 0.5) / W * 2 - 1) * a.tan_half_fov * W / H
fy = ((y + 0.5) / H * 2 - 1) * a.tan_half_fov
fwd, right, up = (np.array(list(v), dtype=np.float64) for v in (a.cam_fwd, a.cam_right, a.cam_up))
d = fwd[None] + fx[:, None] * right[None] + fy[:, None] * up[None]
d /= np.linalg.norm(d, axis=1, keepdims=True)
cam = np.zeros(len(x), dtype=RAY)
cam["o"] = np.array(list(a.cam_pos), dtype=np.float32); cam["d"] = d.astype(np.float32); cam["tmax"] = 1e30
lo = np.minimum.reduce([soup["v0"].min(0), soup["v1"].min(0), soup["v2"].min(0)])
hi = np.maximum.reduce([soup["v0"].max(0), soup["v1"].max(0), soup["v2"].max(0)])
rnd = rt.rays_random(lo, hi, 7, 0, 4 << 20)
with rt.RayTracer(nodes, tris) as t:
    t.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    t.dev.dcr_write(vx.VX_DCR_RT_KEEP_ACTIVE, tune)
    for name, rays in (("camera", cam), ("random", rnd)):
        best = 1e30
        for _ in range(4):
            h = t.trace(rays)
            best = min(best, t.dev.mpm_query(vx.VX_CSR_RT_ELAPSED_NS))
        t.dev.dcr_write(vx.VX_DCR_RT_STATS, 1); t.trace(rays); c = t.counters(); t.dev.dcr_write(vx.VX_DCR_RT_STATS, 0)
        print(f"variant {variant} tune {tune} {name}: {len(rays)} rays, {best/1e6:.3f} ms, {len(rays)/best*1e3:.0f} Mrays/s, "
              f"hit {np.mean(h['prim_id'] != 0xFFFFFFFF):.2f}, nodes/ray {c['nodes']/len(rays):.1f}, tris/ray {c['tris']/len(rays):.1f}")
