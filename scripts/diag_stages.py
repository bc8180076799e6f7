"""Diagnostic: cost of each ray population of the C3 path (primary / shadow / bounce / bounce shadow) when
traced as separate, stage-sorted streams by the explicit-ray kernels -- the upper bound a wavefront split
could reach.  Populations are rebuilt in numpy from GPU hit records (approximate arithmetic, timing only).
usage: python scripts/diag_stages.py [variant] [tune]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from skybox_rt_b200 import raytrace as rt, vortex as vx
from skybox_rt_b200.formats import RAY

variant = int(sys.argv[1]) if len(sys.argv) > 1 else 2
tune = int(sys.argv[2]) if len(sys.argv) > 2 else 0
soup = rt.scene_terrain(1000, 500)
nodes, tris = rt.build_bvh2(soup)
W, H = 1920, 1080
a = rt.make_arg(W, H, 0, scene="terrain")
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
light = np.array(list(a.light_pos), dtype=np.float64)
rec_of_prim = np.empty(len(tris), dtype=np.int64); rec_of_prim[tris["prim_id"]] = np.arange(len(tris))
rng = np.random.default_rng(1)


def surface(rays, hits):
    m = hits["prim_id"] != 0xFFFFFFFF
    r, h = rays[m], hits[m]
    t = tris[rec_of_prim[h["prim_id"]]]
    n = np.cross(t["e1"].astype(np.float64), t["e2"].astype(np.float64)); n /= np.linalg.norm(n, axis=1, keepdims=True)
    dd = r["d"].astype(np.float64)
    n[np.sum(n * dd, axis=1) > 0] *= -1
    po = r["o"] + dd * h["t"][:, None] + n * a.ray_eps
    return po, n


def shadow_rays(po, n):
    lv = light[None] - po; dist = np.linalg.norm(lv, axis=1); ld = lv / dist[:, None]
    m = np.sum(n * ld, axis=1) > 0
    s = np.zeros(int(m.sum()), dtype=RAY)
    s["o"] = po[m]; s["d"] = ld[m]; s["tmax"] = dist[m]
    return s


with rt.RayTracer(nodes, tris) as t:
    t.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    t.dev.dcr_write(vx.VX_DCR_RT_KEEP_ACTIVE, tune)

    def timed(name, rays):
        best = 1e30
        for _ in range(3):
            h = t.trace(rays)
            best = min(best, t.dev.mpm_query(vx.VX_CSR_RT_ELAPSED_NS))
        t.dev.dcr_write(vx.VX_DCR_RT_STATS, 1); t.trace(rays); c = t.counters(); t.dev.dcr_write(vx.VX_DCR_RT_STATS, 0)
        print(f"variant {variant} {name:14s}: {len(rays):8d} rays {best/1e6:.3f} ms {len(rays)/best*1e3:6.0f} Mrays/s "
              f"nodes/ray {c['nodes']/len(rays):.1f} tris/ray {c['tris']/len(rays):.1f}")
        return h, best

    tot = 0
    h1, ns = timed("primary", cam); tot += ns
    po, n = surface(cam, h1)
    sh1 = shadow_rays(po, n); _, ns = timed("shadow", sh1); tot += ns
    v = rng.normal(size=po.shape); v /= np.linalg.norm(v, axis=1, keepdims=True)
    bd = n + v; bd /= np.linalg.norm(bd, axis=1, keepdims=True)
    b = np.zeros(len(po), dtype=RAY); b["o"] = po; b["d"] = bd; b["tmax"] = 1e30
    h2, ns = timed("bounce", b); tot += ns
    po2, n2 = surface(b, h2)
    sh2 = shadow_rays(po2, n2); _, ns = timed("bounce shadow", sh2); tot += ns
    nrays = len(cam) + len(sh1) + len(b) + len(sh2)
    print(f"variant {variant} all four streams: {nrays} rays, {tot/1e6:.3f} ms, {nrays/tot*1e3:.0f} Mrays/s (closest-hit for all)")
