"""Render a few frames (or trace a ray list) of one BASELINE config with one kernel variant -- the short command
that ncu captures are taken from (profiles/).  No torch, no oracle.

    python scripts/frame_once.py --workload c3 --variant 4 --frames 3 [--tune N] [--builder device]
Prints the device time of every frame (VX_CSR_RT_ELAPSED_NS) and the ray counts of the last one.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from skybox_rt_b200 import raytrace as rt  # noqa: E402
from skybox_rt_b200 import vortex as vx  # noqa: E402

CONFIGS = {
    "c2": dict(quads=(250, 200), w=1024, h=1024, mode=0, spp=1),
    "c3": dict(quads=(1000, 500), w=1920, h=1080, mode=2, spp=1),
    "c4": dict(quads=(2500, 2000), w=3840, h=2160, mode=2, spp=8),
    "c5": dict(quads=(2500, 2000), rays=1 << 24),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--tune", type=int, default=0)
    ap.add_argument("--stats", type=int, default=0)
    ap.add_argument("--chains", type=int, default=0, help="VX_DCR_RT_CHAINS (0 = default)")
    ap.add_argument("--rotate", type=int, default=0, help="VX_DCR_RT_LBVH_ROTATE for --builder device (0 default, 1 none, k = k-1 passes)")
    ap.add_argument("--unfused", type=int, default=0, help="VX_DCR_RT_FUSE_BOUNCE (1 = separate K4 / K5 launches)")
    ap.add_argument("--sort", type=int, default=0, help="c5 experiment: order the rays on the host by (origin cell of a "
                    "G^3 grid, direction octant) before the upload, G = this value (0 = as generated)")
    ap.add_argument("--builder", default="host", choices=["host", "device"])
    ap.add_argument("--rays", type=int, default=None)
    ap.add_argument("--ranks", type=int, default=1, help="render rank 0's strip of a frame split over this many ranks")
    args = ap.parse_args()
    cfg = CONFIGS[args.workload]
    soup = rt.scene_terrain(*cfg["quads"])
    if args.builder == "device":
        dev0 = vx.Device()
        dev0.dcr_write(vx.VX_DCR_RT_LBVH_ROTATE, args.rotate)
        tracer = rt.RayTracer(soup=soup, device=dev0)
    else:
        tracer = rt.RayTracer(*rt.build_bvh2(soup))
    dev = tracer.dev
    dev.dcr_write(vx.VX_DCR_RT_VARIANT, args.variant)
    dev.dcr_write(vx.VX_DCR_RT_KEEP_ACTIVE, args.tune)
    dev.dcr_write(vx.VX_DCR_RT_STATS, args.stats)
    dev.dcr_write(vx.VX_DCR_RT_CHAINS, args.chains)
    dev.dcr_write(vx.VX_DCR_RT_FUSE_BOUNCE, args.unfused)
    times = []
    if "rays" in cfg:
        n = args.rays or cfg["rays"]
        lo = np.minimum.reduce([soup["v0"].min(0), soup["v1"].min(0), soup["v2"].min(0)]).astype(np.float32)
        hi = np.maximum.reduce([soup["v0"].max(0), soup["v1"].max(0), soup["v2"].max(0)]).astype(np.float32)
        rays = rt.rays_random(lo, hi, 0x5EED5EED, 0, n)
        if args.sort:
            g = args.sort
            cell = np.clip(((rays["o"] - lo) / np.maximum(hi - lo, 1e-20) * g).astype(np.int64), 0, g - 1)
            octant = (rays["d"][:, 0] < 0) * 1 + (rays["d"][:, 1] < 0) * 2 + (rays["d"][:, 2] < 0) * 4
            key = ((cell[:, 0] * g + cell[:, 1]) * g + cell[:, 2]) * 8 + octant
            rays = rays[np.argsort(key, kind="stable")]
        for _ in range(args.frames):
            tracer.trace(rays)
            times.append(tracer.counters()["elapsed_ns"] * 1e-6)
    else:
        for i in range(args.frames):
            arg = rt.make_arg(cfg["w"], cfg["h"], cfg["mode"], spp=cfg["spp"], seed=1234, scene="terrain",
                              rank=0, num_ranks=args.ranks, rows_per_block=8)
            tracer.start(arg)
            dev.ready_wait()
            times.append(tracer.counters()["elapsed_ns"] * 1e-6)
    c = tracer.counters()
    rays = c["rays_primary"] + c["rays_shadow"] + c["rays_bounce"]
    print(f"{args.workload} variant {args.variant}: frames ms {[round(t, 3) for t in times]}  rays {rays}  "
          f"-> {rays / (min(times) * 1e-3) / 1e6:.0f} Mrays/s (L2 warm)  nodes {c['nodes']} tris {c['tris']}")
    if args.stats:
        wn, wt = dev.mpm_query(vx.VX_CSR_RT_WARP_NODE_STEPS), dev.mpm_query(vx.VX_CSR_RT_WARP_TRI_STEPS)
        print(f"  per ray: {c['nodes'] / rays:.2f} node steps, {c['tris'] / rays:.2f} triangle tests; lanes per warp step: "
              f"node {c['nodes'] / max(wn, 1):.1f}, triangle {c['tris'] / max(wt, 1):.1f} ({wn} + {wt} warp steps)")
    tracer.close()


if __name__ == "__main__":
    main()
