"""Host pre-pass vs device pre-pass (vx_cuda_draw3d_binning) on synthetic meshes of growing size.

A displaced UV grid of n quads (2n triangles) seen by an orthographic camera at 4096x4096, wide tile headers.
Host: srt_draw3d_binning (one core, as the reference's graphics::Binning runs).  Device: vertices / indices
resident, wall clock around the call (it waits for the output sizes) + the queued tail.  Prints one table;
run on the GPU box:  python profiles/binning_scaling.py
"""
import ctypes as C
import os
import struct
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skybox_rt_b200 import draw3d as d3  # noqa: E402
from skybox_rt_b200 import vortex as vx  # noqa: E402


def mesh(nx, ny, seed=3):
    rng = np.random.default_rng(seed)
    xs, ys = np.meshgrid(np.linspace(-0.95, 0.95, nx + 1), np.linspace(-0.95, 0.95, ny + 1), indexing="xy")
    v = np.zeros(((nx + 1) * (ny + 1), 10), np.float32)
    v[:, 0], v[:, 1] = xs.ravel(), ys.ravel()
    v[:, 2] = rng.uniform(-0.5, 0.5, len(v))
    v[:, 3] = 1.0
    v[:, 4:8] = rng.uniform(0, 1, (len(v), 4))
    v[:, 8:10] = v[:, 0:2] * 4
    i = (np.arange(ny)[:, None] * (nx + 1) + np.arange(nx)[None, :]).ravel().astype(np.uint32)
    p = np.concatenate([np.stack([i, i + 1, i + nx + 2], 1), np.stack([i, i + nx + 2, i + nx + 1], 1)])
    return v, np.ascontiguousarray(p, np.uint32)


def trace_blob(v, p):
    states = [1, 5, 0xF] + [0] * 21
    return (b"SKBXTRC1" + struct.pack("<II", 1, 0) + struct.pack("<24I", *states) + struct.pack("<4f", 0, 0, 0, 0) +
            struct.pack("<I", 0) + struct.pack("<4i", 0, 0, 0, 0) + struct.pack("<2f", 0.0, 1.0) +
            struct.pack("<II", len(v), len(p)) + v.tobytes() + p.tobytes())


def main():
    W = H = 4096
    print(f"{'triangles':>10} {'tiles':>7} {'pairs':>9} {'host ms':>9} {'device ms':>10} {'speed-up':>9}  identical")
    with vx.Device() as dev:
        L = vx.lib()
        sizes = ((50, 50), (160, 160), (500, 500), (1000, 1000), (2000, 1500))
        if len(sys.argv) == 3:  # one size only (for an ncu launch list)
            sizes = ((int(sys.argv[1]), int(sys.argv[2])),)
        for nx, ny in sizes:
            v, p = mesh(nx, ny)
            t = d3.Trace(blob=trace_blob(v, p))
            t0 = time.perf_counter()
            want = t.binning(0, W, H, fmt=d3.TILEBUF_WIDE)
            host_ms = (time.perf_counter() - t0) * 1e3
            vb, pb = dev.upload(v), dev.upload(p)
            best = 1e9
            for it in range(4):
                tb, prb, info = C.c_void_p(), C.c_void_p(), d3.BinningInfo()
                dev.ready_wait()
                t0 = time.perf_counter()
                rc = L.vx_cuda_draw3d_binning(dev.handle, vb.handle, len(v), pb.handle, len(p), 0.0, 1.0, W, H, 5,
                                              d3.TILEBUF_WIDE, 0, 1, C.byref(tb), C.byref(prb), C.byref(info))
                dev.ready_wait()
                ms = (time.perf_counter() - t0) * 1e3
                assert rc == 0
                if it:
                    best = min(best, ms)
                if it == 3:
                    tile = np.zeros(info.tilebuf_size, np.uint8)
                    prim = np.zeros(info.primbuf_size, np.uint8)
                    vx.Buffer(dev, tb, info.tilebuf_size).read(tile)
                    vx.Buffer(dev, prb, info.primbuf_size).read(prim)
                    same = tile.tobytes() == want[1] and prim.tobytes() == want[2]
                L.vx_mem_free(tb)
                L.vx_mem_free(prb)
            vb.free()
            pb.free()
            print(f"{len(p):>10} {want[0]:>7} {want[3]:>9} {host_ms:>9.2f} {best:>10.3f} {host_ms / best:>8.1f}x  {same}")


if __name__ == "__main__":
    main()
