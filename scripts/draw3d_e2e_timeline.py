"""Host-side timeline of the pipelined draw3d e2e loop (bench_draw3d.py's `e2e`): per frame, how long the host spends
waiting for the previous read-back, clearing, rendering (bin + upload + start) and queueing the read-back, plus the
bare device-to-host copy time of one frame for comparison.  Diagnostic, not a bench value.
    python scripts/draw3d_e2e_timeline.py [--frames 24] [--size 4096] [--trace mouse]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=24)
    ap.add_argument("--size", type=int, default=4096)
    ap.add_argument("--trace", default="mouse")
    ap.add_argument("--apps", type=int, default=2, help="app objects (render targets) alternated")
    args = ap.parse_args()
    import numpy as np
    import torch

    import _oracle_draw3d as od
    from skybox_rt_b200 import draw3d as d3
    from skybox_rt_b200 import vortex as vx
    W = H = args.size
    trace = d3.Trace(blob=od.trace_blob(args.trace))
    dev = vx.Device()
    n = args.apps
    apps = [d3.Draw3D(W, H, flags=7 | d3.WIDE_TILES | d3.AUTO, device=dev) for _ in range(n)]
    host = [torch.empty(H * W, dtype=torch.int32).pin_memory() for _ in range(n)]
    # bare copy: one frame device -> pinned host, blocking
    apps[0].clear(); apps[0].render(trace); dev.ready_wait()
    for _ in range(2):
        t0 = time.perf_counter(); apps[0].read(out=host[0]); t1 = time.perf_counter()
    print(f"blocking read of {W * H * 4 >> 20} MiB: {(t1 - t0) * 1e3:.3f} ms = {W * H * 4 / (t1 - t0) / 1e9:.1f} GB/s")
    t0 = time.perf_counter(); tk = apps[0].read_async(host[0]); apps[0].read_wait(tk); t1 = time.perf_counter()
    print(f"async read + wait: {(t1 - t0) * 1e3:.3f} ms")
    # does a read-back overlap the next frame's rendering?  A: four reads back to back; C: one frame rendered, no read;
    # B: a read of app 0 queued, then a frame rendered into app 1 -- overlap gives max(A / 4, C), none gives the sum
    def timed(fn, reps=5):
        best = 1e9
        for _ in range(reps):
            dev.ready_wait()
            t0 = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t0)
        return best * 1e3

    def exp_a():
        tk = [apps[0].read_async(host[0]) for _ in range(4)]
        for t in tk:
            apps[0].read_wait(t)

    def exp_c():
        apps[1].clear(); apps[1].render(trace); dev.ready_wait()

    def exp_b():
        tk = apps[0].read_async(host[0])
        apps[1].clear(); apps[1].render(trace); dev.ready_wait()
        apps[0].read_wait(tk)

    def exp_d():  # the same with the frame first and the read of the OTHER target second
        apps[1].clear(); apps[1].render(trace)
        tk = apps[0].read_async(host[0])
        dev.ready_wait(); apps[0].read_wait(tk)
    def exp_e():  # B with every upload of the frame queued before its first kernel (prepare + submit)
        tk = apps[0].read_async(host[0])
        apps[1].clear(); apps[1].prepare(trace); apps[1].submit(); dev.ready_wait()
        apps[0].read_wait(tk)

    def exp_f():  # C likewise
        apps[1].clear(); apps[1].prepare(trace); apps[1].submit(); dev.ready_wait()
    apps[1].clear(); apps[1].prepare(trace); dev.ready_wait()
    blob = np.zeros(3 << 20, np.uint8)
    up = dev.alloc(blob.nbytes)

    def exp_g():  # read || the prepared frame's kernels only (no upload at all)
        tk = apps[0].read_async(host[0])
        apps[1].submit(); dev.ready_wait()
        apps[0].read_wait(tk)

    def exp_g0():
        apps[1].submit(); dev.ready_wait()

    def exp_h():  # read || one 3 MB upload only
        tk = apps[0].read_async(host[0])
        up.write(blob); dev.ready_wait()
        apps[0].read_wait(tk)

    def exp_h0():
        up.write(blob); dev.ready_wait()

    def exp_i():  # read || the device-side clear only
        tk = apps[0].read_async(host[0])
        apps[1].clear(); dev.ready_wait()
        apps[0].read_wait(tk)

    def exp_i0():
        apps[1].clear(); dev.ready_wait()
    print(f"G read || kernels only: {timed(exp_g):.3f} ms (kernels alone {timed(exp_g0):.3f})   H read || 3 MB upload: {timed(exp_h):.3f} ms "
          f"(upload alone {timed(exp_h0):.3f})   I read || clear: {timed(exp_i):.3f} ms (clear alone {timed(exp_i0):.3f})")
    a, c, b, d = timed(exp_a), timed(exp_c), timed(exp_b), timed(exp_d)
    print(f"E read || (all uploads, then all draws): {timed(exp_e):.3f} ms   F the same without the read: {timed(exp_f):.3f} ms")
    print(f"A four reads: {a:.3f} ms ({a / 4:.3f} each)   C one frame: {c:.3f} ms   B read || frame: {b:.3f} ms   "
          f"D frame, then read of the other target: {d:.3f} ms   (overlap: {max(a / 4, c):.3f}, none: {a / 4 + c:.3f})")
    # steady-state period of the pipelined loop with parts of the frame left out
    def loop(kind, frames=16):
        for a in apps:
            a.clear(); a.prepare(trace)
        dev.ready_wait()
        tk = [None] * n
        t0 = time.perf_counter()
        for i in range(frames):
            a = apps[i % n]
            if tk[i % n] is not None:
                a.read_wait(tk[i % n])
            if kind in ("clear", "clear+kernels"):
                a.clear()
            if kind in ("kernels", "clear+kernels"):
                a.submit()
            if kind == "upload":
                up.write(blob)
            if kind == "host work":  # the host pre-pass of one frame, nothing sent to the device
                for d in range(trace.num_drawcalls):
                    trace.binning(d, W, H)
            tk[i % n] = a.read_async(host[i % n])
        for j, t in enumerate(tk):
            if t is not None:
                apps[j].read_wait(t)
        return (time.perf_counter() - t0) * 1e3 / frames
    print("period of read-back loops (ms per frame): " + "   ".join(
        f"{k}: {loop(k):.3f}" for k in ("reads only", "clear", "kernels", "clear+kernels", "upload", "host work")))
    for rep in range(2):
        tickets = [None] * n
        rows = []
        begin = time.perf_counter()
        for i in range(args.frames):
            a = apps[i % n]
            t0 = time.perf_counter()
            if tickets[i % n] is not None:
                a.read_wait(tickets[i % n])
            t1 = time.perf_counter()
            a.clear()
            t2 = time.perf_counter()
            a.render(trace)
            t3 = time.perf_counter()
            tickets[i % n] = a.read_async(host[i % n])
            t4 = time.perf_counter()
            st = a.stats()
            rows.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3, st["bin_ms"], st["upload_ms"], st["run_ms"]))
        for j, t in enumerate(tickets):
            if t is not None:
                apps[j].read_wait(t)
        end = time.perf_counter()
        if rep == 1:
            print("frame  wait_prev  clear  render  read_async | app stats: bin  upload  run   (ms)")
            for i, r in enumerate(rows):
                print(f"{i:4d}  {r[0] * 1e3:8.3f} {r[1] * 1e3:6.3f} {r[2] * 1e3:7.3f} {r[3] * 1e3:7.3f}     | {r[4]:6.3f} {r[5]:6.3f} {r[6]:6.3f}")
            a = np.array(rows[4:])
            print("mean of frames 4.. :", " ".join(f"{x:.3f}" for x in (a[:, :4].mean(0) * 1e3)), "| host busy per frame",
                  f"{a[:, 1:4].sum(1).mean() * 1e3:.3f} ms")
            print(f"total {args.frames} frames: {(end - begin) * 1e3:.2f} ms = {(end - begin) * 1e3 / args.frames:.3f} ms per frame")


if __name__ == "__main__":
    main()
