#!/usr/bin/env python
"""bench.py -- the hot path of BASELINE.json measured on B200 through the vx_* C ABI.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl reference]

A "step" is one frame of the workload (or one pass over the ray list for c5).  Prints ONE JSON line.
  value   Mrays/s with the scene resident in HBM, device time from CUDA events on the launch stream
          (L2 flushed before every timed step), max over ranks.
  e2e     the same metric through the public host flow with HOST buffers: upload kernel_arg_t ->
          vx_start -> vx_ready_wait -> vx_copy_from_dev(framebuffer) into pinned memory, wall clock.
  --impl reference : the CPU oracle (oracle/, this repo's restatement -- the reference has no RT
          code) on all host threads, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

# BASELINE.json configs
WORKLOADS = {
    "c1": dict(scene="cornell", quads=None, w=64, h=64, mode=2, spp=1,
               name="C1 Cornell-style box 36 tris, 64x64, primary+shadow+1 bounce, 1 spp"),
    "c2": dict(scene="terrain", quads=(250, 200), w=1024, h=1024, mode=0, spp=1,
               name="C2 100K-tri displaced grid, 1024x1024 primary rays"),
    "c3": dict(scene="terrain", quads=(1000, 500), w=1920, h=1080, mode=2, spp=1,
               name="C3 1M-tri displaced grid, 1920x1080 primary+shadow+1-bounce diffuse, 1 spp"),
    "c4": dict(scene="terrain", quads=(2500, 2000), w=3840, h=2160, mode=2, spp=8,
               name="C4 10M-tri displaced grid, 3840x2160 x 8 spp, primary+shadow+1-bounce"),
    "c5": dict(scene="terrain", quads=(2500, 2000), rays=64 << 20, mode=3,
               name="C5 64Mi incoherent rays vs 10M-tri BVH, closest hit"),
}
ROWS_PER_BLOCK = 8  # multi-GPU row interleave granularity
L2_FLUSH_BYTES = 256 << 20


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS) + ["draw3d"],
                    help="c1..c5: the RT configs of BASELINE.json; draw3d: the raster path (bench_draw3d.py)")
    ap.add_argument("--trace", default="mouse", help="draw3d: which reference trace (tests/golden/raster/<name>.trc.gz)")
    ap.add_argument("--size", type=int, default=4096, help="draw3d: render target is size x size")
    ap.add_argument("--split-samples", type=int, default=0,
                    help="diagnostic: VX_DCR_RT_SPLIT_SAMPLES (0 = split multi-sample frames into per-sample jobs, 1 = never)")
    ap.add_argument("--builder", default="host", choices=["host", "device"],
                    help="RT: BVH from the host binned-SAH builder (default) or built on the device (vx_cuda_rt_build_lbvh)")
    ap.add_argument("--device-clear", action="store_true",
                    help="draw3d: e2e leg clears the targets on the device (vx_cuda_fill32) instead of the reference's host upload")
    ap.add_argument("--gpu-binning", action="store_true",
                    help="draw3d: e2e leg bins on the device (vx_cuda_draw3d_binning) instead of the host pre-pass")
    ap.add_argument("--ppt", type=int, default=0, help="draw3d diagnostic: force pixels per thread (4 or 1; 0 = by grid size)")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--variant", type=int, default=0, help="VX_DCR_RT_VARIANT: 0 default (frames: wavefront pipeline, ray lists: vote-scheduled persistent kernel), "
                         "1 baseline BVH2 megakernel, 2 / 7 Node4 megakernel while-while / vote, 3 ray pool, 4-6 wavefront")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N>1: weak = N x the pixels of the same view (per-GPU work fixed), strong = the same frame split N ways")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true",
                    help="skip the strong-scaling legs (C4 frame and C5 ray list split over the N GPUs) of the default run")
    ap.add_argument("--no-variants", action="store_true", help="skip the untimed A/B steps of the other kernel variants")
    ap.add_argument("--no-flush", action="store_true", help="diagnostic: keep L2 warm between steps")
    ap.add_argument("--keep-active", type=int, default=0, help="diagnostic: refill threshold DCR (0 = default)")
    ap.add_argument("--prefetch", type=int, default=0, help="diagnostic: VX_DCR_RT_PREFETCH (2 = no L2 warm-up)")
    ap.add_argument("--unfused", type=int, default=0, help="diagnostic: VX_DCR_RT_FUSE_BOUNCE (1 = separate K4 / K5 launches)")
    ap.add_argument("--chains", type=int, default=0, help="diagnostic: VX_DCR_RT_CHAINS (0 = default)")
    ap.add_argument("--mode", type=int, default=None, help="diagnostic: override the workload's shading mode (0/1/2)")
    ap.add_argument("--spp", type=int, default=None, help="diagnostic: override the workload's samples per pixel")
    return ap.parse_args()


class ClockSampler:
    """SM clock + throttle reasons while the timed region runs (B200_PROFILING.md).  NVML is polled from a thread every
    ~4 ms (on rank 0 only: its record is the one that is printed, and eight processes polling the driver at once
    showed up as launch gaps in an 8-GPU run) with a time stamp per sample (the timed region of a 20-step run is ~30 ms: `nvidia-smi -lms 100`, used in
    round 1, returned no sample inside it at N = 8); the record describes the samples that fall inside the window
    the caller marks with `window(t0, t1)` (time.perf_counter values around the timed loop).  Falls back to
    nvidia-smi when NVML is not importable."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc, self.th = gpu_index, [], None, None
        self.t0 = self.t1 = None
        self.nvml = None
        self.stop_flag = False

    def window(self, t0, t1):
        self.t0, self.t1 = t0, t1

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(self.idx).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:  # noqa: BLE001 -- no torch / no uuid: physical index (no CUDA_VISIBLE_DEVICES remapping)
                self.h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            self.nvml = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:  # noqa: BLE001
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i",
                 str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        names = (("hw_slowdown", n.nvmlClocksThrottleReasonHwSlowdown), ("hw_thermal_slowdown", n.nvmlClocksThrottleReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", n.nvmlClocksThrottleReasonSwThermalSlowdown), ("sw_power_cap", n.nvmlClocksThrottleReasonSwPowerCap))
        while not self.stop_flag:
            try:
                t = time.perf_counter()
                sm = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((t, sm, [k for k, bit in names if mask & bit]))
            except Exception:  # noqa: BLE001
                break
            time.sleep(0.004)

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.th.join(timeout=2)
            rows = self.rows
            inside = [r for r in rows if self.t0 is not None and self.t0 <= r[0] <= self.t1]
            used, scope = (inside, "inside the timed region") if inside else (rows, "whole run (no sample fell inside the timed region)")
            sm = [r[1] for r in used]
            reasons = sorted({k for r in used for k in r[2]})
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "samples": len(sm),
                    "samples_total": len(rows), "scope": scope, "source": "NVML polled every ~4 ms", "reasons": reasons}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.th.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for _, r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        busy = sorted(sm)[len(sm) // 2:] if sm else []  # upper half = samples under load
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "source": "nvidia-smi -lms 20", "reasons": sorted(reasons)}


def weak_frame(w0, h0, n):
    """Frame of n x the pixels of w0 x h0 at (nearly) the same aspect: the height a multiple of n row blocks so
    that every rank owns the same number of rows, the width a multiple of the 8-pixel job tile."""
    unit = ROWS_PER_BLOCK * n
    h = max(unit, int(round(h0 * n ** 0.5 / unit)) * unit)
    w = max(8, int(round(w0 * h0 * n / h / 8)) * 8)
    return w, h


def profile_inputs(args, wl_key):
    """Counters read from the committed ncu captures: profiles/roofline_inputs.json maps "<workload>:v<variant>" to
    {source, issue_active, lanes_per_inst, dram_bytes} (written by profiles/make_roofline_inputs.py from the named
    summary files).  None for any configuration that was not captured -- nothing is assumed to carry over."""
    if args.gpus != 1 or args.builder != "host" or args.spp is not None or args.mode is not None:
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_inputs.json")) as f:
            return json.load(f).get(f"{wl_key}:v{args.variant}")
    except (OSError, ValueError):
        return None


def roofline_record(args, wl, r, dev, tracer, vx, cpu):
    """Roofline of the frame's traversal work.
    Algorithmic bytes = the ORACLE's per-ray visit counts on the identical (sub)sample x the oracle's record sizes
    (BVH2 node 32 B, triangle record 48 B), scaled to the rays of the step, + the framebuffer / ray I/O.  The bound is
    chosen by working set: a scene that fits the L2 is measured against the L2 -> SM read bandwidth probed on this
    box (vx_cuda_probe_read_bandwidth), an HBM-resident one against MEASURED_PEAKS.json; the issue roofline
    (148 SM x 4 schedulers x SM clock) is reported next to it with the counters of the committed capture."""
    peak_hbm, peak_src = load_peaks()
    # the kernels' own counts, from an extra untimed step of the counting instantiations
    dev.dcr_write(vx.VX_DCR_RT_STATS, 1)
    dev.start(tracer.kernel, tracer.args)
    dev.ready_wait()
    st = tracer.counters()
    wn, wt = dev.mpm_query(vx.VX_CSR_RT_WARP_NODE_STEPS), dev.mpm_query(vx.VX_CSR_RT_WARP_TRI_STEPS)
    dev.dcr_write(vx.VX_DCR_RT_STATS, 0)
    node_bytes = dev.mpm_query(vx.VX_CSR_RT_NODE_BYTES)
    world = r["world"]
    rays_rank = r["rays_total"] / world
    fb_bytes = 0 if wl["mode"] == 3 else wl["w"] * wl["h"] * 4 // world
    ray_io = rays_rank * 48 if wl["mode"] == 3 else 0
    kernel_bytes = st["nodes"] * node_bytes + st["tris"] * 48 + fb_bytes + ray_io
    kern_s = r["kern_ms"] * 1e-3
    rec = {"kernel": "whole vx_start pipeline (trace + shade)", "kernel_ms": r["kern_ms"], "unit": "GB/s",
           "kernel_counts": {"nodes_fetched": st["nodes"], "node_bytes": int(node_bytes), "tris_tested": st["tris"],
                             "bytes": int(kernel_bytes),
                             "lanes_per_node_step": round(st["nodes"] / wn, 2) if wn else None,
                             "lanes_per_tri_step": round(st["tris"] / wt, 2) if wt else None}}
    if cpu is not None:
        oc = cpu["counters"]
        scale = rays_rank / max(1, cpu["rays_per_step"])
        alg = (oc["nodes_visited"] * 32 + oc["tris_tested"] * 48) * scale + fb_bytes + ray_io
        rec.update(counts_source="oracle", algorithmic_bytes=int(alg),
                   oracle_counts={"nodes_visited": oc["nodes_visited"], "tris_tested": oc["tris_tested"],
                                  "rays": cpu["rays_per_step"], "sample": cpu["sample"], "scaled_by": round(scale, 3),
                                  "node_bytes": 32})
    else:
        alg = kernel_bytes
        rec.update(counts_source="kernel (no oracle run in this invocation)", algorithmic_bytes=int(alg))
    achieved = alg / kern_s / 1e9 if kern_s > 0 else 0.0
    # working set: the wide-node image + triangle records the kernels read
    ws = r["n_tris"] * 48 + (r["n_nodes"] * 32 if args.variant == vx.RT_VARIANT_BASELINE else dev.node4_image_size())
    l2_bytes = r["torch"].cuda.get_device_properties(r["torch"].cuda.current_device()).L2_cache_size
    try:
        l2_gbs = dev.read_bandwidth_gbs(64 << 20, 20)
        hbm_gbs = dev.read_bandwidth_gbs(2 << 30, 4)
    except Exception:  # noqa: BLE001 -- the probe is a report, not the product
        l2_gbs = hbm_gbs = None
    resident = ws < 0.8 * l2_bytes
    if resident and l2_gbs:
        rec.update(bound="l2", peak=l2_gbs, peak_source="L2 -> SM read bandwidth probed on this box: 64 MiB buffer, "
                   "16-byte read-only loads, 20 passes (vx_cuda_probe_read_bandwidth)")
    else:
        rec.update(bound="hbm", peak=peak_hbm, peak_source=peak_src)
    rec.update(achieved=achieved, frac=achieved / rec["peak"], hbm_frac=achieved / peak_hbm,
               l2_frac=achieved / l2_gbs if l2_gbs else None, working_set_bytes=int(ws),
               probed={"l2_read_gbs": l2_gbs, "hbm_read_gbs": hbm_gbs})
    sm_mhz = (r["clocks"] or {}).get("sm_mhz") or 1965.0
    issue_peak = 148 * 4 * sm_mhz * 1e6
    pi = profile_inputs(args, args.workload)
    rec["issue"] = {"peak_warp_inst_per_s": issue_peak, "sm_mhz": sm_mhz}
    if pi:
        part = pi.get("step_fraction", 1.0)  # the capture covers this share of the bench step (c5: 16 Mi of 64 Mi rays)
        rec["traffic"] = int(pi["dram_bytes"] / part) if pi.get("dram_bytes") is not None else None
        rec["issue"].update(issue_active=pi.get("issue_active"), lanes_per_inst=pi.get("lanes_per_inst"),
                            warp_inst_per_step=int(pi["warp_inst"] / part) if pi.get("warp_inst") else None,
                            source=pi.get("source"), capture_note=pi.get("partial"))
        if pi.get("warp_inst"):
            # share of the issue slots of the step that carried an instruction, and the share that carried useful
            # lanes: the two factors of the issue roofline
            rec["issue"]["frac"] = pi["warp_inst"] / part / kern_s / issue_peak
            if pi.get("lanes_per_inst"):
                rec["issue"]["lane_weighted_frac"] = rec["issue"]["frac"] * pi["lanes_per_inst"] / 32.0
    else:
        rec["traffic"] = None
    return rec


def variants_ms(args, r, dev, tracer, vx):
    """Device ms per frame of every shipped kernel variant on this workload (3 frames each after one warm-up, L2
    flushed before each, untimed as far as `value` is concerned): the A/B every optimisation is argued from."""
    torch, stream, flush = r["torch"], r["stream"], r["flush"]
    res = {}
    for v in sorted(vx.RT_VARIANTS):
        if v == vx.RT_VARIANT_POOL:
            continue  # explicit ray lists only
        dev.dcr_write(vx.VX_DCR_RT_VARIANT, v)
        ts = []
        for i in range(4):
            with torch.cuda.stream(stream):
                if not args.no_flush:
                    flush.zero_()
            stream.synchronize()
            dev.start(tracer.kernel, tracer.args)
            dev.ready_wait()
            if i:
                ts.append(dev.mpm_query(vx.VX_CSR_RT_ELAPSED_NS) * 1e-6)
        res[str(v)] = round(sum(ts) / len(ts), 4)
    dev.dcr_write(vx.VX_DCR_RT_VARIANT, args.variant)
    return res


def build_scene(wl, rt):
    if wl["scene"] == "cornell":
        soup = rt.scene_cornell()
    else:
        soup = rt.scene_terrain(*wl["quads"])
    t0 = time.perf_counter()
    nodes, tris = rt.build_bvh2(soup)
    return soup, nodes, tris, time.perf_counter() - t0


def scene_box(soup):
    lo = np.minimum.reduce([soup["v0"].min(0), soup["v1"].min(0), soup["v2"].min(0)])
    hi = np.maximum.reduce([soup["v0"].max(0), soup["v1"].max(0), soup["v2"].max(0)])
    return lo.astype(np.float32), hi.astype(np.float32)


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle on the host cores, on a bounded sample of the workload
# ------------------------------------------------------------------------------------------------
def cpu_sample_run(wl, soup, nodes, tris, steps, warmup, threads=None, shrink=1):
    """Times the CPU oracle.  Sample = the rows a rank would own if the frame were split over
    `div` ranks (interleaved blocks of 8 rows => representative of the whole frame), or the first
    1/div of the ray list for c5; `shrink` makes the sample that many times smaller again (single-thread leg).
    Returns a dict: Mrays/s, sample description, threads, seconds per step, the oracle's counters, and what it
    rendered (`fb` + the rank split, or `hits` + the rays) so that the GPU frame can be checked against it."""
    import _oracle as o
    threads = threads or (os.cpu_count() or 1)
    out = {"threads": threads}
    if wl["mode"] == 3:
        n = min(wl["rays"], (1 << 20) // shrink)
        lo, hi = scene_box(soup)
        rays = o.rays_random(lo, hi, 0x5EED5EED, 0, n)

        def run():
            h, c = o.trace(rays, nodes, tris, threads=threads, num_tasks=threads * 16)
            out["hits"] = h
            return c
        out["rays_in"] = rays
        sample = f"first {n} of {wl['rays']} rays"
    else:
        paths = wl["w"] * wl["h"] * wl["spp"]
        div = 1
        while paths // div > 300_000 // shrink and div < 256:
            div *= 2
        a = o.make_arg(wl["w"], wl["h"], wl["mode"], spp=wl["spp"], seed=1234, scene=wl["scene"],
                       num_tasks=wl["h"])
        a.num_ranks, a.rank_id, a.rows_per_block = div, 0, ROWS_PER_BLOCK
        out["div"] = div

        def run():
            fb, c = o.render(a, nodes, tris, threads=threads)
            out["fb"] = fb
            return c
        sample = (f"rows of rank 0 of {div} (interleaved blocks of {ROWS_PER_BLOCK} rows) of the "
                  f"{wl['w']}x{wl['h']}x{wl['spp']}spp frame" if div > 1 else "the whole frame")
    for _ in range(warmup):
        run()
    t0 = time.perf_counter()
    rays_total = 0
    for _ in range(steps):
        c = run()
        rays_total += c["rays_primary"] + c["rays_shadow"] + c["rays_bounce"]
    dt = time.perf_counter() - t0
    out.update(value=rays_total / dt / 1e6, sample=sample, sec_per_step=dt / steps, counters=c,
               rays_per_step=rays_total // steps)
    return out


def parity_check(wl, tracer, cpu):
    """The frame the bench times against what the oracle just rendered: the GPU renders the same rows (same rank
    split, same seed) once, untimed, and every pixel and ray counter must match bit for bit; for c5 the first hit
    records are compared as raw bytes.  (tests/regression/draw3d/main.cpp:505-514: compare what you rendered.)"""
    from skybox_rt_b200 import raytrace as rt
    if wl["mode"] == 3:
        hits = tracer.trace(cpu["rays_in"])
        ref = cpu["hits"]
        bad = int((hits.view(np.uint8).reshape(len(hits), -1) != ref.view(np.uint8).reshape(len(ref), -1)).any(axis=1).sum())
        return {"rays": int(len(ref)), "mismatches": bad, "against": "oracle hit records, raw bytes"}
    a = rt.make_arg(wl["w"], wl["h"], wl["mode"], spp=wl["spp"], seed=1234, scene=wl["scene"], rank=0,
                    num_ranks=cpu["div"], rows_per_block=ROWS_PER_BLOCK)
    fb = tracer.render(a).copy()
    c = tracer.counters()
    ref, oc = cpu["fb"], cpu["counters"]
    bad = int((fb != ref).sum()) if fb.shape == ref.shape else int(ref.size)
    counters_equal = all(c[k] == oc[k] for k in ("rays_primary", "rays_shadow", "rays_bounce"))
    return {"rows": int(ref.shape[0]), "pixels": int(ref.size), "mismatches": bad, "ray_counters_equal": counters_equal,
            "against": f"oracle frame, rank 0 of {cpu['div']}, A8R8G8B8 words"}


def reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from skybox_rt_b200 import raytrace as rt  # host pre-pass only (plain C, no CUDA)
    soup, nodes, tris, _ = build_scene(wl, rt)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    r = cpu_sample_run(wl, soup, nodes, tris, steps, warmup)
    v = r["value"]
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s (primary+shadow+bounce rays traced)", "value": v, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["sec_per_step"] * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"]},
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": r["threads"], "kind": "port", "sample": r["sample"],
                         "tree": "binned-SAH BVH2 from this repo's host builder (libskybox_rt_host.so), the one the GPU "
                                 "arm uploads: a better tree than the oracle's own spatial-median builder, so this arm "
                                 "is not understated"},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference repo has no RT path (SURVEY.md section 0); this is the repo's plain-C oracle",
    }))


# ------------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------------
class DevMem:
    """A vx buffer's device memory exposed to torch (zero copy) via __cuda_array_interface__."""

    def __init__(self, address, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<u4", "data": (address, False),
                                         "version": 2, "strides": None}


def native_arm(args, wl, steps, solo=False, sample_clocks=True):
    """One workload measured on this process' GPU (all ranks call it; `solo`: rank 0 alone runs the whole workload,
    the others return None -- the 1-GPU leg of a strong-scaling pair measured inside an N-GPU run)."""
    import torch
    import torch.distributed as dist

    from skybox_rt_b200 import raytrace as rt
    from skybox_rt_b200 import vortex as vx
    from skybox_rt_b200.formats import HIT, RAY, rank_rows

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if solo:
        if rank != 0:
            return None
        world = 1

    soup, nodes, tris, build_s = build_scene(wl, rt)
    device_build_ms = None
    if args.builder == "device":
        tracer = rt.RayTracer(soup=soup)  # warm: module load, pool growth
        tracer.close()
        t0 = time.perf_counter()
        tracer = rt.RayTracer(soup=soup)
        tracer.dev.ready_wait()
        device_build_ms = (time.perf_counter() - t0) * 1e3  # device open + 40 MB soup upload + build
    else:
        tracer = rt.RayTracer(nodes, tris)
    dev = tracer.dev
    dev.dcr_write(vx.VX_DCR_RT_VARIANT, args.variant)
    dev.dcr_write(vx.VX_DCR_RT_KEEP_ACTIVE, args.keep_active)
    dev.dcr_write(vx.VX_DCR_RT_PREFETCH, args.prefetch)
    dev.dcr_write(vx.VX_DCR_RT_SPLIT_SAMPLES, args.split_samples)
    dev.dcr_write(vx.VX_DCR_RT_FUSE_BOUNCE, args.unfused)
    dev.dcr_write(vx.VX_DCR_RT_CHAINS, args.chains)
    stream = torch.cuda.ExternalStream(dev.cuda_stream(), device=torch.device("cuda", local))
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.int32, device="cuda")
    extra = []  # buffers / communicators to release with the tracer

    rays_mode = wl["mode"] == 3
    gather = None
    if rays_mode:
        n_total = wl["rays"]
        n_local = n_total // world
        first = rank * n_local
        lo, hi = scene_box(soup)
        host_rays = torch.empty(n_local * RAY.itemsize, dtype=torch.uint8).pin_memory()
        rt.rays_random(lo, hi, 0x5EED5EED, first, n_local, out=host_rays.numpy().view(RAY))
        host_hits = torch.empty(n_local * HIT.itemsize, dtype=torch.uint8).pin_memory()
        ray_buf = dev.alloc(n_local * RAY.itemsize, vx.VX_MEM_READ)
        hit_buf = dev.alloc(n_local * HIT.itemsize)
        extra += [ray_buf, hit_buf]
        ray_buf.write(host_rays)
        from skybox_rt_b200.formats import KernelArg
        arg = KernelArg()
        arg.num_tasks, arg.mode = 1, 3
        arg.ray_addr, arg.hit_addr, arg.num_rays = ray_buf.address, hit_buf.address, n_local
        arg.node_addr, arg.tri_addr = tracer.node_buf.address, tracer.tri_buf.address
        arg.num_nodes, arg.num_tris = tracer.num_nodes, tracer.num_tris
        args_buf = dev.upload(arg)
        tracer.args = args_buf
        h2d, d2h = n_local * RAY.itemsize + 192, n_local * HIT.itemsize
        pipeline_depth = 1

        def launch():
            dev.start(tracer.kernel, args_buf)

        def e2e_serial(i):
            ray_buf.write(host_rays)
            args_buf.write(arg)
            dev.start(tracer.kernel, args_buf)
            dev.ready_wait()
            hit_buf.read(host_hits)

        def e2e_loop(k):  # a ray list has no next frame to overlap with: the loop is the serial one
            for i in range(k):
                e2e_serial(i)
    else:
        W, H = wl["w"], wl["h"]
        arg = rt.make_arg(W, H, wl["mode"], spp=wl["spp"], seed=1234, scene=wl["scene"], rank=rank,
                          num_ranks=world, rows_per_block=ROWS_PER_BLOCK)
        rows = tracer.start(arg)  # binds cbuf + args, first (cold) frame
        dev.ready_wait()
        host_frames = [torch.empty(H * W, dtype=torch.int32).pin_memory() for _ in range(2)]
        h2d, d2h = 192, (H * W * 4 if rank == 0 else 0)
        pipeline_depth = 2
        if world > 1:
            # one NCCL exchange of the strips per frame, queued on the render stream by the C entry
            # vx_cuda_frame_gather; torch.distributed only carried the communicator id
            from skybox_rt_b200.multi import NcclFrameGather
            fg = NcclFrameGather(dev, W, H, ROWS_PER_BLOCK)
            frames = [fg.frame, dev.alloc(H * W * 4) if rank == 0 else None]
            extra += [fg] + ([frames[1]] if rank == 0 else [])

            def gather(slot=0):
                fg.comm.frame_gather(tracer.cbuf, frames[slot], W, H, ROWS_PER_BLOCK, 0)
            targets = None
        else:
            targets = [tracer.cbuf, dev.alloc(max(rows, 1) * W * 4)]  # double-buffered render target
            extra.append(targets[1])

        def launch():
            dev.start(tracer.kernel, tracer.args)

        def e2e_serial(i):
            arg.seed = 1234 + i  # a new kernel_arg_t every frame, from the host
            tracer.args.write(arg)
            dev.start(tracer.kernel, tracer.args)
            if gather:
                gather(0)
                dev.ready_wait()
                if rank == 0:  # one D2H copy of the assembled frame
                    frames[0].read(host_frames[0])
            else:
                dev.ready_wait()
                tracer.cbuf.read(host_frames[0])

        def e2e_loop(k):
            """The render loop an application runs: frame i+1 is queued while frame i travels to the host
            (vx_cuda_copy_from_dev_async, two render targets / two host frames).  Every frame's kernel_arg_t upload
            and read-back are inside the timed region; the loop ends when the last frame is on the host."""
            tickets = [None, None]
            for i in range(k):
                slot = i & 1
                if tickets[slot] is not None:
                    dev.copy_wait(tickets[slot])  # the app is done with frame i-2; its buffers are free again
                arg.seed = 1234 + i
                if targets:
                    arg.cbuf_addr = targets[slot].address
                tracer.args.write(arg)
                dev.start(tracer.kernel, tracer.args)
                if gather:
                    gather(slot)
                    if rank == 0:
                        tickets[slot] = frames[slot].read_async(host_frames[slot])
                else:
                    tickets[slot] = targets[slot].read_async(host_frames[slot])
            for t in tickets:
                if t is not None:
                    dev.copy_wait(t)
            dev.ready_wait()
            if targets:
                arg.cbuf_addr = targets[0].address
                tracer.args.write(arg)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_rays():
        c = tracer.counters()
        return c["rays_primary"] + c["rays_shadow"] + c["rays_bounce"], c

    sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
    if sampler:
        sampler.start()  # polling runs through the warm-up as well: it is in steady state when the timed region begins
    # ---- warm-up --------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        launch()
        if gather:
            gather()
    barrier()
    rays_per_step, ctr = step_rays()

    # ---- timed: K steps, device time on the launch stream, L2 flushed before each ------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    t_wall = time.perf_counter()
    with torch.cuda.stream(stream):
        for a, b in ev:
            if not args.no_flush:
                flush.zero_()
            a.record(stream)
            launch()
            if gather:
                gather()
            b.record(stream)
    barrier()
    t_end = time.perf_counter()
    if sampler:
        sampler.window(t_wall, t_end)
    t_wall = t_end - t_wall
    clocks = sampler.stop() if sampler else None
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    kern_ns = dev.mpm_query(vx.VX_CSR_RT_ELAPSED_NS)  # vx_start..end of the last step, without gather

    # ---- e2e: host buffers, copies inside the timed region, wall clock ---------------------------
    e2e_loop(2)
    barrier()
    t0 = time.perf_counter()
    e2e_loop(steps)
    barrier()
    e2e_s = time.perf_counter() - t0
    for i in range(2):
        e2e_serial(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        e2e_serial(i)
    barrier()
    e2e_serial_s = time.perf_counter() - t0

    t = torch.tensor([dev_ms, e2e_s, float(rays_per_step), kern_ns * 1e-6, e2e_serial_s], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, e2e_s, kern_ms, e2e_serial_s = mx[0].item(), mx[1].item(), mx[3].item(), mx[4].item()
        rays_total = sm[2].item()
    else:
        rays_total, kern_ms = float(rays_per_step), kern_ns * 1e-6
    return dict(rank=rank, world=world, dev_ms=dev_ms, e2e_s=e2e_s, e2e_serial_s=e2e_serial_s, rays_total=rays_total,
                kern_ms=kern_ms, clocks=clocks, h2d=h2d, d2h=d2h, ctr=ctr, build_s=build_s, steps=steps,
                device_build_ms=device_build_ms, t_wall=t_wall, soup=soup, nodes=nodes, pipeline_depth=pipeline_depth,
                tris=tris, tracer=tracer, n_nodes=len(nodes), n_tris=len(tris), stream=stream, flush=flush, launch=launch,
                torch=torch, arg=arg, extra=extra)


def release(r):
    if r is None:
        return
    for x in r["extra"]:
        if hasattr(x, "close"):
            x.close()
        else:
            x.free()
    r["tracer"].close()


def strong_legs(args):
    """north_star: "image tiles partitioned across the GPUs of one box ... scaling >= 7x on 8 GPUs" is STRONG scaling
    of one fixed job.  For C4 (the 4K x 8 spp frame) and C5 (the 64 Mi-ray list), in this same run: the job split over
    all N ranks (device time incl. the frame gather, max over ranks) and the whole job on rank 0 alone, with a clock
    record each.  At N = 1 the pair collapses to the single-GPU figure."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    out = {}
    for key in ("c4", "c5"):
        wl = dict(WORKLOADS[key])
        r = native_arm(args, wl, steps=5)
        K = r["steps"]
        leg = None
        if r["rank"] == 0:
            ms = r["dev_ms"] / K
            leg = {"workload": wl["name"], "n_gpus": world, "steps": K, "ms_per_step": ms,
                   "value": r["rays_total"] / (ms * 1e-3) / 1e6, "unit": "Mrays/s",
                   "e2e_ms_per_step": r["e2e_s"] / K * 1e3, "e2e_value": r["rays_total"] / (r["e2e_s"] / K) / 1e6,
                   "e2e_serial_ms_per_step": r["e2e_serial_s"] / K * 1e3,
                   "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"], "clocks": r["clocks"]}
        release(r)
        if world > 1:
            dist.barrier()
            r1 = native_arm(args, wl, steps=3, solo=True)
            if r1 is not None:
                ms1 = r1["dev_ms"] / r1["steps"]
                leg.update(n1_ms_per_step=ms1, n1_value=r1["rays_total"] / (ms1 * 1e-3) / 1e6, n1_clocks=r1["clocks"],
                           n1_e2e_ms_per_step=r1["e2e_s"] / r1["steps"] * 1e3,
                           speedup_vs_n1=ms1 / leg["ms_per_step"],
                           e2e_speedup_vs_n1=(r1["e2e_s"] / r1["steps"] * 1e3) / leg["e2e_ms_per_step"])
                release(r1)
            dist.barrier()
        if leg is not None:
            out[key] = leg
    return out


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md)"


def main():
    args = parse()
    if args.workload == "draw3d":
        import bench_draw3d
        if args.impl == "reference":
            bench_draw3d.reference_arm(args)
        else:
            bench_draw3d.native_arm(args, ClockSampler, load_peaks)
        return
    wl = dict(WORKLOADS[args.workload])
    if args.mode is not None and wl["mode"] != 3:
        wl["mode"] = args.mode
        wl["name"] += f" [diagnostic: mode {args.mode}]"
    if args.spp is not None and wl["mode"] != 3:
        wl["spp"] = args.spp
        wl["name"] += f" [diagnostic: spp {args.spp}]"
    if args.gpus > 1 and args.scaling == "weak" and wl["mode"] != 3:
        # weak scaling: the same view at sqrt(N) x the resolution in each direction, rows interleaved over the
        # N ranks -- pixels (= jobs), samples and so rays per GPU stay those of the 1-GPU frame while the whole
        # job grows N-fold.  The height is rounded so that every rank owns the same number of row blocks.
        n = args.gpus
        wl["w"], wl["h"] = weak_frame(wl["w"], wl["h"], n)
        wl["name"] += f" [weak scaling: {n} x the pixels, {wl['w']}x{wl['h']}]"
    if args.impl == "reference":
        reference_arm(args, wl)
        return
    from skybox_rt_b200 import vortex as vx
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ["VX_CUDA_DEVICE"] = str(local)
    torch.cuda.set_device(local)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)"
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    r = native_arm(args, wl, steps=args.steps)
    tracer, dev = r["tracer"], r["tracer"].dev
    out = None
    if r["rank"] == 0:
        K = args.steps
        ms_per_step = r["dev_ms"] / K
        value = r["rays_total"] / (ms_per_step * 1e-3) / 1e6
        e2e = r["rays_total"] / (r["e2e_s"] / K) / 1e6
        out = {
            "metric": "Mrays/s (primary+shadow+bounce rays traced)", "value": value, "unit": "Mrays/s",
            "n_gpus": r["world"], "steps": K, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["name"], "triangles": r["n_tris"], "bvh2_nodes": r["n_nodes"],
                       "rays_per_step": int(r["rays_total"]), "kernel_variant": args.variant,
                       "l2": "kept warm (--no-flush)" if args.no_flush else "flushed before every timed step (256 MiB memset)",
                       "partition": f"rows interleaved in blocks of {ROWS_PER_BLOCK} over {r['world']} GPU(s), "
                                    "BVH replicated, one NCCL gather per frame" if r["world"] > 1 else "single GPU"},
            "rays": {k: r["ctr"][k] for k in ("rays_primary", "rays_shadow", "rays_bounce")},
            "e2e": {"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                    "ms_per_step": r["e2e_s"] / K * 1e3, "frames_in_flight": r["pipeline_depth"],
                    "loop": "kernel_arg_t upload -> vx_start -> (frame gather) -> vx_cuda_copy_from_dev_async into pinned "
                            "host memory, frame i+1 queued while frame i travels; ends when the last frame is on the host"
                            if r["pipeline_depth"] > 1 else "rays upload -> vx_start -> vx_ready_wait -> hits read-back",
                    "serial_ms_per_step": r["e2e_serial_s"] / K * 1e3,
                    "serial_value": r["rays_total"] / (r["e2e_serial_s"] / K) / 1e6},
            "gpu_launches": int(dev.mpm_query(vx.VX_CSR_RT_LAUNCHES)) * K,
            "clocks": r["clocks"],
            "host_prepass_s": {"bvh2_build": r["build_s"]},
            "builder": args.builder, "device_open_upload_build_ms": r.get("device_build_ms"),
        }
        cpu = None
        if r["world"] == 1 and not args.no_cpu_baseline:
            cpu = cpu_sample_run(wl, r["soup"], r["nodes"], r["tris"], steps=3, warmup=1)
            one = cpu_sample_run(wl, r["soup"], r["nodes"], r["tris"], steps=1, warmup=0, threads=1, shrink=8)
            out["cpu_baseline"] = {"value": cpu["value"], "unit": "Mrays/s", "cores": cpu["threads"], "kind": "port",
                                   "sample": cpu["sample"],
                                   "single_thread": {"value": one["value"], "unit": "Mrays/s", "cores": 1, "sample": one["sample"]}}
            out["parity_check"] = parity_check(wl, tracer, cpu)
            tracer.args.write(r["arg"])  # back to the kernel_arg_t of the timed step (the check rendered a rank strip)
        out["roofline"] = roofline_record(args, wl, r, dev, tracer, vx, cpu)
        if r["world"] == 1 and wl["mode"] != 3 and not args.no_variants:
            out["variants_ms"] = variants_ms(args, r, dev, tracer, vx)
    release(r)
    if not args.no_strong and args.workload == "c3" and args.scaling == "weak":
        legs = strong_legs(args)
        if out is not None:
            out["strong"] = legs
    if r["world"] > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))
        pc = out.get("parity_check")
        if pc and (pc["mismatches"] != 0 or pc.get("ray_counters_equal") is False):
            sys.exit(f"bench.py: the timed frame differs from the oracle: {pc}")


if __name__ == "__main__":
    main()
