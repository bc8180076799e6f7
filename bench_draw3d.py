"""bench_draw3d.py -- the raster hot path (group R) measured on B200.

Reached through `python bench.py --workload draw3d [--trace mouse] [--size 4096] ...`; same JSON contract
as bench.py.  A "step" is one frame: every draw call of one of the reference's own traces
(tests/regression/draw3d/*.cgltrace, re-encoded under tests/golden/raster/) at --size x --size.

  value  Mfragments/s (fragments shaded and sent to the output merger), tile/primitive/texture buffers
         resident in HBM, all draws of the frame submitted back to back (srt_draw3d_submit); device time
         from CUDA events on the launch stream, L2 flushed and the frame cleared on the device before
         every timed step, max over ranks.
  e2e    the reference app's own loop with HOST buffers (apps/draw3d/draw3d_app.c: srt_draw3d_clear +
         srt_draw3d_render + srt_draw3d_read): per draw binning on the host, upload of tile / primitive /
         texture buffers and kernel_arg_t, vx_start, vx_ready_wait; colour buffer read back; wall clock.
  cpu_baseline / --impl reference: the reference's own code compiled natively (oracle/_ref, kind
         "reference") when that library travelled with the snapshot, else the plain-C oracle ("port"),
         on all host threads, whole frames.
N > 1: tile rows are dealt round-robin to the ranks (tile_y % N), buffers replicated, and the strips
are assembled on rank 0 with one NCCL gather per frame.  Weak scaling renders an N-times taller target
(--size x N*--size): the work per GPU stays that of the 1-GPU frame.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "Mfragments/s (fragments shaded by the draw3d kernel)"
ALL_SW = 7
TILE = 32


def workload_name(trace, w, h):
    return f"draw3d: reference trace '{trace}' (tests/regression/draw3d/{trace}.cgltrace), {w}x{h}, all draw calls"


def frame_size(args, world):
    h = args.size * (world if args.scaling == "weak" else 1)
    return args.size, h


def count_fragments(d3, trace, w, h):
    """Fragments per draw, counted by the kernel itself in an untimed pass (VX_DCR_RT_STATS = 1)."""
    from skybox_rt_b200 import vortex as vx
    frags = tests = 0
    with d3.Draw3D(w, h, flags=ALL_SW | d3.WIDE_TILES) as app:
        app.dev.dcr_write(vx.VX_DCR_RT_STATS, 1)
        app.clear()
        for d in range(trace.num_drawcalls):
            app.render(trace, d, d)
            if app.stats()["draws_submitted"]:
                frags += app.dev.mpm_query(d3.VX_CSR_RASTER_FRAGMENTS)
                tests += app.dev.mpm_query(d3.VX_CSR_RASTER_QUAD_TESTS)
    return frags, tests


def cpu_frames(blob, w, h, steps, warmup):
    import _oracle_draw3d as od
    lib, kind = (od.reference(), "reference") if od.reference() is not None else (od.oracle(), "port")
    threads = os.cpu_count() or 1
    for _ in range(warmup):
        od.render(lib, blob, w, h, threads=threads, depth=False)
    t0 = time.perf_counter()
    ksec = 0.0
    for _ in range(steps):
        _, _, k = od.render(lib, blob, w, h, threads=threads, depth=False)
        ksec += k
    wall = time.perf_counter() - t0
    return kind, threads, wall / steps, ksec / steps


def reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import _oracle_draw3d as od
    from skybox_rt_b200 import draw3d as d3  # host pre-pass only (binning totals); no CUDA call
    w, h = frame_size(args, args.gpus)
    blob = od.trace_blob(args.trace)
    steps, warmup = max(1, min(args.steps, 5)), max(0, min(args.warmup, 1))
    kind, threads, sec, ksec = cpu_frames(blob, w, h, steps, warmup)
    frags = fragments_cpu(blob, w, h)  # untimed counting run of the oracle
    v = frags / sec / 1e6
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mfragments/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "i32 fixed point (16.16 / 8.24) + f32 gradients", "data": "reference trace",
        "config": {"workload": workload_name(args.trace, w, h), "fragments_per_step": frags},
        "cpu_baseline": {"value": v, "unit": "Mfragments/s", "cores": threads, "kind": kind,
                         "sample": f"whole frame, {steps} frames, binning + upload-free host flow included "
                                   f"(kernel-only: {ksec * 1e3:.1f} ms/frame)"},
        "e2e": {"value": v, "unit": "Mfragments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def fragments_cpu(blob, w, h):
    """Fragment count of a frame without a GPU: an (untimed) counting run of the plain-C oracle."""
    import ctypes as C

    import _oracle_draw3d as od
    L = od.oracle()
    L.ref_draw3d_count_fragments.restype = C.c_uint64
    L.ref_draw3d_count_fragments.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
    return int(L.ref_draw3d_count_fragments(blob, len(blob), w, h, os.cpu_count() or 1))


class DevMem:
    """A vx buffer's device memory exposed to torch (zero copy) via __cuda_array_interface__."""

    def __init__(self, address, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<i4", "data": (address, False),
                                         "version": 2, "strides": None}


def native_arm(args, ClockSampler, load_peaks):
    import torch
    import torch.distributed as dist

    import _oracle_draw3d as od  # fixtures (trace blobs) + the cpu_baseline leg only
    from skybox_rt_b200 import draw3d as d3
    from skybox_rt_b200 import vortex as vx

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ["VX_CUDA_DEVICE"] = str(local)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)"
    W, H = frame_size(args, world)
    blob = od.trace_blob(args.trace)
    trace = d3.Trace(blob=blob)
    K, warm = args.steps, max(args.warmup, 3)

    # The device outlives every torch tensor that touched its stream (NCCL records its buffers on the stream it ran
    # on; freeing them after vx_dev_close destroyed that stream aborts at interpreter exit), so it is opened here and
    # left to process teardown.
    dev = vx.Device()
    app = d3.Draw3D(W, H, flags=ALL_SW | d3.WIDE_TILES, rank=rank, num_ranks=world, device=dev)
    stream = torch.cuda.ExternalStream(dev.cuda_stream(), device=torch.device("cuda", local))
    flush = torch.empty((256 << 20) // 4, dtype=torch.int32, device="cuda")
    caddr, zaddr = app.buffer_addresses()
    cbuf = torch.as_tensor(DevMem(caddr, W * H * 4), device="cuda").view(H, W)
    zbuf = torch.as_tensor(DevMem(zaddr, W * H * 4), device="cuda").view(H, W)
    dev.dcr_write(0x806, args.ppt)  # VX_DCR_RASTER_EXT_PPT
    app.prepare(trace)
    prep = app.stats()
    draws = prep["draws_submitted"]

    if world > 1:
        # one NCCL exchange per frame through the C entry vx_cuda_frame_gather_inplace: every rank renders its tile rows
        # at their final positions of a full-size colour buffer, the root receives the other ranks' 32-row blocks
        # straight into its own buffer; torch.distributed only carries the communicator id
        ident = [vx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        comm = vx.Comm(dev, ident[0], rank, world)

        def gather(a=None):
            comm.frame_gather_inplace((a or app).color_buffer_handle(), W, H, TILE, 0)
    else:
        gather = None

    def clear_dev():  # device-side equivalent of the host-upload clear (main.cpp:478-490), untimed plumbing
        cbuf.fill_(-16777216)  # 0xff000000
        zbuf.fill_(-1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None  # rank 0's record is the one that is printed
    if sampler:
        sampler.start()  # polls through the warm-up too; the record describes the samples inside the timed window
    with torch.cuda.stream(stream):
        for _ in range(warm):
            clear_dev()
            app.submit()
            if gather:
                gather()
    barrier()

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(K)]
    barrier()
    t_begin = time.perf_counter()
    with torch.cuda.stream(stream):
        for a, b, c in ev:
            clear_dev()
            flush.zero_()
            a.record(stream)
            app.submit()
            b.record(stream)
            if gather:
                gather()
            c.record(stream)
    barrier()
    clocks = None
    if sampler:
        sampler.window(t_begin, time.perf_counter())
        clocks = sampler.stop()
    dev_ms = sum(a.elapsed_time(c) for a, b, c in ev)
    kern_ms = sum(a.elapsed_time(b) for a, b, c in ev) / K  # the draws of one frame, without the gather

    # ---- e2e: the app's loop with host buffers, wall clock ---------------------------------------------
    # Default: the loop a host written for this back-end runs (SRT_DRAW3D_AUTO: device-side clear, device binning for
    # large draws, no wait between the draws of a frame) with two app objects, so that frame i+1 is binned, uploaded
    # and drawn while frame i travels to pinned host memory (srt_draw3d_read_async).  Every frame's uploads, draws
    # and read-back are inside the timed region; it ends when the last frame is on the host.
    # `serial` = the reference app's own flow verbatim (host-upload clear, host binning, wait per draw, blocking read).
    host_frames = [torch.empty(H * W, dtype=torch.int32).pin_memory() for _ in range(2)]
    host_frame = host_frames[0]
    device_clear = getattr(args, "device_clear", False)
    e2e_steps = max(2, min(K, 24))  # enough frames for the pipeline's fill and drain (two frames) not to dominate
    e2e_loop = "n/a"
    if gather is None:
        apps = [d3.Draw3D(W, H, flags=ALL_SW | d3.WIDE_TILES | d3.AUTO, rank=rank, num_ranks=world, device=dev) for _ in range(2)]

        def e2e_loop_run(k):
            tickets = [None, None]
            for i in range(k):
                a = apps[i & 1]
                if tickets[i & 1] is not None:
                    a.read_wait(tickets[i & 1])
                a.clear()
                a.render(trace)
                tickets[i & 1] = a.read_async(host_frames[i & 1])
            for j, t in enumerate(tickets):
                if t is not None:
                    apps[j].read_wait(t)
        e2e_loop_run(2)
        barrier()
        t0 = time.perf_counter()
        e2e_loop_run(e2e_steps)
        barrier()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        st2 = apps[0].stats()
        h2d_auto = st2["h2d_bytes"]
        for a in apps:
            a.close()
        e2e_loop = ("SRT_DRAW3D_AUTO (device clear, host pre-pass on a thread pool, the frame's buffers packed into one upload, "
                    "draws started back to back; device binning for draws >= 10^4 triangles), two app objects, frame i+1 "
                    "prepared and drawn while frame i is read back asynchronously")
    app2 = d3.Draw3D(W, H, flags=ALL_SW | d3.WIDE_TILES | (d3.GPU_BINNING if getattr(args, "gpu_binning", False) else 0) |
                     (d3.DEVICE_CLEAR if device_clear else 0),
                     rank=rank, num_ranks=world, device=dev)
    c2addr, _ = app2.buffer_addresses()
    cbuf2 = torch.as_tensor(DevMem(c2addr, W * H * 4), device="cuda").view(H, W)

    def e2e_step():
        app2.clear()
        app2.render(trace)
        if gather:
            gather(app2)  # queued behind the draws; rank 0's colour buffer then holds the whole frame
            if rank == 0:
                app2.read(out=host_frame)
            else:
                dev.ready_wait()
        else:
            app2.read(out=host_frame)
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    serial_s = (time.perf_counter() - t0) / e2e_steps
    st2 = app2.stats()
    h2d_serial = st2["h2d_bytes"] + (0 if device_clear else 2 * W * H * 4)  # per-draw uploads + the two clear uploads
    if gather is None:
        h2d = h2d_auto
    else:
        e2e_s, h2d = serial_s, h2d_serial
    d2h = W * H * 4 if rank == 0 else 0
    app2.close()

    # ---- units of work: counted by the kernel in an untimed pass ----------------------------------
    frags = tests = 0
    dev.dcr_write(vx.VX_DCR_RT_STATS, 1)
    app3 = d3.Draw3D(W, H, flags=ALL_SW | d3.WIDE_TILES, rank=rank, num_ranks=world, device=dev)
    app3.clear()
    tex_frags = 0
    for d in range(trace.num_drawcalls):
        app3.render(trace, d, d)
        if app3.stats()["draws_submitted"]:
            f = dev.mpm_query(d3.VX_CSR_RASTER_FRAGMENTS)
            frags += f
            tests += dev.mpm_query(d3.VX_CSR_RASTER_QUAD_TESTS)
    check = app3.read()
    app3.close()
    dev.dcr_write(vx.VX_DCR_RT_STATS, 0)
    launches = draws * K

    t = torch.tensor([dev_ms, e2e_s, float(frags), kern_ms, float(prep["tile_prims"]), float(tests)], dtype=torch.float64,
                     device="cuda")
    if world > 1:
        mx, sm = t.clone(), t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, e2e_s, kern_ms = mx[0].item(), mx[1].item(), mx[3].item()
        frags_total, pairs_total, tests_total = sm[2].item(), sm[4].item(), sm[5].item()
    else:
        frags_total, pairs_total, tests_total = float(frags), float(prep["tile_prims"]), float(tests)

    out = None
    if rank == 0:
        ms_per_step = dev_ms / K
        peak, peak_src = load_peaks()
        # algorithmic bytes of one frame on this rank: per (tile, primitive) pair the
        # id + the 120-byte record; per fragment the colour word written and the depth word read+written
        # when the draw tests depth; framebuffer words are counted once per fragment (the reference's
        # per-fragment read-modify-write), texels as 4 taps x bytes per texel for textured fragments
        alg = prep["tile_prims"] * (4 + 120) + frags * (4 + 8) + frags * 4 * 2
        achieved = alg / (kern_ms * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": frags_total / (ms_per_step * 1e-3) / 1e6, "unit": "Mfragments/s",
            "n_gpus": world, "steps": K, "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "i32 fixed point (16.16 / 8.24) + f32 gradients",
            "data": "reference trace",
            "config": {"workload": workload_name(args.trace, W, H), "draw_calls": int(draws),
                       "primitives": int(prep["prims"]), "tiles": int(prep["tiles"]),
                       "tile_primitive_pairs": int(pairs_total), "fragments_per_step": int(frags_total),
                       "quad_tests_per_step": int(tests_total),
                       "l2": "flushed before every timed step (256 MiB memset), frame cleared on the device before the flush",
                       "partition": (f"tile rows dealt round-robin over {world} GPUs (tile_y % N), buffers replicated, "
                                     "one NCCL gather per frame") if world > 1 else "single GPU"},
            "e2e": {"value": frags_total / e2e_s / 1e6, "unit": "Mfragments/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_s * 1e3,
                    "loop": e2e_loop,
                    "serial": {"ms_per_step": serial_s * 1e3, "value": frags_total / serial_s / 1e6, "h2d_bytes_per_step": int(h2d_serial),
                               "host_ms": {"binning": st2["bin_ms"], "upload": st2["upload_ms"], "start+wait": st2["run_ms"]},
                               "binning": "device" if getattr(args, "gpu_binning", False) else "host",
                               "clear": "device fill" if device_clear else "host upload (the reference app's flow)"}},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         # dram bytes read + written by the frame's 10 launches in the committed ncu --set full capture
                         # of this configuration (profiles/r01_draw3d_ncu_summary.txt: 99.0 + 55.6 MB)
                         "traffic": 154599424 if (args.trace == "mouse" and args.size == 4096 and world == 1) else None,
                         "peak_source": peak_src, "kernel": "raster_tile_kernel (all draws of a frame)",
                         "kernel_ms": kern_ms, "algorithmic_bytes": int(alg),
                         "note": "issue-bound integer path; see DESIGN.md section 9 for the issue-slot roofline"},
        }
        if world == 1 and not args.no_cpu_baseline:
            kind, threads, sec, ksec = cpu_frames(blob, W, H, steps=2, warmup=1)
            fo, _, _ = od.render(od.oracle(), blob, W, H, threads=threads, depth=False)
            out["parity_check"] = {"pixels_differing_from_oracle": int((check != fo).sum())}
            out["cpu_baseline"] = {"value": frags_total / sec / 1e6, "unit": "Mfragments/s", "cores": threads, "kind": kind,
                                   "sample": f"whole frame x 2, host flow included (kernel-only {ksec * 1e3:.1f} ms/frame)"}
    app.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))
