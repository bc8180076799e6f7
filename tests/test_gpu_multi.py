"""Multi-GPU frame assembly through the C ABI (include/vortex_cuda.h: vx_cuda_comm_*, vx_cuda_frame_gather,
vx_cuda_copy_from_dev_async): NCCL send/receive group + de-interleave kernel, no torch in the data path.

The reference is single-device, so the checker is the product's own single-GPU frame (itself bit-exact against the
oracle, tests/test_gpu_rt*.py): splitting the rows over ranks and gathering them must not change a byte.
One-GPU boxes run the communicator-of-one and asynchronous-copy cases; the two-rank cases need `gpurun --gpus 2`."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,rpb", [(64, 40, 8), (50, 21, 4)])  # 16-byte rows / ragged rows (4-byte path)
def test_frame_gather_on_a_communicator_of_one_copies_the_strip(native, w, h, rpb):
    from skybox_rt_b200 import raytrace as rt
    from skybox_rt_b200 import vortex as vx
    nodes, tris = rt.build_bvh2(rt.scene_cornell())
    with rt.RayTracer(nodes, tris) as tracer:
        dev = tracer.dev
        comm = vx.Comm(dev, vx.comm_unique_id(), 0, 1)
        frame = dev.alloc(w * h * 4)
        try:
            fb = tracer.render(rt.make_arg(w, h, 2, spp=1, seed=5))
            comm.frame_gather(tracer.cbuf, frame, w, h, rpb, 0)
            dev.ready_wait()
            out = np.empty((h, w), np.uint32)
            frame.read(out)
            assert np.array_equal(out, fb)
            # argument checks: a frame buffer that is too small, a root outside the communicator
            small = dev.alloc(w * h * 4 - 4)
            with pytest.raises(vx.VortexError):
                comm.frame_gather(tracer.cbuf, small, w, h, rpb, 0)
            with pytest.raises(vx.VortexError):
                comm.frame_gather(tracer.cbuf, frame, w, h, rpb, 1)
            small.free()
        finally:
            frame.free()
            comm.close()


@pytest.mark.gpu
def test_async_read_back_equals_blocking_read_and_overlaps_frames(native):
    import torch
    from skybox_rt_b200 import raytrace as rt
    nodes, tris = rt.build_bvh2(rt.scene_terrain(60, 40))
    w, h = 256, 144
    with rt.RayTracer(nodes, tris) as tracer:
        dev = tracer.dev
        ref = [tracer.render(rt.make_arg(w, h, 2, spp=1, seed=s, scene="terrain")).copy() for s in (1, 2, 3)]
        targets = [tracer.cbuf, dev.alloc(w * h * 4)]
        host = [torch.empty(w * h, dtype=torch.int32).pin_memory() for _ in range(2)]
        got, tickets = [], [None, None]
        for i, s in enumerate((1, 2, 3)):
            slot = i & 1
            if tickets[slot] is not None:
                dev.copy_wait(tickets[slot])
                got.append(host[slot].numpy().view(np.uint32).reshape(h, w).copy())
            arg = rt.make_arg(w, h, 2, spp=1, seed=s, scene="terrain")
            tracer._bind(arg, w * h * 4)
            arg.cbuf_addr = targets[slot].address
            tracer.args.write(arg)  # queued behind the previous frame without waiting for it
            dev.start(tracer.kernel, tracer.args)
            tickets[slot] = targets[slot].read_async(host[slot])
        for slot in (1, 0):  # frames 2 and 3 (seeds 2, 3) are in slots 1 and 0
            dev.copy_wait(tickets[slot])
            got.append(host[slot].numpy().view(np.uint32).reshape(h, w).copy())
        targets[1].free()
        with pytest.raises(Exception):
            dev.copy_wait(10 ** 6)  # a ticket that was never issued
    for g, r in zip(got, ref):
        assert np.array_equal(g, r)


@pytest.mark.gpu
@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("size,mode,spp", [((320, 200), 2, 1), ((250, 131), 2, 2)])
def test_c_app_two_ranks_equal_single_gpu_png(native, tmp_path, size, mode, spp):
    """bin/raytrace -g 2 forks one process per GPU; rank 0's PNG of the gathered frame is compared by the app itself
    (-r, tol 0) with the single-GPU PNG."""
    exe = os.path.join(ROOT, "skybox_rt_b200", "bin", "raytrace")
    one, two = str(tmp_path / "one.png"), str(tmp_path / "two.png")
    common = ["-s", "terrain", "-n", "60", "-q", "40", "-w", str(size[0]), "-h", str(size[1]), "-m", str(mode), "-p", str(spp)]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r1 = subprocess.run([exe] + common + ["-o", one], capture_output=True, text=True, timeout=180, env=env)
    assert r1.returncode == 0, r1.stdout + r1.stderr
    r2 = subprocess.run([exe] + common + ["-g", "2", "-o", two, "-r", one], capture_output=True, text=True, timeout=180, env=env)
    assert r2.returncode == 0, r2.stdout + r2.stderr
    assert "PASSED!" in r2.stdout and "rank 0 of 2 joined" in r2.stdout


@pytest.mark.gpu
@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("kind", ["rt", "raster"])
def test_two_ranks_gathered_through_the_c_abi_equal_the_single_gpu_frame(native, kind):
    """tests/_multi_driver.py under torch.distributed.run: vx_cuda_frame_gather (RT strips) and
    vx_cuda_frame_gather_inplace (raster tile rows) against the frame rank 0 renders alone."""
    import sys
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29641" if kind == "rt" else "29642",
                        os.path.join(ROOT, "tests", "_multi_driver.py"), kind], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "0 words differ" in r.stdout


@pytest.mark.gpu
def test_staged_uploads_keep_their_bytes_when_the_host_runs_far_ahead(native):
    """vx_copy_to_dev of up to 4 MiB copies into a page-locked ring (32 MiB, handed out by the byte, 256 transfers in
    flight) and returns before the transfer ran.  Far more bytes than the ring holds are uploaded back to back behind a
    slow kernel, sizes from 1 byte to the staging limit and just above it: every buffer must read back exactly what
    was passed, i.e. no bytes of the ring are reused before the transfer that reads them has completed."""
    from skybox_rt_b200 import raytrace as rt
    from skybox_rt_b200 import vortex as vx
    rng = np.random.default_rng(7)
    sizes = [1, 3, 255, 256, 257, 4096, 4097, 70000, (4 << 20) - 1, 4 << 20, 3 << 20, 1 << 20]
    sizes += [int(s) for s in rng.integers(1, 600000, 330)] + [4 << 20] * 12 + [int(s) for s in rng.integers(1, 5000, 300)]
    sizes += [(4 << 20) + 4]  # above the staging limit: the blocking path, last because it drains the stream
    nodes, tris = rt.build_bvh2(rt.scene_terrain(120, 100))
    with rt.RayTracer(nodes, tris) as tracer:
        dev = tracer.dev
        # something slow in front of the uploads so that they queue up instead of completing at once
        tracer.render(rt.make_arg(1920, 1080, 2, spp=1, seed=1, scene="terrain"))  # allocates the target and the arguments
        arg = rt.make_arg(1920, 1080, 2, spp=4, seed=1, scene="terrain")
        tracer._bind(arg, 1920 * 1080 * 4)
        tracer.args.write(arg)
        dev.start(tracer.kernel, tracer.args)
        bufs, want = [], []
        for i, n in enumerate(sizes):
            data = rng.integers(0, 256, n, dtype=np.uint8)
            b = dev.alloc(n, vx.VX_MEM_READ)
            b.write(data)
            data_copy = data.copy()
            data[:] = 0xEE  # the caller's buffer is free on return: scribbling on it must not reach the device
            bufs.append(b)
            want.append(data_copy)
        assert sum(sizes) > 3 * (32 << 20)
        for b, w in zip(bufs, want):
            got = np.empty(len(w), np.uint8)
            b.read(got)
            assert np.array_equal(got, w), len(w)
            b.free()
