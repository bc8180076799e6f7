"""world_size-2 (and 3) CPU test of the N>1 path's host logic over gloo: each rank renders ITS rows of the
frame (with the CPU oracle standing in for the GPU kernel -- test infrastructure), the product's
FrameGather assembles them on rank 0, and the result must equal the single-rank frame bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, w, h, rpb, out_path):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import _oracle as o
    from skybox_rt_b200.multi import FrameGather
    nodes, tris = o.bvh2_build(o.scene_cornell())
    a = o.make_arg(w, h, 2, spp=2, seed=77)
    a.num_ranks, a.rank_id, a.rows_per_block = world, rank, rpb
    strip, _ = o.render(a, nodes, tris, threads=1)
    g = FrameGather(w, h, rpb, torch.device("cpu"))
    assert g.rows == strip.shape[0]
    frame = g.gather(torch.from_numpy(strip.view(np.int32)))
    if rank == 0:
        np.save(out_path, frame.numpy().view(np.uint32))
    else:
        assert frame is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,rpb,port", [(2, 40, 30, 4, 29531), (3, 33, 17, 8, 29532), (2, 16, 3, 8, 29533)])
def test_strips_gathered_over_gloo_equal_single_rank_frame(tmp_path, world, w, h, rpb, port):
    import _oracle as o
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world, port, w, h, rpb, out), nprocs=world, join=True)
    nodes, tris = o.bvh2_build(o.scene_cornell())
    full, _ = o.render(o.make_arg(w, h, 2, spp=2, seed=77), nodes, tris, threads=1)
    assert np.array_equal(np.load(out), full)


def _raster_worker(rank, world, port, w, h, out_path):
    """Raster path, N>1 host logic: the product's per-rank binning decides which tile rows a rank draws; the
    oracle stands in for the GPU kernel (it renders the frame, the rank keeps only the rows its own tile
    list covers); FrameGather with 32-row blocks puts the strips together on rank 0."""
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import _oracle_draw3d as od
    from skybox_rt_b200 import draw3d as d3
    from skybox_rt_b200.formats import rank_row_index
    from skybox_rt_b200.multi import FrameGather
    blob = od.trace_blob("vase")
    t = d3.Trace(blob=blob)
    full, _, _ = od.render(od.oracle(), blob, w, h, threads=1)
    mine = np.full_like(full, 0xFF000000)  # cleared target
    for d in range(t.num_drawcalls):
        n, tb, _, _ = t.binning(d, w, h, fmt=d3.TILEBUF_WIDE, rank=rank, num_ranks=world)
        hdr = np.frombuffer(tb[:n * 16], np.uint32).reshape(n, 4)
        for ty in set(int(x >> 16) for x in hdr[:, 0]):
            assert ty % world == rank
            mine[ty * 32:(ty + 1) * 32] = full[ty * 32:(ty + 1) * 32]
    rows = rank_row_index(h, 32, world, rank)
    g = FrameGather(w, h, 32, torch.device("cpu"))
    frame = g.gather(torch.from_numpy(mine.view(np.int32))[torch.from_numpy(rows)])
    if rank == 0:
        np.save(out_path, frame.numpy().view(np.uint32))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,port", [(2, 160, 128, 29541), (3, 96, 200, 29542)])
def test_raster_tile_rows_gathered_over_gloo(tmp_path, native, world, w, h, port):
    import _oracle_draw3d as od
    out = str(tmp_path / "frame.npy")
    mp.spawn(_raster_worker, args=(world, port, w, h, out), nprocs=world, join=True)
    full, _, _ = od.render(od.oracle(), od.trace_blob("vase"), w, h, threads=1)
    assert np.array_equal(np.load(out), full)


def test_weak_scaling_frames_keep_the_work_per_rank():
    """bench.py's weak scaling: N x the pixels of the 1-GPU frame, the same rows for every rank."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    from skybox_rt_b200.formats import rank_rows
    for n in (2, 3, 4, 8):
        w, h = bench.weak_frame(1920, 1080, n)
        rows = [rank_rows(h, bench.ROWS_PER_BLOCK, n, r) for r in range(n)]
        assert len(set(rows)) == 1 and sum(rows) == h
        assert abs(w * h / (1920 * 1080 * n) - 1) < 0.005 and w % 8 == 0
        assert abs((w / h) / (1920 / 1080) - 1) < 0.03
