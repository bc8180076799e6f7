"""Multi-GPU frame assembly: one process per GPU, BVH replicated, image rows interleaved in blocks over
the ranks, ONE gather of the per-rank strip buffers per frame.

The reference is single-device (no collective anywhere in /root/reference); the nearest analogue is its
striping of raster tiles over raster units (sim/simx/raster_unit.cpp:109-111).

Product path: `NcclFrameGather` -- the C entry vx_cuda_frame_gather (csrc/vx_nccl.cu: NCCL send/receive group +
one de-interleave kernel on the launch stream); torch.distributed only carries the 128-byte communicator id to
the ranks at start-up, no torch op runs between vx_start and the host frame.
`FrameGather` is the same assembly written with torch.distributed collectives: it runs on gloo, so the CPU tests
(tests/test_multi_gloo.py) can check the row dealing / de-interleave logic with world_size 2 and 3.
"""
import torch
import torch.distributed as dist

from .formats import rank_row_index, rank_rows


class NcclFrameGather:
    """One communicator per device; gather() enqueues the frame assembly behind the frame's kernels."""

    def __init__(self, dev, width, height, rows_per_block, dst=0, group=None):
        from . import vortex as vx
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        ident = [vx.comm_unique_id() if self.rank == dst else None]
        dist.broadcast_object_list(ident, src=dst, group=group)  # start-up plumbing: 128 bytes
        self.comm = vx.Comm(dev, ident[0], self.rank, self.world)
        self.dev, self.w, self.h, self.rpb, self.dst = dev, width, height, rows_per_block, dst
        self.frame = dev.alloc(height * width * 4) if self.rank == dst else None

    def gather(self, strip_buf):
        """strip_buf: this rank's vx Buffer.  Returns the root's frame Buffer (None elsewhere); asynchronous."""
        self.comm.frame_gather(strip_buf, self.frame, self.w, self.h, self.rpb, self.dst)
        return self.frame

    def close(self):
        self.comm.close()
        if self.frame is not None:
            self.frame.free()
            self.frame = None


class FrameGather:
    """Gathers [rows_r, W] uint32 strips (as int32 tensors) to rank `dst` and de-interleaves them.

    All tensors live on `device`; with NCCL the collective is enqueued on the current CUDA stream, so
    wrapping calls in `torch.cuda.stream(render_stream)` orders it after the kernels vx_start queued.
    """

    def __init__(self, width, height, rows_per_block, device, dst=0, group=None):
        self.w, self.h, self.rpb, self.dst, self.group = width, height, rows_per_block, dst, group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.rows = rank_rows(height, rows_per_block, self.world, self.rank)
        self.max_rows = max(rank_rows(height, rows_per_block, self.world, r) for r in range(self.world))
        self.send = torch.zeros(self.max_rows * width, dtype=torch.int32, device=device)
        self.is_dst = self.rank == dst
        if self.is_dst:
            self.recv = [torch.empty(self.max_rows * width, dtype=torch.int32, device=device)
                         for _ in range(self.world)]
            self.row_idx = [torch.as_tensor(rank_row_index(height, rows_per_block, self.world, r), device=device)
                            for r in range(self.world)]
            self.frame = torch.empty((height, width), dtype=torch.int32, device=device)
        else:
            self.recv, self.row_idx, self.frame = None, None, None

    def gather(self, strip):
        """strip: this rank's [rows, W] int32 tensor (or flat).  Returns the frame on dst, None elsewhere."""
        n = self.rows * self.w
        if n:
            self.send[:n].copy_(strip.reshape(-1)[:n], non_blocking=True)
        dist.gather(self.send, self.recv, dst=self.dst, group=self.group)
        if not self.is_dst:
            return None
        for r in range(self.world):
            k = len(self.row_idx[r])
            if k:
                self.frame.index_copy_(0, self.row_idx[r], self.recv[r][:k * self.w].view(k, self.w))
        return self.frame
