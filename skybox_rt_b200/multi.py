"""Multi-GPU frame assembly: one process per GPU, BVH replicated, image rows interleaved in blocks over
the ranks, ONE gather of the per-rank strip buffers per frame.

The reference is single-device (no collective anywhere in /root/reference); the nearest analogue is its
striping of raster tiles over raster units (sim/simx/raster_unit.cpp:109-111).  torch.distributed is
plumbing only: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
import torch
import torch.distributed as dist

from .formats import rank_row_index, rank_rows


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
