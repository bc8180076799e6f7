"""Byte layouts of include/skybox_rt.h as numpy dtypes / ctypes structures (host side)."""
import ctypes as C

import numpy as np

# ---- structs of include/skybox_rt.h as numpy dtypes / ctypes -------------------------------
TRIANGLE = np.dtype([("v0", "<f4", 3), ("v1", "<f4", 3), ("v2", "<f4", 3), ("albedo", "<u4")])
NODE = np.dtype([("bmin", "<f4", 3), ("left_first", "<u4"), ("bmax", "<f4", 3), ("count", "<u4")])
TRIREC = np.dtype([("v0", "<f4", 3), ("prim_id", "<u4"), ("e1", "<f4", 3), ("albedo", "<u4"),
                   ("e2", "<f4", 3), ("pad", "<u4")])
RAY = np.dtype([("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4")])
HIT = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim_id", "<u4")])
assert TRIANGLE.itemsize == 40 and NODE.itemsize == 32 and TRIREC.itemsize == 48
assert RAY.itemsize == 32 and HIT.itemsize == 16

MODE_PRIMARY, MODE_SHADOW, MODE_BOUNCE, MODE_RAYS = 0, 1, 2, 3
MISS = 0xFFFFFFFF


class KernelArg(C.Structure):
    _fields_ = [
        ("num_tasks", C.c_uint32), ("dst_width", C.c_uint32), ("dst_height", C.c_uint32),
        ("cbuf_stride", C.c_uint32), ("cbuf_pitch", C.c_uint32), ("mode", C.c_uint32),
        ("spp", C.c_uint32), ("seed", C.c_uint32),
        ("cbuf_addr", C.c_uint64), ("node_addr", C.c_uint64), ("tri_addr", C.c_uint64),
        ("ray_addr", C.c_uint64), ("hit_addr", C.c_uint64), ("num_rays", C.c_uint64),
        ("num_nodes", C.c_uint32), ("num_tris", C.c_uint32),
        ("cam_pos", C.c_float * 3), ("cam_fwd", C.c_float * 3), ("cam_right", C.c_float * 3),
        ("cam_up", C.c_float * 3), ("tan_half_fov", C.c_float),
        ("light_pos", C.c_float * 3), ("ambient", C.c_float), ("sky", C.c_float * 3),
        ("ray_eps", C.c_float),
        ("rank_id", C.c_uint32), ("num_ranks", C.c_uint32), ("rows_per_block", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


class Bvh2(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("tris", C.c_void_p), ("num_nodes", C.c_uint32),
                ("num_tris", C.c_uint32)]


def rank_rows(height, rows_per_block, num_ranks, rank):
    """Rows of the frame owned by `rank` (srt_rank_rows in include/skybox_rt.h)."""
    if num_ranks <= 1:
        return height
    n = 0
    for y in range(0, height, rows_per_block):
        if (y // rows_per_block) % num_ranks == rank:
            n += min(rows_per_block, height - y)
    return n


def rank_row_index(height, rows_per_block, num_ranks, rank):
    """Frame rows owned by `rank`, in strip-buffer order."""
    rows = []
    for y in range(0, height, rows_per_block):
        if num_ranks <= 1 or (y // rows_per_block) % num_ranks == rank:
            rows.extend(range(y, min(y + rows_per_block, height)))
    return np.asarray(rows, dtype=np.int64)


def kernel_image(pipeline=1, version=1):
    """The 32-byte kernel file: vxbin header (min_vma, max_vma) + pipeline descriptor."""
    import struct
    return struct.pack("<QQQII", 0x80000000, 0x80000000 + 64, 0x5452584f42594b53, pipeline, version)
