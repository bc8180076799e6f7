"""ctypes binding of the CPU oracle (oracle/libref_rt.so).  Test infrastructure only.

Builds the library on first use with oracle/Makefile (gcc -O2 -ffp-contract=off).
"""
import ctypes as C
import os
import subprocess

import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from skybox_rt_b200.formats import (HIT, MISS, MODE_BOUNCE, MODE_PRIMARY, MODE_RAYS, MODE_SHADOW, NODE, RAY,  # noqa: E402,F401
                                   TRIANGLE, TRIREC, Bvh2, KernelArg, rank_rows)


class Counters(C.Structure):
    _fields_ = [("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_bounce", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)
        _lib = C.CDLL(os.path.join(ORACLE_DIR, "libref_rt.so"))
        _lib.ort_render.argtypes = [C.POINTER(KernelArg), C.c_uint32, C.POINTER(Counters)]
        _lib.ort_bvh2_build.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(Bvh2)]
        _lib.ort_bvh2_free.argtypes = [C.POINTER(Bvh2)]
        _lib.ort_scene_cornell.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
        _lib.ort_scene_terrain.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_uint32)]
        _lib.ort_camera_cornell.argtypes = [C.POINTER(KernelArg)]
        _lib.ort_camera_terrain.argtypes = [C.POINTER(KernelArg)]
        _lib.ort_rays_random.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p]
        _lib.ort_closest_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.ort_any_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.ort_any_hit.restype = C.c_int
        _lib.ort_closest_hit_brute.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        _lib.ort_any_hit_brute.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        _lib.ort_any_hit_brute.restype = C.c_int
        _lib.free = C.CDLL(None).free
        _lib.free.argtypes = [C.c_void_p]
    return _lib


def _take(ptr, count, dtype):
    """Copy `count` records from a malloc'ed C array into numpy and free the C array."""
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=count).copy()
    lib().free(ptr)
    return arr


def scene_cornell():
    p, n = C.c_void_p(), C.c_uint32()
    assert lib().ort_scene_cornell(C.byref(p), C.byref(n)) == 0
    return _take(p.value, n.value, TRIANGLE)


def scene_terrain(qx, qy, seed=0x5EED5EED):
    p, n = C.c_void_p(), C.c_uint32()
    assert lib().ort_scene_terrain(qx, qy, seed, C.byref(p), C.byref(n)) == 0
    return _take(p.value, n.value, TRIANGLE)


def bvh2_build(soup):
    soup = np.ascontiguousarray(soup, dtype=TRIANGLE)
    b = Bvh2()
    assert lib().ort_bvh2_build(soup.ctypes.data, len(soup), C.byref(b)) == 0
    nodes = _take(b.nodes, b.num_nodes, NODE)
    tris = _take(b.tris, b.num_tris, TRIREC)
    return nodes, tris


def make_arg(width, height, mode, spp=1, seed=1234, scene="cornell", num_tasks=None):
    a = KernelArg()
    a.num_tasks = num_tasks or height
    a.dst_width, a.dst_height = width, height
    a.cbuf_stride, a.cbuf_pitch = 4, width * 4
    a.mode, a.spp, a.seed = mode, spp, seed
    a.num_ranks, a.rank_id, a.rows_per_block = 1, 0, 1
    if scene == "cornell":
        lib().ort_camera_cornell(C.byref(a))
    else:
        lib().ort_camera_terrain(C.byref(a))
    return a


def render(arg, nodes, tris, threads=None):
    """Run the oracle kernel; returns (framebuffer uint32 [rows, W], counters dict)."""
    threads = threads or min(os.cpu_count() or 1, 64)
    rows = rank_rows(arg.dst_height, arg.rows_per_block, arg.num_ranks, arg.rank_id)
    fb = np.zeros((rows, arg.dst_width), dtype=np.uint32)
    arg.cbuf_addr = fb.ctypes.data
    arg.node_addr, arg.tri_addr = nodes.ctypes.data, tris.ctypes.data
    arg.num_nodes, arg.num_tris = len(nodes), len(tris)
    ctr = Counters()
    assert lib().ort_render(C.byref(arg), threads, C.byref(ctr)) == 0
    return fb, ctr.as_dict()


def trace(rays, nodes, tris, threads=None, num_tasks=256):
    """Closest hit for an explicit ray list (mode 3); returns (hits, counters)."""
    threads = threads or min(os.cpu_count() or 1, 64)
    rays = np.ascontiguousarray(rays, dtype=RAY)
    hits = np.zeros(len(rays), dtype=HIT)
    a = KernelArg()
    a.num_tasks = max(1, min(num_tasks, len(rays)))
    a.mode = MODE_RAYS
    a.ray_addr, a.hit_addr, a.num_rays = rays.ctypes.data, hits.ctypes.data, len(rays)
    a.node_addr, a.tri_addr = nodes.ctypes.data, tris.ctypes.data
    a.num_nodes, a.num_tris = len(nodes), len(tris)
    ctr = Counters()
    assert lib().ort_render(C.byref(a), threads, C.byref(ctr)) == 0
    return hits, ctr.as_dict()


def rays_random(lo, hi, seed, first, count):
    lo = np.asarray(lo, dtype=np.float32)
    hi = np.asarray(hi, dtype=np.float32)
    rays = np.zeros(count, dtype=RAY)
    lib().ort_rays_random(lo.ctypes.data, hi.ctypes.data, seed, first, count, rays.ctypes.data)
    return rays


def closest_brute(rays, tris):
    rays = np.ascontiguousarray(rays, dtype=RAY)
    hits = np.zeros(len(rays), dtype=HIT)
    for i in range(len(rays)):
        lib().ort_closest_hit_brute(tris.ctypes.data, len(tris), rays[i:i + 1].ctypes.data,
                                    hits[i:i + 1].ctypes.data)
    return hits


def any_hit(rays, nodes, tris):
    rays = np.ascontiguousarray(rays, dtype=RAY)
    return np.array([lib().ort_any_hit(nodes.ctypes.data, tris.ctypes.data, rays[i:i + 1].ctypes.data, None)
                     for i in range(len(rays))], dtype=np.int32)


def any_hit_brute(rays, tris):
    rays = np.ascontiguousarray(rays, dtype=RAY)
    return np.array([lib().ort_any_hit_brute(tris.ctypes.data, len(tris), rays[i:i + 1].ctypes.data)
                     for i in range(len(rays))], dtype=np.int32)


def save_png(path, fb):
    """Save an A8R8G8B8 framebuffer bottom-up, like draw3d/main.cpp:383-386 (negative pitch)."""
    from PIL import Image
    h, w = fb.shape
    rgba = np.zeros((h, w, 4), dtype=np.uint8)
    rgba[..., 0] = (fb >> 16) & 255
    rgba[..., 1] = (fb >> 8) & 255
    rgba[..., 2] = fb & 255
    rgba[..., 3] = (fb >> 24) & 255
    Image.fromarray(rgba[::-1].copy(), "RGBA").save(path)
