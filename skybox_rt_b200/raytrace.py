"""Host side of the `raytrace` regression app, in Python, over the vx_* C ABI.

Follows the host flow of the reference's graphics apps step for step
(tests/regression/draw3d/main.cpp:392-517): open device -> check caps -> build the flat device
buffers on the host (here: soup -> BVH2, where draw3d runs graphics::Binning, main.cpp:190) ->
vx_mem_alloc + vx_copy_to_dev each buffer (:198-211) -> fill kernel_arg_t with device addresses ->
vx_upload_bytes(kernel_arg) (:346) -> vx_start (:353) -> vx_ready_wait (:357) -> vx_copy_from_dev
of the colour buffer (:383).  The C twin of this file is apps/raytrace/main.c.

The compute is the CUDA back-end behind libvortex.so; the host pre-pass is libskybox_rt_host.so.
Nothing here falls back to a CPU implementation.
"""
import ctypes as C
import os

import numpy as np

from . import vortex as vx
from .formats import (HIT, MODE_RAYS, NODE, RAY, TRIANGLE, TRIREC, Bvh2, KernelArg, kernel_image, rank_rows)

PKG = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(PKG, "lib", "libskybox_rt_host.so")
HOST_EXPORTS = ["srt_bvh2_build", "srt_bvh2_free", "srt_scene_cornell", "srt_scene_terrain",
                "srt_camera_cornell", "srt_camera_terrain", "srt_rays_random"]

_host = None


def hostlib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise vx.VortexError(f"{HOST_LIB_PATH} is missing: run `python -m skybox_rt_b200.build`")
        L = C.CDLL(HOST_LIB_PATH)
        L.srt_bvh2_build.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(Bvh2)]
        L.srt_bvh2_free.argtypes = [C.POINTER(Bvh2)]
        L.srt_scene_cornell.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
        L.srt_scene_terrain.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p),
                                        C.POINTER(C.c_uint32)]
        L.srt_camera_cornell.argtypes = [C.POINTER(KernelArg)]
        L.srt_camera_terrain.argtypes = [C.POINTER(KernelArg)]
        L.srt_rays_random.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p]
        L.srt_rays_random.restype = None
        L._free = C.CDLL(None).free
        L._free.argtypes = [C.c_void_p]
        _host = L
    return _host


def _view(ptr, count, dtype):
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


def scene_cornell():
    p, n = C.c_void_p(), C.c_uint32()
    if hostlib().srt_scene_cornell(C.byref(p), C.byref(n)) != 0:
        raise vx.VortexError("srt_scene_cornell failed")
    out = _view(p.value, n.value, TRIANGLE).copy()
    hostlib()._free(p)
    return out


def scene_terrain(quads_x, quads_y, seed=0x5EED5EED):
    p, n = C.c_void_p(), C.c_uint32()
    if hostlib().srt_scene_terrain(quads_x, quads_y, seed, C.byref(p), C.byref(n)) != 0:
        raise vx.VortexError("srt_scene_terrain failed")
    out = _view(p.value, n.value, TRIANGLE).copy()
    hostlib()._free(p)
    return out


def build_bvh2(soup):
    """soup (TRIANGLE records) -> (nodes, tris): the flat buffers the device consumes."""
    soup = np.ascontiguousarray(soup, dtype=TRIANGLE)
    b = Bvh2()
    if hostlib().srt_bvh2_build(soup.ctypes.data, len(soup), C.byref(b)) != 0:
        raise vx.VortexError("srt_bvh2_build failed")
    nodes = _view(b.nodes, b.num_nodes, NODE).copy()
    tris = _view(b.tris, b.num_tris, TRIREC).copy()
    hostlib().srt_bvh2_free(C.byref(b))
    return nodes, tris


def rays_random(lo, hi, seed, first, count, out=None):
    lo = np.ascontiguousarray(lo, dtype=np.float32)
    hi = np.ascontiguousarray(hi, dtype=np.float32)
    rays = out if out is not None else np.zeros(count, dtype=RAY)
    hostlib().srt_rays_random(lo.ctypes.data, hi.ctypes.data, seed, first, count, rays.ctypes.data)
    return rays


def make_arg(width, height, mode, spp=1, seed=1234, scene="cornell", rank=0, num_ranks=1, rows_per_block=4):
    a = KernelArg()
    a.num_tasks = height
    a.dst_width, a.dst_height = width, height
    a.cbuf_stride, a.cbuf_pitch = 4, width * 4
    a.mode, a.spp, a.seed = mode, spp, seed
    a.rank_id, a.num_ranks, a.rows_per_block = rank, num_ranks, rows_per_block
    if scene == "cornell":
        hostlib().srt_camera_cornell(C.byref(a))
    else:
        hostlib().srt_camera_terrain(C.byref(a))
    return a


def build_lbvh_noalloc(dev, soup_buf, num_triangles, node_buf, tri_buf):
    """vx_cuda_rt_build_lbvh_noalloc (include/vortex_cuda.h): the device BVH build into buffers the caller owns, with
    the builder's scratch in a fixed array inside libvortex.so -- no allocation in the call.  No counterpart in the
    reference.  Returns (num_nodes, num_tris); raises VortexError (code -3 above VX_CUDA_LBVH_NOALLOC_MAX_TRIS)."""
    info = (C.c_uint32 * 2)()
    rc = vx.lib().vx_cuda_rt_build_lbvh_noalloc(dev.handle, soup_buf.handle, num_triangles, node_buf.handle,
                                                tri_buf.handle, info)
    if rc != 0:
        raise vx.VortexError(f"vx_cuda_rt_build_lbvh_noalloc failed with code {rc}")
    return info[0], info[1]


class RayTracer:
    """One opened device with one uploaded scene."""

    def __init__(self, nodes=None, tris=None, device=None, soup=None):
        """Either the host builder's (nodes, tris) arrays, or `soup` alone: the triangle soup is uploaded and the
        BVH is built on the device (vx_cuda_rt_build_lbvh)."""
        self.dev = device or vx.Device()
        self._own_dev = device is None
        isa = self.dev.caps(vx.VX_CAPS_ISA_FLAGS)
        if not isa & vx.VX_ISA_EXT_RT:
            raise vx.VortexError("RT extension not supported by the device")  # draw3d/main.cpp:400-417
        self.kernel = self.dev.upload_kernel(kernel_image())
        if soup is not None:
            soup = np.ascontiguousarray(soup, dtype=TRIANGLE)
            sbuf = self.dev.upload(soup)
            nb, tb, info = C.c_void_p(), C.c_void_p(), (C.c_uint32 * 2)()
            rc = vx.lib().vx_cuda_rt_build_lbvh(self.dev.handle, sbuf.handle, len(soup), C.byref(nb), C.byref(tb), info)
            sbuf.free()
            if rc != 0:
                raise vx.VortexError(f"vx_cuda_rt_build_lbvh failed with code {rc}")
            self.num_nodes, self.num_tris = info[0], info[1]
            self.node_buf = vx.Buffer(self.dev, nb, self.num_nodes * NODE.itemsize)
            self.tri_buf = vx.Buffer(self.dev, tb, self.num_tris * TRIREC.itemsize)
        else:
            nodes = np.ascontiguousarray(nodes, dtype=NODE)
            tris = np.ascontiguousarray(tris, dtype=TRIREC)
            self.num_nodes, self.num_tris = len(nodes), len(tris)
            self.node_buf = self.dev.alloc(nodes.nbytes, vx.VX_MEM_READ)
            self.node_buf.write(nodes)
            self.tri_buf = self.dev.alloc(tris.nbytes, vx.VX_MEM_READ)
            self.tri_buf.write(tris)
        self.cbuf = None
        self.args = None
        self.ray_buf = self.hit_buf = None

    def bvh(self):
        """The BVH2 the device holds, read back: (nodes, tris) structured arrays."""
        nodes, tris = np.empty(self.num_nodes, NODE), np.empty(self.num_tris, TRIREC)
        self.node_buf.read(nodes)
        self.tri_buf.read(tris)
        return nodes, tris

    def close(self):
        for b in (self.kernel, self.node_buf, self.tri_buf, self.cbuf, self.args, self.ray_buf, self.hit_buf):
            if b is not None:
                b.free()
        self.kernel = self.node_buf = self.tri_buf = self.cbuf = self.args = self.ray_buf = self.hit_buf = None
        if self._own_dev:
            self.dev.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- frame rendering ---------------------------------------------------------------------
    def _bind(self, arg, nbytes_cbuf):
        if self.cbuf is None or self.cbuf.size < nbytes_cbuf:
            if self.cbuf is not None:
                self.cbuf.free()
            self.cbuf = self.dev.alloc(nbytes_cbuf, vx.VX_MEM_READ_WRITE)
        arg.cbuf_addr = self.cbuf.address
        arg.node_addr, arg.tri_addr = self.node_buf.address, self.tri_buf.address
        arg.num_nodes, arg.num_tris = self.num_nodes, self.num_tris

    def _upload_args(self, arg):
        if self.args is None:
            self.args = self.dev.upload(arg)
        else:
            self.args.write(arg)

    def start(self, arg):
        """Asynchronous: upload kernel_arg_t and vx_start."""
        rows = rank_rows(arg.dst_height, arg.rows_per_block, arg.num_ranks, arg.rank_id)
        self._bind(arg, max(rows, 1) * arg.cbuf_pitch)
        self._upload_args(arg)
        self.dev.start(self.kernel, self.args)
        return rows

    def render(self, arg, out=None):
        """Synchronous frame: returns this rank's strip as uint32 [rows, W] (row 0 = bottom)."""
        rows = self.start(arg)
        self.dev.ready_wait()
        fb = out if out is not None else np.empty((rows, arg.dst_width), dtype=np.uint32)
        if rows:
            self.cbuf.read(fb)
        return fb

    # -- explicit rays ---------------------------------------------------------------------------
    def trace(self, rays, out=None):
        rays = np.ascontiguousarray(rays, dtype=RAY)
        n = len(rays)
        hits = out if out is not None else np.empty(n, dtype=HIT)
        if n == 0:
            return hits
        if self.ray_buf is None or self.ray_buf.size < rays.nbytes:
            for b in (self.ray_buf, self.hit_buf):
                if b is not None:
                    b.free()
            self.ray_buf = self.dev.alloc(rays.nbytes, vx.VX_MEM_READ)
            self.hit_buf = self.dev.alloc(n * HIT.itemsize, vx.VX_MEM_READ_WRITE)
        self.ray_buf.write(rays)
        a = KernelArg()
        a.num_tasks, a.mode = 1, MODE_RAYS
        a.ray_addr, a.hit_addr, a.num_rays = self.ray_buf.address, self.hit_buf.address, n
        a.node_addr, a.tri_addr = self.node_buf.address, self.tri_buf.address
        a.num_nodes, a.num_tris = self.num_nodes, self.num_tris
        self._upload_args(a)
        self.dev.start(self.kernel, self.args)
        self.dev.ready_wait()
        self.hit_buf.read(hits)
        return hits

    def counters(self):
        q = self.dev.mpm_query
        return {
            "rays_primary": q(vx.VX_CSR_RT_RAYS_PRIMARY), "rays_shadow": q(vx.VX_CSR_RT_RAYS_SHADOW),
            "rays_bounce": q(vx.VX_CSR_RT_RAYS_BOUNCE), "nodes": q(vx.VX_CSR_RT_NODES),
            "tris": q(vx.VX_CSR_RT_TRIS), "elapsed_ns": q(vx.VX_CSR_RT_ELAPSED_NS),
        }


def assemble_strips(strips, height, rows_per_block, num_ranks):
    """De-interleave per-rank strip buffers (list of [rows_r, W]) into the full frame."""
    from .formats import rank_row_index
    w = strips[0].shape[1]
    frame = np.empty((height, w), dtype=np.uint32)
    for r, s in enumerate(strips):
        idx = rank_row_index(height, rows_per_block, num_ranks, r)
        frame[idx] = s[:len(idx)]
    return frame


def save_png(path, fb):
    """Bottom-up save, like SaveImage(..., bits + (H-1)*pitch, W, H, -pitch) at draw3d/main.cpp:383-386."""
    from PIL import Image
    h, w = fb.shape
    rgba = np.empty((h, w, 4), dtype=np.uint8)
    rgba[..., 0] = (fb >> 16) & 255
    rgba[..., 1] = (fb >> 8) & 255
    rgba[..., 2] = fb & 255
    rgba[..., 3] = (fb >> 24) & 255
    Image.fromarray(rgba[::-1].copy(), "RGBA").save(path)
