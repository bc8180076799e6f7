"""ctypes binding of the raster (draw3d) path: the C host flow in apps/draw3d/draw3d_app.c and the C
pre-pass in csrc/host/draw3d_{trace,binning}.c.  Mirrors the reference app
(/root/reference tests/regression/draw3d/main.cpp): `Draw3D.render(trace)` is its render() loop.

No fallback: the libraries must have been built (`python -m skybox_rt_b200.build`); rendering needs a
CUDA device because `vx_dev_open` does.
"""
import ctypes as C
import gzip
import os

import numpy as np

from . import vortex as vx

PKG = os.path.dirname(os.path.abspath(__file__))
APP_LIB = os.path.join(PKG, "lib", "libdraw3d_app.so")
HOST_LIB = os.path.join(PKG, "lib", "libskybox_rt_host.so")

SW_RAST, SW_OM, SW_TEX, WIDE_TILES, GPU_BINNING, DEVICE_CLEAR = 1, 2, 4, 8, 16, 32
AUTO = 64  # device clear / device binning where they pay, no wait between the draws of a frame (draw3d_app.h)
TILEBUF_REFERENCE, TILEBUF_WIDE = 0, 1
VX_CSR_RASTER_FRAGMENTS, VX_CSR_RASTER_QUAD_TESTS, VX_CSR_RASTER_TILES = 0x7C8, 0x7C9, 0x7CA
VX_DCR_RASTER_EXT_TBUF_FORMAT = 0x804
VX_DCR_RASTER_EXT_TILE_LOGSIZE = 0x80D  # tile size the buffers were binned with; vx_start refuses anything but 5


class Binning(C.Structure):  # srt_binning_t (include/skybox_draw3d.h)
    _fields_ = [("tilebuf", C.c_void_p), ("tilebuf_size", C.c_uint64), ("primbuf", C.c_void_p),
                ("primbuf_size", C.c_uint64), ("num_tiles", C.c_uint32), ("num_prims", C.c_uint32),
                ("num_tile_prims", C.c_uint64), ("format", C.c_uint32)]


class Stats(C.Structure):  # srt_draw3d_stats_t (apps/draw3d/draw3d_app.h)
    _fields_ = [("draws_submitted", C.c_uint32), ("prims", C.c_uint64), ("tiles", C.c_uint64),
                ("tile_prims", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("bin_ms", C.c_double), ("upload_ms", C.c_double), ("run_ms", C.c_double)]


class BinningInfo(C.Structure):  # srt_binning_info_t (include/vortex_cuda.h)
    _fields_ = [("num_tiles", C.c_uint32), ("num_prims", C.c_uint32), ("num_tile_prims", C.c_uint64),
                ("tilebuf_size", C.c_uint64), ("primbuf_size", C.c_uint64), ("format", C.c_uint32)]


class Drawcall(C.Structure):  # srt_drawcall_t
    _fields_ = [("states", C.c_uint32 * 24), ("texture_envcolor", C.c_float * 4), ("texture_id", C.c_uint32),
                ("viewport", C.c_int32 * 4), ("near", C.c_float), ("far", C.c_float), ("num_vertices", C.c_uint32),
                ("num_primitives", C.c_uint32), ("vertices", C.c_void_p), ("primitives", C.c_void_p)]


class TraceStruct(C.Structure):  # srt_trace_t
    _fields_ = [("num_drawcalls", C.c_uint32), ("num_textures", C.c_uint32), ("drawcalls", C.c_void_p),
                ("textures", C.c_void_p), ("blob", C.c_void_p), ("blob_size", C.c_uint64)]


_host = None
_app = None


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB):
            raise vx.VortexError(f"{HOST_LIB} is missing: run `python -m skybox_rt_b200.build`")
        L = C.CDLL(HOST_LIB)
        L.srt_trace_from_blob.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(TraceStruct)]
        L.srt_trace_load.argtypes = [C.c_char_p, C.POINTER(TraceStruct)]
        L.srt_trace_free.argtypes = [C.POINTER(TraceStruct)]
        L.srt_trace_free.restype = None
        L.srt_draw3d_binning_rank.argtypes = [C.c_void_p] + [C.c_uint32] * 6 + [C.POINTER(Binning)]
        L.srt_binning_free.argtypes = [C.POINTER(Binning)]
        L.srt_binning_free.restype = None
        L.srt_drawcall_at.argtypes = [C.POINTER(TraceStruct), C.c_uint32]
        L.srt_drawcall_at.restype = C.c_void_p
        _host = L
    return _host


def app_lib():
    global _app
    if _app is None:
        if not os.path.exists(APP_LIB):
            raise vx.VortexError(f"{APP_LIB} is missing: run `python -m skybox_rt_b200.build`")
        vx.lib()  # libvortex.so first (libdraw3d_app.so links against it)
        host_lib()
        L = C.CDLL(APP_LIB)
        vp, u32 = C.c_void_p, C.c_uint32
        L.srt_draw3d_open.argtypes = [vp, u32, u32, u32, u32, u32, C.POINTER(vp)]
        L.srt_draw3d_close.argtypes = [vp]
        L.srt_draw3d_close.restype = None
        L.srt_draw3d_clear.argtypes = [vp]
        L.srt_draw3d_set_om_override.argtypes = [vp, C.POINTER(u32)]
        L.srt_draw3d_set_om_override.restype = None
        L.srt_draw3d_render.argtypes = [vp, C.POINTER(TraceStruct), u32, u32]
        L.srt_draw3d_prepare.argtypes = [vp, C.POINTER(TraceStruct)]
        L.srt_draw3d_submit.argtypes = [vp]
        L.srt_draw3d_read.argtypes = [vp, vp, vp]
        L.srt_draw3d_read_async.argtypes = [vp, vp, C.POINTER(C.c_uint64)]
        L.srt_draw3d_read_wait.argtypes = [vp, C.c_uint64]
        L.srt_draw3d_stats.argtypes = [vp]
        L.srt_draw3d_stats.restype = C.POINTER(Stats)
        L.srt_draw3d_color_buffer.argtypes = [vp]
        L.srt_draw3d_color_buffer.restype = vp
        L.srt_draw3d_depth_buffer.argtypes = [vp]
        L.srt_draw3d_depth_buffer.restype = vp
        _app = L
    return _app


class Trace:
    """A draw-call trace (SKBXTRC1 blob, plain or gzip'ed) parsed by the C loader."""

    def __init__(self, path=None, blob=None):
        self.s = TraceStruct()
        if blob is None and str(path).endswith(".cgltrace"):  # the reference's Boost-XML archive, read in C
            rc = host_lib().srt_trace_load(str(path).encode(), C.byref(self.s))
            if rc != 0:
                raise vx.VortexError(f"srt_trace_load({path}) failed with code {rc}")
            self.blob = C.string_at(self.s.blob, self.s.blob_size)
            return
        if blob is None:
            with (gzip.open(path, "rb") if str(path).endswith(".gz") else open(path, "rb")) as f:
                blob = f.read()
        self.blob = bytes(blob)
        rc = host_lib().srt_trace_from_blob(self.blob, len(self.blob), C.byref(self.s))
        if rc != 0:
            raise vx.VortexError(f"srt_trace_from_blob failed with code {rc}")

    @property
    def num_drawcalls(self):
        return self.s.num_drawcalls

    def draw_arrays(self, draw):
        """-> (vertices f32 [nv, 10], primitives u32 [np, 3], near, far) of one draw call (copies)."""
        p = host_lib().srt_drawcall_at(C.byref(self.s), draw)
        if not p:
            raise IndexError(draw)
        dc = Drawcall.from_address(p)
        nv, npr = dc.num_vertices, dc.num_primitives
        v = np.frombuffer(C.string_at(dc.vertices, nv * 40), np.float32).reshape(nv, 10) if nv else np.zeros((0, 10), np.float32)
        t = np.frombuffer(C.string_at(dc.primitives, npr * 12), np.uint32).reshape(npr, 3) if npr else np.zeros((0, 3), np.uint32)
        return v, t, dc.near, dc.far

    def gpu_binning(self, dev, draw, width, height, fmt=TILEBUF_REFERENCE, rank=0, num_ranks=1, tile_logsize=5):
        """vx_cuda_draw3d_binning on `dev` -> same tuple as binning(), read back from the device buffers."""
        return gpu_binning(dev, *self.draw_arrays(draw), width, height, fmt, rank, num_ranks, tile_logsize)

    def binning(self, draw, width, height, fmt=TILEBUF_REFERENCE, rank=0, num_ranks=1, tile_logsize=5):
        """srt_draw3d_binning_rank -> (num_tiles, tilebuf bytes, primbuf bytes, num_tile_prims)."""
        b = Binning()
        dc = host_lib().srt_drawcall_at(C.byref(self.s), draw)
        rc = host_lib().srt_draw3d_binning_rank(dc, width, height, tile_logsize, fmt, rank, num_ranks, C.byref(b))
        if rc != 0:
            raise vx.VortexError(f"srt_draw3d_binning failed with code {rc}")
        tb = C.string_at(b.tilebuf, b.tilebuf_size) if b.tilebuf_size else b""
        pb = C.string_at(b.primbuf, b.primbuf_size) if b.primbuf_size else b""
        out = (b.num_tiles, tb, pb, b.num_tile_prims)
        host_lib().srt_binning_free(C.byref(b))
        return out

    def close(self):
        if self.s.blob:
            host_lib().srt_trace_free(C.byref(self.s))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def gpu_binning(dev, vertices, primitives, near, far, width, height, fmt=TILEBUF_REFERENCE, rank=0, num_ranks=1,
                tile_logsize=5):
    """Triangle setup + tile binning on the device (csrc/raster_binning.cu) from host arrays; returns
    (num_tiles, tilebuf bytes, primbuf bytes, num_tile_prims) like Trace.binning()."""
    vertices = np.ascontiguousarray(vertices, np.float32)
    primitives = np.ascontiguousarray(primitives, np.uint32)
    nv, npr = len(vertices), len(primitives)
    vb = dev.upload(vertices) if nv else None
    pb = dev.upload(primitives) if npr else None
    tb, prb, info = C.c_void_p(), C.c_void_p(), BinningInfo()
    try:
        rc = vx.lib().vx_cuda_draw3d_binning(dev.handle, vb.handle if vb else None, nv, pb.handle if pb else None, npr,
                                             near, far, width, height, tile_logsize, fmt, rank, num_ranks,
                                             C.byref(tb), C.byref(prb), C.byref(info))
        if rc != 0:
            raise vx.VortexError(f"vx_cuda_draw3d_binning failed with code {rc}")
        tile_bytes = np.zeros(info.tilebuf_size, np.uint8)
        prim_bytes = np.zeros(info.primbuf_size, np.uint8)
        if tb:
            vx.Buffer(dev, tb, info.tilebuf_size).read(tile_bytes)
        if prb:
            vx.Buffer(dev, prb, info.primbuf_size).read(prim_bytes)
        return info.num_tiles, tile_bytes.tobytes(), prim_bytes.tobytes(), info.num_tile_prims
    finally:
        for h in (tb, prb):
            if h:
                vx.lib().vx_mem_free(h)
        for b in (vb, pb):
            if b is not None:
                b.free()


class Draw3D:
    """The draw3d app's host object: open -> clear -> render / (prepare, submit) -> read."""

    def __init__(self, width, height, flags=SW_RAST | SW_OM | SW_TEX, rank=0, num_ranks=1, device=None):
        self.own_dev = device is None
        self.dev = device or vx.Device()
        self.width, self.height = width, height
        h = C.c_void_p()
        rc = app_lib().srt_draw3d_open(self.dev.handle, width, height, flags, rank, num_ranks, C.byref(h))
        if rc != 0:
            if self.own_dev:
                self.dev.close()
            raise vx.VortexError(f"srt_draw3d_open failed with code {rc}")
        self.h = h

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        if self.h is not None:
            app_lib().srt_draw3d_close(self.h)
            self.h = None
            if self.own_dev:
                self.dev.close()

    def _ck(self, rc, what):
        if rc != 0:
            raise vx.VortexError(f"{what} failed with code {rc}")

    def set_om_override(self, ov):
        """(blend mode, blend func, blend constant, logic op) DCR values for blending draws, or None
        (srt_draw3d_set_om_override: output-merger state a .cgltrace cannot express)."""
        app_lib().srt_draw3d_set_om_override(self.h, (C.c_uint32 * 4)(*ov) if ov is not None else None)

    def clear(self):
        self._ck(app_lib().srt_draw3d_clear(self.h), "srt_draw3d_clear")

    def render(self, trace, start_draw=0, end_draw=0xFFFFFFFF):
        self._ck(app_lib().srt_draw3d_render(self.h, C.byref(trace.s), start_draw, end_draw), "srt_draw3d_render")

    def prepare(self, trace):
        self._ck(app_lib().srt_draw3d_prepare(self.h, C.byref(trace.s)), "srt_draw3d_prepare")

    def submit(self):
        self._ck(app_lib().srt_draw3d_submit(self.h), "srt_draw3d_submit")

    def wait(self):
        self.dev.ready_wait()

    def read(self, depth=False, out=None):
        """-> colour [H, W] uint32 (row 0 = framebuffer row 0), and the depth/stencil words if asked."""
        c = out if out is not None else np.empty((self.height, self.width), np.uint32)
        z = np.empty((self.height, self.width), np.uint32) if depth else None
        cp = c.data_ptr() if hasattr(c, "data_ptr") else c.ctypes.data
        self._ck(app_lib().srt_draw3d_read(self.h, cp, z.ctypes.data if depth else None), "srt_draw3d_read")
        return (c, z) if depth else c

    def read_async(self, out):
        """Colour read-back that overlaps the next frame (rendered by another Draw3D): returns a ticket."""
        t = C.c_uint64()
        cp = out.data_ptr() if hasattr(out, "data_ptr") else out.ctypes.data
        self._ck(app_lib().srt_draw3d_read_async(self.h, cp, C.byref(t)), "srt_draw3d_read_async")
        return t.value

    def read_wait(self, ticket):
        self._ck(app_lib().srt_draw3d_read_wait(self.h, ticket), "srt_draw3d_read_wait")

    def color_buffer_handle(self):
        """vx_buffer_h of the colour target (for vx_cuda_frame_gather_inplace / vx_cuda_copy_from_dev_async)."""
        return C.c_void_p(app_lib().srt_draw3d_color_buffer(self.h))

    def buffer_addresses(self):
        """(colour, depth) device addresses as vx_mem_address reports them (CUDA device pointers)."""
        out = []
        for fn in (app_lib().srt_draw3d_color_buffer, app_lib().srt_draw3d_depth_buffer):
            a = C.c_uint64()
            self._ck(vx.lib().vx_mem_address(fn(self.h), C.byref(a)), "vx_mem_address")
            out.append(a.value)
        return tuple(out)

    def stats(self):
        s = app_lib().srt_draw3d_stats(self.h).contents
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    def frame(self, trace, depth=False):
        """clear + render + read: what one run of the reference app produces."""
        self.clear()
        self.render(trace)
        return self.read(depth=depth)
