"""ctypes binding of the vx_* C ABI (include/vortex.h) exported by skybox_rt_b200/lib/libvortex.so.

Same names, argument order and error convention (0 = success) as the reference host API
(/root/reference runtime/include/vortex.h:74-139).  The thin `Device` / `Buffer` classes below only
turn non-zero return codes into exceptions, the way the reference apps wrap every call in RT_CHECK
(tests/regression/draw3d/main.cpp:23-31).

There is no fallback: if the native library is missing this module raises at import of `lib()`.
"""
import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SKYBOX_RT_LIB") or os.path.join(PKG, "lib", "libvortex.so")  # env: A/B builds only

# caps ids / flags (include/vortex.h)
VX_CAPS_VERSION, VX_CAPS_NUM_THREADS, VX_CAPS_NUM_WARPS, VX_CAPS_NUM_CORES = 0, 1, 2, 3
VX_CAPS_CACHE_LINE_SIZE, VX_CAPS_GLOBAL_MEM_SIZE, VX_CAPS_LOCAL_MEM_SIZE, VX_CAPS_ISA_FLAGS = 4, 5, 6, 7
VX_ISA_EXT_TEX, VX_ISA_EXT_RASTER, VX_ISA_EXT_OM, VX_ISA_EXT_RT = 1 << 38, 1 << 39, 1 << 40, 1 << 44
VX_MEM_READ, VX_MEM_WRITE, VX_MEM_READ_WRITE = 1, 2, 3
VX_MAX_TIMEOUT = 24 * 60 * 60 * 1000

# counters (include/skybox_rt.h)
VX_CSR_MCYCLE, VX_CSR_MINSTRET = 0xB00, 0xB02
VX_CSR_RT_RAYS_PRIMARY, VX_CSR_RT_RAYS_SHADOW, VX_CSR_RT_RAYS_BOUNCE = 0x7C0, 0x7C1, 0x7C2
VX_CSR_RT_NODES, VX_CSR_RT_TRIS, VX_CSR_RT_ELAPSED_NS = 0x7C3, 0x7C4, 0x7C5
VX_CSR_RT_NODE_BYTES, VX_CSR_RT_LAUNCHES = 0x7C6, 0x7C7
VX_CSR_RT_WARP_NODE_STEPS, VX_CSR_RT_WARP_TRI_STEPS = 0x7C8, 0x7C9

# back-end DCRs (csrc/vx_internal.h)
VX_DCR_RT_VARIANT, VX_DCR_RT_STATS, VX_DCR_RT_KEEP_ACTIVE, VX_DCR_RT_PREFETCH = 0x800, 0x801, 0x802, 0x803
VX_DCR_RT_CHAINS = 0x809  # wavefront pipeline: chains per frame (0 = default, 1..4)
VX_DCR_RT_PASS_JOBS = 0x80A  # wavefront pipeline: jobs per pass (0 = default 2^27)
VX_DCR_RT_FUSE_BOUNCE = 0x80B  # wavefront pipeline: 0 = K4 / K5 folded into K3 (default), 1 = separate launches
VX_DCR_RT_LBVH_ROTATE = 0x80C  # device BVH build: rotation passes (0 = default 2, k = k - 1 passes)
VX_DCR_RT_REPACK = 0x808  # 0 = BVH2 -> Node4 on the device, 1 = on the host (A/B)
VX_DCR_RT_SPLIT_SAMPLES = 0x807  # 0 = (pixel, sample) jobs when spp > 1, 1 = never, 2 = always
RT_VARIANT_DEFAULT, RT_VARIANT_BASELINE, RT_VARIANT_WHILE, RT_VARIANT_POOL = 0, 1, 2, 3
# wavefront pipeline: trace kernels + shading passes (csrc/rt_wave.cuh), trace kernels scheduled while-while / vote / vote + deferred triangle
RT_VARIANT_WAVE, RT_VARIANT_WAVE_VOTE, RT_VARIANT_WAVE_DEFER = 4, 5, 6
RT_VARIANT_MEGA_VOTE = 7  # persistent megakernel, vote scheduler (the round-1 default); 0 now picks 4 for frames, 7 for ray lists
RT_VARIANTS = (RT_VARIANT_DEFAULT, RT_VARIANT_BASELINE, RT_VARIANT_WHILE, RT_VARIANT_POOL, RT_VARIANT_WAVE,
               RT_VARIANT_WAVE_VOTE, RT_VARIANT_WAVE_DEFER, RT_VARIANT_MEGA_VOTE)  # every shipped kernel variant

EXPORTS = [
    "vx_dev_open", "vx_dev_close", "vx_dev_caps", "vx_mem_alloc", "vx_mem_reserve", "vx_mem_free",
    "vx_mem_access", "vx_mem_address", "vx_mem_info", "vx_copy_to_dev", "vx_copy_from_dev", "vx_start",
    "vx_ready_wait", "vx_dcr_read", "vx_dcr_write", "vx_mpm_query", "vx_upload_kernel_bytes",
    "vx_upload_kernel_file", "vx_upload_bytes", "vx_upload_file", "vx_check_occupancy", "vx_dump_perf",
    "vx_dev_init",
]
CUDA_EXPORTS = ["vx_cuda_stream", "vx_cuda_device", "vx_cuda_host_register", "vx_cuda_host_unregister",
                "vx_cuda_perf_totals", "vx_cuda_draw3d_binning", "vx_cuda_rt_build_lbvh",
                "vx_cuda_rt_node4_image", "vx_cuda_fill32", "vx_cuda_probe_read_bandwidth",
                "vx_cuda_comm_unique_id", "vx_cuda_comm_init", "vx_cuda_comm_destroy", "vx_cuda_frame_gather",
                "vx_cuda_gather_bytes", "vx_cuda_frame_gather_inplace", "vx_cuda_copy_from_dev_async", "vx_cuda_copy_wait",
                "vx_cuda_rt_build_lbvh_noalloc"]  # include/vortex_cuda.h
VX_CUDA_LBVH_NOALLOC_MAX_TRIS = 65536
VX_CUDA_COMM_ID_BYTES = 128

_lib = None


class VortexError(RuntimeError):
    pass


def lib():
    """Load libvortex.so (never builds, never falls back)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VortexError(
                f"{LIB_PATH} is missing: run `python -m skybox_rt_b200.build` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, u32, u64, i = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
        sig = {
            "vx_dev_open": [C.POINTER(vp)], "vx_dev_close": [vp],
            "vx_dev_caps": [vp, u32, C.POINTER(u64)],
            "vx_mem_alloc": [vp, u64, i, C.POINTER(vp)],
            "vx_mem_reserve": [vp, u64, u64, i, C.POINTER(vp)],
            "vx_mem_free": [vp], "vx_mem_access": [vp, u64, u64, i],
            "vx_mem_address": [vp, C.POINTER(u64)],
            "vx_mem_info": [vp, C.POINTER(u64), C.POINTER(u64)],
            "vx_copy_to_dev": [vp, vp, u64, u64], "vx_copy_from_dev": [vp, vp, u64, u64],
            "vx_start": [vp, vp, vp], "vx_ready_wait": [vp, u64],
            "vx_dcr_read": [vp, u32, C.POINTER(u32)], "vx_dcr_write": [vp, u32, u32],
            "vx_mpm_query": [vp, u32, u32, C.POINTER(u64)],
            "vx_upload_kernel_bytes": [vp, vp, u64, C.POINTER(vp)],
            "vx_upload_kernel_file": [vp, C.c_char_p, C.POINTER(vp)],
            "vx_upload_bytes": [vp, vp, u64, C.POINTER(vp)],
            "vx_upload_file": [vp, C.c_char_p, C.POINTER(vp)],
            "vx_check_occupancy": [vp, u32, C.POINTER(u32)],
            "vx_cuda_stream": [vp, C.POINTER(vp)], "vx_cuda_device": [vp, C.POINTER(i)],
            "vx_cuda_perf_totals": [vp, C.POINTER(u64), u32],
            "vx_cuda_rt_build_lbvh": [vp, vp, u32, C.POINTER(vp), C.POINTER(vp), vp],
            "vx_cuda_rt_build_lbvh_noalloc": [vp, vp, u32, vp, vp, vp],
            "vx_cuda_rt_node4_image": [vp, vp, u64, C.POINTER(u64)],
            "vx_cuda_fill32": [vp, u64, u64, u32],
            "vx_cuda_probe_read_bandwidth": [vp, u64, u32, C.POINTER(C.c_float)],
            "vx_cuda_copy_from_dev_async": [vp, vp, u64, u64, C.POINTER(u64)], "vx_cuda_copy_wait": [vp, u64],
            "vx_cuda_comm_unique_id": [vp], "vx_cuda_comm_init": [vp, vp, u32, u32, C.POINTER(vp)],
            "vx_cuda_comm_destroy": [vp], "vx_cuda_frame_gather": [vp, vp, vp, vp, u32, u32, u32, u32],
            "vx_cuda_gather_bytes": [vp, vp, vp, u64, vp, C.POINTER(u64), C.POINTER(u64), u32],
            "vx_cuda_frame_gather_inplace": [vp, vp, vp, u32, u32, u32, u32],
            "vx_cuda_draw3d_binning": [vp, vp, u32, vp, u32, C.c_float, C.c_float, u32, u32, u32, u32, u32, u32,
                                       C.POINTER(vp), C.POINTER(vp), vp],
        }
        for name, args in sig.items():
            if name.startswith("vx_cuda_") and not hasattr(L, name):
                # SKYBOX_RT_LIB names the reference's own stub (oracle/_ref/libvortex.so, tests/test_gpu_ref_stub.py):
                # it forwards the 22 reference functions to libvortex-cuda.so and knows nothing of the extensions
                continue
            fn = getattr(L, name)
            fn.argtypes = args
            fn.restype = i
        _lib = L
    return _lib


def _check(rc, what):
    if rc != 0:
        raise VortexError(f"{what} failed with code {rc}")


def _host_ptr(arr):
    if isinstance(arr, (bytes, bytearray)):
        return C.cast(C.c_char_p(bytes(arr)), C.c_void_p), len(arr)
    if isinstance(arr, C.Structure):
        return C.cast(C.pointer(arr), C.c_void_p), C.sizeof(arr)
    if isinstance(arr, np.ndarray):
        assert arr.flags["C_CONTIGUOUS"]
        return C.c_void_p(arr.ctypes.data), arr.nbytes
    if hasattr(arr, "data_ptr"):  # torch CPU tensor (pinned or pageable)
        assert arr.device.type == "cpu" and arr.is_contiguous()
        return C.c_void_p(arr.data_ptr()), arr.numel() * arr.element_size()
    raise TypeError(type(arr))


class Buffer:
    def __init__(self, dev, handle, size):
        self.dev, self.handle, self.size = dev, handle, size

    @property
    def address(self):
        a = C.c_uint64()
        _check(lib().vx_mem_address(self.handle, C.byref(a)), "vx_mem_address")
        return a.value

    def write(self, src, offset=0):
        p, n = _host_ptr(src)
        _check(lib().vx_copy_to_dev(self.handle, p, offset, n), "vx_copy_to_dev")

    def read(self, dst, offset=0):
        p, n = _host_ptr(dst)
        _check(lib().vx_copy_from_dev(p, self.handle, offset, n), "vx_copy_from_dev")
        return dst

    def read_async(self, dst, offset=0):
        """vx_cuda_copy_from_dev_async: returns a ticket for Device.copy_wait (dst should be pinned)."""
        p, n = _host_ptr(dst)
        t = C.c_uint64()
        _check(lib().vx_cuda_copy_from_dev_async(p, self.handle, offset, n, C.byref(t)), "vx_cuda_copy_from_dev_async")
        return t.value

    def fill32(self, value, offset=0, size=None):
        """Device-side fill with a 32-bit pattern (vx_cuda_fill32)."""
        _check(lib().vx_cuda_fill32(self.handle, offset, self.size - offset if size is None else size, value), "vx_cuda_fill32")

    def free(self):
        if self.handle is not None:
            _check(lib().vx_mem_free(self.handle), "vx_mem_free")
            self.handle = None


def comm_unique_id():
    """VX_CUDA_COMM_ID_BYTES bytes naming a new communicator (vx_cuda_comm_unique_id): make it on one rank, hand it
    to the others."""
    buf = (C.c_ubyte * VX_CUDA_COMM_ID_BYTES)()
    _check(lib().vx_cuda_comm_unique_id(buf), "vx_cuda_comm_unique_id")
    return bytes(buf)


class Comm:
    """This device's membership in a group of num_ranks devices (vx_cuda_comm_init; collective)."""

    def __init__(self, dev, comm_id, rank, num_ranks):
        assert len(comm_id) == VX_CUDA_COMM_ID_BYTES
        h = C.c_void_p()
        idbuf = (C.c_ubyte * VX_CUDA_COMM_ID_BYTES).from_buffer_copy(comm_id)
        _check(lib().vx_cuda_comm_init(dev.handle, idbuf, rank, num_ranks, C.byref(h)), "vx_cuda_comm_init")
        self.dev, self.handle, self.rank, self.num_ranks = dev, h, rank, num_ranks

    def frame_gather(self, strip, frame, width, height, rows_per_block, root=0):
        """Asynchronous on the launch stream: strips -> the root's `frame` buffer (None on the other ranks)."""
        _check(lib().vx_cuda_frame_gather(self.dev.handle, self.handle, strip.handle, frame.handle if frame else None,
                                          width, height, rows_per_block, root), "vx_cuda_frame_gather")

    def frame_gather_inplace(self, frame_handle, width, height, rows_per_block, root=0):
        """Row blocks of full-size per-rank frame buffers -> the root's buffer, in place (vx_cuda_frame_gather_inplace);
        frame_handle: a vx_buffer_h (e.g. srt_draw3d_color_buffer)."""
        _check(lib().vx_cuda_frame_gather_inplace(self.dev.handle, self.handle, frame_handle, width, height, rows_per_block,
                                                  root), "vx_cuda_frame_gather_inplace")

    def gather_bytes(self, src, nbytes, dst=None, offsets=None, sizes=None, root=0):
        n = self.num_ranks
        off = (C.c_uint64 * n)(*offsets) if offsets is not None else None
        siz = (C.c_uint64 * n)(*sizes) if sizes is not None else None
        _check(lib().vx_cuda_gather_bytes(self.dev.handle, self.handle, src.handle, nbytes, dst.handle if dst else None,
                                          off, siz, root), "vx_cuda_gather_bytes")

    def close(self):
        if self.handle is not None:
            lib().vx_cuda_comm_destroy(self.handle)
            self.handle = None


class Device:
    """vx_dev_open .. vx_dev_close as a context manager."""

    def __init__(self):
        h = C.c_void_p()
        _check(lib().vx_dev_open(C.byref(h)), "vx_dev_open")
        self.handle = h

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        if self.handle is not None:
            _check(lib().vx_dev_close(self.handle), "vx_dev_close")
            self.handle = None

    def caps(self, caps_id):
        v = C.c_uint64()
        _check(lib().vx_dev_caps(self.handle, caps_id, C.byref(v)), "vx_dev_caps")
        return v.value

    def alloc(self, size, flags=VX_MEM_READ_WRITE):
        h = C.c_void_p()
        _check(lib().vx_mem_alloc(self.handle, size, flags, C.byref(h)), "vx_mem_alloc")
        return Buffer(self, h, size)

    def upload(self, src):
        p, n = _host_ptr(src)
        h = C.c_void_p()
        _check(lib().vx_upload_bytes(self.handle, p, n, C.byref(h)), "vx_upload_bytes")
        return Buffer(self, h, n)

    def upload_kernel(self, image_bytes):
        p, n = _host_ptr(image_bytes)
        h = C.c_void_p()
        _check(lib().vx_upload_kernel_bytes(self.handle, p, n, C.byref(h)), "vx_upload_kernel_bytes")
        return Buffer(self, h, n - 16)

    def start(self, kernel, args):
        _check(lib().vx_start(self.handle, kernel.handle, args.handle), "vx_start")

    def ready_wait(self, timeout_ms=VX_MAX_TIMEOUT):
        _check(lib().vx_ready_wait(self.handle, timeout_ms), "vx_ready_wait")

    def copy_wait(self, ticket):
        _check(lib().vx_cuda_copy_wait(self.handle, ticket), "vx_cuda_copy_wait")

    def dcr_write(self, addr, value):
        _check(lib().vx_dcr_write(self.handle, addr, value), "vx_dcr_write")

    def dcr_read(self, addr):
        v = C.c_uint32()
        _check(lib().vx_dcr_read(self.handle, addr, C.byref(v)), "vx_dcr_read")
        return v.value

    def mpm_query(self, addr, core_id=0):
        v = C.c_uint64()
        _check(lib().vx_mpm_query(self.handle, addr, core_id, C.byref(v)), "vx_mpm_query")
        return v.value

    def node4_image_size(self):
        """Size in bytes of the wide-node image the last ray-tracing run traversed (0 before the first one)."""
        n = C.c_uint64()
        return n.value if lib().vx_cuda_rt_node4_image(self.handle, None, 0, C.byref(n)) == 0 else 0

    def node4_image(self):
        """Bytes of the Node4 image the last ray-tracing run traversed (vx_cuda_rt_node4_image)."""
        n = C.c_uint64()
        _check(lib().vx_cuda_rt_node4_image(self.handle, None, 0, C.byref(n)), "vx_cuda_rt_node4_image")
        out = np.zeros(n.value, np.uint8)
        _check(lib().vx_cuda_rt_node4_image(self.handle, out.ctypes.data, n.value, C.byref(n)), "vx_cuda_rt_node4_image")
        return out.tobytes()

    def perf_totals(self):
        """Accumulated performance-monitor totals, slots VX_PERF_* of include/vortex_cuda.h."""
        out = (C.c_uint64 * 32)()
        _check(lib().vx_cuda_perf_totals(self.handle, out, 32), "vx_cuda_perf_totals")
        return list(out)

    def read_bandwidth_gbs(self, nbytes, repeats=20):
        """GB/s of the read probe over a private buffer of nbytes (vx_cuda_probe_read_bandwidth)."""
        ms = C.c_float()
        _check(lib().vx_cuda_probe_read_bandwidth(self.handle, nbytes, repeats, C.byref(ms)), "vx_cuda_probe_read_bandwidth")
        return nbytes * repeats / (ms.value * 1e-3) / 1e9

    def cuda_stream(self):
        """cudaStream_t (as int) that vx_start launches on -- wrap with torch.cuda.ExternalStream."""
        p = C.c_void_p()
        _check(lib().vx_cuda_stream(self.handle, C.byref(p)), "vx_cuda_stream")
        return p.value or 0

    def cuda_device(self):
        n = C.c_int()
        _check(lib().vx_cuda_device(self.handle, C.byref(n)), "vx_cuda_device")
        return n.value

    def mem_info(self):
        f, u = C.c_uint64(), C.c_uint64()
        _check(lib().vx_mem_info(self.handle, C.byref(f), C.byref(u)), "vx_mem_info")
        return f.value, u.value
