"""vx_cuda_rt_build_lbvh_noalloc (include/vortex_cuda.h, csrc/rt_lbvh.cu): the device BVH build for callers that
cannot allocate -- outputs owned by the caller, scratch in a fixed device array inside libvortex.so.

No counterpart in the reference, so the checker is the existing entry
point vx_cuda_rt_build_lbvh, itself held against the oracle in tests/test_gpu_lbvh.py: the same soup must give the
same node and triangle-record BYTES (tolerance 0, the builder is integer / exact-fp32 work), on calls in a row that
share the scratch array -- a large scene first, so the smaller ones that follow find its leftovers there.

Collected last on purpose (the file name): a failure here must not hide the other tests under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200 import vortex as vx
from skybox_rt_b200.formats import NODE, TRIANGLE, TRIREC

pytestmark = pytest.mark.gpu
MAX_TRIS = vx.VX_CUDA_LBVH_NOALLOC_MAX_TRIS


def random_soup(n, seed):
    """n small triangles scattered in a box; ids and albedos arbitrary."""
    rng = np.random.default_rng(seed)
    soup = np.zeros(n, TRIANGLE)
    centre = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    for k in ("v0", "v1", "v2"):
        soup[k] = centre + rng.uniform(-0.2, 0.2, (n, 3)).astype(np.float32)
    soup["albedo"] = rng.integers(0, 1 << 24, n, dtype=np.uint32)
    return soup


def num_nodes_of(n):
    return 1 if n == 1 else 2 * n - 1


def build_with_allocation(dev, sbuf, n):
    """The existing entry point: new buffers, stream-ordered scratch."""
    nb, tb, info = C.c_void_p(), C.c_void_p(), (C.c_uint32 * 2)()
    rc = vx.lib().vx_cuda_rt_build_lbvh(dev.handle, sbuf.handle, n, C.byref(nb), C.byref(tb), info)
    assert rc == 0
    nodes, tris = np.empty(info[0], NODE), np.empty(info[1], TRIREC)
    node_buf, tri_buf = vx.Buffer(dev, nb, nodes.nbytes), vx.Buffer(dev, tb, tris.nbytes)
    node_buf.read(nodes)
    tri_buf.read(tris)
    node_buf.free()
    tri_buf.free()
    return nodes, tris


def build_without_allocation(dev, sbuf, n, node_buf, tri_buf):
    """The new entry point into the caller's buffers (pre-filled with a pattern it must overwrite completely)."""
    junk_n = np.full(num_nodes_of(n) * NODE.itemsize // 4, 0xDEADBEEF, np.uint32)
    junk_t = np.full(n * TRIREC.itemsize // 4, 0xDEADBEEF, np.uint32)
    node_buf.write(junk_n)
    tri_buf.write(junk_t)
    num_nodes, num_tris = rt.build_lbvh_noalloc(dev, sbuf, n, node_buf, tri_buf)
    assert (num_nodes, num_tris) == (num_nodes_of(n), n)
    nodes, tris = np.empty(num_nodes, NODE), np.empty(num_tris, TRIREC)
    node_buf.read(nodes)
    tri_buf.read(tris)
    return nodes, tris


def same_bytes(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and a.tobytes() == b.tobytes()


def test_no_alloc_build_equals_the_allocating_build_on_calls_in_a_row(native):
    soups = [rt.scene_terrain(110, 100),   # 22 000 triangles: fills most of the scratch first
             rt.scene_cornell(),           # a few dozen: runs inside the leftovers of the call before
             random_soup(1, 3),            # the single-leaf tree
             random_soup(5000, 4),
             rt.scene_cornell()]           # the same scene again: the second call must equal the first
    with vx.Device() as dev:
        node_buf = dev.alloc(num_nodes_of(MAX_TRIS) * NODE.itemsize)
        tri_buf = dev.alloc(MAX_TRIS * TRIREC.itemsize)
        for i, soup in enumerate(soups):
            soup = np.ascontiguousarray(soup, dtype=TRIANGLE)
            n = len(soup)
            sbuf = dev.upload(soup)
            want_nodes, want_tris = build_with_allocation(dev, sbuf, n)
            got_nodes, got_tris = build_without_allocation(dev, sbuf, n, node_buf, tri_buf)
            again_nodes, again_tris = build_without_allocation(dev, sbuf, n, node_buf, tri_buf)
            sbuf.free()
            assert same_bytes(got_tris, want_tris), f"call {i} ({n} triangles): triangle records differ"
            assert same_bytes(got_nodes, want_nodes), f"call {i} ({n} triangles): nodes differ"
            assert same_bytes(again_tris, want_tris) and same_bytes(again_nodes, want_nodes), f"call {i}: repeat differs"
        node_buf.free()
        tri_buf.free()


def test_no_alloc_build_at_the_maximum_size_and_its_frame(native):
    """Exactly VX_CUDA_LBVH_NOALLOC_MAX_TRIS triangles use the whole array; a frame rendered from the tree it
    built equals the frame from the allocating build's tree."""
    soup = random_soup(MAX_TRIS, 11)
    with vx.Device() as dev:
        sbuf = dev.upload(soup)
        node_buf = dev.alloc(num_nodes_of(MAX_TRIS) * NODE.itemsize)
        tri_buf = dev.alloc(MAX_TRIS * TRIREC.itemsize)
        want_nodes, want_tris = build_with_allocation(dev, sbuf, MAX_TRIS)
        got_nodes, got_tris = build_without_allocation(dev, sbuf, MAX_TRIS, node_buf, tri_buf)
        assert same_bytes(got_tris, want_tris) and same_bytes(got_nodes, want_nodes)
        for b in (sbuf, node_buf, tri_buf):
            b.free()
    arg = dict(spp=1, seed=9)
    with rt.RayTracer(got_nodes, got_tris) as a, rt.RayTracer(want_nodes, want_tris) as b:
        fa = a.render(rt.make_arg(96, 64, 2, **arg)).copy()
        fb = b.render(rt.make_arg(96, 64, 2, **arg)).copy()
    assert np.array_equal(fa, fb)


def test_no_alloc_build_rejects_what_does_not_fit(native, capfd):
    soup = random_soup(64, 5)
    with vx.Device() as dev:
        sbuf = dev.upload(soup)
        node_buf = dev.alloc(num_nodes_of(64) * NODE.itemsize)
        tri_buf = dev.alloc(64 * TRIREC.itemsize)
        info = (C.c_uint32 * 2)()
        call = vx.lib().vx_cuda_rt_build_lbvh_noalloc
        # beyond the fixed scratch: its own code and a message that names the limit
        assert call(dev.handle, sbuf.handle, MAX_TRIS + 1, node_buf.handle, tri_buf.handle, info) == -3
        assert str(MAX_TRIS) in capfd.readouterr().err
        with pytest.raises(vx.VortexError):
            rt.build_lbvh_noalloc(dev, sbuf, MAX_TRIS + 1, node_buf, tri_buf)
        # bad arguments: nothing to build, null handles, outputs too small, aliased outputs, more triangles than the soup holds
        assert call(dev.handle, sbuf.handle, 0, node_buf.handle, tri_buf.handle, info) == -1
        assert call(dev.handle, None, 64, node_buf.handle, tri_buf.handle, info) == -1
        assert call(dev.handle, sbuf.handle, 64, node_buf.handle, tri_buf.handle, None) == -1
        small_nodes = dev.alloc(num_nodes_of(64) * NODE.itemsize - 4)
        small_tris = dev.alloc(64 * TRIREC.itemsize - 4)
        assert call(dev.handle, sbuf.handle, 64, small_nodes.handle, tri_buf.handle, info) == -1
        assert call(dev.handle, sbuf.handle, 64, node_buf.handle, small_tris.handle, info) == -1
        assert call(dev.handle, sbuf.handle, 64, node_buf.handle, node_buf.handle, info) == -1
        assert call(dev.handle, sbuf.handle, 65, node_buf.handle, tri_buf.handle, info) == -1
        assert (info[0], info[1]) == (0, 0)
        # and the buffers that do fit still build after the refusals
        assert rt.build_lbvh_noalloc(dev, sbuf, 64, node_buf, tri_buf) == (127, 64)
        for b in (sbuf, node_buf, tri_buf, small_nodes, small_tris):
            b.free()
