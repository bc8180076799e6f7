"""BVH build on the device (vx_cuda_rt_build_lbvh, csrc/rt_lbvh.cu).

The tree differs from the host builder's (linear BVH vs binned SAH), so the checks are the ones the formats
state (include/skybox_rt.h) and the one that matters: everything rendered or traced from it is bit-identical
to the host-built tree and to the oracle -- closest hit is tree-independent by construction (min t, ties to the
lowest primitive id, identical (v0, e1, e2) operands)."""
import os

import numpy as np
import pytest

import _oracle as o
from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200 import vortex as vx
from skybox_rt_b200.formats import RAY, TRIANGLE

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VARIANTS = [vx.RT_VARIANT_DEFAULT, vx.RT_VARIANT_BASELINE, vx.RT_VARIANT_WHILE, vx.RT_VARIANT_POOL]


def check_tree(nodes, tris, soup):
    """The srt_bvh2_node_t / srt_tri_rec_t contracts: every triangle in exactly one leaf, leaves <= 4, children
    adjacent, boxes = exact min / max of the contained vertices, records = (v0, v1 - v0, v2 - v0, id, albedo)."""
    n = len(soup)
    assert len(tris) == n and sorted(tris["prim_id"].tolist()) == list(range(n))
    src = soup[tris["prim_id"]]
    assert np.array_equal(tris["v0"], src["v0"]) and np.array_equal(tris["albedo"], src["albedo"])
    assert np.array_equal(tris["e1"], src["v1"] - src["v0"]) and np.array_equal(tris["e2"], src["v2"] - src["v0"])
    verts = np.stack([src["v0"], src["v1"], src["v2"]], 1)  # [n, 3, 3] in record order
    lo, hi = verts.min(1), verts.max(1)
    seen = np.zeros(n, bool)
    depth = 0

    def walk(i, d):  # -> (bmin, bmax) of the subtree, checked against the stored box on the way up
        nonlocal depth
        depth = max(depth, d)
        nd = nodes[i]
        if nd["count"] > 0:
            f, c = int(nd["left_first"]), int(nd["count"])
            assert c <= 4 and f + c <= n and not seen[f:f + c].any()
            seen[f:f + c] = True
            b = lo[f:f + c].min(0), hi[f:f + c].max(0)
        else:
            l = int(nd["left_first"])
            assert 0 < l and l + 1 < len(nodes)
            a, c = walk(l, d + 1), walk(l + 1, d + 1)
            b = np.minimum(a[0], c[0]), np.maximum(a[1], c[1])
        assert np.array_equal(nd["bmin"], b[0]) and np.array_equal(nd["bmax"], b[1]), f"box of node {i}"
        return b

    import sys
    sys.setrecursionlimit(10000)
    walk(0, 1)
    assert seen.all()
    return depth


@pytest.fixture(scope="module")
def terrain(native):
    soup = rt.scene_terrain(120, 90)  # 21 600 triangles
    with rt.RayTracer(soup=soup) as lb, rt.RayTracer(*rt.build_bvh2(soup)) as sah:
        yield soup, lb, sah, o.bvh2_build(soup)


def test_tree_is_a_valid_bvh2(terrain):
    soup, lb, _, _ = terrain
    nodes, tris = lb.bvh()
    assert len(nodes) == 2 * len(soup) - 1
    depth = check_tree(nodes, tris, soup)
    assert depth < 64


def test_build_is_deterministic(terrain):
    soup, lb, _, _ = terrain
    a = lb.bvh()
    with rt.RayTracer(soup=soup) as again:
        b = again.bvh()
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("w,h,mode,spp", [(320, 180, 2, 1), (201, 97, 2, 3), (64, 33, 1, 1)])
def test_frames_equal_host_built_tree_and_oracle(terrain, variant, w, h, mode, spp):
    _, lb, sah, (onodes, otris) = terrain
    arg = lambda: rt.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain")
    for t in (lb, sah):
        t.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    a, b = lb.render(arg()), sah.render(arg())
    ref, ctr = o.render(o.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"), onodes, otris)
    assert np.array_equal(a, b), f"{(a != b).sum()} pixels differ between the two trees"
    assert np.array_equal(a, ref)
    c = lb.counters()
    assert [c[k] for k in ("rays_primary", "rays_shadow", "rays_bounce")] == \
           [ctr[k] for k in ("rays_primary", "rays_shadow", "rays_bounce")]


@pytest.mark.parametrize("variant", VARIANTS)
def test_hit_records_equal_host_built_tree(terrain, variant):
    soup, lb, sah, _ = terrain
    rng = np.random.default_rng(3)
    n = 20000
    verts = np.concatenate([soup["v0"], soup["v1"], soup["v2"]])
    lo, hi = verts.min(0), verts.max(0)
    rays = np.zeros(n, RAY)
    rays["o"] = rng.uniform(lo - 0.2 * (hi - lo), hi + 0.2 * (hi - lo), (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    rays["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays["tmin"], rays["tmax"] = 0.0, np.float32(3.0e38)
    for t in (lb, sah):
        t.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    a, b = lb.trace(rays), sah.trace(rays)
    assert a.tobytes() == b.tobytes()
    assert (a["prim_id"] != 0xFFFFFFFF).sum() > n // 20


def test_cornell_goldens_from_the_device_built_tree(native):
    with rt.RayTracer(soup=rt.scene_cornell()) as t:
        for name, mode, spp in [("cornell_64_primary", 0, 1), ("cornell_64_bounce_spp4", 2, 4)]:
            fb = t.render(rt.make_arg(64, 64, mode, spp=spp, seed=1234))
            assert np.array_equal(fb, np.load(os.path.join(GOLDEN, name + ".npy")))
        check_tree(*t.bvh(), np.ascontiguousarray(rt.scene_cornell(), TRIANGLE))


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 7, 33, 2047, 2048, 2049, 5000])
def test_small_and_chunk_boundary_sizes(native, n):
    rng = np.random.default_rng(n)
    soup = np.zeros(n, TRIANGLE)
    c = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    for k in ("v0", "v1", "v2"):
        soup[k] = c + rng.uniform(-0.05, 0.05, (n, 3)).astype(np.float32)
    soup["albedo"] = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    with rt.RayTracer(soup=soup) as t:
        nodes, tris = t.bvh()
        assert len(nodes) == max(1, 2 * n - 1)
        check_tree(nodes, tris, soup)


def test_duplicate_and_degenerate_triangles(native):
    """All triangles identical (every Morton code equal: the index half of the key splits them), zero-area
    triangles, and a flat scene (zero extent in z)."""
    n = 300
    soup = np.zeros(n, TRIANGLE)
    soup["v0"], soup["v1"], soup["v2"] = (0, 0, 0), (1, 0, 0), (0, 1, 0)
    soup["v1"][100:200] = (0, 0, 0)  # points
    with rt.RayTracer(soup=soup) as t:
        nodes, tris = t.bvh()
        depth = check_tree(nodes, tris, soup)
        assert depth <= 12  # balanced on the index bits
        rays = np.zeros(1, RAY)
        rays["o"], rays["d"], rays["tmax"] = (0.25, 0.25, 1.0), (0, 0, -1), 10.0
        hit = t.trace(rays)
    assert hit["prim_id"][0] == 0 and hit["t"][0] == 1.0  # ties go to the lowest primitive id


def test_errors(native):
    with vx.Device() as dev:
        buf = dev.alloc(400)
        import ctypes as C
        nb, tb, info = C.c_void_p(), C.c_void_p(), (C.c_uint32 * 2)()
        L = vx.lib()
        assert L.vx_cuda_rt_build_lbvh(dev.handle, buf.handle, 0, C.byref(nb), C.byref(tb), info) == -1   # no triangles
        assert L.vx_cuda_rt_build_lbvh(dev.handle, buf.handle, 11, C.byref(nb), C.byref(tb), info) == -1  # past the buffer
        assert L.vx_cuda_rt_build_lbvh(dev.handle, None, 1, C.byref(nb), C.byref(tb), info) == -1
        buf.free()
        assert dev.mem_info()[1] == 0


# ---------------------------------------------------------------- BVH2 -> Node4 repack on the device
def _node4_images(make_tracer, w=48, h=32):
    out = []
    for mode in (1, 0, 2):  # host repack (rt_repack.cpp), device repack in one cooperative launch, device repack per level
        with make_tracer() as t:
            t.dev.dcr_write(vx.VX_DCR_RT_REPACK, mode)
            fb = t.render(rt.make_arg(w, h, 2, spp=1, seed=3, scene="terrain"))
            out.append((t.dev.node4_image(), fb))
    return out


@pytest.mark.parametrize("builder", ["sah", "lbvh"])
@pytest.mark.parametrize("quads", [(1, 1), (3, 2), (40, 30), (250, 200)])
def test_device_repack_equals_host_repack(native, builder, quads):
    """The Node4 image built level by level on the device is the host routine's, byte for byte."""
    soup = rt.scene_terrain(*quads)
    if builder == "sah":
        nodes, tris = rt.build_bvh2(soup)
        make = lambda: rt.RayTracer(nodes, tris)
    else:
        make = lambda: rt.RayTracer(soup=soup)
    (host_img, host_fb), (dev_img, dev_fb), (lvl_img, lvl_fb) = _node4_images(make)
    assert len(host_img) == len(dev_img) and len(host_img) % 128 == 0 and len(host_img) > 0
    assert host_img == dev_img, f"{sum(a != b for a, b in zip(host_img, dev_img))} bytes differ"
    assert host_img == lvl_img
    assert np.array_equal(host_fb, dev_fb) and np.array_equal(host_fb, lvl_fb)


def test_device_repack_refuses_what_the_host_repack_refuses(native):
    soup = rt.scene_terrain(20, 20)
    nodes, tris = rt.build_bvh2(soup)
    bad = []
    n = nodes.copy(); n["left_first"][0] = len(nodes) + 5; bad.append(n)          # child out of range
    n = nodes.copy(); n["left_first"][0] = 0; bad.append(n)                        # the root as its own child
    leaf = int(np.nonzero(nodes["count"] > 0)[0][0])
    n = nodes.copy(); n["count"][leaf] = 5; bad.append(n)                          # leaf larger than SRT_MAX_LEAF_TRIS
    n = nodes.copy(); n["left_first"][leaf] = len(tris); bad.append(n)             # triangle range out of bounds
    inner = int(np.nonzero(nodes["count"] == 0)[0][3])
    n = nodes.copy(); n["left_first"][inner] = 1; bad.append(n)                    # a cycle
    for mode in (0, 1, 2):
        for k, n in enumerate(bad):
            with rt.RayTracer(n, tris) as t:
                t.dev.dcr_write(vx.VX_DCR_RT_REPACK, mode)
                with pytest.raises(vx.VortexError):
                    t.render(rt.make_arg(16, 16, 0))
