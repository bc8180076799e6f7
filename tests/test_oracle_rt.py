"""CPU tests: the RT oracle pinned against known answers, brute force and its frozen goldens; the
product's host pre-pass (scenes, BVH2 builder) checked against the oracle's independent restatement.

The reference has no RT fixtures, so the pins are: analytic cases, an O(N)
all-triangle search, and tests/golden/*.npy (made by tests/golden/make_golden.py).
"""
import os

import numpy as np
import pytest

import _oracle as o
from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200.formats import HIT, MISS, RAY, TRIANGLE

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def one_tri(v0, v1, v2, albedo=0xffffffff):
    t = np.zeros(1, dtype=TRIANGLE)
    t["v0"], t["v1"], t["v2"], t["albedo"] = v0, v1, v2, albedo
    return t


def mkray(o_, d, tmin=0.0, tmax=1e30):
    r = np.zeros(1, dtype=RAY)
    r["o"], r["d"], r["tmin"], r["tmax"] = o_, d, tmin, tmax
    return r


# ---------------------------------------------------------------- known answers
def test_single_triangle_known_answer():
    nodes, tris = o.bvh2_build(one_tri((0, 0, 5), (4, 0, 5), (0, 4, 5)))
    hits, _ = o.trace(mkray((1, 1, 0), (0, 0, 1)), nodes, tris)
    assert hits["prim_id"][0] == 0
    assert hits["t"][0] == np.float32(5.0)
    assert hits["u"][0] == np.float32(0.25) and hits["v"][0] == np.float32(0.25)
    # outside the triangle, behind the origin, beyond tmax, before tmin
    assert o.trace(mkray((3, 3, 0), (0, 0, 1)), nodes, tris)[0]["prim_id"][0] == MISS
    assert o.trace(mkray((1, 1, 6), (0, 0, 1)), nodes, tris)[0]["prim_id"][0] == MISS
    assert o.trace(mkray((1, 1, 0), (0, 0, 1), tmax=4.0), nodes, tris)[0]["prim_id"][0] == MISS
    assert o.trace(mkray((1, 1, 0), (0, 0, 1), tmin=5.5), nodes, tris)[0]["prim_id"][0] == MISS
    miss = o.trace(mkray((1, 1, 0), (0, 0, 1), tmax=4.0), nodes, tris)[0]
    assert miss["t"][0] == np.float32(4.0) and miss["u"][0] == 0 and miss["v"][0] == 0


def test_edges_are_inclusive_and_parallel_rays_miss():
    nodes, tris = o.bvh2_build(one_tri((0, 0, 1), (1, 0, 1), (0, 1, 1)))
    for p in [(0, 0, 0), (1, 0, 0), (0, 1, 0), (0.5, 0.5, 0), (0.5, 0, 0)]:
        assert o.trace(mkray(p, (0, 0, 1)), nodes, tris)[0]["prim_id"][0] == 0, p
    assert o.trace(mkray((0.2, 0.2, 0), (1, 0, 0)), nodes, tris)[0]["prim_id"][0] == MISS  # det == 0


def test_equal_t_goes_to_lowest_prim_id():
    a = one_tri((0, 0, 2), (4, 0, 2), (0, 4, 2))
    soup = np.concatenate([a, a, a])
    nodes, tris = o.bvh2_build(soup)
    assert o.trace(mkray((1, 1, 0), (0, 0, 1)), nodes, tris)[0]["prim_id"][0] == 0
    assert o.closest_brute(mkray((1, 1, 0), (0, 0, 1)), tris)["prim_id"][0] == 0


def test_axis_aligned_box_known_t():
    # unit cube centred at (0,0,5): a ray down +z enters at t = 4.5
    soup = o.scene_cornell()[10:22].copy()  # the tall box: 12 triangles, axis aligned
    nodes, tris = o.bvh2_build(soup)
    h = o.trace(mkray((-0.3, 0.0, -3.0), (0, 0, 1)), nodes, tris)[0]
    assert h["prim_id"][0] != MISS
    assert h["t"][0] == np.float32(3.05)  # front face at z = 0.05


# ---------------------------------------------------------------- BVH vs brute force
@pytest.mark.parametrize("scene", ["cornell", "terrain"])
def test_bvh_traversal_equals_brute_force(scene):
    soup = o.scene_cornell() if scene == "cornell" else o.scene_terrain(24, 20)
    nodes, tris = o.bvh2_build(soup)
    lo = np.minimum.reduce([soup["v0"].min(0), soup["v1"].min(0), soup["v2"].min(0)])
    hi = np.maximum.reduce([soup["v0"].max(0), soup["v1"].max(0), soup["v2"].max(0)])
    rays = o.rays_random(lo, hi, 7, 0, 3000)
    hits, _ = o.trace(rays, nodes, tris)
    brute = o.closest_brute(rays, tris)
    assert hits.tobytes() == brute.tobytes()
    assert (hits["prim_id"] != MISS).mean() > 0.3
    rays["tmax"] = 3.0 if scene == "cornell" else 20.0
    assert np.array_equal(o.any_hit(rays, nodes, tris), o.any_hit_brute(rays, tris))


def test_closest_hit_is_independent_of_tree_shape():
    """Oracle traversal over the oracle's median-split tree == over the product's SAH tree."""
    soup = o.scene_terrain(40, 30)
    n1, t1 = o.bvh2_build(soup)
    n2, t2 = rt.build_bvh2(soup)
    assert len(n1) != len(n2) or n1.tobytes() != n2.tobytes()
    lo = np.array([-70, -10, -50], np.float32)
    hi = np.array([70, 40, 50], np.float32)
    rays = o.rays_random(lo, hi, 11, 0, 20000)
    h1, _ = o.trace(rays, n1, t1)
    h2, _ = o.trace(rays, n2, t2)
    assert h1.tobytes() == h2.tobytes()
    a = o.make_arg(96, 54, o.MODE_BOUNCE, spp=2, scene="terrain")
    f1, c1 = o.render(a, n1, t1)
    f2, c2 = o.render(a, n2, t2)
    assert np.array_equal(f1, f2)
    assert [c1[k] for k in ("rays_primary", "rays_shadow", "rays_bounce")] == \
           [c2[k] for k in ("rays_primary", "rays_shadow", "rays_bounce")]


# ---------------------------------------------------------------- frozen goldens
@pytest.mark.parametrize("name,mode,spp", [("cornell_64_primary", 0, 1), ("cornell_64_shadow", 1, 1),
                                           ("cornell_64_bounce", 2, 1), ("cornell_64_bounce_spp4", 2, 4)])
def test_oracle_reproduces_golden(name, mode, spp):
    nodes, tris = o.bvh2_build(o.scene_cornell())
    fb, ctr = o.render(o.make_arg(64, 64, mode, spp=spp, seed=1234), nodes, tris)
    gold = np.load(os.path.join(GOLDEN, name + ".npy"))
    assert np.array_equal(fb, gold)
    assert ctr["rays_primary"] == 64 * 64 * spp
    assert (fb >> 24 == 0xff).all()


def test_render_is_independent_of_task_and_thread_count():
    nodes, tris = o.bvh2_build(o.scene_cornell())
    ref, _ = o.render(o.make_arg(40, 30, 2, spp=2, num_tasks=30), nodes, tris, threads=1)
    for tasks, threads in [(1, 1), (7, 3), (64, 8)]:
        fb, _ = o.render(o.make_arg(40, 30, 2, spp=2, num_tasks=tasks), nodes, tris, threads=threads)
        assert np.array_equal(fb, ref)


def test_rank_strips_reassemble_to_the_full_frame():
    nodes, tris = o.bvh2_build(o.scene_cornell())
    full, _ = o.render(o.make_arg(33, 30, 1), nodes, tris)
    for ranks, rpb in [(2, 4), (3, 4), (4, 8), (8, 1)]:
        strips = []
        for r in range(ranks):
            a = o.make_arg(33, 30, 1)
            a.num_ranks, a.rank_id, a.rows_per_block = ranks, r, rpb
            strips.append(o.render(a, nodes, tris)[0])
        assert sum(len(s) for s in strips) == 30
        assert np.array_equal(rt.assemble_strips(strips, 30, rpb, ranks), full)


# ---------------------------------------------------------------- product host pre-pass vs oracle
def test_host_scene_generators_match_oracle_bytes():
    assert rt.scene_cornell().tobytes() == o.scene_cornell().tobytes()
    for qx, qy, seed in [(1, 1, 1), (17, 9, 0x5EED5EED), (64, 48, 42)]:
        assert rt.scene_terrain(qx, qy, seed).tobytes() == o.scene_terrain(qx, qy, seed).tobytes()
    lo, hi = np.array([-1, -2, -3], np.float32), np.array([4, 5, 6], np.float32)
    assert rt.rays_random(lo, hi, 9, 1 << 33, 1000).tobytes() == o.rays_random(lo, hi, 9, 1 << 33, 1000).tobytes()
    d = rt.rays_random(lo, hi, 9, 0, 1000)["d"].astype(np.float64)
    assert np.allclose((d * d).sum(1), 1.0, atol=1e-6)


def test_host_camera_presets_match_oracle():
    for scene in ("cornell", "terrain"):
        a, b = rt.make_arg(8, 8, 2, scene=scene), o.make_arg(8, 8, 2, scene=scene)
        for f in ("cam_pos", "cam_fwd", "cam_right", "cam_up", "light_pos", "sky"):
            assert list(getattr(a, f)) == list(getattr(b, f)), f
        assert (a.tan_half_fov, a.ambient, a.ray_eps) == (b.tan_half_fov, b.ambient, b.ray_eps)


def test_host_bvh2_builder_invariants():
    soup = rt.scene_terrain(50, 40)
    nodes, tris = rt.build_bvh2(soup)
    assert sorted(tris["prim_id"].tolist()) == list(range(len(soup)))
    src = soup[tris["prim_id"]]
    assert np.array_equal(tris["v0"], src["v0"])
    assert np.array_equal(tris["e1"], src["v1"] - src["v0"]) and np.array_equal(tris["e2"], src["v2"] - src["v0"])
    assert np.array_equal(tris["albedo"], src["albedo"])
    leaves = nodes[nodes["count"] > 0]
    assert leaves["count"].max() <= 4 and leaves["count"].sum() == len(soup)
    assert len(nodes) == 2 * len(leaves) - 1
    # every node's box is exactly the min/max of what it contains, children inside parents
    lo = np.minimum(np.minimum(src["v0"], src["v1"]), src["v2"])
    hi = np.maximum(np.maximum(src["v0"], src["v1"]), src["v2"])
    for n in leaves[:200]:
        s = slice(int(n["left_first"]), int(n["left_first"] + n["count"]))
        assert np.array_equal(n["bmin"], lo[s].min(0)) and np.array_equal(n["bmax"], hi[s].max(0))
    inner = nodes[nodes["count"] == 0]
    for side in (0, 1):
        ch = nodes[inner["left_first"] + side]
        assert (ch["bmin"] >= inner["bmin"]).all() and (ch["bmax"] <= inner["bmax"]).all()
    # deterministic
    n2, t2 = rt.build_bvh2(soup)
    assert n2.tobytes() == nodes.tobytes() and t2.tobytes() == tris.tobytes()


def test_host_bvh2_degenerate_inputs():
    # all centroids identical -> object-median fallback must still terminate with valid leaves
    soup = np.repeat(one_tri((0, 0, 1), (1, 0, 1), (0, 1, 1)), 37)
    nodes, tris = rt.build_bvh2(soup)
    assert nodes[nodes["count"] > 0]["count"].sum() == 37
    h, _ = o.trace(mkray((0.2, 0.2, 0), (0, 0, 1)), nodes, tris)
    assert h["prim_id"][0] == 0
    with pytest.raises(Exception):
        rt.build_bvh2(np.zeros(0, dtype=TRIANGLE))
    assert HIT.itemsize == 16
