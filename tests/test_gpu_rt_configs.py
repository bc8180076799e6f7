"""GPU parity on the configurations BASELINE.json quotes its numbers on (C2..C5), not on toy scenes.

Every frame / hit list the bench times is compared here with the CPU oracle bit for bit: C2 and C3 as whole
frames, C4 and C5 on a stated subsample against the full 10 M-triangle tree (one rank-of-64 row strip at 8 spp;
the first 1 Mi rays).  This is where deep trees, the traversal-stack spill, the device repack of 1 M / 10 M
nodes and an HBM-resident BVH meet a checker (reference analogue: every regression app compares what it
rendered, tests/regression/draw3d/main.cpp:505-514).  Oracle tree = oracle builder (spatial median), product
tree = host SAH or device LBVH: closest hit is tree-independent by construction, so equality is the claim.
"""
import numpy as np
import pytest

import _oracle as o
from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200 import vortex as vx
from skybox_rt_b200.formats import MISS

pytestmark = pytest.mark.gpu
VARIANTS = sorted(vx.RT_VARIANTS)  # every kernel variant the back-end ships
RPB = 8


def ray_counts(c):
    return [c[k] for k in ("rays_primary", "rays_shadow", "rays_bounce")]


def scene_box(soup):
    lo = np.minimum.reduce([soup["v0"].min(0), soup["v1"].min(0), soup["v2"].min(0)])
    hi = np.maximum.reduce([soup["v0"].max(0), soup["v1"].max(0), soup["v2"].max(0)])
    return lo.astype(np.float32), hi.astype(np.float32)


class Scene:
    def __init__(self, qx, qy):
        self.soup = rt.scene_terrain(qx, qy)
        self.nodes, self.tris = rt.build_bvh2(self.soup)
        self.onodes, self.otris = o.bvh2_build(self.soup)
        self.tracer = rt.RayTracer(self.nodes, self.tris)
        self._refs = {}

    def oracle_frame(self, w, h, mode, spp, seed, ranks=1, rank=0):
        key = (w, h, mode, spp, seed, ranks, rank)
        if key not in self._refs:
            a = o.make_arg(w, h, mode, spp=spp, seed=seed, scene="terrain", num_tasks=h)
            a.num_ranks, a.rank_id, a.rows_per_block = ranks, rank, RPB
            self._refs[key] = o.render(a, self.onodes, self.otris)
        return self._refs[key]

    def close(self):
        self.tracer.close()


@pytest.fixture(scope="module")
def c2(native):
    s = Scene(250, 200)  # 100 000 triangles
    yield s
    s.close()


@pytest.fixture(scope="module")
def c3(native):
    s = Scene(1000, 500)  # 1 000 000 triangles
    yield s
    s.close()


@pytest.fixture(scope="module")
def c45(native):
    s = Scene(2500, 2000)  # 10 000 000 triangles
    yield s
    s.close()


@pytest.mark.parametrize("variant", VARIANTS)
def test_c2_full_frame(c2, variant):
    c2.tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    fb = c2.tracer.render(rt.make_arg(1024, 1024, 0, spp=1, seed=1234, scene="terrain"))
    ref, ctr = c2.oracle_frame(1024, 1024, 0, 1, 1234)
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(c2.tracer.counters()) == ray_counts(ctr)


@pytest.mark.parametrize("variant", VARIANTS)
def test_c3_full_frame(c3, variant):
    """The frame bench.py times by default."""
    c3.tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    fb = c3.tracer.render(rt.make_arg(1920, 1080, 2, spp=1, seed=1234, scene="terrain"))
    ref, ctr = c3.oracle_frame(1920, 1080, 2, 1, 1234)
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(c3.tracer.counters()) == ray_counts(ctr)


def test_c3_full_frame_device_built_tree(c3):
    ref, ctr = c3.oracle_frame(1920, 1080, 2, 1, 1234)
    with rt.RayTracer(soup=c3.soup) as tracer:
        for variant in VARIANTS:
            tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
            fb = tracer.render(rt.make_arg(1920, 1080, 2, spp=1, seed=1234, scene="terrain"))
            assert np.array_equal(fb, ref), f"variant {variant}: {(fb != ref).sum()} pixels differ"
            assert ray_counts(tracer.counters()) == ray_counts(ctr)


def test_c3_weak_scaling_strip(c3):
    """One rank's strip of the N = 8 weak-scaling frame of bench.py (5400 x 3072, rows dealt in blocks of 8)."""
    ref, ctr = c3.oracle_frame(5400, 3072, 2, 1, 1234, ranks=8, rank=3)
    for variant in VARIANTS:
        c3.tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
        fb = c3.tracer.render(rt.make_arg(5400, 3072, 2, spp=1, seed=1234, scene="terrain", rank=3, num_ranks=8,
                                          rows_per_block=RPB))
        assert np.array_equal(fb, ref), f"variant {variant}: {(fb != ref).sum()} pixels differ"
        assert ray_counts(c3.tracer.counters()) == ray_counts(ctr)


@pytest.mark.parametrize("variant", VARIANTS)
def test_c4_strip_of_rank_0_of_64(c45, variant):
    """3840 x 2160 x 8 spp against the 10 M-triangle tree: the rows rank 0 of 64 owns (40 rows, 1.2 M paths)."""
    c45.tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    fb = c45.tracer.render(rt.make_arg(3840, 2160, 2, spp=8, seed=1234, scene="terrain", rank=0, num_ranks=64,
                                       rows_per_block=RPB))
    ref, ctr = c45.oracle_frame(3840, 2160, 2, 8, 1234, ranks=64, rank=0)
    assert fb.shape == ref.shape and fb.shape[0] >= 32
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(c45.tracer.counters()) == ray_counts(ctr)


@pytest.mark.parametrize("variant", VARIANTS)
def test_c5_first_mebi_rays(c45, variant):
    """The first 2^20 of C5's 64 Mi random rays (tmax = 1e30, closest hit), hit records as raw bytes."""
    c45.tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    lo, hi = scene_box(c45.soup)
    rays = rt.rays_random(lo, hi, 0x5EED5EED, 0, 1 << 20)
    hits = c45.tracer.trace(rays)
    ref, _ = o.trace(rays, c45.onodes, c45.otris)
    assert hits.tobytes() == ref.tobytes(), f"{(hits != ref).sum()} hit records differ"
    assert 0.05 < (hits["prim_id"] != MISS).mean() < 1.0


def test_c45_device_built_tree(c45):
    """10 M triangles through the device LBVH build + device repack, C5 rays and a C4 strip."""
    lo, hi = scene_box(c45.soup)
    rays = rt.rays_random(lo, hi, 0x5EED5EED, 0, 1 << 18)
    ref_hits, _ = o.trace(rays, c45.onodes, c45.otris)
    ref_fb, ctr = c45.oracle_frame(3840, 2160, 2, 8, 1234, ranks=64, rank=0)
    with rt.RayTracer(soup=c45.soup) as tracer:
        assert tracer.trace(rays).tobytes() == ref_hits.tobytes()
        fb = tracer.render(rt.make_arg(3840, 2160, 2, spp=8, seed=1234, scene="terrain", rank=0, num_ranks=64,
                                       rows_per_block=RPB))
        assert np.array_equal(fb, ref_fb)
        assert ray_counts(tracer.counters()) == ray_counts(ctr)


# ---------------------------------------------------------------- regressions for ADVICE.md (round 1)
def test_unbounded_rays_tmax_inf(c2):
    """tmax = +inf: empty Node4 slots must stay unreachable (they are inverted infinite boxes)."""
    lo, hi = scene_box(c2.soup)
    rays = rt.rays_random(lo, hi, 77, 0, 100_000)
    rays["tmax"] = np.float32(np.inf)
    # directions with all three components positive and small make 3e38 * (1/d) overflow: the old encoding's hole
    rays["d"][::3] = np.abs(rays["d"][::3]) * np.float32(0.5) + np.float32(1e-3)
    ref, _ = o.trace(rays, c2.onodes, c2.otris)
    for variant in VARIANTS:
        c2.tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
        hits = c2.tracer.trace(rays)
        assert hits.tobytes() == ref.tobytes(), f"variant {variant}: {(hits != ref).sum()} hit records differ"


def test_scene_freed_and_rebuilt_on_one_device(native):
    """vx_mem_free + a new scene of the same size on the same device: the pool hands back the same addresses, and
    the Node4 image derived from the freed buffers must not be reused."""
    soup_a = rt.scene_terrain(60, 40, seed=1)
    soup_b = rt.scene_terrain(60, 40, seed=2)
    assert len(soup_a) == len(soup_b)
    arg = lambda: rt.make_arg(160, 90, 2, spp=1, seed=3, scene="terrain")  # noqa: E731
    refs = []
    for soup in (soup_a, soup_b):
        on, ot = o.bvh2_build(soup)
        refs.append(o.render(o.make_arg(160, 90, 2, spp=1, seed=3, scene="terrain"), on, ot)[0])
    assert not np.array_equal(refs[0], refs[1])
    with vx.Device() as dev:
        for builder in ("host", "device"):
            for soup, ref in ((soup_a, refs[0]), (soup_b, refs[1]), (soup_a, refs[0])):
                tracer = rt.RayTracer(*rt.build_bvh2(soup), device=dev) if builder == "host" else \
                    rt.RayTracer(soup=soup, device=dev)
                for variant in VARIANTS:
                    dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
                    assert np.array_equal(tracer.render(arg()), ref), (builder, variant)
                tracer.close()
