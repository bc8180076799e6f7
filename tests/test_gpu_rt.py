"""GPU parity tests (run on the B200 box): the CUDA path, called through the vx_* C ABI, against the
CPU oracle on the same seeded inputs and against the frozen goldens.  Bar: bit-exact pixels, hit
records (t, u, v, prim_id as raw bits) and ray counts."""
import ctypes as C
import os

import numpy as np
import pytest

import _oracle as o
from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200 import vortex as vx
from skybox_rt_b200.formats import HIT, MISS, RAY

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VARIANTS = sorted(vx.RT_VARIANTS)


@pytest.fixture(scope="module")
def cornell(native):
    soup = rt.scene_cornell()
    nodes, tris = rt.build_bvh2(soup)
    with rt.RayTracer(nodes, tris) as tracer:
        yield tracer, o.bvh2_build(soup)


@pytest.fixture(scope="module")
def terrain(native):
    soup = rt.scene_terrain(120, 90)  # 21 600 triangles
    nodes, tris = rt.build_bvh2(soup)
    with rt.RayTracer(nodes, tris) as tracer:
        yield tracer, o.bvh2_build(soup), soup


def ray_counts(c):
    return [c[k] for k in ("rays_primary", "rays_shadow", "rays_bounce")]


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("name,mode,spp", [("cornell_64_primary", 0, 1), ("cornell_64_shadow", 1, 1),
                                           ("cornell_64_bounce", 2, 1), ("cornell_64_bounce_spp4", 2, 4)])
def test_cornell_golden_images(cornell, variant, name, mode, spp):
    tracer, (onodes, otris) = cornell
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    fb = tracer.render(rt.make_arg(64, 64, mode, spp=spp, seed=1234))
    gold = np.load(os.path.join(GOLDEN, name + ".npy"))
    assert np.array_equal(fb, gold), f"{(fb != gold).sum()} pixels differ"
    _, ctr = o.render(o.make_arg(64, 64, mode, spp=spp, seed=1234), onodes, otris)
    assert ray_counts(tracer.counters()) == ray_counts(ctr)


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("w,h,mode,spp", [(320, 180, 2, 1), (201, 97, 2, 3), (64, 33, 1, 1), (7, 5, 0, 2)])
def test_terrain_frames_match_oracle(terrain, variant, w, h, mode, spp):
    tracer, (onodes, otris), _ = terrain
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    fb = tracer.render(rt.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"))
    ref, ctr = o.render(o.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"), onodes, otris)
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(tracer.counters()) == ray_counts(ctr)


@pytest.mark.parametrize("chains", [1, 2, 3, 4])
@pytest.mark.parametrize("w,h,mode,spp", [(320, 180, 2, 1), (201, 97, 2, 3), (64, 33, 1, 1), (7, 5, 0, 2), (40, 24, 2, 1)])
def test_wavefront_chains_do_not_change_the_frame(terrain, chains, w, h, mode, spp):
    """VX_DCR_RT_CHAINS splits a frame into independent chains on separate streams (tail overlap); 0 picks 2 only for
    large frames, so the split is forced here on small ones, including frames with fewer 256-job groups than chains."""
    tracer, (onodes, otris), _ = terrain
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, vx.RT_VARIANT_WAVE)
    tracer.dev.dcr_write(vx.VX_DCR_RT_CHAINS, chains)
    try:
        fb = tracer.render(rt.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"))
        c = tracer.counters()
    finally:
        tracer.dev.dcr_write(vx.VX_DCR_RT_CHAINS, 0)
    ref, ctr = o.render(o.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"), onodes, otris)
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(c) == ray_counts(ctr)


@pytest.mark.parametrize("variant", [vx.RT_VARIANT_WAVE, vx.RT_VARIANT_WAVE_VOTE, vx.RT_VARIANT_WAVE_DEFER])
@pytest.mark.parametrize("w,h,mode,spp", [(320, 180, 2, 1), (201, 97, 2, 3), (64, 33, 1, 1)])
def test_wavefront_with_separate_bounce_shading_launches(terrain, variant, w, h, mode, spp):
    """VX_DCR_RT_FUSE_BOUNCE = 1: the bounce hit is shaded by its own pass (K4) and its shadow rays traced by their own
    launch (K5) instead of inside the shadow + bounce trace launch -- the A/B form of the default; same frame."""
    tracer, (onodes, otris), _ = terrain
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    tracer.dev.dcr_write(vx.VX_DCR_RT_FUSE_BOUNCE, 1)
    try:
        fb = tracer.render(rt.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"))
        c = tracer.counters()
    finally:
        tracer.dev.dcr_write(vx.VX_DCR_RT_FUSE_BOUNCE, 0)
    ref, ctr = o.render(o.make_arg(w, h, mode, spp=spp, seed=99, scene="terrain"), onodes, otris)
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(c) == ray_counts(ctr)


@pytest.mark.parametrize("pass_jobs", [1, 45000, 70000])  # 1 / 2 / 3 samples per pass of the 20 800-job plane
@pytest.mark.parametrize("chains", [1, 2])
def test_wavefront_sample_passes_keep_the_sample_order(terrain, pass_jobs, chains):
    """A frame whose pixels x samples exceed the queue budget runs as several passes with a running per-pixel sum;
    the additions must stay in sample order (bit-identical frame).  VX_DCR_RT_PASS_JOBS shrinks the budget."""
    tracer, (onodes, otris), _ = terrain
    dev = tracer.dev
    dev.dcr_write(vx.VX_DCR_RT_VARIANT, vx.RT_VARIANT_WAVE)
    dev.dcr_write(vx.VX_DCR_RT_CHAINS, chains)
    dev.dcr_write(vx.VX_DCR_RT_PASS_JOBS, pass_jobs)
    try:
        fb = tracer.render(rt.make_arg(201, 97, 2, spp=5, seed=41, scene="terrain"))
        c = tracer.counters()
    finally:
        dev.dcr_write(vx.VX_DCR_RT_CHAINS, 0)
        dev.dcr_write(vx.VX_DCR_RT_PASS_JOBS, 0)
    ref, ctr = o.render(o.make_arg(201, 97, 2, spp=5, seed=41, scene="terrain"), onodes, otris)
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} pixels differ"
    assert ray_counts(c) == ray_counts(ctr)


@pytest.mark.parametrize("variant", VARIANTS)
def test_explicit_rays_hit_records_bit_exact(terrain, variant):
    tracer, (onodes, otris), soup = terrain
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    lo = np.minimum.reduce([soup["v0"].min(0), soup["v1"].min(0), soup["v2"].min(0)])
    hi = np.maximum.reduce([soup["v0"].max(0), soup["v1"].max(0), soup["v2"].max(0)])
    rays = rt.rays_random(lo, hi, 5, 0, 200_000)
    rays["tmin"][::7] = 0.5
    rays["tmax"][::5] = 30.0
    hits = tracer.trace(rays)
    ref, _ = o.trace(rays, onodes, otris)
    assert hits.tobytes() == ref.tobytes(), f"{(hits != ref).sum()} hit records differ"
    assert 0.2 < (hits["prim_id"] != MISS).mean() < 1.0
    # a slice of them against the O(N) search as well
    assert hits[:300].tobytes() == o.closest_brute(rays[:300], otris).tobytes()


@pytest.mark.parametrize("variant", VARIANTS)
def test_ragged_and_empty_inputs(cornell, variant):
    tracer, (onodes, otris) = cornell
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    assert len(tracer.trace(np.zeros(0, dtype=RAY))) == 0
    one = np.zeros(1, dtype=RAY)
    one["o"], one["d"], one["tmax"] = (0, 0, -3), (0, 0, 1), 1e30
    assert tracer.trace(one).tobytes() == o.trace(one, onodes, otris)[0].tobytes()
    for w, h in [(1, 1), (3, 2), (9, 31), (33, 1)]:
        fb = tracer.render(rt.make_arg(w, h, 2, spp=2))
        ref, _ = o.render(o.make_arg(w, h, 2, spp=2), onodes, otris)
        assert np.array_equal(fb, ref), (w, h)


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("ranks,rpb", [(2, 4), (3, 8), (8, 4)])
def test_rank_strips_reassemble_to_single_gpu_frame(cornell, variant, ranks, rpb):
    tracer, _ = cornell
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    full = tracer.render(rt.make_arg(70, 50, 2, spp=2)).copy()
    strips = [tracer.render(rt.make_arg(70, 50, 2, spp=2, rank=r, num_ranks=ranks, rows_per_block=rpb)).copy()
              for r in range(ranks)]
    assert np.array_equal(rt.assemble_strips(strips, 50, rpb, ranks), full)


def test_single_triangle_scene_and_degenerate_triangles(native):
    soup = np.zeros(3, dtype=rt.TRIANGLE)
    soup["v0"], soup["v1"], soup["v2"] = (0, 0, 5), (4, 0, 5), (0, 4, 5)
    soup["v1"][1] = soup["v0"][1]  # zero-area triangle
    soup["v2"][1] = soup["v0"][1]
    soup["albedo"] = 0xff808080
    nodes, tris = rt.build_bvh2(soup)
    rays = rt.rays_random(np.float32([-1, -1, 0]), np.float32([5, 5, 1]), 3, 0, 5000)
    rays["d"] = (0, 0, 1)
    with rt.RayTracer(nodes, tris) as tracer:
        for variant in VARIANTS:
            tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
            hits = tracer.trace(rays)
            assert hits.tobytes() == o.trace(rays, nodes, tris)[0].tobytes()
            assert set(np.unique(hits["prim_id"])) <= {0, 2, MISS} and (hits["prim_id"] == 0).any()


# ---------------------------------------------------------------- boundary behaviour on a live device
def test_vx_api_error_conventions(native):
    L = vx.lib()
    with vx.Device() as dev:
        assert dev.caps(vx.VX_CAPS_NUM_THREADS) == 32
        assert dev.caps(vx.VX_CAPS_NUM_CORES) >= 100
        assert dev.caps(vx.VX_CAPS_ISA_FLAGS) & vx.VX_ISA_EXT_RT
        v = C.c_uint64()
        assert L.vx_dev_caps(dev.handle, 0x99, C.byref(v)) != 0
        buf = dev.alloc(1024)
        assert buf.address % 64 == 0                       # reference block alignment
        data = np.arange(256, dtype=np.uint32)
        buf.write(data)
        assert np.array_equal(buf.read(np.zeros(256, np.uint32)), data)
        assert L.vx_copy_to_dev(buf.handle, data.ctypes.data, 8, 1024) != 0      # offset+size > size
        assert L.vx_copy_from_dev(data.ctypes.data, buf.handle, 1020, 8) != 0
        h = C.c_void_p()
        assert L.vx_mem_alloc(dev.handle, 0, 3, C.byref(h)) != 0
        assert L.vx_mem_free(None) == 0
        dev.dcr_write(0x001, 0x1234)                        # VX_DCR_BASE_STARTUP_ADDR0
        assert dev.dcr_read(0x001) == 0x1234
        assert L.vx_dcr_write(dev.handle, 0x700, 1) != 0    # outside every DCR range
        junk = dev.upload(np.zeros(64, np.uint8))
        assert L.vx_start(dev.handle, junk.handle, junk.handle) != 0   # not a pipeline descriptor
        dev.ready_wait(1000)                                # idle device: returns at once
        free_b, used_b = dev.mem_info()
        assert used_b >= 1024 and free_b > 0
        buf.free()
        # leaked `junk` is released by vx_dev_close, like the per-draw args buffers of draw3d


def test_start_is_asynchronous_and_wait_times_out(terrain):
    tracer, _, _ = terrain
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, vx.RT_VARIANT_DEFAULT)
    arg = rt.make_arg(2048, 2048, 2, spp=4, scene="terrain")
    tracer.start(arg)
    rc = vx.lib().vx_ready_wait(tracer.dev.handle, 0)      # 0 ms: still running -> -1, or already done
    assert rc in (0, -1)
    tracer.dev.ready_wait()
    c = tracer.counters()
    assert c["rays_primary"] == 2048 * 2048 * 4 and c["elapsed_ns"] > 0
    assert HIT.itemsize == 16


@pytest.mark.parametrize("variant", [vx.RT_VARIANT_DEFAULT, vx.RT_VARIANT_WHILE])
@pytest.mark.parametrize("w,h,mode,spp,ranks", [(201, 97, 2, 3, 1), (64, 64, 2, 8, 1), (130, 70, 1, 2, 3), (33, 9, 2, 1, 1)])
def test_sample_split_jobs_render_the_same_frame(terrain, variant, w, h, mode, spp, ranks):
    """Multi-sample frames run as (pixel, sample) jobs + an ordered resolve (VX_DCR_RT_SPLIT_SAMPLES); the
    additions are the pixel loop's, in its order, so the frame is the one-job-per-pixel frame bit for bit."""
    tracer, (onodes, otris), _ = terrain
    tracer.dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
    try:
        for r in range(ranks):
            frames, counts = [], []
            for split in (1, 0, 2):  # never / when spp > 1 / always (also for spp = 1)
                tracer.dev.dcr_write(vx.VX_DCR_RT_SPLIT_SAMPLES, split)
                frames.append(tracer.render(rt.make_arg(w, h, mode, spp=spp, seed=5, scene="terrain", rank=r,
                                                        num_ranks=ranks, rows_per_block=8)))
                counts.append(ray_counts(tracer.counters()))
            assert np.array_equal(frames[0], frames[1]) and np.array_equal(frames[0], frames[2])
            assert counts[0] == counts[1] == counts[2]
        if ranks == 1:
            ref, _ = o.render(o.make_arg(w, h, mode, spp=spp, seed=5, scene="terrain"), onodes, otris)
            assert np.array_equal(frames[0], ref)
    finally:
        tracer.dev.dcr_write(vx.VX_DCR_RT_SPLIT_SAMPLES, 0)
