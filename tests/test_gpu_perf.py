"""The performance monitor behind vx_mpm_query / vx_dump_perf.

Reference: class selection by VX_DCR_BASE_MPM_CLASS / VORTEX_PROFILING (runtime/stub/vortex.cpp:142-149), the
class CSRs (hw/rtl/VX_types.vh:66-247) and the PERF: line formats (runtime/stub/utils.cpp:159-805).  The counts
themselves are checked against what the host pre-pass and the oracle say the same frame contains."""
import os
import re
import subprocess

import numpy as np
import pytest

import _oracle as o
import _oracle_draw3d as od
from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200 import vortex as vx

pytestmark = pytest.mark.gpu
ALL_SW = 7
MPM_CLASS, CLS_CORE, CLS_MEM, CLS_TEX, CLS_RASTER, CLS_OM = 0x005, 1, 2, 3, 4, 5


@pytest.fixture(scope="module")
def d3(native):
    from skybox_rt_b200 import draw3d
    return draw3d


def test_raster_class_counters_accumulate_over_the_frame(d3):
    blob = od.trace_blob("vase")
    t = d3.Trace(blob=blob)
    nd = od.num_draws(blob)
    expect_reads, per_draw = 0, []
    for d in range(nd):
        ntiles, _, _, pairs = t.binning(d, 128, 128)
        expect_reads += ntiles + 2 * pairs  # tile header + (pid word + primitive record) per (tile, primitive) pair
    with d3.Draw3D(128, 128, flags=ALL_SW) as app:
        dev = app.dev
        dev.dcr_write(MPM_CLASS, CLS_RASTER)
        app.clear()
        for d in range(nd):
            app.render(t, d, d)  # [start, end] inclusive, main.cpp:180
            per_draw.append(dev.mpm_query(d3.VX_CSR_RASTER_FRAGMENTS))  # last-run CSR, still per draw
        cores = dev.caps(vx.VX_CAPS_NUM_CORES)
        assert dev.mpm_query(0xB03) == expect_reads == dev.mpm_query(0xB06)       # RASTER_READS, RCACHE_READS
        assert all(dev.mpm_query(0xB03, c) == expect_reads for c in (1, cores - 1))  # every core reports the total
        assert dev.mpm_query(0xB04) == 0 and dev.mpm_query(0xB07) == 0            # latency / misses: no source
        # MINSTRET: fragments started on each SM; over the SMs they are the frame's fragments
        per_sm = [dev.mpm_query(vx.VX_CSR_MINSTRET, c) for c in range(cores)]
        assert sum(per_sm) == sum(per_draw) > 0
        assert dev.mpm_query(vx.VX_CSR_MCYCLE) > 0
        tot = dev.perf_totals()
        assert tot[2] == nd and tot[10] == sum(per_draw)                          # VX_PERF_RUNS, VX_PERF_FRAGMENTS
        with pytest.raises(vx.VortexError):
            dev.mpm_query(vx.VX_CSR_MINSTRET, cores)                              # no such core
        fb = app.read()
    assert np.array_equal(fb, od.golden_framebuffer(os.path.join(od.GOLDEN_DIR, "vase_ref_128.npy")))


def test_tex_and_om_class_counters(d3):
    blob = od.trace_blob("box")  # 2 draws, bilinear texture, depth test on
    t = d3.Trace(blob=blob)
    with d3.Draw3D(128, 128, flags=ALL_SW) as app:
        dev = app.dev
        dev.dcr_write(MPM_CLASS, CLS_TEX)
        app.clear()
        app.render(t)
        frags = dev.perf_totals()[10]
        texels = dev.mpm_query(0xB03)
        assert texels == 4 * frags > 0  # every fragment of this trace takes one bilinear sample
        # switching the class re-labels the class CSRs; the accumulators keep running
        dev.dcr_write(MPM_CLASS, CLS_OM)
        rd, wr = dev.mpm_query(0xB03), dev.mpm_query(0xB04)
        tiles = sum(t.binning(d, 128, 128)[0] for d in range(od.num_draws(blob)))
        assert rd == tiles * 1024           # depth word of every pixel of every tile; colour is not read (no blend)
        assert 0 < wr <= 2 * min(frags, tiles * 1024)
        assert dev.mpm_query(0xB07) == rd and dev.mpm_query(0xB08) == wr
        dev.dcr_write(MPM_CLASS, CLS_MEM)
        assert dev.mpm_query(0xB18) >= rd + texels and dev.mpm_query(0xB19) == wr


def test_class_zero_reports_the_last_run_only(d3):
    t = d3.Trace(blob=od.trace_blob("triangle"))
    with d3.Draw3D(64, 64, flags=ALL_SW) as app:
        app.dev.dcr_write(vx.VX_DCR_RT_STATS, 1)
        app.frame(t)
        a = app.dev.mpm_query(vx.VX_CSR_MINSTRET)
        app.frame(t)
        assert app.dev.mpm_query(vx.VX_CSR_MINSTRET) == a > 0      # not accumulated
        assert app.dev.mpm_query(vx.VX_CSR_MINSTRET, 1) == 0       # unattributed: core 0 carries the total
        assert app.dev.mpm_query(0xB03) == 0
        assert app.dev.perf_totals()[2] == 0


@pytest.mark.parametrize("variant", [vx.RT_VARIANT_DEFAULT, vx.RT_VARIANT_BASELINE, vx.RT_VARIANT_WHILE])
def test_rt_rays_are_attributed_to_sms(native, variant):
    soup = rt.scene_cornell()
    nodes, tris = rt.build_bvh2(soup)
    onodes, otris = o.bvh2_build(soup)
    _, ctr = o.render(o.make_arg(64, 64, 2, spp=2, seed=7), onodes, otris)
    rays = ctr["rays_primary"] + ctr["rays_shadow"] + ctr["rays_bounce"]
    with rt.RayTracer(nodes, tris) as tracer:
        dev = tracer.dev
        dev.dcr_write(vx.VX_DCR_RT_VARIANT, variant)
        dev.dcr_write(MPM_CLASS, CLS_MEM)
        tracer.render(rt.make_arg(64, 64, 2, spp=2, seed=7))
        tracer.render(rt.make_arg(64, 64, 2, spp=2, seed=7))
        cores = dev.caps(vx.VX_CAPS_NUM_CORES)
        assert sum(dev.mpm_query(vx.VX_CSR_MINSTRET, c) for c in range(cores)) == 2 * rays
        c = tracer.counters()
        assert dev.mpm_query(0xB18) == 2 * (c["nodes"] + c["tris"]) > 0   # MEM_READS: node + triangle records
        assert dev.mpm_query(0xB19) == 2 * 64 * 64                        # MEM_WRITES: pixels


def _perf_lines(text):
    return [l for l in text.splitlines() if l.startswith("PERF:")]


def test_dump_perf_prints_the_reference_lines(native, tmp_path):
    """VORTEX_PROFILING=<class> as the reference's scripts use it (ci/blackbox.sh --perf=N): vx_dev_close dumps."""
    bindir = os.path.join(os.path.dirname(native["vortex"]), "..", "bin")
    trc = os.path.join(od.GOLDEN_DIR, "vase.trc.gz")
    want = {
        CLS_RASTER: [r"raster memory reads=(\d+)", r"raster memory latency=0 cycles", r"rcache reads=(\d+)",
                     r"rcache read misses=0 \(hit ratio=100%\)"],
        CLS_TEX: [r"tex memory reads=(\d+)", r"tcache reads=(\d+)", r"tcache mshr stalls=0"],
        CLS_OM: [r"om memory reads=(\d+)", r"om memory writes=(\d+)", r"ocache writes=(\d+)"],
        CLS_MEM: [r"memory requests=(\d+) \(reads=(\d+), writes=(\d+)\)", r"memory latency=0 cycles"],
        CLS_CORE: [r"loads=(\d+)", r"stores=(\d+)"],
    }
    for cls, pats in want.items():
        env = dict(os.environ, VORTEX_PROFILING=str(cls))
        p = subprocess.run([os.path.join(bindir, "draw3d"), "-t", trc, "-w", "128", "-h", "128", "-x", "-y", "-u", "-o",
                            str(tmp_path / "o.png")], capture_output=True, text=True, timeout=300, env=env)
        assert p.returncode == 0, p.stdout + p.stderr
        lines = _perf_lines(p.stdout)
        for pat in pats:
            m = [re.search("PERF: " + pat, l) for l in lines]
            m = [x for x in m if x]
            assert m, f"class {cls}: no line matching {pat!r} in\n" + "\n".join(lines)
            assert all(int(g) > 0 for g in m[0].groups())
        tot = re.search(r"PERF: instrs=(\d+), cycles=(\d+), IPC=([\d.]+)", lines[-1])  # the line every script greps
        assert tot and int(tot.group(1)) > 0 and int(tot.group(2)) > 0
        core = [re.match(r"PERF: core(\d+): instrs=(\d+), cycles=(\d+), IPC=", l) for l in lines]
        assert sum(int(m.group(2)) for m in core if m) == int(tot.group(1))
        assert any(l.startswith("PERF: raster: tiles=") for l in lines)
    # without the variable: the totals line only, describing the last run
    env = {k: v for k, v in os.environ.items() if k != "VORTEX_PROFILING"}
    p = subprocess.run([os.path.join(bindir, "raytrace"), "-w", "64", "-h", "64"], capture_output=True, text=True,
                       timeout=300, env=env)
    assert p.returncode == 0, p.stdout + p.stderr
    p = subprocess.run([os.path.join(bindir, "raytrace"), "-w", "64", "-h", "64"], capture_output=True, text=True,
                       timeout=300, env=dict(env, VORTEX_PROFILING="2"))
    assert p.returncode == 0, p.stdout + p.stderr
    lines = _perf_lines(p.stdout)
    runs = int([re.match(r"PERF: runs=(\d+)", l) for l in lines if l.startswith("PERF: runs=")][0].group(1))
    assert runs >= 1
    assert any(re.match(rf"PERF: rt: rays_primary={4096 * runs}, ", l) for l in lines), "\n".join(lines)
    assert any(re.match(rf"PERF: memory requests=\d+ \(reads=\d+, writes={4096 * runs}\)", l) for l in lines)
