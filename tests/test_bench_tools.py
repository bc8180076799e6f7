"""CPU tests of the measurement tooling: bench.py's helpers (weak-scaling frame, clock sampler window, roofline
inputs) and the profile post-processing (profiles/make_roofline_inputs.py, profiles/ncu_lines.py) -- the parts of
the record a reader recomputes.  No GPU, no ncu."""
import json
import os
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "profiles"))

import bench  # noqa: E402


@pytest.mark.parametrize("n", [2, 3, 4, 8])
def test_weak_frame_keeps_pixels_per_rank_and_row_blocks_equal(n):
    w, h = bench.weak_frame(1920, 1080, n)
    assert h % (bench.ROWS_PER_BLOCK * n) == 0 and w % 8 == 0
    assert abs(w * h / (1920 * 1080 * n) - 1.0) < 0.02
    from skybox_rt_b200.formats import rank_rows
    rows = [rank_rows(h, bench.ROWS_PER_BLOCK, n, r) for r in range(n)]
    assert len(set(rows)) == 1 and sum(rows) == h


def test_clock_sampler_reports_only_samples_inside_the_window():
    s = bench.ClockSampler(0)

    class FakeNvml:  # stands where pynvml stands once start() succeeded
        pass
    s.nvml, s.max_mhz, s.stop_flag = FakeNvml(), 1965.0, True
    s.th = type("T", (), {"join": lambda self, timeout=None: None})()
    t0 = time.perf_counter()
    s.rows = [(t0 - 1.0, 300.0, []), (t0 + 0.001, 1965.0, []), (t0 + 0.002, 1950.0, ["sw_power_cap"]), (t0 + 5.0, 210.0, ["hw_slowdown"])]
    s.window(t0, t0 + 0.01)
    rec = s.stop()
    assert rec["samples"] == 2 and rec["samples_total"] == 4
    assert rec["sm_mhz"] == pytest.approx(1957.5) and rec["reasons"] == ["sw_power_cap"]
    assert rec["scope"] == "inside the timed region"
    s.window(t0 + 100, t0 + 101)  # nothing inside: the record says so instead of inventing a clock
    rec = s.stop()
    assert rec["samples"] == 4 and rec["scope"].startswith("whole run")


def test_roofline_inputs_file_matches_its_sources():
    import make_roofline_inputs as m
    with open(os.path.join(ROOT, "profiles", "roofline_inputs.json")) as f:
        data = json.load(f)
    assert "c3:v0" in data and "c5:v0" in data and "c4:v0" in data and "c2:v0" in data
    for key, rec in data.items():
        path = os.path.join(ROOT, rec["source"])
        assert os.path.exists(path), f"{key}: {rec['source']} is not committed"
        ks = m.parse(path)
        assert len(ks) == rec["kernels"]
        assert int(sum(k["smsp__inst_executed.sum"] for k in ks)) == rec["warp_inst"]
        assert 0.0 < rec["issue_active"] <= 1.0 and 1.0 <= rec["lanes_per_inst"] <= 32.0
    # the partial captures say what share of the bench step they cover
    assert data["c5:v0"]["step_fraction"] == 0.25 and "partial" in data["c5:v0"]


def test_profile_inputs_are_refused_for_configurations_never_captured():
    class A:
        gpus, builder, spp, mode, variant = 1, "host", None, None, 0
    assert bench.profile_inputs(A, "c3")["source"].startswith("profiles/")
    A.gpus = 2
    assert bench.profile_inputs(A, "c3") is None
    A.gpus, A.builder = 1, "device"
    assert bench.profile_inputs(A, "c3") is None
    A.builder, A.variant = "host", 6
    assert bench.profile_inputs(A, "c3") is None  # variant 6 was never captured with --set full


def test_ncu_lines_function_ranges_find_the_traversal_phases():
    import ncu_lines
    src = open(os.path.join(ROOT, "skybox_rt_b200", "csrc", "rt_wide.cuh")).read().split("\n")
    ranges = {name: (a, b) for name, a, b in ncu_lines.function_ranges({n: t for n, t in enumerate(src, 1)})}
    for fn in ("node_step", "tri_test", "tri_step", "take_jobs", "run_warp", "pop_next"):
        assert fn in ranges, fn
    a, b = ranges["node_step"]
    body = "\n".join(src[a - 1:b])
    assert "SRT_CHILD(w, t3, h3)" in body and "take_jobs" not in body
    phases = {(f, n): p for f, n, p in ncu_lines.PHASES}
    assert phases[("rt_wide.cuh", "node_step")] == "node" and phases[("rt_wide.cuh", "run_warp")] == "scheduler"


def test_workload_table_is_the_baseline_configs():
    with open(os.path.join(ROOT, "BASELINE.json")) as f:
        base = json.load(f)
    assert len(base["configs"]) == 5 and sorted(bench.WORKLOADS) == ["c1", "c2", "c3", "c4", "c5"]
    assert bench.WORKLOADS["c3"]["w"] == 1920 and bench.WORKLOADS["c3"]["h"] == 1080 and bench.WORKLOADS["c3"]["quads"] == (1000, 500)
    assert bench.WORKLOADS["c4"]["spp"] == 8 and bench.WORKLOADS["c5"]["rays"] == 64 << 20
