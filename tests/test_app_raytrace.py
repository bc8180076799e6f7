"""The C regression app (skybox_rt_b200/apps/raytrace/main.c) end to end: host C -> vx_* C ABI -> CUDA,
output PNG compared by the app itself against the frozen oracle renders (tol 0), exit code = #errors --
the way the reference runs tests/regression/draw3d (-r reference.png, draw3d/main.cpp:505-514)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def app(native):
    exe = os.path.join(ROOT, "skybox_rt_b200", "bin", "raytrace")
    assert os.path.exists(exe), "python -m skybox_rt_b200.build did not produce bin/raytrace"
    return exe


@pytest.mark.gpu
@pytest.mark.parametrize("name,mode,spp", [("cornell_64_primary", 0, 1), ("cornell_64_shadow", 1, 1),
                                           ("cornell_64_bounce", 2, 1), ("cornell_64_bounce_spp4", 2, 4)])
def test_app_reproduces_golden_png(native, tmp_path, name, mode, spp):
    out = str(tmp_path / "output.png")
    r = subprocess.run([app(native), "-w", "64", "-h", "64", "-m", str(mode), "-p", str(spp), "-e", "1234", "-o", out,
                        "-r", os.path.join(GOLDEN, name + ".png")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "PASSED!" in r.stdout


@pytest.mark.gpu
def test_app_reports_mismatch_as_exit_code(native, tmp_path):
    out = str(tmp_path / "output.png")
    r = subprocess.run([app(native), "-w", "64", "-h", "64", "-m", "0", "-o", out, "-r",
                        os.path.join(GOLDEN, "cornell_64_bounce.png")], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "FAILED!" in r.stdout


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="only meaningful without a GPU")
def test_app_fails_loudly_without_a_gpu(native, tmp_path):
    r = subprocess.run([app(native), "-o", str(tmp_path / "o.png")], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0
    assert "vx_dev_open" in r.stdout


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="only meaningful without a GPU")
def test_multi_rank_app_reports_a_failed_rank_as_its_exit_code(native, tmp_path):
    """-g 2 forks its ranks; when they cannot open a device the parent must come back with a non-zero code instead
    of waiting for a communicator that will never form."""
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r = subprocess.run([app(native), "-g", "2", "-o", str(tmp_path / "o.png")], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode != 0
    assert "failed (exit code" in r.stdout  # the parent names the rank: ranks other than 0 print nothing themselves
