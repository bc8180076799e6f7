"""CPU test of the draw3d app's host flows against a recording stand-in for the back-end (tests/native/
draw3d_flow_check.c): the SRT_DRAW3D_AUTO flow -- host pre-pass on the app's thread pool, the frame packed into one
arena upload, draws started back to back -- must hand every vx_start what the reference's serial flow hands it (same
DCR file and kernel_arg_t mirrors, same tile headers, primitive lists, primitive records and texels), and must make
every back-end call from the calling thread.  Run plain and, where the tool chain has it, under ThreadSanitizer: the
race check SURVEY.md section 5 asks of anything threaded, here without a GPU."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SOURCES = [os.path.join(ROOT, p) for p in (
    "tests/native/draw3d_flow_check.c", "skybox_rt_b200/apps/draw3d/draw3d_app.c", "skybox_rt_b200/csrc/host/draw3d_binning.c",
    "skybox_rt_b200/csrc/host/draw3d_trace.c", "skybox_rt_b200/csrc/host/draw3d_cgltrace.c")]
TRACES = ["mouse", "vase", "scene", "tekkaman", "evilskull"]


def compile_check(cc, out, extra):
    cmd = [cc, "-O1", "-g", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"), "-o", out] + extra + SOURCES + [
        "-pthread", "-lz", "-lm"]
    return subprocess.run(cmd, capture_output=True, text=True)


@pytest.fixture(scope="module")
def plain(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("flows") / "flow_check")
    r = compile_check(shutil.which("gcc") or "cc", out, ["-Wall", "-Wextra", "-Werror"])
    assert r.returncode == 0, r.stderr
    return out


@pytest.fixture(scope="module")
def tsan(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("flows_tsan") / "flow_check_tsan")
    for cc in ("/usr/bin/gcc", shutil.which("gcc"), shutil.which("clang")):
        if cc and compile_check(cc, out, ["-fsanitize=thread"]).returncode == 0:
            return out
    pytest.skip("no compiler here links -fsanitize=thread")


def run(exe, trace, size, frames, env=None):
    path = os.path.join(ROOT, "tests", "golden", "raster", trace + ".trc.gz")
    return subprocess.run([exe, path, str(size), str(frames)], capture_output=True, text=True, timeout=600,
                          env=dict(os.environ, **(env or {})))


@pytest.mark.parametrize("trace", TRACES)
@pytest.mark.parametrize("size", [256, 2048])
def test_auto_flow_hands_every_launch_what_the_serial_flow_does(plain, trace, size):
    r = run(plain, trace, size, 7)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "PASSED!" in r.stdout
    serial, auto = [ln for ln in r.stdout.splitlines() if ln.startswith("flags")]
    # the packed frame is a handful of uploads (arena pieces + one argument block per draw), not three per draw;
    # a trace of one draw (scene) has nothing to pack and takes the serial path
    n_auto, n_serial = (int(ln.split(" uploads")[0].split()[-1]) for ln in (auto, serial))
    assert n_auto <= n_serial and (n_auto < n_serial or trace == "scene")


@pytest.mark.parametrize("trace", ["mouse", "vase", "tekkaman"])
def test_auto_flow_is_race_free_under_thread_sanitizer(tsan, trace):
    r = run(tsan, trace, 1024, 16, env={"TSAN_OPTIONS": "halt_on_error=1 exitcode=66"})
    assert r.returncode == 0 and "PASSED!" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
    assert "ThreadSanitizer" not in r.stderr
