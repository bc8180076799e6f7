"""The plug-in route under the REFERENCE's own stub.

oracle/_ref/libvortex.so is runtime/stub/vortex.cpp + utils.cpp compiled from /root/reference by oracle/Makefile
(prebuilt file on the GPU box).  With VORTEX_DRIVER=cuda its vx_dev_open dlopen()s "libvortex-cuda.so" -- this repo's
back-end under the reference's plug-in name -- and calls vx_dev_init (runtime/stub/vortex.cpp:58-97).  A ray-traced
frame and a raster draw then run through the stub's forwarders, its dcr_initialize, its upload helpers and its
vx_dump_perf, and must reproduce the goldens."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB = os.path.join(ROOT, "oracle", "_ref", "libvortex.so")


def run_driver(extra_env=None):
    if not os.path.exists(STUB):
        if os.path.isdir("/root/reference"):
            subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
        else:
            pytest.fail(f"{STUB} is missing: it is built where /root/reference exists (make -C oracle) and travels "
                        "to the GPU box as a prebuilt file")
    env = dict(os.environ)
    libdir = os.path.join(ROOT, "skybox_rt_b200", "lib")
    env.update(SKYBOX_RT_LIB=STUB, VORTEX_DRIVER="cuda",
               LD_LIBRARY_PATH=libdir + os.pathsep + env.get("LD_LIBRARY_PATH", ""))
    env.update(extra_env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "tests", "_ref_stub_driver.py")], env=env,
                          capture_output=True, text=True, timeout=600)


def test_frames_through_the_reference_stub(native):
    p = run_driver()
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "STUB-OK raytrace" in p.stdout and "STUB-OK draw3d" in p.stdout and "STUB-OK closed" in p.stdout
    # the stub's vx_dev_close dumps perf through vx_mpm_query on every core (runtime/stub/vortex.cpp:99-104)
    assert "PERF: instrs=" in p.stdout and "PERF: core0: instrs=" in p.stdout


def test_reference_stub_with_vortex_profiling(native):
    """VORTEX_PROFILING=4 (raster class): the stub writes VX_DCR_BASE_MPM_CLASS before every vx_start
    (stub/vortex.cpp:142-149) and prints the class's PERF: lines from the back-end's CSRs."""
    p = run_driver({"VORTEX_PROFILING": "4"})
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "PERF: raster memory reads=" in p.stdout and "PERF: instrs=" in p.stdout


def test_unknown_driver_fails_like_the_reference(native):
    p = run_driver({"VORTEX_DRIVER": "nosuchdriver"})
    assert p.returncode != 0 and "Cannot open library" in p.stderr
