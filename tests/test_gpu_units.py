"""The golden images of the reference's three sub-unit regression apps (tests/regression/{tex,raster,om}; SURVEY.md
section 2 #12, section 8c) against the sm_100a raster path, through the vx_* C ABI.  Bar: tolerance 0, as the apps
compare (tex/main.cpp:268, raster/main.cpp:272, om/main.cpp:295).

The product has one raster kernel (the draw3d path), not three unit test kernels, so the units are reached through
draws that isolate them (tests/_oracle_draw3d.py: quad_trace): a viewport-filling textured quad with texture
environment REPLACE samples exactly where the tex app samples; a white quad is the om app's constant-colour box; the
coverage of the triangle trace is what the raster app writes.  tests/test_oracle_units.py holds the same images
against the oracle and the compiled reference on the CPU."""
import numpy as np
import pytest

import _oracle_draw3d as od

pytestmark = pytest.mark.gpu
ALL_SW = 7  # -x -y -u: unit state from the kernel_arg mirrors; 0: from the device DCR file ("hardware" units)
UNIT_CASES = od.tex_unit_cases()


@pytest.fixture(scope="module")
def d3(native):
    from skybox_rt_b200 import draw3d
    from skybox_rt_b200 import vortex
    vortex.lib()  # raises if libvortex.so is missing: no fallback
    return draw3d


@pytest.mark.parametrize("case", UNIT_CASES, ids=[c[0] for c in UNIT_CASES])
def test_tex_unit_goldens_and_reference_rendered_vectors(d3, case):
    """19 golden PNGs (7 formats, point / bilinear) + the level-0 vectors the reference's kernel rendered in the build
    container (clamp / repeat, magnification, all formats): the frame equals the stored image; sizes whose coordinates
    the raster path computes differently from the tex app (not a power of two) are held against the oracle's frame."""
    name, blob, w, h, want = case
    with d3.Draw3D(w, h, flags=ALL_SW) as app:
        fb = app.frame(d3.Trace(blob=blob))
    if want is None:
        want = od.render(od.oracle(), blob, w, h, threads=4)[0]
    assert fb.shape == want.shape and int((fb != want).sum()) == 0, name


@pytest.mark.parametrize("name", ["toad_ref_f0", "toad_ref_f3", "soccer_ref_g1", "palette4_ref_g2"])
def test_tex_unit_goldens_with_unit_state_from_the_device_dcrs(d3, name):
    _, blob, w, h, want = next(c for c in UNIT_CASES if c[0] == name)
    with d3.Draw3D(w, h, flags=0) as app:
        fb = app.frame(d3.Trace(blob=blob))
    assert int((fb != want).sum()) == 0


@pytest.mark.parametrize("size", [8, 16, 32, 64, 128])
@pytest.mark.parametrize("flags", [ALL_SW, 0])
def test_raster_unit_golden_is_the_coverage_of_the_triangle_frame(d3, flags, size):
    """raster/kernel.cpp stores white for every pixel vx_rast reports over a 0xff000000 clear; the app saves bottom-up."""
    with d3.Draw3D(size, size, flags=flags) as app:
        fb = app.frame(d3.Trace(blob=od.trace_blob("triangle")))
    gold = od.unit_fixtures("raster")[f"triangle_ref_{size}"][::-1]
    assert np.array_equal(fb != 0xFF000000, gold == 0xFFFFFFFF)


@pytest.mark.parametrize("size", [8, 16, 32, 64, 128])
@pytest.mark.parametrize("flags", [ALL_SW, 0])
def test_om_unit_golden_is_a_white_quad(d3, flags, size):
    """om/kernel.cpp sends 0xffffffff for every pixel through the default output-merger state -> whitebox_<n>.png."""
    with d3.Draw3D(size, size, flags=flags) as app:
        fb = app.frame(d3.Trace(blob=od.quad_trace(size, size)))
    assert np.array_equal(fb, od.unit_fixtures("om")[f"whitebox_{size}"])
