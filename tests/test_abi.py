"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/vortex.h and include/skybox_rt.h declare.  No compute calls (there is no GPU here)."""
import ctypes as C
import os
import re

import pytest

from skybox_rt_b200 import raytrace as rt
from skybox_rt_b200 import vortex as vx
from skybox_rt_b200.formats import KernelArg, kernel_image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header, prefix):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(?:int|void)\s+(%s\w+)\s*\(" % prefix, src)))


def test_libvortex_exports_every_declared_entry_point(native):
    names = declared("vortex.h", "vx_")
    assert len(names) == 22, names  # the 22 functions of reference runtime/include/vortex.h:74-139
    L = C.CDLL(native["vortex"])
    for n in names + ["vx_dev_init"]:
        assert hasattr(L, n), n
    assert sorted(names + ["vx_dev_init"]) == sorted(vx.EXPORTS)
    ext = declared("vortex_cuda.h", "vx_")
    assert sorted(ext) == sorted(vx.CUDA_EXPORTS)
    for n in ext:
        assert hasattr(L, n), n
    assert os.path.exists(os.path.join(os.path.dirname(native["vortex"]), "libvortex-cuda.so"))


def test_host_library_exports_every_declared_entry_point(native):
    names = [n for n in declared("skybox_rt.h", "srt_") if n != "srt_rank_rows"]
    L = C.CDLL(native["host"])
    for n in names:
        assert hasattr(L, n), n
    assert sorted(names) == sorted(rt.HOST_EXPORTS)


def test_plugin_entry_fills_all_sixteen_callbacks(native):
    L = C.CDLL(native["vortex"])
    table = (C.c_void_p * 16)()
    assert L.vx_dev_init(table) == 0
    assert all(table[i] for i in range(16))
    assert L.vx_dev_init(None) != 0


def test_argument_checks_that_need_no_device(native):
    L = vx.lib()
    assert L.vx_dev_open(None) != 0
    assert L.vx_dev_close(None) != 0
    assert L.vx_mem_free(None) == 0          # reference callbacks.inc:100-103
    assert L.vx_mem_alloc(None, 16, 3, C.byref(C.c_void_p())) != 0
    assert L.vx_copy_to_dev(None, None, 0, 0) != 0
    assert L.vx_start(None, None, None) != 0
    assert L.vx_ready_wait(None, 0) != 0
    assert L.vx_upload_bytes(None, None, 0, None) != 0


def test_layouts_match_the_c_header(native):
    out = os.path.join(ROOT, "tests", ".sizeof_kernel_arg")
    src = '#include <stdio.h>\n#include "skybox_rt.h"\nint main(){printf("%zu %zu %zu", sizeof(kernel_arg_t), ' \
          'sizeof(srt_kernel_image_t), sizeof(srt_tri_rec_t));return 0;}\n'
    import subprocess
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", out], input=src.encode(),
                   check=True)
    sizes = subprocess.run([out], capture_output=True, check=True).stdout.decode().split()
    os.remove(out)
    assert [int(s) for s in sizes] == [C.sizeof(KernelArg), 32, 48]
    img = kernel_image()
    assert len(img) == 32 and img[16:24] == b"SKYBOXRT"


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="only meaningful without a GPU")
def test_open_fails_loudly_without_a_gpu(native):
    h = C.c_void_p()
    assert vx.lib().vx_dev_open(C.byref(h)) != 0
    with pytest.raises(vx.VortexError):
        vx.Device()
