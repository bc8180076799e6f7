"""In-tree native build of the B200 back-end.

    python -m skybox_rt_b200.build            # build what is stale
    python -m skybox_rt_b200.build --force

Outputs (git-ignored, shipped to the GPU box with the repo snapshot):
    skybox_rt_b200/lib/libvortex.so           vx_* C ABI + CUDA kernels (sm_100a), nvcc
    skybox_rt_b200/lib/libvortex-cuda.so      same file under the reference's plugin name
    skybox_rt_b200/lib/libskybox_rt_host.so   host pre-pass (scenes, BVH2 builder), gcc, no CUDA
    skybox_rt_b200/bin/raytrace               the regression app (C host program)
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "lib")
BIN = os.path.join(PKG, "bin")
INC = os.path.join(ROOT, "include")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CC = os.environ.get("CC", "gcc")

# -fmad=false etc.: the bit-exactness contract of csrc/rt_device.cuh
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-Xcompiler", "-Wall",
    "-I", INC, "-I", CSRC,
]
NCCL_LINK = []  # filled in when the NCCL frame gather is built in (csrc/vx_nccl.cu)
HOST_CFLAGS = ["-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-Wall", "-Wextra", "-I", INC]


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _glob(d, exts):
    out = []
    for base, _, files in os.walk(d):
        for f in sorted(files):
            if f.endswith(exts):
                out.append(os.path.join(base, f))
    return out


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)


def build_experiment(tag, defines, verbose=False):
    """A/B helper: build the CUDA back-end alone with extra -D macros into lib/exp_<tag>/libvortex.so
    (select it at run time with SKYBOX_RT_LIB=<path>).  Not part of the product build."""
    out_dir = os.path.join(LIB, "exp_" + tag)
    os.makedirs(out_dir, exist_ok=True)
    cu_src = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cpp"))]
    out = os.path.join(out_dir, "libvortex.so")
    _run([NVCC] + NVCC_FLAGS + ["-D" + d for d in defines] + ["-shared", "-o", out] + cu_src, verbose)
    return out


def build(force=False, verbose=False):
    os.makedirs(LIB, exist_ok=True)
    os.makedirs(BIN, exist_ok=True)
    headers = _glob(INC, (".h",)) + _glob(CSRC, (".h", ".cuh"))

    # 1. host pre-pass library (plain C)
    host_src = _glob(os.path.join(CSRC, "host"), (".c",))
    host_so = os.path.join(LIB, "libskybox_rt_host.so")
    if force or _stale(host_so, host_src + headers):
        _run([CC] + HOST_CFLAGS + ["-shared", "-o", host_so] + host_src + ["-lm", "-lz"], verbose)

    # 2. the CUDA back-end
    cu_src = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cpp"))]
    vx_so = os.path.join(LIB, "libvortex.so")
    if force or _stale(vx_so, cu_src + headers):
        # one object per translation unit, compiled side by side (objects live in the git-ignored build/)
        from concurrent.futures import ThreadPoolExecutor
        obj_dir = os.path.join(PKG, "build")
        os.makedirs(obj_dir, exist_ok=True)
        objs = [os.path.join(obj_dir, os.path.basename(f) + ".o") for f in cu_src]
        todo = [(f, o) for f, o in zip(cu_src, objs) if force or _stale(o, [f] + headers)]
        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
            list(ex.map(lambda fo: _run([NVCC] + NVCC_FLAGS + ["-c", fo[0], "-o", fo[1]], verbose), todo))
        _run([NVCC, "-shared", "-o", vx_so] + objs + NCCL_LINK, verbose)
    plugin = os.path.join(LIB, "libvortex-cuda.so")
    if force or _stale(plugin, [vx_so]):
        shutil.copyfile(vx_so, plugin)

    # 3. the regression app
    app_dir = os.path.join(PKG, "apps", "raytrace")
    app_src = _glob(app_dir, (".c",)) if os.path.isdir(app_dir) else []
    if app_src:
        exe = os.path.join(BIN, "raytrace")
        if force or _stale(exe, app_src + headers + [vx_so, host_so]):
            _run([CC, "-O2", "-ffp-contract=off", "-Wall", "-Wextra", "-I", INC, "-o", exe] + app_src +
                 ["-L", LIB, "-lvortex", "-lskybox_rt_host", "-lz", "-lm", "-Wl,-rpath,$ORIGIN/../lib"], verbose)
    # 4. the draw3d app: its host flow as a library (ctypes-loadable) and the command-line program
    d3_dir = os.path.join(PKG, "apps", "draw3d")
    d3_lib_src = [os.path.join(d3_dir, "draw3d_app.c")]
    d3_so = os.path.join(LIB, "libdraw3d_app.so")
    link = ["-pthread", "-L", LIB, "-lvortex", "-lskybox_rt_host", "-lz", "-lm"]
    if force or _stale(d3_so, d3_lib_src + headers + [os.path.join(d3_dir, "draw3d_app.h"), vx_so, host_so]):
        _run([CC] + HOST_CFLAGS + ["-shared", "-o", d3_so] + d3_lib_src + link + ["-Wl,-rpath,$ORIGIN"], verbose)
    exe = os.path.join(BIN, "draw3d")
    d3_src = d3_lib_src + [os.path.join(d3_dir, "main.c"), os.path.join(PKG, "apps", "raytrace", "png_io.c")]
    if force or _stale(exe, d3_src + headers + [vx_so, host_so]):
        _run([CC, "-O2", "-ffp-contract=off", "-Wall", "-Wextra", "-I", INC, "-o", exe] + d3_src + link +
             ["-Wl,-rpath,$ORIGIN/../lib"], verbose)
    # 5. kernel images: what an unmodified reference app passes to vx_upload_kernel_file("kernel.vxbin").
    #    16-byte .vxbin header (kernel/scripts/vxbin.py:54-75) + the 16-byte pipeline descriptor of skybox_rt.h
    import struct
    for name, pipeline in (("raytrace", 1), ("draw3d", 2)):
        path = os.path.join(LIB, name + ".vxbin")
        image = struct.pack("<QQQII", 0x80000000, 0x80000000 + 64, 0x5452584F42594B53, pipeline, 1)
        if force or not os.path.exists(path) or open(path, "rb").read() != image:
            with open(path, "wb") as f:
                f.write(image)
    return {"vortex": vx_so, "host": host_so, "draw3d": d3_so}


if __name__ == "__main__":
    if "--exp" in sys.argv:  # python -m skybox_rt_b200.build --exp TAG -DNAME=VALUE ...
        i = sys.argv.index("--exp")
        build_experiment(sys.argv[i + 1], [a[2:] for a in sys.argv[i + 2:] if a.startswith("-D")], verbose=True)
    else:
        build(force="--force" in sys.argv, verbose=True)
