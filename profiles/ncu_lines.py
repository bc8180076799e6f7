"""Per-source-line view of an .ncu-rep captured with --import-source on (kernels built with -lineinfo):
share of warp instructions, average active threads per instruction and stall samples per CUDA line, for ONE kernel
(the source page of a multi-kernel report lists every kernel; -k picks which).
usage: python profiles/ncu_lines.py gpurun_out/x.ncu-rep [top_n] [-k kernel-regex] [--skip n] [--budget rays] > profiles/x_lines.txt
--budget N: also print the thread-instruction budget per ray by phase (node step / triangle test / scheduler /
            advance + job fetch / other), N = rays the captured launch traced."""
import csv
import re
import subprocess
import sys

# which phase of the traversal loop a source line belongs to (file, first line, last line) -> phase; first match wins
PHASES = [
    ("rt_wide.cuh", "node_step", "node"), ("rt_wide.cuh", "pop_next", "node"), ("rt_wide.cuh", "Stack", "node"),
    ("rt_wide.cuh", "sts64", "node"), ("rt_wide.cuh", "lds64", "node"),
    ("rt_wide.cuh", "tri_test", "triangle"), ("rt_wide.cuh", "tri_step", "triangle"),
    ("rt_device.cuh", "intersect_tri", "triangle"), ("rt_device.cuh", "consider", "triangle"),
    ("rt_wide.cuh", "run_warp", "scheduler"), ("rt_wide.cuh", "take_jobs", "advance"), ("rt_wide.cuh", "Cursor", "scheduler"),
    # vector helpers are inlined: inside a trace kernel they belong to the ray/triangle test
    ("rt_device.cuh", "mk", "triangle"), ("rt_device.cuh", "dot", "triangle"), ("rt_device.cuh", "V3", "triangle"),
    ("rt_device.cuh", "safe_inv", "advance"), ("rt_device.cuh", "camera_dir", "advance"), ("rt_device.cuh", "ld3", "advance"),
    ("rt_device.cuh", "rnd01", "advance"), ("rt_device.cuh", "make_camera", "advance"),
]


def function_ranges(src):
    """(name, first, last) of the functions / structs at namespace level of a source file.  src: {line: text}.
    The text between two namespace-level declarations is accumulated until a '{' opens a body; the name is the
    identifier in front of the parameter list (or after `struct`)."""
    out, depth, head, start, name = [], 0, "", 0, None
    base = None  # brace depth of namespace level (1 inside `namespace srt {`)
    for n in range(1, (max(src) if src else 0) + 1):
        code = src.get(n, "").split("//")[0]
        if base is None:
            if re.match(r"\s*namespace\s+\w+\s*\{", code):
                base, depth = 1, 1
            continue
        for ch_i, ch in enumerate(code):
            if ch == "{":
                if depth == base and name is None:
                    text = re.sub(r"__launch_bounds__\s*\([^)]*\)", "", head + code[:ch_i])
                    m = re.search(r"struct\s+(\w+)", text) or re.search(r"(\w+)\s*\(", text)
                    name, start = (m.group(1) if m else "?"), (start or n)
                depth += 1
            elif ch == "}":
                depth -= 1
                if depth == base and name is not None:
                    out.append((name, start, n))
                    name, head, start = None, "", 0
            elif ch == ";" and depth == base:
                head, start = "", 0
        if depth == base and name is None:
            if code.strip() and not start:
                start = n
            head += code + " "
    return out


def main(path, top=40, kernel=None, rays=None, skip=0):
    cmd = ["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"]
    if kernel:
        cmd += ["-k", "regex:" + kernel]  # matches the function name without template arguments
    if skip or kernel:
        cmd += ["-s", str(skip), "-c", "1"]  # the skip-th matching launch of the report
    raw = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    fname, fpath, head, lines, kname = None, None, None, {}, None
    paths, srcs = {}, {}
    for r in rows:
        if len(r) >= 2 and r[0] in ("File Name", "File Path"):
            fpath = r[1]
            fname = r[1].split("/")[-1]
            paths[fname] = fpath
            continue
        if len(r) >= 2 and r[0] == "Function Name":
            kname = kname or r[1]
            if r[1] != kname:
                fname = None  # a second kernel of the report: keep the first match only
            continue
        if len(r) > 4 and r[0] == "Line No":
            head = r
            ii = head.index("Instructions Executed")
            it = head.index("Thread Instructions Executed")
            isamp = head.index("# Samples")
            continue
        if head is None or fname is None or len(r) <= it:
            continue
        if r[2] == "-" and r[0] != "":  # a CUDA source line (its SASS rows follow); metrics on it are the line totals
            try:
                key = (fname, int(r[0]))
                srcs.setdefault(fname, {})[int(r[0])] = r[1]
                v = lines.setdefault(key, [0, 0, 0, r[1].strip()])
                v[0] += int(r[ii] or 0); v[1] += int(r[it] or 0); v[2] += int(r[isamp] or 0)
            except ValueError:
                pass
    tot_i = sum(v[0] for v in lines.values()) or 1
    tot_t = sum(v[1] for v in lines.values())
    tot_s = sum(v[2] for v in lines.values()) or 1
    print(f"# {path}: kernel {kname}")
    print(f"# {tot_i} warp instructions, {tot_t / tot_i:.2f} threads/instruction, {tot_s} stall samples")
    if rays:
        # function ranges come from the files on disk: run this right after the capture (same tree), as the GPU scripts do
        import os
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        ranges = {}
        for f in paths:
            try:
                text = open(os.path.join(root, "skybox_rt_b200", "csrc", f)).read().split("\n")
            except OSError:
                text = []
            ranges[f] = function_ranges({n: t for n, t in enumerate(text, 1)})
        phase = {}
        for (f, ln), v in lines.items():
            fn = next((n for n, a, b in ranges.get(f, []) if a <= ln <= b), None)
            ph = next((p for pf, pn, p in PHASES if pf == f and pn == fn), None)
            if ph is None:
                ph = "advance" if f == "rt_wave.cuh" or (f == "rt_wide.cuh" and fn in ("wide_render_body", "rt_wide_rays_kernel")) else \
                     ("shading math" if f == "rt_device.cuh" else "other")
            p = phase.setdefault(ph, [0, 0])
            p[0] += v[0]; p[1] += v[1]
        print(f"# budget per ray ({rays} rays): phase, warp instructions, thread instructions, lanes per instruction")
        for ph, (wi, ti) in sorted(phase.items(), key=lambda kv: -kv[1][1]):
            print(f"#   {ph:14s} {wi / rays:8.1f} {ti / rays:9.1f} {ti / max(wi, 1):6.1f}")
        print(f"#   {'total':14s} {tot_i / rays:8.1f} {tot_t / rays:9.1f} {tot_t / tot_i:6.1f}")
    print("# inst%  thr/inst  samples%  file:line  source")
    for (f, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
        if v[0] == 0:
            break
        print(f"{v[0] / tot_i * 100:6.2f}  {v[1] / v[0]:6.1f}  {v[2] / tot_s * 100:7.2f}  {f}:{ln}  {v[3][:100]}")


if __name__ == "__main__":
    a = sys.argv[1:]
    kern = a[a.index("-k") + 1] if "-k" in a else None
    rays = int(a[a.index("--budget") + 1]) if "--budget" in a else None
    skip = int(a[a.index("--skip") + 1]) if "--skip" in a else 0
    pos = [x for i, x in enumerate(a) if not x.startswith("-") and (i == 0 or a[i - 1] not in ("-k", "--budget", "--skip"))]
    main(pos[0], int(pos[1]) if len(pos) > 1 else 40, kern, rays, skip)
