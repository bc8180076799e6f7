"""Per-source-line view of an .ncu-rep captured with --import-source on (kernels built with -lineinfo):
share of warp instructions, average active threads per instruction and stall samples per CUDA line.
usage: python profiles/ncu_lines.py gpurun_out/x.ncu-rep [top_n] > profiles/x_lines.txt"""
import csv
import subprocess
import sys


def main(path, top=40):
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    fname, head, lines = None, None, {}
    for r in rows:
        if len(r) >= 2 and r[0] in ("File Name", "File Path"):
            fname = r[1].split("/")[-1]
            continue
        if len(r) > 4 and r[0] == "Line No":
            head = r
            ii = head.index("Instructions Executed")
            it = head.index("Thread Instructions Executed")
            isamp = head.index("# Samples")
            continue
        if head is None or len(r) <= it:
            continue
        if r[2] == "-" and r[0] != "":  # a CUDA source line (its SASS rows follow); metrics on it are the line totals
            try:
                key = (fname, int(r[0]))
                lines[key] = [int(r[ii] or 0), int(r[it] or 0), int(r[isamp] or 0), r[1].strip()]
            except ValueError:
                pass
    tot_i = sum(v[0] for v in lines.values()) or 1
    tot_t = sum(v[1] for v in lines.values())
    tot_s = sum(v[2] for v in lines.values()) or 1
    print(f"# {path}: {tot_i} warp instructions, {tot_t / tot_i:.2f} threads/instruction, {tot_s} stall samples")
    print("# inst%  thr/inst  samples%  file:line  source")
    for (f, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
        if v[0] == 0:
            break
        print(f"{v[0] / tot_i * 100:6.2f}  {v[1] / v[0]:6.1f}  {v[2] / tot_s * 100:7.2f}  {f}:{ln}  {v[3][:100]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
