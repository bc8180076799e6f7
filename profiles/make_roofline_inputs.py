"""profiles/roofline_inputs.json from the committed ncu summaries (profiles/*_summary.txt, written by ncu_summary.py).
bench.py reads it for the `roofline.issue` / `roofline.traffic` fields; a configuration without a capture gets none.

A vx_start of the wavefront pipeline is several kernels: their warp instructions and DRAM bytes are summed, issue
activity is weighted by kernel time and lanes per instruction by warp instructions.
usage: python profiles/make_roofline_inputs.py   (rewrites profiles/roofline_inputs.json)"""
import json
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
# "<workload>:v<variant>" -> summary file (one frame / one ray list of that configuration)
CAPTURES = {
    "c3:v0": "r02c_c3_wave_summary.txt", "c3:v4": "r02c_c3_wave_summary.txt", "c3:v7": "r02_c3_mega_vote_summary.txt",
    "c3:v1": "r02_c3_baseline_summary.txt", "c2:v0": "r02c_c2_wave_summary.txt", "c2:v4": "r02c_c2_wave_summary.txt",
    "c5:v0": "r02c_c5_vote_summary.txt", "c5:v7": "r02c_c5_vote_summary.txt", "c5:v1": "r02_c5_baseline_summary.txt",
    "c5:v3": "r02_c5_pool_summary.txt", "c5:v4": "r02_c5_while_summary.txt",
    "c4:v0": "r02c_c4_wave_summary.txt", "c4:v4": "r02c_c4_wave_summary.txt",
}
# captures of a PART of the step (noted in the record): c5 = 16 Mi of the 64 Mi rays; c4 = the trace kernels of one of 8 sample planes
PARTIAL = {"c5": ("16 Mi of the 64 Mi rays (scripts/frame_once.py --workload c5): dram_bytes / warp_inst scale by 4 for the bench step", 0.25),
           "c4": ("the three trace kernels of the frame (all 8 sample planes in one pass); the two shading passes and the "
                  "combine, ~8 % of the frame, are not in the capture", 1.0)}
UNIT = {"us": 1e-6, "ms": 1e-3, "s": 1.0, "ns": 1e-9, "byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "": 1, "inst": 1, "%": 1}


def parse(path):
    kernels, cur = [], None
    for line in open(path):
        if line.startswith("=="):
            cur = {"name": line[2:].split("(")[0].strip()}
            kernels.append(cur)
            continue
        m = re.match(r"\s+(\S+)\s+([-0-9.eE]+)\s*(\S*)", line)
        if m and cur is not None:
            cur[m.group(1)] = float(m.group(2)) * UNIT.get(m.group(3), 1)
    return kernels


def main():
    out = {}
    for key, fname in CAPTURES.items():
        path = os.path.join(HERE, fname)
        if not os.path.exists(path):
            continue
        ks = parse(path)
        t = sum(k["gpu__time_duration.sum"] for k in ks)
        wi = sum(k["smsp__inst_executed.sum"] for k in ks)
        rec = {
            "source": "profiles/" + fname, "kernels": len(ks), "captured_ms": round(t * 1e3, 4),
            "warp_inst": int(wi),
            "issue_active": round(sum(k["smsp__issue_active.avg.pct_of_peak_sustained_active"] * k["gpu__time_duration.sum"] for k in ks) / t / 100, 4),
            "lanes_per_inst": round(sum(k["smsp__thread_inst_executed_per_inst_executed.ratio"] * k["smsp__inst_executed.sum"] for k in ks) / wi, 2),
            "dram_bytes": int(sum(k.get("dram__bytes_read.sum", 0) + k.get("dram__bytes_write.sum", 0) for k in ks)),
            "l2_bytes": int(sum(k.get("lts__t_bytes.sum", 0) for k in ks)),
        }
        wl = key.split(":")[0]
        if wl in PARTIAL:
            rec["partial"], rec["step_fraction"] = PARTIAL[wl]
        out[key] = rec
    with open(os.path.join(HERE, "roofline_inputs.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    for k, v in sorted(out.items()):
        print(k, v)


if __name__ == "__main__":
    main()
