"""Regenerates the frozen oracle renders under tests/golden/ (run from the repo root).

The reference has no RT golden image, so these are THIS repo's goldens:
the CPU oracle (oracle/ref_rt*.c, gcc -O2 -ffp-contract=off) rendering config 1 of BASELINE.json
-- the Cornell-style box at 64x64, 1 spp -- in each shading mode, plus one multi-sample case.
Stored as raw little-endian A8R8G8B8 (row 0 = bottom) in .npy files, and as PNG for eyeballing.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle as o  # noqa: E402

CASES = [("cornell_64_primary", 0, 1), ("cornell_64_shadow", 1, 1), ("cornell_64_bounce", 2, 1),
         ("cornell_64_bounce_spp4", 2, 4)]

if __name__ == "__main__":
    nodes, tris = o.bvh2_build(o.scene_cornell())
    for name, mode, spp in CASES:
        fb, ctr = o.render(o.make_arg(64, 64, mode, spp=spp, seed=1234), nodes, tris, threads=1)
        np.save(os.path.join(HERE, name + ".npy"), fb)
        o.save_png(os.path.join(HERE, name + ".png"), fb)
        print(name, ctr)
