"""One device BVH build (1 M triangles) + the first frame on it (Node4 repack + trace), for an ncu capture.
Run on the GPU box:  ncu --set full ... python profiles/lbvh_once.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skybox_rt_b200 import raytrace as rt  # noqa: E402

soup = rt.scene_terrain(1000, 500)
with rt.RayTracer(soup=soup) as t:
    fb = t.render(rt.make_arg(64, 64, 0, scene="terrain"))
    print("nodes", t.num_nodes, "tris", t.num_tris, "pixels lit", int((fb != 0).sum()))
