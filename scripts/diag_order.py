#!/usr/bin/env python3
"""Rays whose closest hit in the wide walk depends on the order of the triangle tests: the same rays through two contexts
whose warp votes differ (default against PTC_VOTE_NODE = PTC_VOTE_TRI = 0), and twice through the default one (diagnostic)."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import opencl_pathtracer_b200 as pt
from golden_cases import surface_rays

w, h = 1920, 1080
sc = pt.Scene("atrium", w, h)
n = 1 << 23
o, d = surface_rays(sc, n, 11)
a = pt.PathTracerCUDA(w, h, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE)
os.environ["PTC_VOTE_NODE"] = "0"
os.environ["PTC_VOTE_TRI"] = "0"
b = pt.PathTracerCUDA(w, h, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE)
c = pt.PathTracerCUDA(w, h, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE | pt.FLAG_BVH2)
for r in (a, b, c):
    r.upload(sc)
x1, x2, y, z = a.trace_rays(o, d), a.trace_rays(o, d), b.trace_rays(o, d), c.trace_rays(o, d)
print("default twice: differ", int((x1["hit"] != x2["hit"]).sum()), "| default vs votes 0:", int((x1["hit"] != y["hit"]).sum()),
      "| default vs BVH2:", int((x1["hit"] != z["hit"]).sum()), "| votes 0 vs BVH2:", int((y["hit"] != z["hit"]).sum()), "of", n)
i = np.flatnonzero(x1["hit"] != y["hit"])[:12]
np.set_printoptions(precision=9, suppress=False, linewidth=200)
for k in i:
    print("ray", k, "o", o[k], "d", d[k], "\n    default hit", x1["hit"][k], "dist2", x1["dist2"][k], "st", x1["st"][k], "| votes0 hit", y["hit"][k], "dist2", y["dist2"][k], "st", y["st"][k],
          "| bvh2 hit", z["hit"][k], "dist2", z["dist2"][k])
for r in (a, b, c):
    r.close()
