import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import opencl_pathtracer_b200 as pt
sys.path.insert(0, "tests")
from test_gpu_parity import _rel_mismatch, _surface_rays
w, h = 320, 200
name = sys.argv[1] if len(sys.argv) > 1 else "atrium:4"
sc = pt.Scene(name, w, h)
for depth in (1, 2, 3, 10):
    with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_BVH2) as a, \
         pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST) as b, \
         pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_PARITY) as p:
        a.upload(sc); b.upload(sc); p.upload(sc)
        for it in (0, 3):
            ra, rb, rp = a.render_single(it), b.render_single(it), p.render_single(it)
            m = _rel_mismatch(rb, ra)
            print("depth %d it %d: wide vs BVH2 %g | BVH2 vs PARITY %g | wide vs PARITY %g" % (depth, it, m.mean(), _rel_mismatch(ra, rp).mean(), _rel_mismatch(rb, rp).mean()))
            if depth == 1 and m.any():
                i = np.flatnonzero(m)[:6]
                print(" px", i, "\n BVH2", ra.reshape(-1, 4)[i], "\n wide", rb.reshape(-1, 4)[i])
o, d = _surface_rays(sc, 200000, 5)
with pt.PathTracerCUDA(512, 512, mode=pt.MODE_FAST, flags=pt.FLAG_BVH2) as a, pt.PathTracerCUDA(512, 512, mode=pt.MODE_FAST) as b:
    a.upload(sc); b.upload(sc)
    x, y = a.trace_rays(o, d), b.trace_rays(o, d)
    reg = (d[:, :3] != 0).all(1)
    diff = x["hit"] != y["hit"]
    tie = diff & (x["dist2"] == y["dist2"])
    print("regular rays: differ %g ties %g ; axis-parallel rays: differ %g" % (diff[reg].mean(), tie[reg].mean(), diff[~reg].mean()))
    i = np.flatnonzero(diff & reg & ~tie)[:5]
    print(i, x["hit"][i], y["hit"][i], x["dist2"][i], y["dist2"][i], x["st"][i], y["st"][i])
    i = np.flatnonzero(tie & reg)[:5]
    print("ties", i, x["hit"][i], y["hit"][i], x["dist2"][i], x["st"][i], y["st"][i])
