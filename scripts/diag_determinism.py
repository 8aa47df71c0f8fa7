#!/usr/bin/env python3
"""Run-to-run determinism of FAST mode: the extend stage alone (ray probe, twice on the same rays) and whole images, for the
wide walk and for the BVH2 walk (diagnostic)."""
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
o, d = surface_rays(sc, 1 << 20, 7)
for flags, label in ((0, "wide"), (pt.FLAG_BVH2, "bvh2")):
    with pt.PathTracerCUDA(w, h, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE | flags) as r:
        r.upload(sc)
        g1 = r.trace_rays(o, d)
        g2 = r.trace_rays(o, d)
        dh = g1["hit"] != g2["hit"]
        print(label, "probe: hits differ on", int(dh.sum()), "of", len(o), "| dist2 differ", int((g1["dist2"].view(np.uint32) != g2["dist2"].view(np.uint32)).sum()))
        if dh.any():
            i = np.flatnonzero(dh)[:5]
            print("   examples", g1["hit"][i], g2["hit"][i], g1["dist2"][i], g2["dist2"][i])
        imgs = []
        for k in range(3):
            r.reset()
            r.render(0, 2)
            imgs.append(r.read_image()[0])
        print(label, "image: pixels differing run to run:", [int((imgs[0].view(np.uint32) != x.view(np.uint32)).any(axis=2).sum()) for x in imgs[1:]])
        r.reset()
        r.render(0, 1, sync=True)
        c1 = r.counters()
        r.reset()
        r.render(0, 1, sync=True)
        c2 = r.counters()
        print(label, "counters run to run:", {k: (c1[k], c2[k]) for k in ("extension_rays", "shadow_rays", "box_tests", "triangle_tests")})
