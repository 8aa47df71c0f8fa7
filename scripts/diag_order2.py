#!/usr/bin/env python3
"""Images at small path depths: default warp votes against votes 0 (a different, deterministic schedule) (diagnostic)."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import opencl_pathtracer_b200 as pt

w, h = 1920, 1080
sc = pt.Scene("atrium", w, h)
np.set_printoptions(precision=9, linewidth=200)
for depth in (1, 2, 3):
    os.environ.pop("PTC_VOTE_NODE", None)
    os.environ.pop("PTC_VOTE_TRI", None)
    a = pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE)
    os.environ["PTC_VOTE_NODE"] = "0"
    os.environ["PTC_VOTE_TRI"] = "0"
    b = pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE)
    a.upload(sc)
    b.upload(sc)
    for it in (0, 1):
        x1, x2, y = a.render_single(it), a.render_single(it), b.render_single(it)
        d12 = (x1.view(np.uint32) != x2.view(np.uint32)).any(axis=2)
        d1y = (x1.view(np.uint32) != y.view(np.uint32)).any(axis=2)
        print("depth", depth, "it", it, "| default twice differ:", int(d12.sum()), "| default vs votes 0:", int(d1y.sum()))
        if depth == 1 and d1y.any():
            ys, xs = np.nonzero(d1y)
            for k in range(min(6, len(ys))):
                print("   px", xs[k], ys[k], "default", x1[ys[k], xs[k]], "votes0", y[ys[k], xs[k]])
    a.close()
    b.close()
