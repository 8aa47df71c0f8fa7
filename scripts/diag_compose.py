#!/usr/bin/env python3
"""render(0,2) against render(0,1) + render(1,1) at full size, with and without batch graphs (diagnostic)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import opencl_pathtracer_b200 as pt

w, h = 1920, 1080
sc = pt.Scene("atrium", w, h)
for mode in (pt.MODE_PARITY, pt.MODE_FAST):
    with pt.PathTracerCUDA(w, h, mode=mode, flags=pt.FLAG_NO_VARIANCE) as r:
        r.upload(sc)
        r.render(0, 2)
        a, na = r.read_image()
        r.reset()
        r.render(0, 1)
        b1, _ = r.read_image()
        r.render(1, 1)
        a2, n2 = r.read_image()
        r.reset()
        r.render(1, 1)
        b2, _ = r.read_image()
        r.reset()
        r.render(0, 2)
        a3, _ = r.read_image()
        print("mode", mode, "graphs", os.environ.get("PTC_GRAPHS"), "| a vs a2 differ:", int((a.view(np.uint32) != a2.view(np.uint32)).any(axis=2).sum()),
              "| a vs a3:", int((a.view(np.uint32) != a3.view(np.uint32)).any(axis=2).sum()),
              "| b1+b2 vs a:", int(((b1 + b2).view(np.uint32) != a.view(np.uint32)).any(axis=2).sum()),
              "| counts", float(na.min()), float(n2.min()), float(n2.max()), "| zero px in a2", int((a2[..., :3].sum(axis=2) == 0).sum()),
              "zero px in a", int((a[..., :3].sum(axis=2) == 0).sum()))
