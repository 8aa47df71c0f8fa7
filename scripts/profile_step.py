#!/usr/bin/env python3
"""One short render of the bench workload for ncu captures: atrium 1080p, FAST mode, `spp` iterations in one batch."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import opencl_pathtracer_b200 as pt

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 2
mode = pt.MODE_PARITY if (len(sys.argv) > 2 and sys.argv[2] == "parity") else pt.MODE_FAST
scene_name = sys.argv[3] if len(sys.argv) > 3 else "atrium"
W, H = 1920, 1080
sc = pt.Scene(scene_name, W, H)
with pt.PathTracerCUDA(W, H, mode=mode, flags=pt.FLAG_NO_VARIANCE, iterations_per_batch=spp) as r:
    r.upload(sc)
    r.render(0, spp, sync=True)
    c = r.counters()
    print({k: c[k] for k in ("paths", "extension_rays", "shadow_rays", "box_tests", "triangle_tests", "kernel_launches")})
