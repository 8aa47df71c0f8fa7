#!/usr/bin/env python3
"""Pixels whose depth-2 radiance differs between two warp-vote settings of the wide walk: the reference path of the pixel
(oracle path debug) and its rays through the extend probe of three contexts (diagnostic)."""
import ctypes as C
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import opencl_pathtracer_b200 as pt
from oracle import ptref

w, h, depth = 1920, 1080, 3
sc = pt.Scene("atrium", w, h)
np.set_printoptions(precision=9, linewidth=220)
a = pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE)
os.environ["PTC_VOTE_NODE"] = "0"
os.environ["PTC_VOTE_TRI"] = "0"
b = pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE)
c = pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE | pt.FLAG_BVH2)
for r in (a, b, c):
    r.upload(sc)
L = ptref.lib()
rs = ptref.ref_scene(sc)
found = 0
for it in range(8):
    x, y, z = a.render_single(it), b.render_single(it), c.render_single(it)
    dm = (x.view(np.uint32) != y.view(np.uint32)).any(axis=2)
    ys, xs = np.nonzero(dm)
    print("iteration", it, "differing pixels", len(ys))
    for k in range(min(3, len(ys))):
        px, py = int(xs[k]), int(ys[k])
        rays = np.zeros((depth, 8), np.float32)
        ids = np.zeros((depth, 2), np.int32)
        work = np.zeros((depth, 4), np.uint32)
        st = np.zeros((depth, 2), np.float32)
        n = L.ptref_path_debug(C.byref(rs), w, h, depth, px, py, it, rays.ctypes.data_as(C.c_void_p), ids.ctypes.data_as(C.c_void_p),
                               work.ctypes.data_as(C.c_void_p), st.ctypes.data_as(C.c_void_p))
        print(" pixel", px, py, "| default", x[py, px], "| votes0", y[py, px], "| bvh2", z[py, px])
        print("   reference path: segments", n, "ids (tri, material type)", ids[:n].tolist(), "work", work[:n].tolist())
        o = np.ascontiguousarray(rays[:n, 0:4])
        d = np.ascontiguousarray(rays[:n, 4:8])
        for label, r in (("default", a), ("votes0", b), ("bvh2", c)):
            g = r.trace_rays(o, d)
            print("   ", label, "hits", g["hit"].tolist(), "dist2", g["dist2"], "st", g["st"].reshape(-1).tolist())
        for i in range(n):
            print("    ray", i, "o", rays[i, :4], "d", rays[i, 4:])
        found += 1
    if found >= 6:
        break
