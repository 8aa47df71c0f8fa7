#!/usr/bin/env python3
"""Where do PARITY and the reference kernel part? Per iteration: pixels whose 1-spp radiance differs, each walked bounce by bounce
(tests/test_gpu_configs.py classify_divergence) with the details of the first diverging bounce.
usage: diag_mismatch.py [scene] [w] [h] [iterations] [depth]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import opencl_pathtracer_b200 as pt
from oracle import ptref

name = sys.argv[1] if len(sys.argv) > 1 else "cornell"
w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (512, 512)
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 4
depth = int(sys.argv[5]) if len(sys.argv) > 5 else 10
sc = pt.Scene(name, w, h, load_sky=name.startswith(("gallery", "terrain")))
with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_PARITY, iterations_per_batch=1) as r:
    r.upload(sc)
    for it in range(iters):
        ref = ptref.render(sc, w, h, depth, 0, it, it + 1)["imageColor"].reshape(-1, 4)
        got = r.render_single(it).reshape(-1, 4)
        diff = (got.view(np.uint32) != ref.view(np.uint32)).any(axis=1)
        a, b = got.astype(np.float64), ref.astype(np.float64)
        bad = (np.abs(a - b) > 1e-3 * np.maximum(np.abs(a), np.abs(b)) + 1e-6).any(axis=1)
        print("it %d: %d pixels differ in some bit, %d outside 1e-3" % (it, diff.sum(), bad.sum()))
        pix = np.flatnonzero(diff)[:12]
        probes = [r.path_probe(it, b) for b in range(depth + 1)]
        for p in pix:
            rp = ptref.path_debug(sc, w, h, depth, int(p % w), int(p // w), it)
            line = "  px %d (%d,%d) rad gpu %s ref %s | ref path n=%d mats %s" % (p, p % w, p // w, got[p][:3], ref[p][:3], rp["n"], rp["mat"].tolist())
            for b in range(rp["n"]):
                go, gd = probes[b]["org"][p], probes[b]["dir"][p]
                if not (np.array_equal(go.view(np.uint32), rp["org"][b].view(np.uint32)) and np.array_equal(gd.view(np.uint32), rp["dir"][b].view(np.uint32))):
                    line += " | ray %d differs: max|ddir| %.3g max|dorg| %.3g (gpu refl %d)" % (b, np.abs(gd - rp["dir"][b]).max(), np.abs(go - rp["org"][b]).max(), probes[b]["refl"][p])
                    break
                if b + 1 <= depth:
                    gh = probes[b + 1]["hit"][p]
                    if int(gh) != int(rp["hit"][b]) or (gh >= 0 and not np.array_equal(probes[b + 1]["st"][p].view(np.uint32), rp["st"][b].view(np.uint32))):
                        line += " | hit %d differs: gpu %d st %s ref %d st %s" % (b, gh, probes[b + 1]["st"][p], rp["hit"][b], rp["st"][b])
                        break
            else:
                line += " | all %d rays and hits identical" % rp["n"]
            print(line)
