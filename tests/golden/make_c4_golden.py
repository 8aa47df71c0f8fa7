#!/usr/bin/env python3
"""Generates tests/golden/c4_primary_hash.json: hashes of the reference kernel's primary hits on BASELINE.json configs[3]'s scene
(terrain + flattened instances, 10.1 M triangles, sky) at 1920x1080, iteration 0 — from oracle/_ref (the reference's own kernel
source compiled for the CPU; needs /root/reference to build it). The GPU test compares the CUDA path with these hashes, so that
parity on the 10 M-triangle scene is pinned even where the oracle is absent."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import opencl_pathtracer_b200 as pt  # noqa: E402
from oracle import ptref  # noqa: E402

W, H = 1920, 1080
sc = pt.Scene("terrain", W, H, load_sky=True)
ref = ptref.primary(sc, W, H, sampler=0, iteration=0)
sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
out = {"scene": "terrain", "image": [W, H], "iteration": 0, "triangles": int(sc.info.triangulationSize), "bvh_nodes": int(sc.info.bvhSize),
       "hit_fraction": float((ref["hit"] >= 0).mean()),
       "hit_sha256": sha(ref["hit"]), "st_sha256": sha(ref["st"].view(np.uint32)), "counts_sha256": sha(ref["counts"]),
       "dist2_sha256": sha(ref["dist2"].view(np.uint32))}
with open(os.path.join(ROOT, "tests", "golden", "c4_primary_hash.json"), "w") as f:
    json.dump(out, f, indent=1)
print(out)
