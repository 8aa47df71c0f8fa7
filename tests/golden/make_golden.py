#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref = the reference kernel + BVH builder compiled
from /root/reference by oracle/build_ref.py). Run in the build container (the reference tree is not on the GPU box):

    python tests/golden/make_golden.py

The fixtures pin (a) the oracle build against silent changes and (b) the GPU path on machines where oracle/_ref is
absent. The reference ships no golden vectors of its own."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import opencl_pathtracer_b200 as pt  # noqa: E402
from oracle import ptref  # noqa: E402
sys.path.insert(0, os.path.dirname(HERE))
from golden_cases import surface_rays  # noqa: E402

CASES = [("cornell", 64, 48, False, 10), ("gallery", 80, 60, True, 6), ("atrium:8", 64, 40, False, 5)]


def bvh_digest(scene):
    n = scene.nodes()
    h = hashlib.sha256()
    for f in ("cutAxis", "triangleStartIndex", "nbTriangles", "son1Id", "son2Id", "isLeaf"):
        h.update(np.ascontiguousarray(n[f]).tobytes())
    h.update(np.ascontiguousarray(scene.triangles()["id"]).tobytes())
    return h.hexdigest()


def main():
    out = {}
    # RNG known answers: InitializeRandomSeed + 8 draws
    kat = []
    for (x, y, it, W, H) in [(0, 0, 0, 512, 512), (1, 0, 0, 512, 512), (511, 511, 0, 512, 512), (0, 0, 1, 512, 512),
                             (1919, 1079, 1023, 1920, 1080), (7, 3, 15, 1920, 1080)]:
        s0, seeds, us = ptref.rng(x, y, it, W, H, 8)
        kat.append(np.concatenate([[x, y, it, W, H, s0], seeds]).astype(np.int64))
        out["rng_u_%d" % (len(kat) - 1)] = us
    out["rng_kat"] = np.stack(kat)
    # cosine-sample KAT (trig sensitive)
    cs = []
    for seed, n in [(12345, (0.0, 0.0, 1.0, 0.0)), (99991, (0.6, 0.0, 0.8, 0.0)), (31337, (-0.48, 0.6, -0.64, 0.0))]:
        v, after = ptref.cosine_sample(seed, n)
        cs.append(np.concatenate([v, [np.float32(after).view(np.float32)]]))
        out["cos_after_%d" % (len(cs) - 1)] = np.array([after], np.int64)
        out["cos_v_%d" % (len(cs) - 1)] = v
    for name, w, h, sky, depth in CASES:
        key = name.replace(":", "_")
        sc = pt.Scene(name, w, h, load_sky=sky)
        out[key + "_bvh_sha256"] = np.frombuffer(bvh_digest(sc).encode(), np.uint8)
        i = sc.info
        out[key + "_bvh_meta"] = np.array([i.triangulationSize, i.bvhSize, i.bvhMaxDepth], np.int64)
        p = ptref.primary(sc, w, h, 0, 0)
        for k in ("hit", "st", "dist2", "dir", "seed_after", "counts"):
            out[key + "_primary_" + k] = p[k]
        out[key + "_radiance_it0"] = ptref.render(sc, w, h, depth, 0, 0, 1)["imageColor"]
        o, d = surface_rays(sc, 4000, 21)  # rays that start on surfaces, as the bounces shoot them
        tr = ptref.trace_rays(sc, o, d)
        out[key + "_rays_origin"], out[key + "_rays_direction"] = o, d
        for k in ("hit", "st", "dist2", "dir", "counts"):
            out[key + "_rays_" + k] = tr[k]
        acc = ptref.render(sc, w, h, depth, 0, 0, 4)
        for k in ("imageColor", "imageRayNb", "rayDepths", "rayIntersectedBBx", "rayIntersectedTri"):
            out[key + "_acc4_" + k] = acc[k]
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote", os.path.join(HERE, "reference_vectors.npz"), os.path.getsize(os.path.join(HERE, "reference_vectors.npz")), "bytes")


if __name__ == "__main__":
    main()
