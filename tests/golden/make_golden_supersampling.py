#!/usr/bin/env python3
"""Generates tests/golden/supersampling_vectors.npz from the REFERENCE ITSELF built with its `-D SUPER_SAMPLING` option
(oracle/_ref, namespace refk_jittered_ss of oracle/ref_glue.cpp) and the table of oracle/shim/X2inv.cl — the reference's own
Kernel/X2inv.cl is an empty file. Run in the build container:

    python tests/golden/make_golden_supersampling.py

Pins the accumulators after 24 iterations of the Cornell box: which (pixel, iteration) paths the criterion dropped shows in
imageRayNb."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import opencl_pathtracer_b200 as pt  # noqa: E402
from oracle import ptref  # noqa: E402

CASE = ("cornell", 64, 48, 5, 24)  # scene, width, height, depth, iterations


def main():
    name, w, h, depth, n_it = CASE
    sc = pt.Scene(name, w, h)
    acc = ptref.render(sc, w, h, depth, ptref.SAMPLER_JITTERED_SUPERSAMPLING, 0, n_it)
    out = {"x2inv": ptref.x2inv(), "case": np.array([w, h, depth, n_it], np.int64)}
    for k in ("imageColor", "imageRayNb", "imageV", "rayDepths", "rayIntersectedBBx", "rayIntersectedTri"):
        out[k] = acc[k]
    path = os.path.join(HERE, "supersampling_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; paths kept %d of %d" % (acc["imageRayNb"].sum(), w * h * n_it))


if __name__ == "__main__":
    main()
