#!/usr/bin/env python3
"""The case compute-sanitizer runs (scripts/gpu_r2_c.sh): cornell_specular 96x64, depth 6, both modes — plain
launches, a replayed batch graph, a pixel shard, the display epilogue and the parity probes. Small on purpose: the tool slows every
kernel down by one to two orders of magnitude. Prints one line per leg; any finding is compute-sanitizer's to report."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import opencl_pathtracer_b200 as pt

w, h, depth = 96, 64, 6
sc = pt.Scene("cornell_specular", w, h)
ref = None
for mode, name in ((pt.MODE_PARITY, "PARITY"), (pt.MODE_FAST, "FAST"), (pt.MODE_FAST, "FAST|BVH2")):
    flags = pt.FLAG_BVH2 if name.endswith("BVH2") else 0
    with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=mode, flags=flags, iterations_per_batch=2) as r:
        r.upload(sc)
        r.trace_primary(0)
        r.render(0, 3)                      # two batches; small batches replay a captured graph from the second call on
        r.render(3, 2)
        r.render(5, 2)
        color, nb = r.read_image()
        rgb = r.read_rgb8()
        stats = r.read_stats()
        r.set_pixel_stripes(8, 1, 3)        # a pixel shard (what one of three GPUs renders of a progressive frame)
        r.reset()
        r.render(0, 1)
        shard = r.read_rgb8()
        r.set_pixel_stripes(0)
        r.reset()
        one = r.render_single(2)
        probe = r.path_probe(2, 3)
    assert np.isfinite(color).all() and (nb == 7).all() and stats[0].sum() == 7 * w * h and np.isfinite(one).all()
    if ref is None:
        ref = color
    print("%-10s ok: mean %.5f, shard rows lit %d, max |d| vs PARITY %.3g" % (name, color[..., :3].mean() / 7, int((shard.sum(axis=(1, 2)) > 0).sum()),
                                                                             float(np.abs(color - ref).max())))
with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_GPU_BVH) as r:  # the device BVH builder
    r.upload(sc)
    r.render(0, 2)
    c2, _ = r.read_image()
print("GPU_BVH    ok: mean %.5f" % (c2[..., :3].mean() / 2))
