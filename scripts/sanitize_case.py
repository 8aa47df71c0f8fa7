#!/usr/bin/env python3
"""The memory-checker case. compute-sanitizer is closed on the GPU pool ("runs under it have left GPUs needing a
reset"), so the kernels are run from the CHECKED build instead (PTC_LIB_DIR=opencl_pathtracer_b200/_lib_checked: every queue entry,
node / triangle / material / texture / texel index, stack level and target pixel is compared with its bound, violations are
reported through ptc_counters.faults): three scenes (untextured box, textured gallery with sky, 16 k-triangle atrium), both modes
and both walks — plain launches, a replayed batch graph, a pixel shard, the display epilogue, the parity probes, the device BVH
builder, and a multi-device context where the box has two GPUs. Exits 1 on any fault. tests/test_checked_build.py runs it."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import opencl_pathtracer_b200 as pt

from opencl_pathtracer_b200 import _native

version = _native.cuda_lib().ptc_version().decode()
print("library:", version, "(%s)" % _native.LIB_DIR)
faults = 0


def check(r, what):
    global faults
    c = r.counters()
    if c["faults"]:
        faults |= c["faults"]
        print("FAULT in %s: mask %#x, first: code %d value %d bound %d" % (what, c["faults"], c["fault_code"], c["fault_value"], c["fault_bound"]))


w, h, depth = 96, 64, 6
for scene_name, sky in (("cornell_specular", False), ("gallery", True), ("atrium:4", False)):
    sc = pt.Scene(scene_name, w, h, load_sky=sky)
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
            check(r, scene_name + " " + name + " render")
            r.set_pixel_stripes(8, 1, 3)        # a pixel shard (what one of three GPUs renders of a progressive frame)
            r.reset()
            r.render(0, 1)
            shard = r.read_rgb8()
            r.set_pixel_stripes(0)
            r.reset()
            one = r.render_single(2)
            probe = r.path_probe(2, 3)
            check(r, scene_name + " " + name + " shard + probes")
        assert np.isfinite(color).all() and (nb == 7).all() and stats[0].sum() == 7 * w * h and np.isfinite(one).all()
        if ref is None:
            ref = color
        print("%-16s %-10s ok: mean %.5f, shard rows lit %d, max |d| vs PARITY %.3g" % (scene_name, name, color[..., :3].mean() / 7, int((shard.sum(axis=(1, 2)) > 0).sum()),
                                                                                 float(np.abs(color - ref).max())))
with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_GPU_BVH) as r:  # the device BVH builder
    r.upload(sc)
    r.render(0, 2)
    c2, _ = r.read_image()
    check(r, "GPU_BVH")
print("GPU_BVH    ok: mean %.5f" % (c2[..., :3].mean() / 2))
if _native.cuda_lib().ptc_device_count() >= 2:
    with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, devices=[0, 1]) as r:
        r.upload(sc)
        r.render(0, 5)
        r.render(5, 1)
        c3, n3 = r.read_image()
        r.read_rgb8()
        check(r, "two devices")
    print("2 devices  ok: mean %.5f" % (c3[..., :3].sum() / n3.sum()))
print("faults: %#x" % faults)
sys.exit(1 if faults else 0)
