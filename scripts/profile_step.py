#!/usr/bin/env python3
"""One short render of the bench workload for ncu captures and tuning sweeps: atrium 1080p, `spp` iterations in one batch.
usage: profile_step.py [spp] [fast|parity] [scene] [repeat]"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import opencl_pathtracer_b200 as pt

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 2
mode = pt.MODE_PARITY if (len(sys.argv) > 2 and sys.argv[2] == "parity") else pt.MODE_FAST
scene_name = sys.argv[3] if len(sys.argv) > 3 else "atrium"
repeat = int(sys.argv[4]) if len(sys.argv) > 4 else 1
W, H = 1920, 1080
gpu_bvh = os.environ.get("PTC_GPU_BVH") == "1"
if scene_name.startswith("terrain"):
    W, H = 3840, 2160
t0 = time.perf_counter()
sc = pt.Scene(scene_name, W, H, load_sky=scene_name.startswith("terrain"), build_bvh=not gpu_bvh)
t_scene = time.perf_counter() - t0
with pt.PathTracerCUDA(W, H, mode=mode, flags=pt.FLAG_NO_VARIANCE | pt.FLAG_STAGE_TIMING | (pt.FLAG_GPU_BVH if gpu_bvh else 0),
                       iterations_per_batch=spp) as r:
    t0 = time.perf_counter()
    r.upload(sc)
    t_up = time.perf_counter() - t0
    if mode == pt.MODE_FAST:
        nodes, order, depth, ms = r.wide_bvh()
        print("scene (import%s) %.2f s | upload %.2f s | wide BVH: %d nodes, depth %d, device build %.2f ms" %
              ("" if gpu_bvh else " + BVH_Create", t_scene, t_up, len(nodes), depth, ms))
    r.render(0, spp, sync=True)
    r.reset()
    t0 = time.perf_counter()
    for i in range(repeat):
        r.render(spp * (i + 1), spp)
    r.synchronize()
    dt = time.perf_counter() - t0
    c = r.counters()
    rays = c["extension_rays"] + c["shadow_rays"]
    env = " ".join("%s=%s" % (k[4:], v) for k, v in sorted(os.environ.items()) if k.startswith("PTC_"))
    print("env %s | %.1f Msamples/s %.0f Mrays/s | ms: ext %.1f con %.1f shade %.1f acc %.1f | box/path %.0f tri/path %.0f" % (
        env, c["paths"] / dt / 1e6, rays / dt / 1e6,
        c["ms_extend"], c["ms_connect"], c["ms_shade"], c["ms_accumulate"], c["box_tests"] / c["paths"], c["triangle_tests"] / c["paths"]))
    try:
        import ctypes as C
        from opencl_pathtracer_b200 import _native as N
        fn = N.cuda_lib().ptc_debug_lane_stats  # only in a -DPTC_LANE_STATS build
        buf = (C.c_ulonglong * 16)()
        fn(buf, 1)
        for k, name in ((0, "extend"), (8, "connect")):
            ns, ln, idle, fin, blk, ts, lt, wn = [buf[k + i] for i in range(8)]
            print("lanes %-7s | node steps %.2f M: stepping %.1f, no ray %.1f, finished %.1f, blocked on triangles %.1f | triangle steps %.2f M: testing %.1f, waiting for a node %.1f"
                  % (name, ns / 1e6, ln / max(ns, 1), idle / max(ns, 1), fin / max(ns, 1), blk / max(ns, 1), ts / 1e6, lt / max(ts, 1), wn / max(ts, 1)))
    except AttributeError:
        pass
