"""Second half of bench.py: the CUDA arm (see bench.py's docstring for the contract)."""
import json
import os
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


class _DevArray:
    """__cuda_array_interface__ view of a device pointer owned by the PathTracer_CUDA context (zero-copy for NCCL)."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist

    import opencl_pathtracer_b200 as pt
    from bench import DEPTH, WORKLOADS, ClockSampler, QuietStdout, host_threads, load_peaks

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this benchmark has no CPU fallback (use --impl reference for the CPU reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    quiet = QuietStdout()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.workload]
    W, H = wl["size"]
    S, K, WU = (args.spp_per_step or wl["spp"]), args.steps, args.warmup
    mode = pt.MODE_FAST if args.mode == "fast" else pt.MODE_PARITY
    n_pix = W * H
    t0 = time.perf_counter()
    scene = pt.Scene(wl["scene"], W, H, load_sky=wl["sky"], build_bvh=False)
    t_import = time.perf_counter() - t0
    t0 = time.perf_counter()
    scene.build_bvh()  # BVH_Create (the reference's host builder); the roofline / PARITY legs need it even with --bvh device
    t_bvh = time.perf_counter() - t0
    main_flags = pt.FLAG_NO_VARIANCE | (pt.FLAG_GPU_BVH if (args.bvh == "device" and args.mode == "fast") else 0)
    info = scene.info
    scene_mb = (info.triangulationSize * 64 + (info.bvhSize // 2) * 64) / 1e6  # DevTri + DevNode (one per inner node)
    stream = torch.cuda.Stream(device=dev)

    def first_iter(step):
        return (step * world + rank) * S  # disjoint iteration ranges per rank

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    with torch.cuda.stream(stream):
        r = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=mode, device=local_rank, flags=main_flags, stream=stream.cuda_stream)
        t0 = time.perf_counter()
        r.upload(scene)
        t_upload = time.perf_counter() - t0
        setup = {"import_s": t_import, "host_bvh_create_s": t_bvh, "upload_s": t_upload, "bvh": args.bvh}
        if mode == pt.MODE_FAST:
            wn, wo, wdepth, wms = r.wide_bvh()
            setup.update({"wide_nodes": int(len(wn)), "wide_depth": wdepth, "device_build_ms": wms})
            del wn, wo
        bufs = r.device_buffers()
        color = torch.as_tensor(_DevArray(bufs.image_color, n_pix * 4), device=dev)
        count = torch.as_tensor(_DevArray(bufs.image_ray_nb, n_pix), device=dev)

        # ---------------- device-timed region: K steps + one NCCL reduce of the accumulators
        for i in range(WU):
            r.render(first_iter(i), S)
        r.synchronize()
        r.reset()
        c0 = r.counters()
        clocks = ClockSampler(local_rank)
        barrier()
        if rank == 0:
            clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(K):
            r.render(first_iter(WU + i), S)
        if world > 1:
            dist.reduce(color, dst=0, op=dist.ReduceOp.SUM)
            dist.reduce(count, dst=0, op=dist.ReduceOp.SUM)
        e1.record(stream)
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        clk = clocks.stop() if rank == 0 else None
        c1 = r.counters()
        launches = c1["kernel_launches"] - c0["kernel_launches"]
        rays = sum_over_ranks((c1["extension_rays"] - c0["extension_rays"]) + (c1["shadow_rays"] - c0["shadow_rays"]))
        paths = sum_over_ranks(c1["paths"] - c0["paths"])
        assert paths == float(K) * S * n_pix * world, (paths, K, S, n_pix, world)
        if rank == 0:
            mean_rgb = float((color.view(-1, 4)[:, :3].sum(0) / count.sum()).mean().item())
        value = paths / (ms * 1e-3)
        # the reference's own ray count (Controleur/PathTracer.cpp:238-243, numberOfProcessedRays): every path that ended at
        # depth i is charged (1 + lights) * (i + 1) rays, whether or not its shadow rays were traced
        depths = r.read_stats()[0].astype(np.float64)
        ref_rays = sum_over_ranks(float(sum((1 + info.lightsSize) * (i + 1) * depths[i] for i in range(min(DEPTH, len(depths))))))

        # ---------------- e2e: reference-facing call sequence with host buffers, host wall clock
        r.reset()
        host_c = torch.empty(n_pix * 4, dtype=torch.float32, pin_memory=True)
        host_n = torch.empty(n_pix, dtype=torch.float32, pin_memory=True)
        cam = scene.camera()
        if world > 1:
            tmp_c, tmp_n = torch.empty_like(color), torch.empty_like(count)
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            r.set_camera(cam)  # the step's input: 64 bytes host -> device (kernel parameter block)
            r.render(first_iter(i), S)
            if world == 1:
                r.read_image_into(host_c.data_ptr(), host_n.data_ptr())  # device -> pinned host, synchronises
            else:
                tmp_c.copy_(color)
                tmp_n.copy_(count)
                dist.reduce(tmp_c, dst=0, op=dist.ReduceOp.SUM)
                dist.reduce(tmp_n, dst=0, op=dist.ReduceOp.SUM)
                if rank == 0:
                    host_c.copy_(tmp_c, non_blocking=True)
                    host_n.copy_(tmp_n, non_blocking=True)
                stream.synchronize()
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e_value = K * S * n_pix * world / e2e_s
        r.close()

        # ---------------- roofline of the dominant traversal kernel (rank 0; separate context with per-stage CUDA events)
        roof = None
        extra = {}
        if rank == 0:
            peaks, peak_kind = load_peaks()
            rp = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=mode, device=local_rank, flags=main_flags | pt.FLAG_STAGE_TIMING,
                                   stream=stream.cuda_stream)
            rp.upload(scene)
            rp.render(0, S, sync=True)
            rp.reset()
            rp.render(S, S, sync=True)
            c = rp.counters()
            rp.close()
            # Algorithmic bytes per traced ray: 64 B per inner-node visit + 64 B per triangle test of the
            # reference BVH2 walked with a correct distance cull (FAST | FLAG_BVH2), counted by the kernels themselves, per
            # extension ray and per shadow ray. The kernel being timed may need fewer bytes (wide BVH): `fetched_bytes_per_launch`
            # is what ITS walk asks for in its own layout (80 B per 8-wide node, 64 B per triangle record).
            wide = args.mode == "fast"
            if wide:
                rb = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=pt.MODE_FAST, device=local_rank,
                                       flags=pt.FLAG_NO_VARIANCE | pt.FLAG_BVH2, stream=stream.cuda_stream)
                rb.upload(scene)
                rb.render(S, S, sync=True)
                cb = rb.counters()
                rb.close()
            else:
                cb = c
            per_ext = 64.0 * (cb["ext_box_tests"] / 2 + cb["ext_triangle_tests"]) / max(cb["extension_rays"], 1)
            per_con = 64.0 * ((cb["box_tests"] - cb["ext_box_tests"]) / 2 + (cb["triangle_tests"] - cb["ext_triangle_tests"])) / max(cb["shadow_rays"], 1)
            node_b, fan = (80.0, 8) if wide else (64.0, 2)
            own_ext = node_b * c["ext_box_tests"] / fan + 64.0 * c["ext_triangle_tests"]
            own_con = node_b * (c["box_tests"] - c["ext_box_tests"]) / fan + 64.0 * (c["triangle_tests"] - c["ext_triangle_tests"])
            kern = {"k_extend": (per_ext * c["extension_rays"], c["ms_extend"], c["launches_extend"], c["extension_rays"], own_ext, per_ext),
                    "k_connect": (per_con * c["shadow_rays"], c["ms_connect"], c["launches_connect"], c["shadow_rays"], own_con, per_con)}
            dom = max(kern, key=lambda k: kern[k][1])
            traffic = None
            tp = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tp):
                with open(tp) as f:
                    traffic = json.load(f).get(args.workload, {}).get(dom)
            # measured read bandwidths with the traversal's own load instruction: 32 MB (L2 hits) and 4 GB (HBM)
            rm = pt.PathTracerCUDA(64, 64, mode=pt.MODE_FAST, device=local_rank, flags=pt.FLAG_NO_VARIANCE, stream=stream.cuda_stream)
            l2_read = rm.measure_read_bandwidth(32 << 20, 200)
            hbm_read = rm.measure_read_bandwidth(4 << 30, 3)
            rm.close()

            def roofline(name):
                b, t_ms, n, nrays, own, per_ray = kern[name]
                ach = b / (t_ms * 1e-3) / 1e9 if t_ms > 0 else 0.0
                return {"kernel": name, "bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
                        "peak_kind": peak_kind + " HBM copy bandwidth (%s)" % ("working set is L2-resident: see DESIGN.md" if scene_mb < 100 else "node + triangle records exceed the L2"),
                        "algorithmic_bytes_per_launch": b / max(n, 1), "algorithmic_bytes_per_ray": per_ray,
                        "fetched_bytes_per_launch": own / max(n, 1), "fetched_gbs": own / (t_ms * 1e-3) / 1e9 if t_ms > 0 else 0.0,
                        "avg_launch_ms": t_ms / max(n, 1), "launches": n,
                        "mrays_per_s": nrays / (t_ms * 1e-3) / 1e6 if t_ms > 0 else 0.0,
                        "l2_read_gbs_measured": l2_read, "hbm_read_gbs_measured": hbm_read,
                        "frac_of_l2_read": (own / (t_ms * 1e-3) / 1e9 / l2_read) if (t_ms > 0 and l2_read > 0) else None}

            roof = roofline(dom)
            roof["traffic"] = traffic
            other = "k_connect" if dom == "k_extend" else "k_extend"
            extra["roofline_other"] = roofline(other)
            tot = c["ms_generate"] + c["ms_extend"] + c["ms_connect"] + c["ms_shade"] + c["ms_accumulate"]
            extra["stage_ms_per_step"] = {k: c["ms_" + k] for k in ("generate", "extend", "connect", "shade", "accumulate")}
            extra["stage_share"] = {k: (c["ms_" + k] / tot if tot else 0.0) for k in ("generate", "extend", "connect", "shade", "accumulate")}
            extra["work_per_path"] = {"box_tests": c["box_tests"] / c["paths"], "triangle_tests": c["triangle_tests"] / c["paths"],
                                      "rays": (c["extension_rays"] + c["shadow_rays"]) / c["paths"]}

            # PARITY mode beside it (bit-exact arithmetic + the reference's culling quirks), 2 steps
            if world == 1 and args.mode == "fast" and args.workload != "cornell_area":  # PARITY has no area lights (an extension)
                rq = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=pt.MODE_PARITY, device=local_rank, flags=pt.FLAG_NO_VARIANCE,
                                       stream=stream.cuda_stream)
                rq.upload(scene)
                rq.render(0, S, sync=True)
                q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                q0.record(stream)
                rq.render(S, 2 * S)
                q1.record(stream)
                rq.synchronize()
                cq = rq.counters()
                rq.close()
                qms = q0.elapsed_time(q1)
                extra["parity_mode"] = {"value": 2 * S * n_pix / (qms * 1e-3), "unit": "samples/s", "ms_per_step": qms / 2,
                                        "box_tests_per_path": cq["box_tests"] / cq["paths"]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.workload != "cornell_area":  # nor has the reference kernel
        try:
            from oracle import ptref
            rows = (0, H)  # the whole frame once: ~10-15 s on 16 host cores
            t0 = time.perf_counter()
            ptref.render(scene, W, H, DEPTH, 0, 0, 1, nthreads=host_threads(), rows=rows)
            dt = time.perf_counter() - t0
            cpu = {"value": (rows[1] - rows[0]) * W / dt, "unit": "samples/s", "cores": host_threads(), "kind": "reference",
                   "sample": "rows %d..%d of the %dx%d frame, 1 spp (%d paths), %.1f s; oracle/_ref = the reference kernel source compiled with g++ -O2 + OpenMP"
                             % (rows[0], rows[1], W, H, (rows[1] - rows[0]) * W, dt)}
        except Exception as e:  # the oracle is test infrastructure: its absence must not hide the GPU number
            cpu = {"value": None, "unit": "samples/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % e}

    if world > 1:
        dist.barrier()
    if rank == 0:
        line = {"metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": WU,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "mrays_per_s": rays / (ms * 1e-3) / 1e6,
                "mrays_per_s_reference_convention": ref_rays / (ms * 1e-3) / 1e6,
                "config": {"workload": wl["desc"], "image": [W, H], "ray_max_depth": DEPTH, "sampler": "jittered", "mode": args.mode,
                           "spp_per_step_per_gpu": S, "triangles": int(info.triangulationSize), "bvh_nodes": int(info.bvhSize),
                           "parallelism": "iteration-range sharding x%d, one NCCL reduce of the accumulators per run" % world,
                           "l2": "not flushed: each step streams ~%.1f GB of path state (> L2); scene records %.0f MB (%s the 126 MB L2)"
                                 % (S * n_pix * 112 / 1e9, scene_mb, "inside" if scene_mb < 100 else "larger than")},
                "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": n_pix * 20,
                        "ms_per_step": 1000.0 * e2e_s / K},
                "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "cpu_baseline": cpu, "mean_rgb": mean_rgb, "setup": setup}
        line.update(extra)
        quiet.emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_progressive(args, rank=0, local_rank=0, world=1):
    """BASELINE.json configs[4] (C5): progressive 1 spp/frame 1080p with a moving camera (the Maya viewport mode).
    Per frame, through the C ABI with host buffers: ptc_set_camera -> ptc_reset_accumulation -> ptc_render(1 iteration) ->
    ptc_read_rgb8 (display epilogue on the device, 3 bytes/pixel back to the host). Reports p50 / p99 frame latency.
    On N > 1 GPUs a 1-spp frame cannot be split by iterations: every rank renders interleaved 8-row stripes
    (ptc_set_pixel_stripes), converts them to RGB8 on the device, and one uint8 NCCL reduce composes the frame on rank 0,
    which copies it to pinned host memory; latency = rank 0's wall clock per frame."""
    import numpy as np
    import torch

    import opencl_pathtracer_b200 as pt
    from bench import DEPTH, WORKLOADS, QuietStdout

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this benchmark has no CPU fallback")
    import torch.distributed as dist
    from opencl_pathtracer_b200.sharding import STRIPE_ROWS, compose_stripes, display_tensor
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    quiet = QuietStdout()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    wl = WORKLOADS[args.workload]
    W, H = wl["size"]
    scene = pt.Scene(wl["scene"], W, H, load_sky=wl["sky"])
    tris = scene.triangles()
    lo = np.array([min(tris[k][a].min() for k in ("S1", "S2", "S3")) for a in "xyz"], dtype=np.float64)
    hi = np.array([max(tris[k][a].max() for k in ("S1", "S2", "S3")) for a in "xyz"], dtype=np.float64)
    centre = 0.5 * (lo + hi)
    cam0 = scene.camera().astype(np.float64).reshape(4, 4)  # position, direction, right, up
    axis = cam0[3, :3] / np.linalg.norm(cam0[3, :3])

    def rotated(angle):
        c, s = np.cos(angle), np.sin(angle)

        def rot(v):  # Rodrigues
            return v * c + np.cross(axis, v) * s + axis * np.dot(axis, v) * (1 - c)

        cam = cam0.copy()
        cam[0, :3] = centre + rot(cam0[0, :3] - centre)
        for r in (1, 2, 3):
            cam[r, :3] = rot(cam0[r, :3])
        return cam.astype(np.float32).reshape(16)

    mode = pt.MODE_FAST if args.mode == "fast" else pt.MODE_PARITY
    frames, warm = args.progressive, max(3, args.warmup)
    lat = []
    with pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=mode, device=local_rank, stream=stream.cuda_stream, flags=pt.FLAG_NO_VARIANCE,
                           iterations_per_batch=1) as r:
        r.upload(scene)
        host = torch.empty(H * W * 3, dtype=torch.uint8, pin_memory=True)  # the display frame lands in pinned host memory
        rgb = host.numpy().reshape(H, W, 3)
        if world > 1:
            r.set_pixel_stripes(STRIPE_ROWS, rank, world)
            frame = display_tensor(r, dev)
        for f in range(warm + frames):
            if f == warm:
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                t_all = time.perf_counter()
            t0 = time.perf_counter()
            r.set_camera(rotated(np.radians(0.5) * f))
            r.reset()
            r.render(f, 1)
            if world > 1:
                r.convert_rgb8()
                with torch.cuda.stream(stream):
                    compose_stripes(frame, dst=0)
                    if rank == 0:
                        host.copy_(frame, non_blocking=True)
                stream.synchronize()
            else:
                r._ck(r._lib.ptc_read_rgb8(r._h, rgb.ctypes.data))
            if f >= warm:
                lat.append((time.perf_counter() - t0) * 1e3)
        total = time.perf_counter() - t_all
        # reset zeroes the launch counter with the other counters: count the launches of one frame instead
        r.reset()
        r.render(0, 1, sync=True)
        per_frame = r.counters()["kernel_launches"] + 1  # + the RGB8 epilogue kernel
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    lat.sort()
    p50, p99 = lat[len(lat) // 2], lat[min(len(lat) - 1, int(len(lat) * 0.99))]
    per = "ptc_set_camera + ptc_reset_accumulation + ptc_render(1) + ptc_read_rgb8" if world == 1 else \
        "per rank: ptc_set_camera + ptc_reset_accumulation + ptc_render(1) of its 8-row stripes + ptc_convert_rgb8; uint8 NCCL reduce to rank 0; copy to pinned host"
    line = {"metric": "frame latency p50 (progressive 1 spp/frame, moving camera)", "value": p50, "unit": "ms", "n_gpus": world, "steps": frames,
            "warmup": warm, "ms_per_step": 1e3 * total / frames, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "p99_ms": p99, "frames_per_s": frames / total, "samples_per_s": frames * W * H / total,
            "config": {"workload": "C5: " + wl["desc"] + "; camera orbits 0.5 deg/frame, accumulation reset every frame", "image": [W, H],
                       "ray_max_depth": DEPTH, "mode": args.mode, "per_frame": per,
                       "parallelism": "whole frame on one GPU" if world == 1 else "interleaved %d-row stripes x%d, one uint8 NCCL reduce per frame" % (STRIPE_ROWS, world)},
            "e2e": {"value": p50, "unit": "ms", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": W * H * 3},
            "gpu_launches": int(per_frame) * frames * world, "mean_rgb8": float(rgb.mean())}
    quiet.emit(json.dumps(line))
