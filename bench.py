#!/usr/bin/env python3
"""bench.py — benchmark of the path-tracing hot path (DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--mode fast|parity]
                  [--total-spp S] [--progressive FRAMES]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

Default workload = BASELINE.json configs[2] (the config the 1/2/4/8-GPU samples/s + traversal roofline is quoted on): procedural
atrium, ~262 k triangles, diffuse + varnished + glass, 2 point + 1 directional light, 1920x1080, depth 10, jittered sampler.

One STEP = `--spp-per-step` (default 8) iterations of the whole frame on every GPU ("weak" scaling: the work per GPU is fixed);
GPUs render disjoint iteration ranges (a path's random stream depends on the iteration number only), the accumulators and the three
histograms are summed to rank 0 with one NCCL reduce inside the timed region. `--total-spp S` is the STRONG form BASELINE.json words
its configs with ("256 spp at 1/2/4/8 GPUs"): one step = S iterations split over the N GPUs, the reduce and the read-back of the
image into host memory inside every step.

Two ways onto N > 1 GPUs, same numbers asked of both:
  * under torchrun (what the driver launches): one process per GPU, torch.distributed over NCCL on tensors aliasing the library's
    device buffers;
  * plain `python bench.py --gpus N` (no torchrun): ONE process, one host thread, ptc_create_multi — the multi-GPU context of the
    C ABI (ncclCommInitAll inside the library), what the reference's own caller gets.

Prints ONE JSON line on rank 0. `value` = samples/s (paths/s), scene resident in HBM, device-timed; `e2e` = the same through the
reference-facing call sequence with host buffers; `roofline` = the dominant traversal kernel against the memory level its node and
triangle records live in; `cpu_baseline` = the reference kernel itself (oracle/_ref) on the host cores. `--impl reference` times
that CPU reference alone, without loading the CUDA library.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEPTH = 10
WORKLOADS = {
    "atrium": dict(scene="atrium", size=(1920, 1080), sky=False, spp=8,
                   desc="C3: procedural atrium ~262k tris, diffuse+varnished+glass, 3 lights, 1920x1080, depth 10, jittered"),
    "atrium_1m": dict(scene="atrium@2", size=(1920, 1080), sky=False, spp=8,
                      desc="north-star target: the atrium at twice the tessellation, ~1.05M tris, diffuse+varnished+glass, 3 lights, 1920x1080, depth 10, jittered"),
    "cornell": dict(scene="cornell", size=(1920, 1080), sky=False, spp=8,
                    desc="C2: Cornell box 30 tris, diffuse + point light, 1920x1080, depth 10, jittered"),
    "cornell512": dict(scene="cornell", size=(512, 512), sky=False, spp=16,
                       desc="C1: Cornell box 30 tris, 512x512, 16 spp per step, depth 10, jittered"),
    "cornell_area": dict(scene="cornell_area", size=(1920, 1080), sky=False, spp=8,
                         desc="C2 as BASELINE.json words it: Cornell box 30 tris, diffuse + 130x105 area light (PTC_LIGHT_AREA, an extension: "
                              "the reference has no area lights, so no reference arm exists for this workload), 1920x1080, depth 10, jittered"),
    "terrain": dict(scene="terrain", size=(3840, 2160), sky=True, spp=1,
                    desc="C4: procedural terrain + 64x64 flattened rock instances ~10.1M tris, 16 textures 1024^2, sun + sky cube map, 3840x2160, depth 10, jittered"),
    "terrain_1080p": dict(scene="terrain", size=(1920, 1080), sky=True, spp=4,
                          desc="C4 scene (~10.1M tris, textures, sun + sky) at 1920x1080, depth 10, jittered"),
    "atrium_small": dict(scene="atrium:4", size=(1920, 1080), sky=False, spp=8,
                         desc="atrium at 1/4 tessellation (~16k tris), 1920x1080 (smoke only)"),
}
L2_BYTES = 126e6  # B200 L2; a scene whose node + triangle records fit is served from it (ncu: DRAM at 3-5 % of peak on the atrium)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "MEASURED_PEAKS.json"
    return {"hbm_gbs": 6650.0}, "fallback of B200_PROFILING.md"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "power_w_max": max(pw) if pw else None,
                "samples": len(sm), "reasons": sorted(reasons)}


class QuietStdout:
    """The bench contract is ONE JSON line on stdout. Libraries write there too (NCCL prints its version banner to stdout when the
    box exports NCCL_DEBUG=VERSION): while the run is in progress file descriptor 1 points at stderr; emit() restores it for the line."""

    def __init__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def emit(self, text):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)
        print(text, flush=True)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_rows(step, rows_per_step, H):
    if rows_per_step >= H:
        return 0, H
    r0 = (H // 4 + step * rows_per_step) % (H - rows_per_step)
    return r0, r0 + rows_per_step


# ------------------------------------------------------------------------------------------------ reference arm
def reference_scene(wl):
    """The workload's scene for the CPU reference: generated by the scene library (libpathtracer_scene.so, no CUDA code), its BVH
    built by the REFERENCE's own BVH_Create (oracle/_ref, Controleur/PathTracer_BVH.cpp compiled where it lies)."""
    from opencl_pathtracer_b200.scene import Scene
    from oracle import ptref
    W, H = wl["size"]
    scene = Scene(wl["scene"], W, H, load_sky=wl["sky"], build_bvh=False)
    t0 = time.perf_counter()
    nodes, tris, size, depth = ptref.bvh_create(scene.triangles())
    scene.triangles()[:] = tris  # the builder permutes the triangles: install its order together with its nodes
    scene.set_bvh(nodes, size, depth)
    return scene, time.perf_counter() - t0


def run_reference(args, rank):
    """--impl reference: the reference's own kernel (oracle/_ref) on the host cores, bounded row samples of the same frame. The CUDA
    library is never loaded by this arm."""
    if rank != 0:
        return
    from oracle import ptref
    wl = WORKLOADS[args.workload]
    if args.workload == "cornell_area":
        print(json.dumps({"impl": "reference", "unavailable": "the reference kernel has no area lights (PTC_LIGHT_AREA is an extension); use --workload cornell"}))
        return
    W, H = wl["size"]
    scene, t_bvh = reference_scene(wl)
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 to its workers, which would leave the reference
    # kernel on one core — the thread count is passed explicitly (num_threads clause)
    cores = host_threads()
    rows = min(args.ref_rows, H)
    times = []
    for step in range(args.warmup + args.steps):
        r0, r1 = reference_rows(step, rows, H)
        t0 = time.perf_counter()
        ptref.render(scene, W, H, DEPTH, 0, 0, 1, nthreads=cores, rows=(r0, r1))
        dt = time.perf_counter() - t0
        if step >= args.warmup:
            times.append(dt)
    total = sum(times)
    paths = rows * W * args.steps
    v = paths / total
    sample = "%d image rows x %d px x 1 spp per step (= %d paths) of the %dx%d frame" % (rows, W, rows * W, W, H)
    loaded = sorted({os.path.basename(l.split()[-1]) for l in open("/proc/self/maps") if "pathtracer" in l or "ptref" in l})
    line = {"impl": "reference", "metric": "samples/s", "value": v, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "image": [W, H], "ray_max_depth": DEPTH, "sampler": "jittered", "sample": sample,
                       "bvh": "reference BVH_Create (oracle/_ref), %.2f s" % t_bvh, "libraries_mapped": loaded},
            "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ roofline (rank 0)
def roofline_block(pt, args, scene, W, H, S, mode, main_flags, device, stream_ptr):
    """Per-stage device times (a context with PTC_FLAG_STAGE_TIMING: CUDA events on the stream the kernels run on around every
    launch) -> the dominant traversal kernel against the memory level that holds its node + triangle records."""
    peaks, peak_src = load_peaks()
    info = scene.info
    rp = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=mode, device=device, flags=main_flags | pt.FLAG_STAGE_TIMING, stream=stream_ptr)
    rp.upload(scene)
    rp.render(0, S, sync=True)
    rp.reset()
    rp.render(S, S, sync=True)
    c = rp.counters()
    rp.close()
    # Algorithmic bytes per traced ray: 64 B per inner-node visit + 64 B per triangle test of the reference BVH2
    # walked with a correct distance cull (FAST | FLAG_BVH2), counted by the kernels themselves, per extension ray and per shadow
    # ray. The kernel being timed may need fewer bytes (wide BVH): `fetched_*` is what ITS walk asks for in its own layout (96 B per
    # padded 8-wide node, 64 B per triangle record).
    wide = args.mode == "fast"
    if wide:
        rb = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=pt.MODE_FAST, device=device, flags=pt.FLAG_NO_VARIANCE | pt.FLAG_BVH2, stream=stream_ptr)
        rb.upload(scene)
        rb.render(S, S, sync=True)
        cb = rb.counters()
        rb.close()
    else:
        cb = c
    per_ext = 64.0 * (cb["ext_box_tests"] / 2 + cb["ext_triangle_tests"]) / max(cb["extension_rays"], 1)
    per_con = 64.0 * ((cb["box_tests"] - cb["ext_box_tests"]) / 2 + (cb["triangle_tests"] - cb["ext_triangle_tests"])) / max(cb["shadow_rays"], 1)
    node_b, fan = (96.0, 8) if wide else (64.0, 2)
    own_ext = node_b * c["ext_box_tests"] / fan + 64.0 * c["ext_triangle_tests"]
    own_con = node_b * (c["box_tests"] - c["ext_box_tests"]) / fan + 64.0 * (c["triangle_tests"] - c["ext_triangle_tests"])
    kern = {"k_extend": (per_ext * c["extension_rays"], c["ms_extend"], c["launches_extend"], c["extension_rays"], own_ext, per_ext),
            "k_connect": (per_con * c["shadow_rays"], c["ms_connect"], c["launches_connect"], c["shadow_rays"], own_con, per_con)}
    dom = max(kern, key=lambda k: kern[k][1])
    ncu = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            ncu = json.load(f).get(args.workload, {})
    # measured read bandwidths with the traversal's own load instruction: 32 MB (L2 hits) and 4 GB (HBM)
    rm = pt.PathTracerCUDA(64, 64, mode=pt.MODE_FAST, device=device, flags=pt.FLAG_NO_VARIANCE, stream=stream_ptr)
    l2_read = rm.measure_read_bandwidth(32 << 20, 200)
    hbm_read = rm.measure_read_bandwidth(4 << 30, 3)
    rm.close()
    # working set: the flattened BVH2 + triangle records (the wide walk's own records are about the same size)
    scene_bytes = info.triangulationSize * 64 + (info.bvhSize // 2) * 64
    resident = scene_bytes < 0.8 * L2_BYTES

    def roofline(name):
        b, t_ms, n, nrays, own, per_ray = kern[name]
        sec = t_ms * 1e-3
        alg = b / sec / 1e9 if sec > 0 else 0.0
        fetched = own / sec / 1e9 if sec > 0 else 0.0
        k = ncu.get(name, {}) if isinstance(ncu.get(name), dict) else {"dram_bytes_per_launch": ncu.get(name)}
        dram = k.get("dram_bytes_per_launch")
        launch_ms = t_ms / max(n, 1)
        out = {"kernel": name, "unit": "GB/s", "achieved": alg,
               "algorithmic_bytes_per_launch": b / max(n, 1), "algorithmic_bytes_per_ray": per_ray,
               "fetched_bytes_per_launch": own / max(n, 1), "fetched_gbs": fetched,
               "avg_launch_ms": launch_ms, "launches": n, "mrays_per_s": nrays / sec / 1e6 if sec > 0 else 0.0,
               "l2_read_gbs_measured": l2_read, "hbm_read_gbs_measured": hbm_read, "hbm_copy_gbs": peaks["hbm_gbs"],
               "traffic": dram}
        if resident:
            # node + triangle records live in the L2 (ncu: DRAM 3-5 % of peak): the memory roofline of this kernel is the L2 read
            # rate, measured here with the kernel's own 256-bit loads; what actually limits it is instruction issue (see `limiter`)
            out.update({"bound": "l2", "peak": l2_read, "frac": alg / l2_read if l2_read else None,
                        "peak_kind": "L2 read bandwidth measured in this run (ptc_measure_read_bandwidth, 32 MB buffer, the traversal's own LDG.E.256); "
                                     "scene records %.0f MB are L2-resident" % (scene_bytes / 1e6),
                        "fetched_frac": fetched / l2_read if l2_read else None,
                        "algorithmic_over_hbm_copy": alg / peaks["hbm_gbs"]})
        else:
            out.update({"bound": "hbm", "peak": peaks["hbm_gbs"], "frac": alg / peaks["hbm_gbs"],
                        "peak_kind": "%s HBM copy bandwidth; scene records %.0f MB exceed the L2" % (peak_src, scene_bytes / 1e6),
                        "fetched_frac": fetched / peaks["hbm_gbs"]})
        if dram:
            # ncu's per-launch DRAM bytes were captured on the larger launches of a step (cold caches, serialised): scaled by time
            out["dram_gbs_ncu"] = k.get("dram_gbs")
            if k.get("dram_gbs"):
                # what the DRAM itself moved for this kernel, as a fraction of the HBM copy rate: the algorithmic fraction above counts
                # every node / triangle visit, and the L2 serves most of them (top of the tree, coherent camera rays), so it can pass 1
                out["dram_frac_ncu"] = k["dram_gbs"] / peaks["hbm_gbs"]
                out["l2_hit_pct_ncu"] = k.get("l2_hit_pct")
                if not resident:
                    out["frac_note"] = ("frac = algorithmic bytes (64 B per BVH2 node visit and triangle test of the reference walk) / time / HBM copy rate; "
                                        "the L2 (hit rate %.0f %% under ncu) serves repeated visits, so DRAM itself runs at dram_frac_ncu = %.2f of the copy rate "
                                        "and the kernel is issue bound like the L2-resident scenes (see limiter)" % (k.get("l2_hit_pct") or 0.0, out["dram_frac_ncu"]))
        if k.get("issue_active_pct") is not None:
            out["limiter"] = "instruction issue + SIMT efficiency (ncu %s: issue slots %.0f %% busy, %.1f of 32 lanes per instruction, %s spill instructions)" % (
                k.get("source", "profiles/"), k["issue_active_pct"], k["lanes_per_inst"], k.get("spill_insts", "?"))
        return out

    roof = roofline(dom)
    extra = {"roofline_other": roofline("k_connect" if dom == "k_extend" else "k_extend")}
    tot = sum(c["ms_" + k] for k in ("generate", "extend", "connect", "shade", "accumulate"))
    extra["stage_ms_per_step"] = {k: c["ms_" + k] for k in ("generate", "extend", "connect", "shade", "accumulate")}
    extra["stage_share"] = {k: (c["ms_" + k] / tot if tot else 0.0) for k in ("generate", "extend", "connect", "shade", "accumulate")}
    extra["work_per_path"] = {"box_tests": c["box_tests"] / c["paths"], "triangle_tests": c["triangle_tests"] / c["paths"],
                              "rays": (c["extension_rays"] + c["shadow_rays"]) / c["paths"]}
    # the north star's "extension throughput": closest-hit rays per second inside k_extend, and over the whole step
    extra["extension_grays_per_s"] = {"in_kernel": c["extension_rays"] / (c["ms_extend"] * 1e-3) / 1e9 if c["ms_extend"] else None,
                                      "whole_step": c["extension_rays"] / (tot * 1e-3) / 1e9 if tot else None}
    return roof, extra


# ------------------------------------------------------------------------------------------------ CUDA arm, one process per GPU
def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist

    import opencl_pathtracer_b200 as pt
    from opencl_pathtracer_b200.sharding import accumulator_tensors, iteration_range, reduce_accumulators

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this benchmark has no CPU fallback (use --impl reference for the CPU reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    quiet = QuietStdout()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.workload]
    W, H = wl["size"]
    strong = args.total_spp > 0
    K, WU = args.steps, args.warmup
    S = args.total_spp if strong else (args.spp_per_step or wl["spp"])  # iterations per step: of the whole job (strong) / per GPU (weak)
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

    def step_range(step):
        """iterations this rank renders in `step`"""
        if strong:
            return iteration_range(rank, world, step * S, S)
        return (step * world + rank) * S, S

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def over_ranks(x, op):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=op)
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
        acc = accumulator_tensors(r, dev)  # colour, count, the three histograms: aliases of the library's device buffers
        color, count = acc[0], acc[1]
        host_c = torch.empty(n_pix * 4, dtype=torch.float32, pin_memory=True)
        host_n = torch.empty(n_pix, dtype=torch.float32, pin_memory=True)
        tmp = [torch.empty_like(a) for a in acc] if world > 1 else None  # a reduce inside a loop sums copies: the accumulators keep running

        def reduce_and_read_back():
            """sum of all ranks' accumulators so far -> rank 0 -> pinned host memory (what OpenCL_RunKernel's read-back hands the caller)"""
            if world == 1:
                host_c.copy_(color, non_blocking=True)
                host_n.copy_(count, non_blocking=True)
                return
            for t, a in zip(tmp, acc):
                t.copy_(a)
            reduce_accumulators(tmp, dst=0)
            if rank == 0:
                host_c.copy_(tmp[0], non_blocking=True)
                host_n.copy_(tmp[1], non_blocking=True)

        # ---------------- device-timed region. weak: K steps + one reduce. strong: every step = its share of S iterations, then the
        # reduce and the copy of the image to pinned host memory
        for i in range(WU):
            b, k = step_range(i)
            r.render(b, k)
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
            b, k = step_range(WU + i)
            r.render(b, k)
            if strong:
                reduce_and_read_back()
        if not strong:
            reduce_accumulators(acc, dst=0)
        e1.record(stream)
        barrier()
        ms = over_ranks(e0.elapsed_time(e1), dist.ReduceOp.MAX)
        clk = clocks.stop() if rank == 0 else None
        c1 = r.counters()
        launches = c1["kernel_launches"] - c0["kernel_launches"]
        rays = over_ranks((c1["extension_rays"] - c0["extension_rays"]) + (c1["shadow_rays"] - c0["shadow_rays"]), dist.ReduceOp.SUM)
        paths = over_ranks(c1["paths"] - c0["paths"], dist.ReduceOp.SUM)
        expect = float(K) * S * n_pix * (1 if strong else world)
        assert paths == expect, (paths, expect)
        mean_rgb = None
        if rank == 0:
            mean_rgb = float((color.view(-1, 4)[:, :3].sum(0) / count.sum()).mean().item())
        value = paths / (ms * 1e-3)
        # the reference's own ray count (Controleur/PathTracer.cpp:238-243): every path that ended at depth i is charged
        # (1 + lights) * (i + 1) rays, whether or not its shadow rays were traced. The histograms were reduced: rank 0 holds the job's
        depths = r.read_stats()[0].astype(np.float64)
        ref_rays = float(sum((1 + info.lightsSize) * (i + 1) * depths[i] for i in range(min(DEPTH, len(depths)))))
        if not strong and rank == 0:
            assert depths.sum() == paths, "the depth histogram was not reduced over the ranks"
        if strong and world > 1:
            ref_rays = None  # the running histograms of a strong run live in the per-step copies

        # ---------------- e2e: the reference-facing call sequence with host buffers, host wall clock
        r.reset()
        cam = scene.camera()
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            r.set_camera(cam)  # the step's input: 64 bytes host -> device (kernel parameter block)
            b, k = step_range(i)
            r.render(b, k)
            if world == 1:
                r.read_image_into(host_c.data_ptr(), host_n.data_ptr())  # device -> pinned host, synchronises
            else:
                reduce_and_read_back()
                stream.synchronize()
        barrier()
        e2e_s = over_ranks(time.perf_counter() - t0, dist.ReduceOp.MAX)
        e2e_value = expect / e2e_s
        r.close()

        roof, extra = None, {}
        if rank == 0:
            roof, extra = roofline_block(pt, args, scene, W, H, (args.spp_per_step or wl["spp"]), mode, main_flags, local_rank, stream.cuda_stream)
            # PARITY mode beside it (bit-exact arithmetic + the reference's culling quirks), 2 steps
            if world == 1 and args.mode == "fast" and args.workload != "cornell_area":  # PARITY has no area lights (an extension)
                Sq = args.spp_per_step or wl["spp"]
                rq = pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=pt.MODE_PARITY, device=local_rank, flags=pt.FLAG_NO_VARIANCE,
                                       stream=stream.cuda_stream)
                rq.upload(scene)
                rq.render(0, Sq, sync=True)
                q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                q0.record(stream)
                rq.render(Sq, 2 * Sq)
                q1.record(stream)
                rq.synchronize()
                cq = rq.counters()
                rq.close()
                qms = q0.elapsed_time(q1)
                extra["parity_mode"] = {"value": 2 * Sq * n_pix / (qms * 1e-3), "unit": "samples/s", "ms_per_step": qms / 2,
                                        "box_tests_per_path": cq["box_tests"] / cq["paths"]}
                extra["value_parity_mode"] = extra["parity_mode"]["value"]  # the mode the bit-exact claims are made in

    # ---------------- the drop-in DEFAULT (CUDA_Options{}: PARITY, read-back + callback after every iteration), 16 iterations
    if rank == 0 and world == 1 and not args.no_default_leg and args.workload != "cornell_area":
        try:
            d = pt.run_default_drop_in(scene, 16, DEPTH)
            sec = (d["trace_ms"] + d["readback_ms"] + d["display_ms"]) * 1e-3
            extra["e2e_default"] = {"value": 16 * n_pix / sec, "unit": "samples/s", "iterations": 16, "trace_ms": d["trace_ms"], "readback_ms": d["readback_ms"],
                                    "d2h_bytes_per_iteration": n_pix * 20,
                                    "what": "CUDA_SetupContext / CUDA_InitializeMemory / CUDA_RunKernel with CUDA_Options{} untouched: PARITY arithmetic, 1 spp per pass, "
                                            "imageColor + imageRayNb read back after every iteration (reference Controleur/PathTracer_OpenCL.cpp:76-107)"}
        except Exception as e:
            extra["e2e_default"] = {"value": None, "error": str(e)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.workload != "cornell_area":  # nor has the reference kernel
        try:
            from oracle import ptref
            # a bounded sample: the whole frame once where that takes ~10 s (1080p atrium / Cornell); for the 10 M-triangle scene at 4K
            # the 1080p-equivalent number of rows ("1 spp at 1080p, stated")
            rows = (0, H) if n_pix <= 1920 * 1080 else (H // 2 - H // 8, H // 2 + H // 8)
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
        par = ("strong: %d iterations per step split over %d GPU(s) by iteration range, NCCL reduce of accumulators + histograms and image read-back inside every step" % (S, world)) \
            if strong else ("iteration-range sharding x%d (one process per GPU), one NCCL reduce of the accumulators + histograms per run" % world)
        line = {"metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": WU,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "mrays_per_s": rays / (ms * 1e-3) / 1e6,
                "mrays_per_s_reference_convention": ref_rays / (ms * 1e-3) / 1e6 if ref_rays is not None else None,
                "config": {"workload": wl["desc"], "image": [W, H], "ray_max_depth": DEPTH, "sampler": "jittered", "mode": args.mode,
                           "spp_per_step": S, "spp_per_step_is": "whole job" if strong else "per GPU", "triangles": int(info.triangulationSize),
                           "bvh_nodes": int(info.bvhSize), "parallelism": par,
                           "l2": "not flushed: each step streams ~%.1f GB of path state (> L2); scene records %.0f MB (%s the 126 MB L2)"
                                 % ((S // (world if strong else 1)) * n_pix * 112 / 1e9, scene_mb, "inside" if scene_mb < 100 else "larger than")},
                "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": n_pix * 20,
                        "ms_per_step": 1000.0 * e2e_s / K},
                "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "cpu_baseline": cpu, "mean_rgb": mean_rgb, "setup": setup}
        line.update(extra)
        quiet.emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ CUDA arm, ONE process, C ABI
def run_ours_single_process(args):
    """N GPUs through ptc_create_multi from one process and one host thread (no torch.distributed): the multi-GPU context of the C
    ABI. Timing: host wall clock around render + ptc_synchronize of all devices (weak), or render + ptc_read_image into pinned host
    memory (strong; the read sums the devices with ncclReduce inside the library)."""
    import ctypes as C

    import numpy as np

    import opencl_pathtracer_b200 as pt
    from opencl_pathtracer_b200 import _native

    L = _native.cuda_lib()
    if L.ptc_device_count() < args.gpus:
        raise SystemExit("bench.py: %d CUDA devices, --gpus %d" % (L.ptc_device_count(), args.gpus))
    quiet = QuietStdout()
    wl = WORKLOADS[args.workload]
    W, H = wl["size"]
    n_pix = W * H
    N = args.gpus
    strong = args.total_spp > 0
    S = args.total_spp if strong else (args.spp_per_step or wl["spp"]) * N  # iterations per step over ALL devices
    K, WU = args.steps, args.warmup
    mode = pt.MODE_FAST if args.mode == "fast" else pt.MODE_PARITY
    scene = pt.Scene(wl["scene"], W, H, load_sky=wl["sky"])
    info = scene.info
    L.ptc_host_alloc.restype = C.c_void_p
    hc, hn = L.ptc_host_alloc(n_pix * 16), L.ptc_host_alloc(n_pix * 4)
    with pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=mode, flags=pt.FLAG_NO_VARIANCE, devices=list(range(N))) as r:
        t0 = time.perf_counter()
        r.upload(scene)
        t_upload = time.perf_counter() - t0
        for i in range(WU):
            r.render(i * S, S)
        r.read_image_into(hc, hn)
        r.reset()
        c0 = r.counters()
        clocks = ClockSampler(0)
        clocks.start()
        t0 = time.perf_counter()
        for i in range(K):
            r.render((WU + i) * S, S)
            if strong:
                # ncclReduce into device 0 (whose accumulators then hold the running total; the others restart from zero) + copy to
                # pinned host memory
                r.read_image_into(hc, hn)
        if not strong:
            r.read_image_into(hc, hn)
        sec = time.perf_counter() - t0
        clk = clocks.stop()
        c1 = r.counters()
        assert c1["paths"] - c0["paths"] == K * S * n_pix, (c1["paths"], c0["paths"], K, S, n_pix)
        rays = (c1["extension_rays"] - c0["extension_rays"]) + (c1["shadow_rays"] - c0["shadow_rays"])
        kind = r.reduce_kind()
        img = np.ctypeslib.as_array(C.cast(hc, C.POINTER(C.c_float)), shape=(n_pix, 4))
        cnt = np.ctypeslib.as_array(C.cast(hn, C.POINTER(C.c_float)), shape=(n_pix,))
        mean_rgb = float((img[:, :3].sum(0) / cnt.sum()).mean())
        launches = c1["kernel_launches"] - c0["kernel_launches"]
    L.ptc_host_free(hc); L.ptc_host_free(hn)
    value = K * S * n_pix / sec
    line = {"metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": N, "steps": K, "warmup": WU, "ms_per_step": 1000.0 * sec / K,
            "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_per_s": rays / sec / 1e6,
            "config": {"workload": wl["desc"], "image": [W, H], "ray_max_depth": DEPTH, "sampler": "jittered", "mode": args.mode,
                       "spp_per_step": S, "spp_per_step_is": "whole job", "triangles": int(info.triangulationSize), "bvh_nodes": int(info.bvhSize),
                       "parallelism": "ONE process, one host thread, ptc_create_multi over %d devices (iteration ranges per device); accumulators + histograms summed by the "
                                      "library (%s) when the image is read%s" % (N, kind, "; read-back inside every step" if strong else ", once per run"),
                       "timing": "host wall clock, read-back to pinned host memory included"},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": n_pix * 20 if strong else n_pix * 20 // K,
                    "ms_per_step": 1000.0 * sec / K},
            "gpu_launches": int(launches), "clocks": clk, "roofline": None, "cpu_baseline": None, "mean_rgb": mean_rgb,
            "setup": {"upload_s": t_upload, "what": "one flattening + upload to device 0, peer copies to the others"}}
    quiet.emit(json.dumps(line))


# ------------------------------------------------------------------------------------------------ progressive frames (C5)
def run_progressive(args, rank=0, local_rank=0, world=1):
    """BASELINE.json configs[4] (C5): progressive 1 spp/frame 1080p with a moving camera (the Maya viewport mode).
    Per frame, through the C ABI with host buffers: ptc_set_camera -> ptc_reset_accumulation -> ptc_render(1 iteration) ->
    ptc_read_rgb8 (display epilogue on the device, 3 bytes/pixel back to the host). Reports p50 / p99 frame latency, for all frames
    of the orbit and for the frames that look into the box (mean RGB8 > 1: outside it every path ends at its first hit).
    N > 1 GPUs: a 1-spp frame cannot be split by iterations: every GPU renders interleaved 8-row stripes, converts
    them to RGB8 on the device, one uint8 NCCL reduce composes the frame on GPU 0, which copies it to pinned host memory. Under
    torchrun: one process per GPU (ptc_set_pixel_stripes + torch.distributed); without: ONE process, ptc_create_multi."""
    import numpy as np
    import torch

    import opencl_pathtracer_b200 as pt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this benchmark has no CPU fallback")
    import torch.distributed as dist
    from opencl_pathtracer_b200.sharding import STRIPE_ROWS, compose_stripes, display_tensor
    single = world == 1 and args.gpus > 1  # one process, several devices
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
    lat, inside = [], []
    kw = dict(devices=list(range(args.gpus))) if single else dict(device=local_rank, stream=stream.cuda_stream)
    with pt.PathTracerCUDA(W, H, ray_max_depth=DEPTH, mode=mode, flags=pt.FLAG_NO_VARIANCE, iterations_per_batch=1, **kw) as r:
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
                if rank == 0:
                    inside.append(float(rgb[::8, ::8].mean()) > 1.0)
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

    def pct(v, q):
        v = sorted(v)
        return v[min(len(v) - 1, int(len(v) * q))] if v else None

    lat_in = [l for l, i in zip(lat, inside) if i]
    n_gpus = args.gpus if single else world
    if n_gpus == 1:
        per = "ptc_set_camera + ptc_reset_accumulation + ptc_render(1) + ptc_read_rgb8"
        par = "whole frame on one GPU"
    elif single:
        per = "one process, ptc_create_multi: ptc_set_camera + ptc_reset_accumulation + ptc_render(1) (8-row stripes per device) + ptc_read_rgb8 (per-device RGB8, one uint8 reduce inside the library, copy to pinned host)"
        par = "interleaved %d-row stripes x%d inside one C-ABI context" % (STRIPE_ROWS, n_gpus)
    else:
        per = "per rank: ptc_set_camera + ptc_reset_accumulation + ptc_render(1) of its 8-row stripes + ptc_convert_rgb8; uint8 NCCL reduce to rank 0; copy to pinned host"
        par = "interleaved %d-row stripes x%d, one process per GPU, one uint8 NCCL reduce per frame" % (STRIPE_ROWS, n_gpus)
    line = {"metric": "frame latency p50 (progressive 1 spp/frame, moving camera)", "value": pct(lat, 0.5), "unit": "ms", "n_gpus": n_gpus, "steps": frames,
            "warmup": warm, "ms_per_step": 1e3 * total / frames, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "p99_ms": pct(lat, 0.99), "frames_per_s": frames / total, "samples_per_s": frames * W * H / total,
            "interior_frames": {"count": len(lat_in), "p50_ms": pct(lat_in, 0.5), "p99_ms": pct(lat_in, 0.99),
                                "what": "frames whose display image is not black: the camera looks into the box and paths really bounce"},
            "config": {"workload": "C5: " + wl["desc"] + "; camera orbits 0.5 deg/frame, accumulation reset every frame", "image": [W, H],
                       "ray_max_depth": DEPTH, "mode": args.mode, "per_frame": per, "parallelism": par},
            "e2e": {"value": pct(lat, 0.5), "unit": "ms", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": W * H * 3},
            "gpu_launches": int(per_frame) * frames * (world if world > 1 else 1), "mean_rgb8": float(rgb.mean())}
    quiet.emit(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="atrium", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="fast", choices=["fast", "parity"])
    ap.add_argument("--spp-per-step", type=int, default=0, help="weak scaling: iterations per step per GPU (0 = the workload's default)")
    ap.add_argument("--total-spp", type=int, default=0, help="strong scaling: iterations per step for the WHOLE job, split over the GPUs "
                                                             "(BASELINE.json: 256 for the atrium, 1024 for the terrain), reduce + read-back inside every step")
    ap.add_argument("--ref-rows", type=int, default=128)
    ap.add_argument("--bvh", choices=["host", "device"], default="host",
                    help="FAST mode: wide BVH collapsed from the reference's host-built BVH2 (default) or built on the device (PTC_FLAG_GPU_BVH)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-default-leg", action="store_true", help="skip the e2e_default leg (the drop-in with CUDA_Options{} untouched)")
    ap.add_argument("--progressive", type=int, default=0, metavar="FRAMES",
                    help="C5: time FRAMES progressive frames (1 spp each, moving camera, RGB8 read-back) instead of the throughput run")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.progressive:
        run_progressive(args, rank, local_rank, world)
        return
    if world == 1 and args.gpus > 1:
        run_ours_single_process(args)  # no torchrun: the C ABI's own multi-GPU context
        return
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
