#!/usr/bin/env python3
"""bench.py — headline benchmark of the path-tracing hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload atrium|cornell|terrain|terrain_1080p] [--mode fast|parity]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

Workload (BASELINE.json configs[2], the one the metric's 1/2/4/8-GPU samples/s + traversal roofline is quoted on):
procedural atrium, ~262 k triangles, diffuse + varnished + glass, 2 point + 1 directional light, 1920x1080, depth 10,
jittered sampler. One STEP = `--spp-per-step` (default 8) iterations of the whole frame on every rank; ranks render
disjoint iteration ranges (random streams depend on the iteration number only), the per-GPU accumulators are summed to
rank 0 with one NCCL reduce inside the timed region ("weak" scaling: per-GPU work fixed).

Prints ONE JSON line on rank 0. `value` = samples/s (paths/s) with the scene resident in HBM, device-timed;
`e2e` = the same metric through the reference-facing call sequence with host buffers (camera upload, render, image
read-back into pinned host memory each step); `roofline` = dominant traversal kernel vs measured HBM copy bandwidth;
`cpu_baseline` = the reference kernel itself (oracle/_ref, compiled from the reference sources) on the host cores.
`--impl reference` times that CPU reference alone.
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
    "cornell": dict(scene="cornell", size=(1920, 1080), sky=False, spp=8,
                    desc="C2: Cornell box 30 tris, diffuse + point light, 1920x1080, depth 10, jittered"),
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


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


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
    r0 = (H // 4 + step * rows_per_step) % (H - rows_per_step)
    return r0, r0 + rows_per_step


def run_reference(args, rank):
    """--impl reference: the reference's own kernel (oracle/_ref) on the host cores, bounded row samples of the same frame."""
    if rank != 0:
        return
    import opencl_pathtracer_b200 as pt
    from oracle import ptref
    wl = WORKLOADS[args.workload]
    if args.workload == "cornell_area":
        print(json.dumps({"impl": "reference", "unavailable": "the reference kernel has no area lights (PTC_LIGHT_AREA is an extension); use --workload cornell"}))
        return
    W, H = wl["size"]
    scene = pt.Scene(wl["scene"], W, H, load_sky=wl["sky"])
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 to its workers, which would leave the reference
    # kernel on one core (14 k instead of 230 k samples/s) — the thread count is passed explicitly (num_threads clause)
    cores = host_threads()
    rows = args.ref_rows
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
    line = {"impl": "reference", "metric": "samples/s", "value": v, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "image": [W, H], "ray_max_depth": DEPTH, "sampler": "jittered", "sample": sample},
            "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="atrium", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="fast", choices=["fast", "parity"])
    ap.add_argument("--spp-per-step", type=int, default=0, help="iterations per step per GPU (0 = the workload's default)")
    ap.add_argument("--ref-rows", type=int, default=128)
    ap.add_argument("--bvh", choices=["host", "device"], default="host",
                    help="FAST mode: wide BVH collapsed from the reference's host-built BVH2 (default) or built on the device (PTC_FLAG_GPU_BVH)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    run_ours(args, rank, local_rank, world)


from bench_impl import run_ours, run_progressive  # noqa: E402  (second half of this file, split to keep each file small)

if __name__ == "__main__":
    main()
