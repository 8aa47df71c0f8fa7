"""Worker of tests/test_multi_gpu.py::test_torchrun_nccl_reduce_equals_one_gpu (launched by torch.distributed.run, one rank per GPU)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import opencl_pathtracer_b200 as pt  # noqa: E402
from opencl_pathtracer_b200.sharding import accumulator_tensors, iteration_range, reduce_accumulators  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    w, h, depth, first, n = 320, 180, 8, 2, 7
    sc = pt.Scene("atrium:4", w, h)
    stream = torch.cuda.Stream(device=dev)
    msg = "OK"
    with torch.cuda.stream(stream):
        with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_PARITY, device=local, flags=pt.FLAG_NO_VARIANCE, stream=stream.cuda_stream) as r:
            r.upload(sc)
            b, k = iteration_range(rank, world, first, n)
            r.render(b, k)
            reduce_accumulators(accumulator_tensors(r, dev), dst=0)
            stream.synchronize()
            if rank == 0:
                color, nb = r.read_image()
                stats = r.read_stats()
                with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_PARITY, device=local, flags=pt.FLAG_NO_VARIANCE) as one:
                    one.upload(sc)
                    one.render(first, n)
                    c1, n1 = one.read_image()
                    s1 = one.read_stats()
                ok = np.array_equal(nb, n1) and all(np.array_equal(a, b_) for a, b_ in zip(stats, s1))
                mean = c1[..., :3].astype(np.float64).mean()
                rmse = np.sqrt(((c1[..., :3].astype(np.float64) - color[..., :3]) ** 2).mean())
                ok = ok and rmse <= 1e-4 * mean
                msg = "%s world %d: counts+histograms equal %s, RMSE / mean %.3g" % ("OK" if ok else "FAIL", world, ok, rmse / mean)
    dist.barrier()
    if rank == 0:
        with open(sys.argv[1], "w") as f:
            f.write(msg)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
