"""CPU, world_size 2 over gloo: the N>1 path of the render — disjoint iteration ranges per rank + one sum-reduce of the
accumulators — against a single-process run of all iterations. The per-rank "device" is the oracle here (no GPU in
this container); the sharding arithmetic and the reduce are the product's (opencl_pathtracer_b200/sharding.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist

    import opencl_pathtracer_b200 as pt
    from opencl_pathtracer_b200.sharding import iteration_range, reduce_accumulators
    from oracle import ptref
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w, h, depth, total = 48, 32, 6, 5
    sc = pt.Scene("cornell_specular", w, h)
    first, n = iteration_range(rank, world, 0, total)
    acc = ptref.render(sc, w, h, depth, 0, first, first + n, nthreads=2)
    ts = [torch.from_numpy(acc["imageColor"]), torch.from_numpy(acc["imageRayNb"]),
          torch.from_numpy(acc["rayDepths"].astype(np.int32)), torch.from_numpy(acc["rayIntersectedBBx"].astype(np.int32))]
    reduce_accumulators(ts, dst=0)
    if rank == 0:
        np.savez(os.path.join(tmp, "reduced.npz"), color=ts[0].numpy(), nb=ts[1].numpy(), depths=ts[2].numpy(), bbx=ts[3].numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_iteration_sharding(tmp_path, ptref):
    import torch.multiprocessing as mp

    import opencl_pathtracer_b200 as pt
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), "reduced.npz"))
    w, h, depth, total = 48, 32, 6, 5
    sc = pt.Scene("cornell_specular", w, h)
    ref = ptref.render(sc, w, h, depth, 0, 0, total)
    assert np.array_equal(got["nb"], ref["imageRayNb"])
    assert np.array_equal(got["depths"], ref["rayDepths"].astype(np.int32))
    assert np.array_equal(got["bbx"], ref["rayIntersectedBBx"].astype(np.int32))
    # the same paths are traced; only the fp32 summation order differs (RMSE <= 1e-4 of the mean radiance)
    mean = ref["imageColor"][..., :3].mean()
    rmse = np.sqrt(((got["color"] - ref["imageColor"]) ** 2).mean())
    assert rmse <= 1e-4 * mean
