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


def _stripe_runs(rows):
    """[(first, last + 1)] of the contiguous runs in a sorted row list."""
    runs, start = [], None
    for i, y in enumerate(rows):
        if start is None:
            start = y
        if i + 1 == len(rows) or rows[i + 1] != y + 1:
            runs.append((start, y + 1))
            start = None
    return runs


def _stripe_worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist

    import opencl_pathtracer_b200 as pt
    from golden_cases import rgb8_of
    from opencl_pathtracer_b200.sharding import compose_stripes, stripe_rows
    from oracle import ptref
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w, h, depth = 48, 37, 6
    sc = pt.Scene("cornell_specular", w, h)
    acc = None
    for r0, r1 in _stripe_runs(stripe_rows(rank, world, h)):
        acc = ptref.render(sc, w, h, depth, 0, 0, 1, nthreads=2, acc=acc, rows=(r0, r1))
    frame = torch.from_numpy(rgb8_of(acc["imageColor"], acc["imageRayNb"]))
    compose_stripes(frame, dst=0)
    if rank == 0:
        np.save(os.path.join(tmp, "frame.npy"), frame.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_pixel_stripes(tmp_path, ptref):
    """Progressive frames (1 spp): interleaved 8-row stripes per rank, display frames summed as uint8 to rank 0."""
    import torch.multiprocessing as mp

    import opencl_pathtracer_b200 as pt
    from golden_cases import rgb8_of
    from opencl_pathtracer_b200.sharding import stripe_rows
    w, h, depth = 48, 37, 6
    rows = [stripe_rows(r, 2, h) for r in range(2)]
    assert sorted(rows[0] + rows[1]) == list(range(h)) and not set(rows[0]) & set(rows[1])
    assert rows[0][:9] == [0, 1, 2, 3, 4, 5, 6, 7, 16] and rows[1][:8] == list(range(8, 16))
    assert stripe_rows(0, 1, h) == list(range(h)) and stripe_rows(2, 3, 10, 4) == [8, 9]
    with pytest.raises(ValueError):
        stripe_rows(2, 2, h)
    port = 31500 + os.getpid() % 2000
    mp.spawn(_stripe_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), "frame.npy"))
    sc = pt.Scene("cornell_specular", w, h)
    ref = ptref.render(sc, w, h, depth, 0, 0, 1)
    assert np.array_equal(got, rgb8_of(ref["imageColor"], ref["imageRayNb"]))
    assert got.any()
