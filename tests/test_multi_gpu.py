"""Multi-GPU on hardware (needs >= 2 CUDA devices; skipped on a one-GPU box). SURVEY.md 8(e) / BASELINE.md section 4:
the image of N GPUs rendering disjoint shares of the SAME iterations against the image of one GPU —
counts and histograms exactly equal, colours to an RMSE of 1e-4 of the mean.

  * through the C ABI alone (ptc_create_multi: one process, one host thread, ncclCommInitAll / peer copies; ctypes, no torch);
  * through the per-process harness bench.py uses under torchrun (one rank per GPU, torch.distributed reduce over NCCL on
    tensors aliasing the library's device buffers).
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import opencl_pathtracer_b200 as pt
from opencl_pathtracer_b200 import _native

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_devices():
    try:
        return _native.cuda_lib().ptc_device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_n_devices() < 2, reason="needs two CUDA devices")


def _render(sc, w, h, depth, mode, first, n, devices=None, flags=0):
    with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=mode, devices=devices, flags=flags | (0 if devices else pt.FLAG_NO_VARIANCE)) as r:
        r.upload(sc)
        r.render(first, n)
        color, nb = r.read_image()
        stats = r.read_stats()
        rgb = r.read_rgb8()
        c = r.counters()
        kind = r.reduce_kind()
    return color, nb, stats, rgb, c, kind


def _compare(one, many, n_paths, exact_work_hist):
    (c1, n1, s1, g1, k1, _), (cm, nm, sm, gm, km, _) = one, many
    assert np.array_equal(n1, nm), "sample counts differ"
    assert k1["paths"] == km["paths"] == n_paths
    assert k1["extension_rays"] == km["extension_rays"] and k1["shadow_rays"] == km["shadow_rays"]
    assert np.array_equal(s1[0], sm[0]), "depth histograms differ"
    if exact_work_hist:
        assert np.array_equal(s1[1], sm[1]) and np.array_equal(s1[2], sm[2]), "work histograms differ"
    mean = c1[..., :3].astype(np.float64).mean()
    rmse = np.sqrt(((c1[..., :3].astype(np.float64) - cm[..., :3]) ** 2).mean())
    assert rmse <= 1e-4 * mean, "RMSE / mean = %g" % (rmse / mean)
    assert np.abs(g1.astype(np.int16) - gm.astype(np.int16)).max() <= 1
    return rmse / mean


@needs2
@pytest.mark.parametrize("reduce", ["nccl", "p2p"])
@pytest.mark.parametrize("mode", [pt.MODE_PARITY, pt.MODE_FAST])
def test_c_abi_multi_device_equals_one_device(scene_factory, monkeypatch, reduce, mode):
    """ptc_create_multi over all devices of the box, iteration ranges (incl. a ragged split) and a second call that adds to the
    first: the same image as one device rendering the same iterations."""
    monkeypatch.setenv("PTC_MULTI_REDUCE", reduce)
    nd = _n_devices()
    w, h, depth = 192, 128, 8
    sc = scene_factory("gallery", w, h, True)
    first, n = 3, 3 * nd + 1
    one = _render(sc, w, h, depth, mode, first, n)
    many = _render(sc, w, h, depth, mode, first, n, devices=list(range(nd)))
    assert many[5] == reduce
    err = _compare(one, many, w * h * n, exact_work_hist=(mode == pt.MODE_PARITY))
    print("%d devices, %s, mode %d: RMSE / mean %.3g" % (nd, reduce, mode, err))


@needs2
def test_c_abi_multi_device_progressive_stripes(scene_factory):
    """Fewer iterations than devices: pixel stripes. Every pixel is rendered by exactly one device, so the accumulators and the
    display frame are BIT-identical to one device's; repeated reads and a read_image between frames keep them so."""
    nd = _n_devices()
    w, h, depth = 160, 100, 6
    sc = scene_factory("cornell_specular", w, h, False)
    with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, flags=pt.FLAG_NO_VARIANCE) as a, \
            pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_FAST, devices=list(range(nd))) as b:
        a.upload(sc)
        b.upload(sc)
        for frame in range(3):
            for r in (a, b):
                r.reset()
                r.render(frame, 1)
            ga, gb = a.read_rgb8(), b.read_rgb8()
            assert np.array_equal(ga, gb), "frame %d" % frame
            assert np.array_equal(a.read_rgb8(), b.read_rgb8())  # a second read of the same frame
        # accumulate two more frames on top without a reset, reading the float image in between
        for r in (a, b):
            r.render(3, 1)
        (ca, na), (cb, nb) = a.read_image(), b.read_image()
        assert np.array_equal(ca.view(np.uint32), cb.view(np.uint32)) and np.array_equal(na, nb)
        for r in (a, b):
            r.render(4, 1)
        assert np.array_equal(a.read_rgb8(), b.read_rgb8())
        (ca, na), (cb, nb) = a.read_image(), b.read_image()
        assert np.array_equal(ca.view(np.uint32), cb.view(np.uint32)) and (nb == 3).all()
        assert b.counters()["paths"] == a.counters()["paths"] == 3 * w * h
        with pytest.raises(pt.PathTracerError):
            b.set_pixel_stripes(8, 0, 2)
        with pytest.raises(pt.PathTracerError):
            b.read_variance()


@needs2
def test_c_abi_multi_device_bench_scene(scene_factory):
    """BASELINE.json configs[2] (atrium, 262 k triangles, 1080p, depth 10) on all devices of the box through ptc_create_multi
    against one device, 2 iterations per device."""
    nd = _n_devices()
    w, h, depth = 1920, 1080, 10
    sc = scene_factory("atrium", w, h, False)
    n = 2 * nd
    one = _render(sc, w, h, depth, pt.MODE_FAST, 0, n)
    many = _render(sc, w, h, depth, pt.MODE_FAST, 0, n, devices=list(range(nd)))
    err = _compare(one, many, w * h * n, exact_work_hist=False)
    print("C3 on %d devices vs 1: RMSE / mean %.3g" % (nd, err))


@needs2
def test_torchrun_nccl_reduce_equals_one_gpu(tmp_path):
    """The harness path (bench.py under torchrun): world_size 2, each rank renders its iteration_range on its own GPU, the
    accumulators are summed with torch.distributed over NCCL on tensors aliasing the device buffers; rank 0 compares with a
    one-context render of all iterations: counts and all three histograms exactly equal, RMSE <= 1e-4 of the mean."""
    out = tmp_path / "result.txt"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29571",
           os.path.join(ROOT, "tests", "mgpu_torch_worker.py"), str(out)]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    text = out.read_text()
    print(text)
    assert text.startswith("OK"), text
