"""CPU: the oracle (reference kernel + builder compiled from /root/reference) against the committed golden vectors,
and properties of the reference's own algorithm (its in-kernel ASSERTs) restated as tests."""
import hashlib
import os

import numpy as np
import pytest

import opencl_pathtracer_b200 as pt
from golden_cases import CASES, key, load


def test_record_sizes_match_reference(ptref):
    s = ptref.sizes()
    # host structs (Controleur/PathTracer_Structs.h) and kernel structs (header.cl) agree, Ray3D is 128 bytes
    assert list(s[:7]) == [64, 160, 336, 64, 48, 12, 92]
    assert list(s[7:14]) == [64, 160, 336, 64, 48, 12, 92]
    assert s[14] == 128


def test_rng_known_answers(ptref):
    g = load()
    for i, row in enumerate(g["rng_kat"]):
        x, y, it, W, H, s0 = [int(v) for v in row[:6]]
        seed0, seeds, us = ptref.rng(x, y, it, W, H, 8)
        assert seed0 == s0 and np.array_equal(seeds, row[6:].astype(np.int32))
        assert np.array_equal(us.view(np.uint32), g["rng_u_%d" % i].view(np.uint32))
        # restatement of header.cl:246-264 in integer arithmetic (what the CUDA kernel computes)
        s = (x + y * W + it * W * H) & 0xFFFFFFFF
        s = (s * 2011) & 0xFFFFFFFF
        s = (s * s) & 0xFFFFFFFF
        s = s or 1
        s = s - (1 << 32) if s >= (1 << 31) else s
        assert s == seed0
        st = s
        for k in range(8):
            st = (16807 * (st & 0xFFFFFFFF)) & 0x7FFFFFFF
            assert st == int(seeds[k])
            assert np.float32(st) / np.float32(2147483648.0) == us[k]


def test_cosine_sample_known_answers(ptref):
    g = load()
    for i, (seed, n) in enumerate([(12345, (0.0, 0.0, 1.0, 0.0)), (99991, (0.6, 0.0, 0.8, 0.0)), (31337, (-0.48, 0.6, -0.64, 0.0))]):
        v, after = ptref.cosine_sample(seed, n)
        assert after == int(g["cos_after_%d" % i][0])
        assert np.allclose(v, g["cos_v_%d" % i], rtol=0, atol=2e-7)  # libm ulps only
        assert abs(np.linalg.norm(v) - 1) < 1e-5 and np.dot(v[:3], n[:3]) >= 0  # kernel.cl:333


@pytest.mark.parametrize("name,w,h,sky,depth", CASES)
def test_oracle_matches_golden(ptref, scene_factory, name, w, h, sky, depth):
    g, k = load(), key(name)
    sc = scene_factory(name, w, h, sky)
    p = ptref.primary(sc, w, h, 0, 0)
    for f in ("hit", "seed_after", "counts"):
        assert np.array_equal(p[f], g[k + "_primary_" + f]), f
    for f in ("st", "dist2", "dir"):
        assert np.array_equal(p[f].view(np.uint32), g[k + "_primary_" + f].view(np.uint32)), f
    acc = ptref.render(sc, w, h, depth, 0, 0, 4)
    assert np.array_equal(acc["imageRayNb"], g[k + "_acc4_imageRayNb"])
    assert np.array_equal(acc["rayDepths"], g[k + "_acc4_rayDepths"])
    assert np.allclose(acc["imageColor"], g[k + "_acc4_imageColor"], rtol=1e-5, atol=1e-6)
    # properties the reference asserts in-kernel
    assert (acc["imageColor"][..., :3] >= 0).all()                      # L >= 0 (kernel.cl:951)
    assert acc["rayDepths"].sum() == w * h * 4
    d = p["dir"]
    assert np.allclose((d * d).sum(1), 1, atol=1e-5)                     # unit directions
    hit = p["hit"] >= 0
    assert ((p["st"][hit] >= 0).all() and (p["st"][hit].sum(1) <= 1).all())  # barycentrics inside (kernel.cl:561)
    # rays that start on surfaces (the bounces), through BVH_IntersectRay of the reference
    from golden_cases import surface_rays
    o, dd = surface_rays(sc, 4000, 21)
    assert np.array_equal(o, g[k + "_rays_origin"]) and np.array_equal(dd, g[k + "_rays_direction"])
    tr = ptref.trace_rays(sc, o, dd)
    for f in ("hit", "counts"):
        assert np.array_equal(tr[f], g[k + "_rays_" + f]), f
    for f in ("st", "dist2", "dir"):
        assert np.array_equal(tr[f].view(np.uint32), g[k + "_rays_" + f].view(np.uint32)), f
    assert (tr["hit"] >= 0).mean() > 0.3


def test_oracle_threads_do_not_change_result(ptref, scene_factory):
    sc = scene_factory("cornell", 64, 48, False)
    a = ptref.render(sc, 64, 48, 10, 0, 0, 2, nthreads=1)
    b = ptref.render(sc, 64, 48, 10, 0, 0, 2, nthreads=4)
    assert np.array_equal(a["imageColor"].view(np.uint32), b["imageColor"].view(np.uint32))
    assert np.array_equal(a["rayIntersectedBBx"], b["rayIntersectedBBx"])


def test_oracle_iteration_ranges_compose(ptref, scene_factory):
    sc = scene_factory("cornell", 64, 48, False)
    a = ptref.render(sc, 64, 48, 10, 0, 0, 4)
    b = ptref.render(sc, 64, 48, 10, 0, 0, 1)
    b = ptref.render(sc, 64, 48, 10, 0, 1, 4, acc=b)
    assert np.array_equal(a["imageColor"].view(np.uint32), b["imageColor"].view(np.uint32))


def test_row_sample_is_a_subset(ptref, scene_factory):
    sc = scene_factory("cornell", 64, 48, False)
    full = ptref.render(sc, 64, 48, 10, 0, 0, 1)
    part = ptref.render(sc, 64, 48, 10, 0, 0, 1, rows=(10, 20))
    assert np.array_equal(part["imageColor"][10:20].view(np.uint32), full["imageColor"][10:20].view(np.uint32))
    assert part["imageRayNb"].sum() == 10 * 64


def test_supersampling_build_matches_golden(ptref, scene_factory):
    """The reference kernel built with `-D SUPER_SAMPLING` (Controleur/PathTracer_OpenCL.cpp:311; the table of
    oracle/shim/X2inv.cl stands in for the reference's empty Kernel/X2inv.cl) against tests/golden/supersampling_vectors.npz:
    bit-exact accumulators, i.e. the same paths dropped."""
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "supersampling_vectors.npz"))
    w, h, depth, n_it = (int(v) for v in g["case"])
    assert np.array_equal(ptref.x2inv(), g["x2inv"])
    sc = scene_factory("cornell", w, h, False)
    acc = ptref.render(sc, w, h, depth, ptref.SAMPLER_JITTERED_SUPERSAMPLING, 0, n_it)
    for k in ("imageColor", "imageRayNb", "imageV", "rayDepths", "rayIntersectedBBx", "rayIntersectedTri"):
        assert np.array_equal(acc[k], g[k]), k
    assert g["imageRayNb"].min() >= 6 and g["imageRayNb"].sum() < 0.8 * w * h * n_it
