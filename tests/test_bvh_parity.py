"""CPU: host/PathTracer_BVH.cpp (this repo's builder) must emit exactly the reference builder's tree
(Controleur/PathTracer_BVH.cpp:12-356): node numbering, child links, cut axes, leaf ranges, boxes and the triangle
permutation — the kernel's closest-hit tie-break is traversal order."""
import hashlib

import numpy as np
import pytest

import opencl_pathtracer_b200 as pt
from opencl_pathtracer_b200 import scene as S
from golden_cases import CASES, key, load


@pytest.mark.parametrize("name", ["cornell", "cornell_specular", "gallery", "atrium:8", "atrium:2", "terrain:6"])
def test_builder_identical_to_reference(ptref, name):
    sc = pt.Scene(name, 64, 64, build_bvh=False)
    rn, rt, rsize, rdepth = ptref.bvh_create(sc.triangles())
    sc.build_bvh()
    i = sc.info
    assert (i.bvhSize, i.bvhMaxDepth) == (rsize, rdepth)
    a = sc.nodes()
    b = np.frombuffer(rn.tobytes(), dtype=S.NODE)
    for f in ("cutAxis", "triangleStartIndex", "nbTriangles", "son1Id", "son2Id", "isLeaf"):
        assert np.array_equal(a[f], b[f]), f
    for box in ("trianglesAABB", "centroidsAABB"):
        for f in ("pMin", "pMax"):
            for c in "xyz":
                assert np.array_equal(a[box][f][c], b[box][f][c]), (box, f, c)
    leaf = a["isLeaf"] != 0
    assert np.array_equal(a["comments"][leaf], b["comments"][leaf])
    assert np.array_equal(sc.triangles()["id"], rt["id"]), "triangle permutation differs"
    for f in ("S1", "S2", "S3", "N"):
        assert np.array_equal(sc.triangles()[f], rt[f]), f


@pytest.mark.parametrize("name,w,h,sky,depth", CASES)
def test_bvh_golden_digest(name, w, h, sky, depth):
    g, k = load(), key(name)
    sc = pt.Scene(name, w, h, load_sky=sky)
    n = sc.nodes()
    hsh = hashlib.sha256()
    for f in ("cutAxis", "triangleStartIndex", "nbTriangles", "son1Id", "son2Id", "isLeaf"):
        hsh.update(np.ascontiguousarray(n[f]).tobytes())
    hsh.update(np.ascontiguousarray(sc.triangles()["id"]).tobytes())
    assert hsh.hexdigest() == bytes(g[k + "_bvh_sha256"]).decode()
    i = sc.info
    assert [i.triangulationSize, i.bvhSize, i.bvhMaxDepth] == list(g[k + "_bvh_meta"])


@pytest.mark.parametrize("name", ["gallery", "atrium:4"])
def test_bvh_invariants(name):
    sc = pt.Scene(name, 32, 32)
    n, t = sc.nodes(), sc.triangles()
    i = sc.info
    assert i.bvhMaxDepth < 30
    leaf = n["isLeaf"] != 0
    inner = ~leaf
    # depth-first numbering: left child = parent + 1
    idx = np.arange(len(n))
    assert np.array_equal(n["son1Id"][inner], idx[inner] + 1)
    # leaves partition the triangle array
    order = np.argsort(n["triangleStartIndex"][leaf], kind="stable")
    starts, counts = n["triangleStartIndex"][leaf][order], n["nbTriangles"][leaf][order]
    assert starts[0] == 0 and np.array_equal(starts[1:], np.cumsum(counts)[:-1]) and counts.sum() == len(t)
    # every triangle inside its leaf box
    for li in np.nonzero(leaf)[0][:200]:
        s0, c = n["triangleStartIndex"][li], n["nbTriangles"][li]
        for ax in "xyz":
            lo = min(t[v][ax][s0:s0 + c].min() for v in ("S1", "S2", "S3"))
            hi = max(t[v][ax][s0:s0 + c].max() for v in ("S1", "S2", "S3"))
            assert n["trianglesAABB"]["pMin"][ax][li] <= lo and hi <= n["trianglesAABB"]["pMax"][ax][li]
    assert sorted(t["id"].tolist()) == list(range(len(t)))
