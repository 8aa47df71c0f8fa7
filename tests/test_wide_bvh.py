"""CPU checks of the FAST-mode 8-wide compressed BVH (csrc/pt_wide_build.cu, decoded as csrc/pt_wide.cuh does):
structure, conservative boxes, and a scalar restatement of the traversal against brute force."""
import numpy as np
import pytest


def _decode(nodes, i):
    n = nodes[i]
    step = np.ldexp(1.0, n["e"].astype(np.int64) - 127)  # per axis
    lo = n["p"].astype(np.float64)[:, None] + n["qlo"].astype(np.float64) * step[:, None]
    hi = n["p"].astype(np.float64)[:, None] + n["qhi"].astype(np.float64) * step[:, None]
    return n, lo, hi  # lo/hi: [axis, slot]


def _children(n, i):
    """(slot, kind, first, count): inner -> node index, leaf -> first record of the sparse triangle order."""
    out = []
    valid, imask = int(n["valid"]), int(n["imask"])
    assert valid >> 24 == 0
    for s in range(8):
        bits = (valid >> (3 * s)) & 7
        if (imask >> s) & 1:
            assert bits == 0
            rel = bin(imask & ((1 << s) - 1)).count("1")
            out.append((s, "inner", int(n["childBase"]) + rel, 1))
        elif bits:
            assert bits in (1, 3, 7)
            out.append((s, "leaf", 24 * i + 3 * s, bin(bits).count("1")))
    return out


def check_wide_bvh(nodes, order, depth, tris):
    """Structure of a wide BVH over `tris` (scene.TRIANGLE records in the order `order` refers to): every triangle exactly
    once, every quantised child box contains its subtree and is at most one grid step loose, node count and depth."""
    nT = len(tris)
    assert len(order) == 24 * len(nodes)
    assert sorted(order[order != 0xffffffff].tolist()) == list(range(nT)), "every triangle exactly once"
    verts = np.stack([np.stack([tris[k][a] for a in "xyz"], -1) for k in ("S1", "S2", "S3")], 1).astype(np.float64)  # [tri, vertex, axis]
    seen = np.zeros(len(nodes), bool)
    seen_tris, max_depth = 0, 0
    # iterative post-order: (node, depth, expanded)
    bounds = {}
    stack = [(0, 1, False)]
    while stack:
        i, d, expanded = stack.pop()
        n, lo, hi = _decode(nodes, i)
        kids = _children(n, i)
        if not expanded:
            assert not seen[i]
            seen[i] = True
            max_depth = max(max_depth, d)
            stack.append((i, d, True))
            for s, kind, first, count in kids:
                if kind == "inner":
                    stack.append((first, d + 1, False))
            continue
        blo, bhi = np.full(3, np.inf), np.full(3, -np.inf)
        step = np.ldexp(1.0, n["e"].astype(np.int64) - 127)
        for s, kind, first, count in kids:
            if kind == "inner":
                clo, chi = bounds.pop(first)
            else:
                idx = order[first:first + count]
                assert (idx != 0xffffffff).all()
                v = verts[idx].reshape(-1, 3)
                clo, chi = v.min(0), v.max(0)
                seen_tris += count
            assert (lo[:, s] <= clo).all() and (hi[:, s] >= chi).all(), "quantised box must contain the child"
            assert (clo - lo[:, s] < step * 1.0001).all() and (hi[:, s] - chi < step * 1.0001).all()
            blo, bhi = np.minimum(blo, clo), np.maximum(bhi, chi)
        assert (n["p"].astype(np.float64) <= blo).all()
        bounds[i] = (blo, bhi)
    assert seen.all() and seen_tris == nT and max_depth == depth


@pytest.mark.parametrize("name", ["cornell", "gallery", "atrium:8"])
def test_wide_bvh_structure_and_boxes(scene_factory, name):
    sc = scene_factory(name, 64, 48, load_sky=(name == "gallery"))
    nodes, order, depth = sc.wide_bvh()
    check_wide_bvh(nodes, order, depth, sc.triangles())


def _tri_hit(o, d, v):
    e1, e2 = v[1] - v[0], v[2] - v[0]
    p = np.cross(d, e2)
    det = e1 @ p
    if abs(det) < 1e-14:
        return None
    t0 = o - v[0]
    u = (t0 @ p) / det
    q = np.cross(t0, e1)
    w = (d @ q) / det
    t = (e2 @ q) / det
    if u < -1e-9 or w < -1e-9 or u + w > 1 + 1e-9 or t <= 1e-6:
        return None
    return t


def test_wide_bvh_traversal_matches_brute_force(scene_factory):
    """The traversal of pt_wide.cuh restated in scalar Python (same decode rules: octant ordering, hit masks,
    relative child indices, triangle groups) finds the brute-force closest hit."""
    sc = scene_factory("atrium:8", 64, 48)
    nodes, order, _ = sc.wide_bvh()
    tris = sc.triangles()
    verts = np.stack([np.stack([tris[k][a] for a in "xyz"], -1) for k in ("S1", "S2", "S3")], 1).astype(np.float64)
    rng = np.random.default_rng(7)
    lo, hi = verts.reshape(-1, 3).min(0), verts.reshape(-1, 3).max(0)
    visited_total = 0
    for _ in range(120):
        o = lo + (hi - lo) * rng.uniform(0.2, 0.8, 3)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        # brute force
        best, best_t = -1, np.inf
        for i in range(len(verts)):
            t = _tri_hit(o, d, verts[i])
            if t is not None and t < best_t:
                best, best_t = i, t
        # wide traversal
        inv = 1.0 / np.where(np.abs(d) < 1e-30, 1e-30, d)
        octinv = 7 - ((4 if d[0] < 0 else 0) | (2 if d[1] < 0 else 0) | (1 if d[2] < 0 else 0))
        stack, g = [], (0, 0x80000000)
        hit, tmax = -1, np.inf
        while True:
            if g[1] > 0x00ffffff:
                bit = g[1].bit_length() - 1
                slot = (bit - 24) ^ octinv
                rest = g[1] & ~(1 << bit)
                rel = bin(g[1] & 0xff & ((1 << slot) - 1)).count("1")
                if rest > 0x00ffffff:
                    stack.append((g[0], rest))
                node = g[0] + rel
                n, blo, bhi = _decode(nodes, node)
                visited_total += 1
                slot_hit = 0
                for s in range(8):
                    t1, t2 = (blo[:, s] - o) * inv, (bhi[:, s] - o) * inv
                    tn, tf = max(np.minimum(t1, t2).max(), 0.0), min(np.maximum(t1, t2).min(), tmax)
                    if tn <= tf:
                        slot_hit |= 1 << s
                imask = int(n["imask"])
                ih = 0
                for s in range(8):
                    if (slot_hit & imask) >> s & 1:
                        ih |= 1 << (s ^ octinv)
                th = 0
                for s in range(8):
                    if slot_hit >> s & 1:
                        th |= 7 << (3 * s)
                th &= int(n["valid"])
                g = (int(n["childBase"]), (ih << 24) | imask)
                tg = [24 * node, th]
                while tg[1]:
                    b = (tg[1] & -tg[1]).bit_length() - 1
                    tg[1] &= tg[1] - 1
                    k = int(order[tg[0] + b])
                    t = _tri_hit(o, d, verts[k])
                    if t is not None and t < tmax:
                        hit, tmax = k, t
            elif stack:
                g = stack.pop()
            else:
                break
        assert hit == best or (best >= 0 and hit >= 0 and abs(tmax - best_t) < 1e-9 * max(1.0, best_t)), (hit, best, tmax, best_t)
    assert visited_total > 0
