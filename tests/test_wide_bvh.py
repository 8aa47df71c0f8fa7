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
    nodes, order, depth = sc.wide_bvh()
    tris = sc.triangles()
    verts = np.stack([np.stack([tris[k][a] for a in "xyz"], -1) for k in ("S1", "S2", "S3")], 1).astype(np.float64)
    rng = np.random.default_rng(7)
    max_stack = 0
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
                    max_stack = max(max_stack, len(stack))
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
    # the walk stacks the unvisited siblings of the child it descends into and the root has none: never more than depth - 1 groups
    # (ptc_upload_scene sizes the shared-memory stack by this bound: item word + parked group + depth - 1 + one spare level)
    assert 0 < max_stack <= depth - 1, (max_stack, depth)


# ---- the packed half-precision slab test of pt_wide.cuh (slab_test_half2), restated operation by operation in numpy
def _f32(x):
    return np.asarray(x, dtype=np.float32)


def _fma32(a, b, c):
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


def _hfma(q, a16, b16):
    # HFMA2: the exact q * A + B rounded once to half (q is the denormal half q * 2^-24: the 2^-24 is applied exactly here)
    with np.errstate(over="ignore", invalid="ignore"):
        return (np.ldexp(q.astype(np.float64), -24) * a16.astype(np.float64) + b16.astype(np.float64)).astype(np.float16)


def slab_test_half2_restated(p, e, qlo, qhi, o, inv, tmax):
    """p[3] grid origin, e[3] biased step exponents, qlo/qhi [3, 8] bytes, o[3], inv[3] (float32, finite), tmax -> 8 booleans."""
    with np.errstate(over="ignore", invalid="ignore"):
        step = np.ldexp(np.float32(1), e.astype(np.int64) - 127).astype(np.float32)
        a = (step * inv).astype(np.float32)
        b = ((p - o).astype(np.float32) * inv).astype(np.float32)
        t0 = np.float32(max(np.minimum(b, _fma32(255.0, a, b)).max(), 0.0))
        amax = np.abs(a).max()
        E = int(min(max(int(np.float32(amax).view(np.uint32)) >> 23, 14), 240))
        sA, sT = np.float32(np.ldexp(1.0, 268 - E - 127)), np.float32(np.ldexp(1.0, 244 - E - 127))
        A = (a * sA).astype(np.float32)
        B = ((b - t0).astype(np.float32) * sT).astype(np.float32)
        pn = _fma32(np.float32(t0 * sT), 4.76837158e-7, 5.9604645e-8)
        pad = _fma32(np.abs(B), 4.96e-4, _fma32(np.abs(A), 7.52e-9, pn))
        hA = A.astype(np.float16)
        hBl, hBh = (B - pad).astype(np.float32).astype(np.float16), (B + pad).astype(np.float32).astype(np.float16)
        tb = _fma32(np.float32(t0 * sT), -1.001, -5.9604645e-8).astype(np.float16)
        te0 = np.float32((np.float32(tmax) - t0) * sT)
        te = np.float32(_fma32(np.abs(te0), 1.0e-3, te0) + np.float32(5.9604645e-8)).astype(np.float16)
        neg = inv < 0
        qe = np.where(neg[:, None], qhi, qlo)
        qf = np.where(neg[:, None], qlo, qhi)
        te_ = _hfma(qe, hA[:, None], hBl[:, None])  # [axis, slot]
        tf_ = _hfma(qf, hA[:, None], hBh[:, None])
        t_enter = np.fmax(np.fmax(np.fmax(te_[0], te_[1]), te_[2]), tb)  # fmax / fmin: a NaN operand is ignored, as HMNMX2 does
        t_exit = np.fmin(np.fmin(np.fmin(tf_[0], tf_[1]), tf_[2]), te)
        return t_enter <= t_exit


def test_half2_slab_test_is_conservative(scene_factory):
    """Every child box the exact slab test (float64, on the quantised boxes) enters is entered by the packed half-precision
    test the FAST node step runs; the boxes it enters in addition are few (reported)."""
    sc = scene_factory("atrium:8", 64, 48)
    nodes, _, _ = sc.wide_bvh()
    tris = sc.triangles()
    verts = np.stack([np.stack([tris[k][a] for a in "xyz"], -1) for k in ("S1", "S2", "S3")], 1).astype(np.float64).reshape(-1, 3)
    slo, shi = verts.min(0), verts.max(0)
    rng = np.random.default_rng(11)
    exact_hits = half_hits = missed = 0
    plain_exact = plain_half = n_single = 0
    for it in range(6000):
        i = int(rng.integers(len(nodes)))
        n, blo, bhi = _decode(nodes, i)
        # a ray aimed at a point inside (or near) the node, from inside the scene, near or far; some axis-parallel, some with a bound
        nlo, nhi = blo.min(1), bhi.max(1)
        target = nlo + (nhi - nlo) * rng.uniform(-0.2, 1.2, 3)
        o = (slo + (shi - slo) * rng.uniform(0, 1, 3)) if it % 3 else (target + rng.normal(size=3) * (nhi - nlo).max() * 0.7)
        d = target - o
        plain = True
        if it % 7 == 0:
            d[int(rng.integers(3))] = 0.0
            plain = False
        if it % 11 == 0:
            d[int(rng.integers(3))] *= 1e-6
            plain = False
        if it % 13 == 0:
            d[int(rng.integers(3))] *= 2.0 ** -12  # below the kSingle threshold: the packed test still runs, with less resolution
            plain = False
        if not np.any(d):
            continue
        o32 = _f32(o)
        d32 = _f32(d / np.linalg.norm(d))
        with np.errstate(divide="ignore"):
            inv = (np.float32(1) / d32).astype(np.float32)
        inv = np.where(np.abs(inv) < 1e30, inv, np.copysign(np.float32(1e30), d32)).astype(np.float32)  # finite_inverse
        if not np.abs(inv).max() < np.float32(16384) * np.abs(inv).min():
            n_single += 1  # WideLane::kSingle: this ray takes the single-precision test
            continue
        dist = np.linalg.norm(target - o)
        tmax = np.inf if it % 2 else dist * rng.uniform(0.5, 1.5)
        got = slab_test_half2_restated(n["p"].astype(np.float32), n["e"], n["qlo"], n["qhi"], o32, inv, tmax)
        # exact: the same ray (float32 origin, the reciprocal the kernel uses) against the quantised boxes in float64
        o64, inv64 = o32.astype(np.float64), inv.astype(np.float64)
        for s in range(8):
            t1, t2 = (blo[:, s] - o64) * inv64, (bhi[:, s] - o64) * inv64
            tn, tf = max(np.minimum(t1, t2).max(), 0.0), min(np.maximum(t1, t2).min(), tmax)
            ex = tn <= tf
            exact_hits += ex
            half_hits += bool(got[s])
            plain_exact += ex and plain
            plain_half += bool(got[s]) and plain
            if ex and not got[s]:
                missed += 1
    print("half2 slab test: %d exact hits, %d half hits, %d missed (%d rays flagged kSingle left out); rays without a small direction component: %d -> %d (+%.2f %%)"
          % (exact_hits, half_hits, missed, n_single, plain_exact, plain_half, 100.0 * (plain_half - plain_exact) / max(plain_exact, 1)))
    assert missed == 0 and n_single > 500
    # the scale follows the coarsest axis, so a ray close to parallel to a coordinate plane has less resolution on the other two axes:
    # more boxes entered, never fewer (beyond a ratio of 2^14 the kernel switches to single precision). For the bulk of the rays the
    # boxes grow by a fraction of a grid step.
    assert plain_half <= plain_exact * 1.03 and half_hits <= exact_hits * 1.06
