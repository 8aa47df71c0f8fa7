import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")
CASES = [("cornell", 64, 48, False, 10), ("gallery", 80, 60, True, 6), ("atrium:8", 64, 40, False, 5)]


def load():
    return np.load(GOLDEN)


def key(name):
    return name.replace(":", "_")


def surface_rays(scene, n, seed):
    """Rays like the ones a bounce shoots: origins 0.001 off random points of random triangles, random directions."""
    rng = np.random.default_rng(seed)
    t = scene.triangles()
    k = rng.integers(0, len(t), n)
    v = [np.stack([t[name][a][k] for a in "xyzw"], -1).astype(np.float32) for name in ("S1", "S2", "S3")]
    u, w = rng.uniform(0, 1, n).astype(np.float32), rng.uniform(0, 1, n).astype(np.float32)
    flip = u + w > 1
    u[flip], w[flip] = 1 - u[flip], 1 - w[flip]
    d = rng.normal(size=(n, 4)).astype(np.float32)
    d[:, 3] = 0
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    # a share of axis-parallel and grazing directions: zero components take the slow box test in the BVH2 walk
    d[: n // 50, 1] = 0
    d[n // 50: n // 25, 0] = 0
    o = v[0] + u[:, None] * (v[1] - v[0]) + w[:, None] * (v[2] - v[0]) + np.float32(0.001) * d
    return o.astype(np.float32), d.astype(np.float32)
