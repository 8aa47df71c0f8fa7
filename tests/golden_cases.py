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


def rgb8_of(color, nb):
    """numpy restatement of the display conversion (reference Alone/PathTracer_bitmap.cpp:259-279): sum * 255 / count clamped
    to 255 in float32, truncated; a negative red sum marks the pixel pure red; pixels without samples are written as 0 here
    (what ptc_convert_rgb8 does for rows outside a pixel shard)."""
    c = color[..., :3].astype(np.float32)
    n = nb.astype(np.float32)[..., None]
    with np.errstate(divide="ignore", invalid="ignore"):
        v = np.minimum(c * np.float32(255.0) / n, np.float32(255.0))
    out = np.where(n > 0, v, 0).astype(np.int32).astype(np.uint8)
    red = (color[..., 0] < 0) & (nb > 0)
    out[red] = (255, 0, 0)
    return out
