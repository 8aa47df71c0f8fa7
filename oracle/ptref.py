"""TEST INFRASTRUCTURE ONLY — ctypes wrapper around oracle/_ref/libptref.so, the reference's own kernel
(Kernel/PathTracer_FullKernel.cl) and BVH builder (Controleur/PathTracer_BVH.cpp) compiled for the CPU by
oracle/build_ref.py. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product (opencl_pathtracer_b200/) never does.

Parity status: PINNED — the oracle IS the reference source, run here.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libptref.so")


class RefScene(C.Structure):
    _fields_ = [("bvh", C.c_void_p), ("triangulation", C.c_void_p), ("lights", C.c_void_p), ("materiaux", C.c_void_p),
                ("textures", C.c_void_p), ("texturesData", C.c_void_p), ("sky", C.c_void_p), ("camera", C.c_float * 16),
                ("lightsSize", C.c_uint32)]


_lib = None


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            from . import build_ref  # noqa: only possible where /root/reference exists
            build_ref.build("/root/reference", verbose=False)
        L = C.CDLL(SO)
        vp, u32 = C.c_void_p, C.c_uint32
        L.ptref_sizes.argtypes = [vp]
        L.ptref_bvh_create.argtypes = [vp, u32, vp, vp, vp]
        L.ptref_render.argtypes = [C.POINTER(RefScene), u32, u32, u32, C.c_int, u32, u32, vp, vp, vp, vp, vp, vp, C.c_int]
        L.ptref_render_rows.argtypes = [C.POINTER(RefScene), u32, u32, u32, C.c_int, u32, u32, u32, u32, vp, vp, vp, vp, vp, vp, C.c_int]
        L.ptref_primary.argtypes = [C.POINTER(RefScene), u32, u32, C.c_int, u32, vp, vp, vp, vp, vp, vp]
        L.ptref_trace_rays.argtypes = [C.POINTER(RefScene), u32, vp, vp, vp, vp, vp, vp, vp]
        L.ptref_rng.argtypes = [u32, u32, u32, u32, u32, C.c_int, vp, vp, vp]
        L.ptref_cosine_sample.argtypes = [C.c_int32, vp, vp, vp]
        L.ptref_x2inv.argtypes = [vp, C.c_int]
        L.ptref_path_debug.argtypes = [C.POINTER(RefScene), u32, u32, u32, u32, u32, u32, vp, vp, vp, vp]
        _lib = L
    return _lib


def sizes():
    a = np.zeros(16, np.uint32)
    n = lib().ptref_sizes(a.ctypes.data)
    return a[:n]


def bvh_create(triangles):
    """Reference BVH_Create on a COPY of `triangles` (structured array, 336-byte records).
    Returns (nodes[160-byte records as uint8 rows], permuted triangles copy, bvhSize, bvhMaxDepth)."""
    tris = np.array(triangles, copy=True)
    n = len(tris)
    nodes = np.zeros((2 * n - 1, 160), np.uint8)
    size, depth = C.c_uint32(), C.c_uint32()
    rc = lib().ptref_bvh_create(tris.ctypes.data, n, nodes.ctypes.data, C.byref(size), C.byref(depth))
    assert rc == 0
    return nodes[:size.value], tris, size.value, depth.value


def ref_scene(scene):
    """RefScene over the host arrays of an opencl_pathtracer_b200.Scene (the records are byte-identical)."""
    i = scene.info
    r = RefScene()
    r.bvh, r.triangulation, r.lights, r.materiaux = i.bvh, i.triangulation, i.lights, i.materiaux
    r.textures, r.texturesData, r.sky = i.textures, i.texturesData, i.sky
    for k in range(16):
        r.camera[k] = i.camera[k]
    r.lightsSize = i.lightsSize
    return r


def render(scene, W, H, max_depth, sampler, it_begin, it_end, nthreads=0, acc=None, rows=None):
    """Reference launch loop. Returns dict of accumulators (new zeroed ones unless `acc` is passed back in)."""
    if acc is None:
        acc = {"imageColor": np.zeros((H, W, 4), np.float32), "imageRayNb": np.zeros((H, W), np.float32),
               "imageV": np.zeros((H, W, 4), np.float32), "rayDepths": np.zeros(max_depth + 1, np.uint32),
               "rayIntersectedBBx": np.zeros(5000, np.uint32), "rayIntersectedTri": np.zeros(5000, np.uint32)}
    rs = ref_scene(scene)
    r0, r1 = rows if rows is not None else (0, H)
    rc = lib().ptref_render_rows(C.byref(rs), W, H, max_depth, sampler, it_begin, it_end, r0, r1, acc["imageColor"].ctypes.data,
                                 acc["imageRayNb"].ctypes.data, acc["imageV"].ctypes.data, acc["rayDepths"].ctypes.data,
                                 acc["rayIntersectedBBx"].ctypes.data, acc["rayIntersectedTri"].ctypes.data, nthreads)
    assert rc == 0
    return acc


SAMPLER_JITTERED_SUPERSAMPLING = 3  # ref_glue.cpp: the JITTERED kernel built with -D SUPER_SAMPLING


def x2inv():
    """The chi-square table the SUPER_SAMPLING build of the reference kernel indexes (oracle/shim/X2inv.cl)."""
    out = np.zeros(4096, np.float32)
    n = lib().ptref_x2inv(out.ctypes.data, len(out))
    return out[:n].copy()


def primary(scene, W, H, sampler=0, iteration=0):
    n = W * H
    out = {"hit": np.empty(n, np.int32), "st": np.empty((n, 2), np.float32), "dist2": np.empty(n, np.float32),
           "dir": np.empty((n, 4), np.float32), "seed_after": np.empty(n, np.int32), "counts": np.empty((n, 2), np.uint32)}
    rs = ref_scene(scene)
    rc = lib().ptref_primary(C.byref(rs), W, H, sampler, iteration, out["hit"].ctypes.data, out["st"].ctypes.data,
                             out["dist2"].ctypes.data, out["dir"].ctypes.data, out["seed_after"].ctypes.data, out["counts"].ctypes.data)
    assert rc == 0
    return out


def trace_rays(scene, origin, direction):
    """Closest hits of arbitrary rays (origin, direction: float32 [n, 4]; the direction is normalised by Ray3D_Create)."""
    o = np.ascontiguousarray(origin, np.float32)
    d = np.ascontiguousarray(direction, np.float32)
    n = len(o)
    out = {"hit": np.empty(n, np.int32), "st": np.empty((n, 2), np.float32), "dist2": np.empty(n, np.float32),
           "dir": np.empty((n, 4), np.float32), "counts": np.empty((n, 2), np.uint32)}
    rs = ref_scene(scene)
    rc = lib().ptref_trace_rays(C.byref(rs), n, o.ctypes.data, d.ctypes.data, out["hit"].ctypes.data, out["st"].ctypes.data,
                                out["dist2"].ctypes.data, out["dir"].ctypes.data, out["counts"].ctypes.data)
    assert rc == 0
    return out


def rng(x, y, it, W, H, n):
    seed0 = C.c_int32()
    seeds = np.empty(n, np.int32)
    us = np.empty(n, np.float32)
    lib().ptref_rng(x, y, it, W, H, n, C.byref(seed0), seeds.ctypes.data, us.ctypes.data)
    return seed0.value, seeds, us


def cosine_sample(seed, normal4):
    nrm = np.ascontiguousarray(normal4, np.float32)
    out = np.empty(4, np.float32)
    after = C.c_int32()
    lib().ptref_cosine_sample(seed, nrm.ctypes.data, out.ctypes.data, C.byref(after))
    return out, after.value


def path_debug(scene, W, H, max_depth, x, y, iteration):
    """Kernel_Main's path loop for ONE work-item, bounce by bounce, made of the reference's own functions (ref_glue.cpp
    ptref_path_debug). Returns dict(n = bounces logged, org [n,4], dir [n,4], hit [n] (triangle index or -1), mat [n]
    (MaterialType of the hit or -1), work [n,4], st [n,2])."""
    rays = np.zeros((max_depth + 1, 8), np.float32)
    ids = np.zeros((max_depth + 1, 2), np.int32)
    work = np.zeros((max_depth + 1, 4), np.uint32)
    st = np.zeros((max_depth + 1, 2), np.float32)
    rs = ref_scene(scene)
    n = lib().ptref_path_debug(C.byref(rs), W, H, max_depth, x, y, iteration, rays.ctypes.data, ids.ctypes.data, work.ctypes.data, st.ctypes.data)
    return {"n": n, "org": rays[:n, :4].copy(), "dir": rays[:n, 4:].copy(), "hit": ids[:n, 0].copy(), "mat": ids[:n, 1].copy(),
            "work": work[:n].copy(), "st": st[:n].copy()}


def max_threads():
    return lib().ptref_max_threads()
