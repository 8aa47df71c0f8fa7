"""Python face of the PathTracer_CUDA boundary: every call goes through the C ABI (include/pathtracer_cuda.h).

PathTracerCUDA mirrors the life cycle of the reference's device layer
(OpenCL_SetupContext -> OpenCL_InitializeMemory -> OpenCL_RunKernel, Controleur/PathTracer_OpenCL.h:17-19):
construct (= setup), upload(scene) (= initialize memory), render(first, n) + read_image() (= the body of the
run loop), read_stats(), close() (= the release at the end of OpenCL_RunKernel).
"""
import ctypes as C
import os

import numpy as np

from . import _native as N


class PathTracerError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("PathTracer_CUDA error %d: %s" % (code, message))
        self.code = code


class PathTracerCUDA:
    def __init__(self, width, height, ray_max_depth=10, sampler=N.SAMPLER_JITTERED, mode=N.MODE_PARITY, device=0,
                 iterations_per_batch=0, flags=0, stream=None, devices=None):
        """devices: a list of CUDA ordinals = ONE context over several GPUs (ptc_create_multi): iteration ranges (or pixel
        stripes for calls with fewer iterations than devices) per GPU, accumulators summed by NCCL when they are read."""
        self._lib = N.cuda_lib()
        cfg = N.Config()
        cfg.struct_size = C.sizeof(N.Config)
        cfg.device, cfg.image_width, cfg.image_height = device, width, height
        cfg.ray_max_depth, cfg.sampler, cfg.mode = ray_max_depth, sampler, mode
        cfg.iterations_per_batch, cfg.flags = iterations_per_batch, flags
        cfg.stream = stream
        h = C.c_void_p()
        if devices is not None:
            arr = (C.c_int32 * len(devices))(*devices)
            rc = self._lib.ptc_create_multi(C.byref(cfg), arr, len(devices), C.byref(h))
        else:
            rc = self._lib.ptc_create(C.byref(cfg), C.byref(h))
        if rc != N.PTC_OK:
            raise PathTracerError(rc, self._lib.ptc_last_error(None).decode())
        self._h = h
        self.width, self.height, self.ray_max_depth = width, height, ray_max_depth
        self.n_pixels = width * height

    def devices(self):
        a = (C.c_int32 * 64)()
        n = self._lib.ptc_context_devices(self._h, a, 64)
        return list(a[:n])

    def reduce_kind(self):
        """'nccl' / 'p2p' for a multi-device context, 'none' for one device."""
        return self._lib.ptc_multi_reduce_kind(self._h).decode()

    def _ck(self, rc):
        if rc != N.PTC_OK:
            raise PathTracerError(rc, self._lib.ptc_last_error(self._h).decode())

    def upload(self, scene):
        v = scene.view() if hasattr(scene, "view") else scene
        self._ck(self._lib.ptc_upload_scene(self._h, C.byref(v)))

    def set_camera(self, camera16):
        a = np.ascontiguousarray(camera16, dtype=np.float32)
        assert a.size == 16
        self._ck(self._lib.ptc_set_camera(self._h, a.ctypes.data_as(C.POINTER(C.c_float))))

    def reset(self):
        self._ck(self._lib.ptc_reset_accumulation(self._h))

    def set_supersampling(self, x2inv=None, count=None):
        """The reference's -D SUPER_SAMPLING early-out with the caller's chi-square table (None + count: the library's 95 %
        quantiles; count 0: off)."""
        if x2inv is None:
            self._ck(self._lib.ptc_set_supersampling(self._h, None, 0 if count is None else count))
        else:
            t = np.ascontiguousarray(x2inv, dtype=np.float32)
            self._ck(self._lib.ptc_set_supersampling(self._h, t.ctypes.data, len(t)))

    def render(self, first_iteration, n_iterations, sync=False):
        self._ck(self._lib.ptc_render(self._h, first_iteration, n_iterations))
        if sync:
            self.synchronize()

    def synchronize(self):
        self._ck(self._lib.ptc_synchronize(self._h))

    def read_image(self, color=None, ray_nb=None):
        """imageColor (running sum, float32 [H, W, 4]) and imageRayNb (running count, [H, W])."""
        if color is None:
            color = np.empty((self.height, self.width, 4), dtype=np.float32)
        if ray_nb is None:
            ray_nb = np.empty((self.height, self.width), dtype=np.float32)
        self._ck(self._lib.ptc_read_image(self._h, color.ctypes.data, ray_nb.ctypes.data))
        return color, ray_nb

    def read_image_into(self, color_ptr, ray_nb_ptr):
        self._ck(self._lib.ptc_read_image(self._h, color_ptr, ray_nb_ptr))

    def read_variance(self):
        v = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._ck(self._lib.ptc_read_variance(self._h, v.ctypes.data))
        return v

    def read_stats(self):
        d = np.zeros(self.ray_max_depth + 1, dtype=np.uint32)
        b = np.zeros(N.MAX_INTERSECTION_NUMBER, dtype=np.uint32)
        t = np.zeros(N.MAX_INTERSECTION_NUMBER, dtype=np.uint32)
        self._ck(self._lib.ptc_read_stats(self._h, d.ctypes.data, b.ctypes.data, t.ctypes.data))
        return d, b, t

    def read_rgb8(self):
        rgb = np.empty((self.height, self.width, 3), dtype=np.uint8)
        self._ck(self._lib.ptc_read_rgb8(self._h, rgb.ctypes.data))
        return rgb

    def convert_rgb8(self):
        """Enqueue sum / count -> RGB8 into the device display buffer (device_buffers().rgb8); no host copy."""
        self._ck(self._lib.ptc_convert_rgb8(self._h))

    def set_pixel_stripes(self, stripe_rows, rank=0, world=1):
        """Render only the rows y with (y // stripe_rows) % world == rank from now on (0 = whole frame).
        See sharding.stripe_rows / sharding.compose_stripes."""
        self._ck(self._lib.ptc_set_pixel_stripes(self._h, stripe_rows, rank, world))

    def add_image(self, color, ray_nb):
        color = np.ascontiguousarray(color, dtype=np.float32)
        ray_nb = np.ascontiguousarray(ray_nb, dtype=np.float32)
        self._ck(self._lib.ptc_add_image(self._h, color.ctypes.data, ray_nb.ctypes.data))

    def counters(self):
        c = N.Counters()
        self._ck(self._lib.ptc_read_counters(self._h, C.byref(c)))
        return {f[0]: getattr(c, f[0]) for f in N.Counters._fields_}

    def device_buffers(self):
        b = N.DeviceBuffers()
        self._ck(self._lib.ptc_get_device_buffers(self._h, C.byref(b)))
        return b

    def trace_primary(self, iteration=0):
        n = self.n_pixels
        out = {"hit": np.empty(n, np.int32), "st": np.empty((n, 2), np.float32), "dist2": np.empty(n, np.float32),
               "dir": np.empty((n, 4), np.float32), "seed_after": np.empty(n, np.int32), "counts": np.empty((n, 2), np.uint32)}
        self._ck(self._lib.ptc_trace_primary(self._h, iteration, out["hit"].ctypes.data, out["st"].ctypes.data, out["dist2"].ctypes.data,
                                             out["dir"].ctypes.data, out["seed_after"].ctypes.data, out["counts"].ctypes.data))
        return out

    def trace_rays(self, origin, direction):
        """Closest hits of arbitrary rays (float32 [n, 4] each) through the extend stage."""
        o = np.ascontiguousarray(origin, np.float32)
        d = np.ascontiguousarray(direction, np.float32)
        n = len(o)
        assert o.shape == (n, 4) and d.shape == (n, 4)
        out = {"hit": np.empty(n, np.int32), "st": np.empty((n, 2), np.float32), "dist2": np.empty(n, np.float32),
               "dir": np.empty((n, 4), np.float32), "counts": np.empty((n, 2), np.uint32)}
        self._ck(self._lib.ptc_trace_rays(self._h, n, o.ctypes.data, d.ctypes.data, out["hit"].ctypes.data, out["st"].ctypes.data,
                                          out["dist2"].ctypes.data, out["dir"].ctypes.data, out["counts"].ctypes.data))
        return out

    def trace_rays_static(self, origin, direction):
        """trace_rays through ptc_trace_rays_static: no device allocation, scratch in the library's static device array, at most
        _native.STATIC_MAX_RAYS rays per call (no counterpart in the reference; include/pathtracer_cuda.h)."""
        o = np.ascontiguousarray(origin, np.float32)
        d = np.ascontiguousarray(direction, np.float32)
        n = len(o)
        assert o.shape == (n, 4) and d.shape == (n, 4)
        out = {"hit": np.empty(n, np.int32), "st": np.empty((n, 2), np.float32), "dist2": np.empty(n, np.float32),
               "counts": np.empty((n, 2), np.uint32)}
        self._ck(self._lib.ptc_trace_rays_static(self._h, n, o.ctypes.data, d.ctypes.data, out["hit"].ctypes.data, out["st"].ctypes.data,
                                                 out["dist2"].ctypes.data, out["counts"].ctypes.data))
        return out

    def path_probe(self, iteration, bounces):
        """State of every path of `iteration` after `bounces` bounces: dict(org [n,4], dir [n,4], hit [n] int32, st [n,2],
        tpar [n], refl [n] = bounces the path really took)."""
        n = self.n_pixels
        o, d, h = np.empty((n, 4), np.float32), np.empty((n, 4), np.float32), np.empty((n, 4), np.float32)
        m = np.empty(n, np.uint32)
        self._ck(self._lib.ptc_path_probe(self._h, iteration, bounces, o.ctypes.data, d.ctypes.data, h.ctypes.data, m.ctypes.data))
        return {"org": o, "dir": d, "hit": h[:, 0].copy().view(np.int32), "st": h[:, 1:3].copy(), "tpar": h[:, 3].copy(), "refl": (m & 0xff).astype(np.int32)}

    def wide_bvh(self):
        """The wide BVH this context traverses, read back from the device: (nodes as scene.NODE8 records, tri_order[24 n],
        depth, device build time in ms — 0 for the host-built tree)."""
        from .scene import NODE8, _as_array
        w = N.WideBvh()
        self._ck(self._lib.ptc_wide_bvh_read(self._h, C.byref(w)))
        try:
            nodes = _as_array(w.nodes, w.n_nodes, NODE8).copy()
            order = _as_array(w.tri_order, w.n_triangles, np.dtype("<u4")).copy()
            return nodes, order, int(w.depth), w.reserved / 1000.0
        finally:
            self._lib.ptc_wide_bvh_free(C.byref(w))

    def measure_read_bandwidth(self, nbytes, repeats):
        """GB/s of a read-only streaming pass over an nbytes device buffer (L2 hit bandwidth if it fits the L2, else HBM)."""
        g = C.c_double()
        self._ck(self._lib.ptc_measure_read_bandwidth(self._h, nbytes, repeats, C.byref(g)))
        return g.value

    def render_single(self, iteration):
        r = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._ck(self._lib.ptc_render_single(self._h, iteration, r.ctypes.data))
        return r

    def close(self):
        if getattr(self, "_h", None):
            # CHECKED build (PTC_LIB_DIR=.../_lib_checked): a context must end with no bounds-check fault, whatever test drove it —
            # so the whole GPU suite run on that library is a memory-checker pass (PTC_CHECKED_SELFTEST provokes faults on purpose)
            fault = None
            if os.path.basename(N.LIB_DIR.rstrip("/")) == "_lib_checked" and not os.environ.get("PTC_CHECKED_SELFTEST"):
                c = N.Counters()
                if self._lib.ptc_read_counters(self._h, C.byref(c)) == N.PTC_OK and c.faults:
                    fault = "CHECKED build: fault mask %#x, first: code %d value %d bound %d" % (c.faults, c.fault_code, c.fault_value, c.fault_bound)
            self._lib.ptc_destroy(self._h)
            self._h = None
            if fault:
                raise RuntimeError(fault)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def PathTracer_Main(scene, image_width, image_height, numImagesToRender, saveRenderedImages=False, loadSky=False, sampler=0, rayDepth=10,
                    mode=N.MODE_PARITY, readBackEvery=0, device=0, exportFolder=None, seed=20261104, nDevices=0):
    """The reference's whole host pipeline (Controleur/PathTracer.cpp:25-110) through the C++ controller of
    host/PathTracer.cpp: importer -> BVH_Create -> CUDA_SetupContext / CUDA_InitializeMemory / CUDA_RunKernel.
    Returns dict(imageColor, imageRayNb, rayDepths, rayIntersectedBBx, rayIntersectedTri)."""
    L = N.host_lib()
    color = np.zeros((image_height, image_width, 4), np.float32)
    nb = np.zeros((image_height, image_width), np.float32)
    d = np.zeros(rayDepth + 1, np.uint32)
    b = np.zeros(N.MAX_INTERSECTION_NUMBER, np.uint32)
    t = np.zeros(N.MAX_INTERSECTION_NUMBER, np.uint32)
    rc = L.pth_main(scene.encode(), seed, image_width, image_height, numImagesToRender, int(saveRenderedImages), int(loadSky), sampler,
                    rayDepth, mode, readBackEvery or numImagesToRender, device, nDevices, exportFolder.encode() if exportFolder else None,
                    color.ctypes.data, nb.ctypes.data, d.ctypes.data, b.ctypes.data, t.ctypes.data)
    if rc != 0:
        raise RuntimeError("PathTracer_Main failed (see stderr)")
    return {"imageColor": color, "imageRayNb": nb, "rayDepths": d, "rayIntersectedBBx": b, "rayIntersectedTri": t}


def run_default_drop_in(scene, numImagesToRender, rayDepth=10):
    """CUDA_SetupContext / CUDA_InitializeMemory / CUDA_RunKernel with CUDA_Options{} untouched (PARITY, read-back and callback
    after every iteration: the reference's loop, Controleur/PathTracer_OpenCL.cpp:76-107) on an imported scene with its BVH.
    Returns dict(setup_ms, trace_ms, readback_ms, display_ms, paths)."""
    L = N.host_lib()
    t = (C.c_double * 4)()
    d = np.zeros(rayDepth + 1, np.uint32)
    if L.pth_run_default(scene._h, numImagesToRender, rayDepth, t, d.ctypes.data) != 0:
        raise RuntimeError(L.pth_host_last_error().decode())
    return {"setup_ms": t[0], "trace_ms": t[1], "readback_ms": t[2], "display_ms": t[3], "paths": int(d.sum())}
