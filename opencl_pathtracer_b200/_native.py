"""ctypes view of the two native libraries (built in-tree by build.py).

libpathtracer_cuda.so  — the C ABI of include/pathtracer_cuda.h (the drop-in boundary, reference
                         Controleur/PathTracer_OpenCL.h:17-19)
libpathtracer_scene.so — scene side of the host mirror of the reference interface (importers, BVH builder) + pth_scene_*;
                         independent of the CUDA library (bench.py's reference arm loads only this one)
libpathtracer_host.so  — controller side (PathTracer_Main, CUDA_* shims) + pth_main / pth_run_default

Loading never falls back to anything: a missing library raises, and ptc_create fails without a CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.environ.get("PTC_LIB_DIR") or os.path.join(_HERE, "_lib")  # PTC_LIB_DIR: A/B builds of the kernels (scripts/build_variant.sh)

PTC_OK = 0
PTC_ERR_INVALID, PTC_ERR_CUDA, PTC_ERR_NO_DEVICE, PTC_ERR_STATE, PTC_ERR_LIMIT, PTC_ERR_UNSUPPORTED = 1, 2, 3, 4, 5, 6
SAMPLER_JITTERED, SAMPLER_RANDOM, SAMPLER_UNIFORM = 0, 1, 2
MODE_PARITY, MODE_FAST = 0, 1
STATIC_MAX_RAYS, STATIC_SCRATCH_BYTES = 32768, 4 << 20  # ptc_trace_rays_static (include/pathtracer_cuda.h)
FLAG_STAGE_TIMING, FLAG_NO_VARIANCE, FLAG_NO_GRAPH, FLAG_BVH2, FLAG_GPU_BVH, FLAG_RUSSIAN_ROULETTE, FLAG_TEXTURE_BILINEAR, FLAG_WORK_COUNTERS = 1, 2, 4, 8, 16, 32, 64, 128
MAX_INTERSECTION_NUMBER = 5000


class Float4(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("w", C.c_float)]


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("image_width", C.c_uint32), ("image_height", C.c_uint32),
                ("ray_max_depth", C.c_uint32), ("sampler", C.c_uint32), ("mode", C.c_uint32), ("iterations_per_batch", C.c_uint32),
                ("flags", C.c_uint32), ("reserved", C.c_uint32), ("stream", C.c_void_p)]


class SceneView(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("bvh_size", C.c_uint32), ("triangulation_size", C.c_uint32), ("lights_size", C.c_uint32),
                ("materiaux_size", C.c_uint32), ("textures_size", C.c_uint32), ("textures_data_size", C.c_uint32), ("reserved", C.c_uint32),
                ("bvh", C.c_void_p), ("triangulation", C.c_void_p), ("lights", C.c_void_p), ("materiaux", C.c_void_p),
                ("textures", C.c_void_p), ("textures_data", C.c_void_p), ("sky", C.c_void_p),
                ("_align16", C.c_uint64),  # ptc_float4 is 16-byte aligned in C (include/pathtracer_structs.h); ctypes aligns a float struct to 4
                ("camera_position", Float4), ("camera_direction", Float4), ("camera_right", Float4), ("camera_up", Float4)]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("extension_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("box_tests", C.c_uint64),
                ("triangle_tests", C.c_uint64), ("kernel_launches", C.c_uint64), ("ms_generate", C.c_double), ("ms_extend", C.c_double),
                ("ms_connect", C.c_double), ("ms_shade", C.c_double), ("ms_accumulate", C.c_double), ("launches_extend", C.c_uint64),
                ("launches_connect", C.c_uint64), ("scene_bytes_nodes", C.c_uint64), ("scene_bytes_triangles", C.c_uint64),
                ("ext_box_tests", C.c_uint64), ("ext_triangle_tests", C.c_uint64),
                ("faults", C.c_uint32), ("fault_code", C.c_uint32), ("fault_value", C.c_uint32), ("fault_bound", C.c_uint32)]


class DeviceBuffers(C.Structure):
    _fields_ = [("image_color", C.c_void_p), ("image_ray_nb", C.c_void_p), ("image_v", C.c_void_p), ("ray_depths", C.c_void_p),
                ("ray_bbx", C.c_void_p), ("ray_tri", C.c_void_p), ("stream", C.c_void_p), ("rgb8", C.c_void_p)]


class WideBvh(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("tri_order", C.c_void_p), ("n_nodes", C.c_uint32), ("n_triangles", C.c_uint32),
                ("depth", C.c_uint32), ("reserved", C.c_uint32)]


class SceneInfo(C.Structure):
    _fields_ = [("bvhSize", C.c_uint32), ("bvhMaxDepth", C.c_uint32), ("triangulationSize", C.c_uint32), ("lightsSize", C.c_uint32),
                ("materiauxSize", C.c_uint32), ("texturesSize", C.c_uint32), ("texturesDataSize", C.c_uint32), ("imageWidth", C.c_uint32),
                ("imageHeight", C.c_uint32), ("reserved", C.c_uint32),
                ("bvh", C.c_void_p), ("triangulation", C.c_void_p), ("lights", C.c_void_p), ("materiaux", C.c_void_p),
                ("textures", C.c_void_p), ("texturesData", C.c_void_p), ("sky", C.c_void_p), ("camera", C.c_float * 16)]


# every symbol include/pathtracer_cuda.h declares (tests check that the library exports all of them)
PTC_SYMBOLS = ["ptc_create", "ptc_upload_scene", "ptc_set_camera", "ptc_reset_accumulation", "ptc_set_supersampling", "ptc_render", "ptc_synchronize",
               "ptc_read_image", "ptc_read_variance", "ptc_read_stats", "ptc_read_rgb8", "ptc_convert_rgb8", "ptc_set_pixel_stripes", "ptc_add_image", "ptc_read_counters",
               "ptc_get_device_buffers", "ptc_trace_primary", "ptc_trace_rays", "ptc_render_single", "ptc_path_probe", "ptc_host_alloc", "ptc_host_free",
               "ptc_last_error", "ptc_version", "ptc_device_count", "ptc_destroy", "ptc_wide_bvh_build", "ptc_wide_bvh_read", "ptc_wide_bvh_free", "ptc_measure_read_bandwidth", "ptc_create_multi", "ptc_context_devices", "ptc_multi_reduce_kind",
               "ptc_trace_rays_static"]

_cuda = None
_host = None
_scene = None


def _ensure_built():
    from . import build as _build
    _build.build()


def cuda_lib():
    """libpathtracer_cuda.so with argtypes set. Raises if the library is missing (no fallback)."""
    global _cuda
    if _cuda is not None:
        return _cuda
    path = os.path.join(LIB_DIR, "libpathtracer_cuda.so")
    if not os.path.exists(path):
        _ensure_built()
    L = C.CDLL(path, mode=C.RTLD_GLOBAL)
    vp, u32, i32 = C.c_void_p, C.c_uint32, C.c_int32
    P = C.POINTER
    L.ptc_create.argtypes = [P(Config), P(vp)]
    L.ptc_upload_scene.argtypes = [vp, P(SceneView)]
    L.ptc_set_camera.argtypes = [vp, P(C.c_float)]
    L.ptc_reset_accumulation.argtypes = [vp]
    L.ptc_set_supersampling.argtypes = [vp, vp, u32]
    L.ptc_render.argtypes = [vp, u32, u32]
    L.ptc_synchronize.argtypes = [vp]
    L.ptc_read_image.argtypes = [vp, vp, vp]
    L.ptc_read_variance.argtypes = [vp, vp]
    L.ptc_read_stats.argtypes = [vp, vp, vp, vp]
    L.ptc_read_rgb8.argtypes = [vp, vp]
    L.ptc_convert_rgb8.argtypes = [vp]
    L.ptc_set_pixel_stripes.argtypes = [vp, u32, u32, u32]
    L.ptc_add_image.argtypes = [vp, vp, vp]
    L.ptc_read_counters.argtypes = [vp, P(Counters)]
    L.ptc_get_device_buffers.argtypes = [vp, P(DeviceBuffers)]
    L.ptc_wide_bvh_build.argtypes = [P(SceneView), P(WideBvh)]
    L.ptc_wide_bvh_read.argtypes = [vp, P(WideBvh)]
    L.ptc_wide_bvh_free.argtypes = [P(WideBvh)]
    L.ptc_wide_bvh_free.restype = None
    L.ptc_trace_primary.argtypes = [vp, u32, vp, vp, vp, vp, vp, vp]
    L.ptc_trace_rays.argtypes = [vp, u32, vp, vp, vp, vp, vp, vp, vp]
    L.ptc_trace_rays_static.argtypes = [vp, u32, vp, vp, vp, vp, vp, vp]
    L.ptc_render_single.argtypes = [vp, u32, vp]
    L.ptc_path_probe.argtypes = [vp, u32, u32, vp, vp, vp, vp]
    L.ptc_measure_read_bandwidth.argtypes = [vp, C.c_size_t, u32, P(C.c_double)]
    L.ptc_create_multi.argtypes = [P(Config), P(i32), u32, P(vp)]
    L.ptc_context_devices.argtypes = [vp, P(i32), u32]
    L.ptc_multi_reduce_kind.argtypes = [vp]
    L.ptc_multi_reduce_kind.restype = C.c_char_p
    L.ptc_host_alloc.argtypes = [C.c_size_t]
    L.ptc_host_alloc.restype = vp
    L.ptc_host_free.argtypes = [vp]
    L.ptc_host_free.restype = None
    L.ptc_last_error.argtypes = [vp]
    L.ptc_last_error.restype = C.c_char_p
    L.ptc_version.restype = C.c_char_p
    L.ptc_device_count.restype = i32
    L.ptc_destroy.argtypes = [vp]
    L.ptc_destroy.restype = None
    for name in PTC_SYMBOLS:
        f = getattr(L, name)
        if f.restype is C.c_int and name not in ("ptc_device_count",):
            f.restype = i32
    _cuda = L
    return L


def scene_lib():
    """libpathtracer_scene.so: procedural / file importers + BVH_Create. Does NOT load the CUDA library."""
    global _scene
    if _scene is not None:
        return _scene
    path = os.path.join(LIB_DIR, "libpathtracer_scene.so")
    if not os.path.exists(path):
        _ensure_built()
    L = C.CDLL(path, mode=C.RTLD_GLOBAL)
    vp, u32 = C.c_void_p, C.c_uint32
    L.pth_last_error.restype = C.c_char_p
    L.pth_scene_create.argtypes = [C.c_char_p, u32, u32, u32, C.c_int]
    L.pth_scene_create.restype = vp
    L.pth_scene_build_bvh.argtypes = [vp]
    L.pth_scene_set_bvh.argtypes = [vp, vp, u32, u32]
    L.pth_scene_get_info.argtypes = [vp, C.POINTER(SceneInfo)]
    L.pth_scene_get_info.restype = None
    L.pth_scene_get_view.argtypes = [vp, C.POINTER(SceneView)]
    L.pth_scene_get_view.restype = None
    L.pth_scene_import.argtypes = [C.c_char_p, u32, u32, C.c_int]
    L.pth_scene_import.restype = vp
    L.pth_scene_export.argtypes = [vp, C.c_char_p]
    L.pth_scene_free.argtypes = [vp]
    L.pth_scene_free.restype = None
    _scene = L
    return L


def host_lib():
    """libpathtracer_host.so: the controller (PathTracer_Main) and the CUDA_* shims; loads the CUDA and scene libraries."""
    global _host
    if _host is not None:
        return _host
    cuda_lib()
    scene_lib()
    path = os.path.join(LIB_DIR, "libpathtracer_host.so")
    if not os.path.exists(path):
        _ensure_built()
    L = C.CDLL(path)
    vp, u32 = C.c_void_p, C.c_uint32
    L.pth_host_last_error.restype = C.c_char_p
    L.pth_main.argtypes = [C.c_char_p, u32, u32, u32, u32, C.c_int, C.c_int, C.c_int, u32, C.c_int, u32, C.c_int, C.c_int, C.c_char_p,
                           vp, vp, vp, vp, vp]
    L.pth_run_default.argtypes = [vp, u32, u32, C.POINTER(C.c_double), vp]
    _host = L
    return L
