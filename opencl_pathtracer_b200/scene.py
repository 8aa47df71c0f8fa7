"""Host-side scenes: the procedural importer + BVH builder behind the reference's importer interface.

numpy record dtypes below mirror the reference AoS layouts (Controleur/PathTracer_Structs.h:16-128,
include/pathtracer_structs.h) so tests can look inside nodes and triangles.
"""
import ctypes as C

import numpy as np

from . import _native as N

F4 = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("w", "<f4")])
BBOX = np.dtype({"names": ["pMin", "pMax", "centroid", "isEmpty"], "formats": [F4, F4, F4, "i1"], "offsets": [0, 16, 32, 48], "itemsize": 64})
NODE = np.dtype({"names": ["trianglesAABB", "centroidsAABB", "cutAxis", "triangleStartIndex", "nbTriangles", "son1Id", "son2Id", "comments", "isLeaf"],
                 "formats": [BBOX, BBOX, "<u4", "<u4", "<u4", "<u4", "<u4", "<i4", "i1"],
                 "offsets": [0, 64, 128, 132, 136, 140, 144, 148, 152], "itemsize": 160})
F2 = np.dtype([("x", "<f4"), ("y", "<f4")])
TRIANGLE = np.dtype({"names": ["S1", "S2", "S3", "N1", "N2", "N3", "T1", "T2", "T3", "BT1", "BT2", "BT3", "N",
                               "UVP1", "UVP2", "UVP3", "UVN1", "UVN2", "UVN3", "AABB",
                               "materialWithPositiveNormalIndex", "materialWithNegativeNormalIndex", "id"],
                     "formats": [F4] * 13 + [F2] * 6 + [BBOX, "<u4", "<u4", "<u4"],
                     "offsets": [16 * i for i in range(13)] + [208 + 8 * i for i in range(6)] + [256, 320, 324, 328], "itemsize": 336})
LIGHT = np.dtype({"names": ["position", "direction", "color", "power", "cosOfInnerFallOffAngle", "cosOfOuterFallOffAngle", "type"],
                  "formats": [F4, F4, F4, "<f4", "<f4", "<f4", "<i4"], "offsets": [0, 16, 32, 48, 52, 56, 60], "itemsize": 64})
MATERIAL = np.dtype({"names": ["simpleColor", "textureName", "opacity", "textureId", "type", "isSimpleColor", "hasAlphaMap"],
                     "formats": [F4, "<u8", "<f4", "<i4", "<i4", "u1", "u1"], "offsets": [0, 16, 24, 28, 32, 36, 37], "itemsize": 48})


LIGHT_AREA = 4  # include/pathtracer_structs.h PTC_LIGHT_AREA (extension; the reference has no area lights)


def area_light(corner, eu, ev, rgb=(1.0, 1.0, 1.0), power=1.0):
    """One LIGHT record holding a parallelogram emitter, packed as ptc_light_make_area (include/pathtracer_structs.h) packs
    it: corner + a eu + b ev, emitting power * rgb as radiance towards cross(eu, ev). FAST mode only."""
    l = np.zeros(1, dtype=LIGHT)
    l["position"] = (corner[0], corner[1], corner[2], ev[0])
    l["direction"] = (eu[0], eu[1], eu[2], ev[1])
    l["color"] = (rgb[0], rgb[1], rgb[2], 1.0)
    l["power"] = power
    l["cosOfInnerFallOffAngle"] = ev[2]
    l["type"] = LIGHT_AREA
    return l[0]


# csrc/pt_wide.cuh DevNode8
NODE8 = np.dtype([("p", "<f4", 3), ("e", "u1", 3), ("imask", "u1"), ("childBase", "<u4"), ("valid", "<u4"), ("spare", "<u4", 2),
                  ("qlo", "u1", (3, 8)), ("qhi", "u1", (3, 8))])
assert NODE8.itemsize == 80


def _as_array(ptr, count, dtype):
    if not ptr or count == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


class Scene:
    """A GlobalVars-like scene produced by PathTracerProceduralImporter (host/PathTracer_ProceduralImporter.h)."""

    def __init__(self, name, width, height, seed=20261104, load_sky=False, build_bvh=True):
        self._lib = N.scene_lib()
        self._h = self._lib.pth_scene_create(name.encode(), seed, width, height, 1 if load_sky else 0)
        if not self._h:
            raise RuntimeError(self._lib.pth_last_error().decode())
        self.name, self.width, self.height = name, width, height
        if build_bvh:
            self.build_bvh()

    @classmethod
    def from_files(cls, folder, width, height, load_sky=True, build_bvh=True):
        """A scene from the reference's file triple sizes.pth / pointers.pth / textureData.pth (host/PathTracer_FileImporter.h,
        counterpart of reference Controleur/PathTracer_FileImporter.cpp:67-146)."""
        self = cls.__new__(cls)
        self._lib = N.scene_lib()
        self._h = self._lib.pth_scene_import(str(folder).encode(), width, height, 1 if load_sky else 0)
        if not self._h:
            raise RuntimeError(self._lib.pth_last_error().decode())
        self.name, self.width, self.height = "file:%s" % folder, width, height
        if build_bvh:
            self.build_bvh()
        return self

    def export(self, folder):
        """Write the scene as it is now (triangles in BVH order once the BVH is built) — reference :15-63."""
        if self._lib.pth_scene_export(self._h, str(folder).encode()) != 0:
            raise RuntimeError(self._lib.pth_last_error().decode())

    def build_bvh(self):
        if self._lib.pth_scene_build_bvh(self._h) != 0:
            raise RuntimeError(self._lib.pth_last_error().decode())

    def set_bvh(self, nodes, bvh_size, bvh_max_depth):
        nodes = np.ascontiguousarray(nodes)
        self._lib.pth_scene_set_bvh(self._h, nodes.ctypes.data, bvh_size, bvh_max_depth)

    @property
    def info(self):
        i = N.SceneInfo()
        self._lib.pth_scene_get_info(self._h, C.byref(i))
        return i

    def view(self):
        v = N.SceneView()
        self._lib.pth_scene_get_view(self._h, C.byref(v))
        return v

    # numpy views into the host arrays (valid while the Scene lives and the BVH is not rebuilt)
    def nodes(self):
        i = self.info
        return _as_array(i.bvh, i.bvhSize, NODE)

    def triangles(self):
        i = self.info
        return _as_array(i.triangulation, i.triangulationSize, TRIANGLE)

    def lights(self):
        i = self.info
        return _as_array(i.lights, i.lightsSize, LIGHT)

    def materials(self):
        i = self.info
        return _as_array(i.materiaux, i.materiauxSize, MATERIAL)

    def wide_bvh(self):
        """The 8-wide compressed BVH FAST mode traverses (csrc/pt_wide_build.cu): (nodes[n] as NODE8 records,
        tri_order[24 n]: record 24 i + 3 slot + k -> BVH-order triangle index or 0xffffffff, depth)."""
        lib = N.cuda_lib()
        w = N.WideBvh()
        v = self.view()
        if lib.ptc_wide_bvh_build(C.byref(v), C.byref(w)) != N.PTC_OK:
            raise RuntimeError(lib.ptc_last_error(None).decode())
        try:
            nodes = _as_array(w.nodes, w.n_nodes, NODE8).copy()
            order = _as_array(w.tri_order, w.n_triangles, np.dtype("<u4")).copy()
            return nodes, order, int(w.depth)
        finally:
            lib.ptc_wide_bvh_free(C.byref(w))

    def camera(self):
        return np.array(self.info.camera, dtype=np.float32)

    def close(self):
        if self._h:
            self._lib.pth_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
