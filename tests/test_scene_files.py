"""The .pth scene files (host/PathTracer_FileImporter.{h,cpp}; reference Controleur/PathTracer_FileImporter.cpp:15-146):
byte layout, round trip, error behaviour. CPU only."""
import os
import struct

import numpy as np
import pytest

import opencl_pathtracer_b200 as pt
from opencl_pathtracer_b200 import scene as S


def _texels(sc):
    i = sc.info
    return S._as_array(i.texturesData, i.texturesDataSize, np.dtype("u1, u1, u1, u1")).copy()


def test_file_layout_matches_the_reference_format(scene_factory, tmp_path):
    sc = scene_factory("gallery", 64, 48, True)
    sc.export(tmp_path)
    i = sc.info
    # sizes.pth: direction, right, up, position (reference order :27-30), then five counts = 84 bytes
    raw = (tmp_path / "sizes.pth").read_bytes()
    assert len(raw) == 4 * 16 + 5 * 4
    cam = np.frombuffer(raw[:64], "<f4").reshape(4, 4)
    pos, direction, right, up = np.array(i.camera, np.float32).reshape(4, 4)
    assert np.array_equal(cam, np.stack([direction, right, up, pos]))
    assert struct.unpack("<5I", raw[64:]) == (i.triangulationSize, i.lightsSize, i.materiauxSize, i.texturesSize, i.texturesDataSize)
    # pointers.pth: Triangle[336] Light[64] Material[48] Texture[12] Sky[92], no padding between the arrays
    ptr = (tmp_path / "pointers.pth").read_bytes()
    sky_bytes = 6 * 12 + 5 * 4
    assert len(ptr) == 336 * i.triangulationSize + 64 * i.lightsSize + 48 * i.materiauxSize + 12 * i.texturesSize + sky_bytes
    assert ptr[:336 * i.triangulationSize] == sc.triangles().tobytes()
    off = 336 * i.triangulationSize
    assert ptr[off:off + 64 * i.lightsSize] == sc.lights().tobytes()
    off += 64 * i.lightsSize
    mats = np.frombuffer(ptr[off:off + 48 * i.materiauxSize], S.MATERIAL)
    assert (mats["textureName"] == 0).all(), "the host pointer never reaches the file"
    live = sc.materials()
    for f in ("simpleColor", "opacity", "textureId", "type", "isSimpleColor", "hasAlphaMap"):
        if f in S.MATERIAL.names:
            assert np.array_equal(mats[f], live[f]), f
    assert len((tmp_path / "textureData.pth").read_bytes()) == 4 * i.texturesDataSize


@pytest.mark.parametrize("name,sky", [("cornell", False), ("gallery", True), ("atrium:8", False)])
def test_round_trip(scene_factory, tmp_path, name, sky):
    """export -> import gives the same arrays; building the BVH of the imported scene (triangles already in BVH order)
    reproduces a valid tree over the same triangles."""
    sc = scene_factory(name, 96, 64, sky)
    sc.export(tmp_path)
    back = pt.Scene.from_files(tmp_path, 96, 64, load_sky=True, build_bvh=False)
    a, b = sc.info, back.info
    assert (a.triangulationSize, a.lightsSize, a.materiauxSize, a.texturesSize, a.texturesDataSize) == \
           (b.triangulationSize, b.lightsSize, b.materiauxSize, b.texturesSize, b.texturesDataSize)
    assert np.array_equal(np.array(a.camera), np.array(b.camera))
    assert sc.triangles().tobytes() == back.triangles().tobytes()
    assert sc.lights().tobytes() == back.lights().tobytes()
    assert np.array_equal(_texels(sc), _texels(back))
    back.build_bvh()
    assert back.info.bvhSize > 0 and back.info.bvhMaxDepth < 30
    assert sorted(back.triangles()["id"].tolist()) == sorted(sc.triangles()["id"].tolist())
    # a second export of the re-imported, re-built scene is again loadable (idempotence of the format)
    back.export(tmp_path)
    again = pt.Scene.from_files(tmp_path, 96, 64, load_sky=True, build_bvh=False)
    assert again.triangles().tobytes() == back.triangles().tobytes()


def test_unloaded_sky_is_a_black_texel(scene_factory, tmp_path):
    sc = scene_factory("gallery", 64, 48, True)
    sc.export(tmp_path)
    back = pt.Scene.from_files(tmp_path, 64, 48, load_sky=False, build_bvh=False)
    tx = _texels(back)
    i = back.info
    sky = S._as_array(i.sky, 1, np.dtype([("tex", "<u4", (6, 3)), ("f", "<f4", 5)]))[0]
    for face in sky["tex"]:
        w, h, off = int(face[0]), int(face[1]), int(face[2])
        assert (w, h) == (1, 1) and tuple(tx[off]) == (0, 0, 0, 0)
    assert tuple(sky["f"]) == (1.0, 0.0, 0.0, 1.0, 0.0)  # groundScale, exposantX, exposantY, cos, sin (:130-134)


def test_missing_and_truncated_files(scene_factory, tmp_path):
    with pytest.raises(RuntimeError, match="Fail to read the files to import"):
        pt.Scene.from_files(tmp_path / "nowhere", 32, 32)
    sc = scene_factory("cornell", 32, 32, False)
    sc.export(tmp_path)
    p = tmp_path / "pointers.pth"
    p.write_bytes(p.read_bytes()[:1000])
    with pytest.raises(RuntimeError, match="shorter than announced"):
        pt.Scene.from_files(tmp_path, 32, 32)
    os.remove(tmp_path / "textureData.pth")
    with pytest.raises(RuntimeError, match="textureData.pth"):
        pt.Scene.from_files(tmp_path, 32, 32)
