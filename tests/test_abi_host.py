"""CPU: the C-ABI library loads and exports every symbol include/pathtracer_cuda.h declares; host-side mirror of the
reference interface (importer conventions, controller error behaviour, BMP writer, sharding arithmetic)."""
import ctypes as C
import os
import re
import struct
import subprocess

import numpy as np
import pytest

import opencl_pathtracer_b200 as pt
from opencl_pathtracer_b200 import _native as N
from opencl_pathtracer_b200.sharding import iteration_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "pathtracer_cuda.h")).read()
    declared = re.findall(r"PTC_API\s+[\w\s\*]+?\b(ptc_\w+)\s*\(", hdr)
    assert len(declared) >= 20 and sorted(declared) == sorted(N.PTC_SYMBOLS)
    lib = N.cuda_lib()
    for name in declared:
        assert getattr(lib, name) is not None
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(N.LIB_DIR, "libpathtracer_cuda.so")], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (ptc_\w+)", out))
    assert set(declared) <= exported
    assert lib.ptc_version().decode().startswith("PathTracer_CUDA")


def test_abi_struct_sizes(tmp_path):
    """The ctypes mirrors have the sizes the C header gives its structs (the header compiled as plain C)."""
    assert C.sizeof(N.Config) == 48 and C.sizeof(N.SceneView) == 32 + 7 * 8 + 8 + 64
    assert C.sizeof(N.Counters) == 17 * 8 + 4 * 4
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "pathtracer_cuda.h"\nint main(void){printf("%zu %zu %zu %zu %zu\\n", sizeof(ptc_config), sizeof(ptc_scene_view), '
                   'sizeof(ptc_counters), sizeof(ptc_device_buffers), sizeof(ptc_wide_bvh)); return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert got == [C.sizeof(N.Config), C.sizeof(N.SceneView), C.sizeof(N.Counters), C.sizeof(N.DeviceBuffers), C.sizeof(N.WideBvh)], got


def test_no_cpu_fallback():
    lib = N.cuda_lib()
    if lib.ptc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pt.PathTracerError) as e:
        pt.PathTracerCUDA(16, 16)
    assert e.value.code == N.PTC_ERR_NO_DEVICE and "no CPU path" in str(e.value)
    with pytest.raises(RuntimeError):
        pt.PathTracer_Main("cornell", 16, 16, 1)  # the controller returns false, as the reference's does on any device error


def test_importer_conventions():
    """Maya importer conventions restated by the procedural importer (Maya/PathTracer_MayaImporter.cpp:934-1047)."""
    sc = pt.Scene("gallery", 32, 32, load_sky=True)
    t = sc.triangles()
    for v in ("S1", "S2", "S3"):
        assert (t[v]["w"] == 1).all()  # points
    for v in ("N1", "N2", "N3"):
        assert (t[v]["w"] == 0).all()  # vectors
        ln = np.sqrt(t[v]["x"] ** 2 + t[v]["y"] ** 2 + t[v]["z"] ** 2)
        assert np.allclose(ln, 1, atol=1e-5)
    assert (t["N"]["w"] == 1).all()  # host cross() returns w = 1 and normalize() keeps it

    def lex_le(a, b):
        return (a["x"] < b["x"]) | ((a["x"] == b["x"]) & (a["y"] < b["y"])) | ((a["x"] == b["x"]) & (a["y"] == b["y"]) & (a["z"] <= b["z"]))
    assert lex_le(t["S1"], t["S2"]).all() and lex_le(t["S2"], t["S3"]).all()  # vertices stored in lexicographic order
    m = sc.materials()
    assert set(m["type"].tolist()) >= {0, 1, 2, 3, 4}
    lt = sc.lights()
    assert sorted(lt["type"].tolist()) == [0, 1, 2]
    # deterministic
    sc2 = pt.Scene("gallery", 32, 32, load_sky=True)
    assert np.array_equal(sc2.triangles().view(np.uint8), t.view(np.uint8))
    # black sky when loadSky = false: six 1x1 faces over one zero texel (MayaImporter.cpp:703-719)
    assert pt.Scene("cornell", 8, 8).info.texturesDataSize == 1


def test_unknown_scene_and_scene_sizes():
    with pytest.raises(RuntimeError):
        pt.Scene("no_such_scene", 8, 8)
    assert pt.Scene("cornell", 8, 8).info.triangulationSize == 30
    n = pt.Scene("atrium", 8, 8, build_bvh=False).info.triangulationSize
    assert abs(n - 262144) / 262144 < 0.01  # BASELINE.json configs[2]: ~262k triangles


def test_iteration_range_partition():
    for total in (0, 1, 7, 256, 1024):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                b, n = iteration_range(r, world, 5, total)
                seen += list(range(b, b + n))
            assert seen == list(range(5, 5 + total))
    with pytest.raises(ValueError):
        iteration_range(2, 2, 0, 4)


def test_alone_fails_loudly_without_a_device(tmp_path):
    """Without a GPU the stand-alone front-end must exit with an error (exit code 1), never render on the CPU.
    (With a device present this test has nothing to check: the picture it writes is tested by test_alone_writes_the_reference_bmp.)"""
    exe = os.path.join(N.LIB_DIR, "pathtracer_alone")
    assert os.path.exists(exe)
    if N.cuda_lib().ptc_device_count() == 0:
        r = subprocess.run([exe, "cornell", "16", "16", "1", "2", "0", "0", "1", str(tmp_path) + "/"], capture_output=True, text=True)
        assert r.returncode == 1 and "ERROR" in r.stderr


def bmp_bytes(color, ray_nb):
    """ConvertRGBAToBMPBuffer + SaveBMP of the reference (Alone/PathTracer_bitmap.cpp:237-286, :146-201) restated in numpy: rows top to
    bottom as stored, B G R per pixel, rows padded to 4 bytes, value = (int) min(sum * 255.f / count, 255.f) in float32, a negative
    red sum marks the pixel pure red; 14-byte file header + 40-byte info header (24 bit, 0x0ec4 pixels per metre)."""
    import struct

    import numpy as np
    h, w = ray_nb.shape
    c = color.reshape(h, w, 4).astype(np.float32)
    n = ray_nb.astype(np.float32)[..., None]
    with np.errstate(divide="ignore", invalid="ignore"):
        v = np.minimum(c[..., :3] * np.float32(255.0) / n, np.float32(255.0))
    v = np.nan_to_num(v, nan=0.0).astype(np.int32).astype(np.uint8)
    neg = c[..., 0] < 0
    v[neg] = (255, 0, 0)
    psw = (3 * w + 3) & ~3
    rows = np.zeros((h, psw), np.uint8)
    rows[:, :3 * w] = v[..., ::-1].reshape(h, 3 * w)
    hdr = b"BM" + struct.pack("<IHHI", 54 + psw * h, 0, 0, 0x36) + struct.pack("<IiiHHIIiiII", 40, w, h, 1, 24, 0, 0, 0x0ec4, 0x0ec4, 0, 0)
    return hdr + rows.tobytes()


@pytest.mark.gpu
def test_alone_writes_the_reference_bmp(tmp_path):
    """N1's host half on a device: the headless front-end (counterpart of reference Alone/main.cpp:9-34) renders Cornell 66x48, 2
    iterations, PARITY, and PathTracerDialog::PaintWindow saves the picture. The file must equal, byte for byte, the reference's
    conversion (restated above) of the accumulators the library itself returns for those iterations; against the ORACLE's
    accumulators (the reference kernel on the CPU) at most a last-bit handful of channel values may differ by one."""
    import numpy as np
    import opencl_pathtracer_b200 as pt
    from oracle import ptref
    exe = os.path.join(N.LIB_DIR, "pathtracer_alone")
    w, h, spp, depth = 66, 48, 2, 10  # 66 * 3 = 198 bytes per row: 2 bytes of padding
    out = str(tmp_path) + "/"
    r = subprocess.run([exe, "cornell", str(w), str(h), str(spp), str(depth), "0", "0", "1", out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    files = sorted(f for f in os.listdir(out) if f.endswith(".bmp"))
    assert len(files) == 1, files
    got = open(os.path.join(out, files[0]), "rb").read()
    sc = pt.Scene("cornell", w, h, load_sky=True)  # the front-end loads the sky, as reference Alone/main.cpp:27 does
    with pt.PathTracerCUDA(w, h, ray_max_depth=depth, mode=pt.MODE_PARITY) as ctx:
        ctx.upload(sc)
        ctx.render(0, spp)
        color, nb = ctx.read_image()
    assert got == bmp_bytes(color, nb), "the saved picture is not the reference's conversion of the library's accumulators"
    if ptref.available():
        acc = ptref.render(sc, w, h, depth, 0, 0, spp)
        ref = bmp_bytes(acc["imageColor"], acc["imageRayNb"])
        assert got[:54] == ref[:54]
        a, b = np.frombuffer(got[54:], np.uint8).astype(np.int16), np.frombuffer(ref[54:], np.uint8).astype(np.int16)
        assert np.abs(a - b).max() <= 1 and (a != b).mean() <= 1e-3, "picture differs from the oracle's: %g of the bytes" % (a != b).mean()


def test_area_light_record_packing(tmp_path):
    """PTC_LIGHT_AREA (extension): the C helper of include/pathtracer_structs.h, compiled as plain C, the Python helper and the
    procedural importer's "cornell_area" scene pack the emitter into the 64-byte reference Light record the same way."""
    import numpy as np
    import opencl_pathtracer_b200 as pt
    from opencl_pathtracer_b200 import scene as S
    src = tmp_path / "a.c"
    src.write_text('#include <stdio.h>\n#include "pathtracer_structs.h"\nint main(void){ptc_light l; float c[3]={227,213,548.7f},u[3]={0,130,0},v[3]={105,0,0},'
                   'rgb[3]={1,0.96f,0.9f}; ptc_light_make_area(&l,c,u,v,rgb,100.f); fwrite(&l,sizeof l,1,stdout); return 0;}\n')
    exe = tmp_path / "a.out"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    raw = subprocess.run([str(exe)], capture_output=True, check=True).stdout
    assert len(raw) == 64
    from_c = np.frombuffer(raw, dtype=S.LIGHT)[0]
    from_py = pt.area_light((227, 213, 548.7), (0, 130, 0), (105, 0, 0), (1, 0.96, 0.9), 100.0)
    assert from_c.tobytes() == from_py.tobytes()
    sc = pt.Scene("cornell_area", 32, 24)
    lights = sc.lights()
    assert len(lights) == 1 and lights[0].tobytes() == from_py.tobytes()
    assert int(lights[0]["type"]) == S.LIGHT_AREA == 4
