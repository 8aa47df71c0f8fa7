"""Memory-checker evidence without compute-sanitizer (closed on the GPU pool): the kernels of the CHECKED build
(opencl_pathtracer_b200/_lib_checked, device code compiled -DPTC_CHECKED) compare every index they form with its bound and report
violations through ptc_counters.faults. scripts/sanitize_case.py drives three scenes, both modes and walks, graphs, pixel shards,
probes, the device BVH builder and (on two GPUs) a multi-device context through them; it must finish with no fault."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECKED = os.path.join(ROOT, "opencl_pathtracer_b200", "_lib_checked")


@pytest.mark.gpu
def test_kernels_under_bounds_checks():
    if not os.path.exists(os.path.join(CHECKED, "libpathtracer_cuda.so")):
        pytest.skip("_lib_checked not built (python -m opencl_pathtracer_b200.build)")
    env = dict(os.environ, PTC_LIB_DIR=CHECKED)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize_case.py")], env=env, capture_output=True, text=True, timeout=900)
    print(p.stdout[-3000:])
    assert "checked" in p.stdout.splitlines()[0], "the case did not run on the checked build: " + p.stdout[:200]
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "faults: 0x0" in p.stdout


@pytest.mark.gpu
def test_bounds_checks_do_fire():
    """The checks are live: with a deliberately wrong bound (PTC_CHECKED_SELFTEST: path-slot bound 1) the same kernels report
    queue-entry faults — and nothing else."""
    if not os.path.exists(os.path.join(CHECKED, "libpathtracer_cuda.so")):
        pytest.skip("_lib_checked not built")
    code = ("import sys; sys.path.insert(0, %r)\nimport opencl_pathtracer_b200 as pt\nsc = pt.Scene('cornell', 64, 48)\n"
            "r = pt.PathTracerCUDA(64, 48, mode=pt.MODE_FAST)\nr.upload(sc)\nr.render(0, 2)\nc = r.counters()\n"
            "print('FAULTS', c['faults'], c['fault_code'], c['fault_bound'])\n" % ROOT)
    env = dict(os.environ, PTC_LIB_DIR=CHECKED, PTC_CHECKED_SELFTEST="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300).stdout
    line = [l for l in out.splitlines() if l.startswith("FAULTS")][0].split()
    mask = int(line[1])
    assert mask != 0 and mask & ~((1 << 0) | (1 << 5)) == 0, line  # kFaultQueueItem / kFaultQueueAppend only
    assert int(line[3]) == 1


def test_checked_build_exists_and_says_so():
    """CPU: the variant is built by build.py and identifies itself (no device needed for ptc_version)."""
    import ctypes as C
    so = os.path.join(CHECKED, "libpathtracer_cuda.so")
    if not os.path.exists(so):
        pytest.skip("_lib_checked not built")
    code = "import ctypes as C; L = C.CDLL(%r); L.ptc_version.restype = C.c_char_p; print(L.ptc_version().decode())" % so
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True).stdout
    assert out.strip().endswith("checked"), out
