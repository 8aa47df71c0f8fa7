import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        from opencl_pathtracer_b200 import _native
        return _native.cuda_lib().ptc_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def ptref():
    """The parity oracle: the reference kernel + BVH builder compiled for the CPU (oracle/_ref)."""
    from oracle import ptref as m
    if not m.available():
        if os.path.isdir("/root/reference/Kernel"):
            from oracle import build_ref
            build_ref.build("/root/reference", verbose=False)
        else:
            pytest.skip("oracle/_ref/libptref.so not built and /root/reference absent")
    return m


_scene_cache = {}


@pytest.fixture(scope="session")
def scene_factory():
    import opencl_pathtracer_b200 as pt

    def make(name, w, h, load_sky=False):
        key = (name, w, h, load_sky)
        if key not in _scene_cache:
            _scene_cache[key] = pt.Scene(name, w, h, load_sky=load_sky)
        return _scene_cache[key]

    return make
