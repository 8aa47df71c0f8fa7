"""ptc_trace_rays_static: the closest-hit probe without device allocation (include/pathtracer_cuda.h; no counterpart in the
reference). Collected LAST on purpose.

The entry point runs the kernels of ptc_trace_rays with every scratch buffer inside the library's file-scope static device array, so
its outputs are compared with ptc_trace_rays bit for bit — the tolerance the traversal parity tests use for hit, (s, t), distance
and the work counters — on two calls in a row (the array keeps the previous call's contents), in both arithmetic modes, and with
the previous call larger than the next one (stale records beyond n must not leak into the result)."""
import numpy as np
import pytest

import opencl_pathtracer_b200 as pt
from opencl_pathtracer_b200 import _native as N
from golden_cases import surface_rays


def _same(a, b, what):
    for f in ("hit", "counts"):
        assert np.array_equal(a[f], b[f]), (what, f, int((a[f] != b[f]).sum()))
    for f in ("st", "dist2"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), (what, f)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [pt.MODE_PARITY, pt.MODE_FAST])
def test_static_scratch_matches_trace_rays_twice(scene_factory, mode):
    w, h = 256, 256
    sc = scene_factory("atrium:4", w, h, False)
    o1, d1 = surface_rays(sc, N.STATIC_MAX_RAYS, 31)  # the maximum batch
    o2, d2 = surface_rays(sc, 5000, 32)               # then a smaller one over the leftovers of the first
    with pt.PathTracerCUDA(w, h, mode=mode) as r:
        r.upload(sc)
        ref1, ref2 = r.trace_rays(o1, d1), r.trace_rays(o2, d2)
        got1 = r.trace_rays_static(o1, d1)
        got2 = r.trace_rays_static(o2, d2)
        got1_again = r.trace_rays_static(o1, d1)
    assert (ref1["hit"] >= 0).mean() > 0.5
    _same(got1, ref1, "first call")
    _same(got2, ref2, "second call")
    _same(got1_again, ref1, "third call")


@pytest.mark.gpu
def test_static_scratch_shared_by_contexts(scene_factory):
    """Two contexts with different scenes share the one array; calls one after the other do not disturb each other."""
    w, h = 256, 256
    sa, sb = scene_factory("cornell", w, h, False), scene_factory("gallery", w, h, True)
    oa, da = surface_rays(sa, 20000, 3)
    ob, db = surface_rays(sb, 12000, 4)
    with pt.PathTracerCUDA(w, h, mode=pt.MODE_PARITY) as a, pt.PathTracerCUDA(w, h, mode=pt.MODE_FAST) as b:
        a.upload(sa)
        b.upload(sb)
        ra, rb = a.trace_rays(oa, da), b.trace_rays(ob, db)
        for _ in range(2):
            _same(a.trace_rays_static(oa, da), ra, "context a")
            _same(b.trace_rays_static(ob, db), rb, "context b")


@pytest.mark.gpu
def test_static_scratch_limits_and_errors(scene_factory):
    w, h = 64, 64
    sc = scene_factory("cornell", w, h, False)
    lib = N.cuda_lib()
    n = N.STATIC_MAX_RAYS + 1
    o = np.zeros((n, 4), np.float32)
    d = np.ones((n, 4), np.float32)
    hit = np.full(n, 7, np.int32)
    with pt.PathTracerCUDA(w, h) as r:
        rc = lib.ptc_trace_rays_static(r._h, 16, o.ctypes.data, d.ctypes.data, hit.ctypes.data, None, None, None)
        assert rc == N.PTC_ERR_STATE  # no scene yet
        r.upload(sc)
        rc = lib.ptc_trace_rays_static(r._h, n, o.ctypes.data, d.ctypes.data, hit.ctypes.data, None, None, None)
        assert rc == N.PTC_ERR_LIMIT
        msg = lib.ptc_last_error(r._h).decode()
        assert "PTC_STATIC_MAX_RAYS" in msg and str(N.STATIC_MAX_RAYS) in msg, msg
        assert (hit == 7).all(), "a refused call must not write"
        assert lib.ptc_trace_rays_static(r._h, 0, o.ctypes.data, d.ctypes.data, None, None, None, None) == N.PTC_OK
        assert lib.ptc_trace_rays_static(r._h, 16, None, d.ctypes.data, None, None, None, None) == N.PTC_ERR_INVALID
        with pytest.raises(pt.PathTracerError):
            r.trace_rays_static(o, d)
        # any output may be NULL; the others are unaffected
        oo, dd = surface_rays(sc, 1000, 9)
        full = r.trace_rays_static(oo, dd)
        only_hit = np.empty(1000, np.int32)
        assert lib.ptc_trace_rays_static(r._h, 1000, oo.ctypes.data, dd.ctypes.data, only_hit.ctypes.data, None, None, None) == N.PTC_OK
        assert np.array_equal(only_hit, full["hit"])


@pytest.mark.gpu
def test_static_scratch_device_pointers(scene_factory):
    """Inputs and outputs may live in device memory (the caller's own buffers; torch is only the allocator of this test)."""
    torch = pytest.importorskip("torch")
    w, h = 128, 128
    sc = scene_factory("atrium:4", w, h, False)
    o, d = surface_rays(sc, 8192, 17)
    with pt.PathTracerCUDA(w, h, mode=pt.MODE_FAST) as r:
        r.upload(sc)
        ref = r.trace_rays(o, d)
        dev = torch.device("cuda", r.devices()[0])
        to, td = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
        thit = torch.empty(len(o), dtype=torch.int32, device=dev)
        tst = torch.empty((len(o), 2), dtype=torch.float32, device=dev)
        td2 = torch.empty(len(o), dtype=torch.float32, device=dev)
        torch.cuda.synchronize(dev)
        for _ in range(2):
            thit.fill_(-5)
            torch.cuda.synchronize(dev)
            rc = N.cuda_lib().ptc_trace_rays_static(r._h, len(o), to.data_ptr(), td.data_ptr(), thit.data_ptr(), tst.data_ptr(), td2.data_ptr(), None)
            assert rc == N.PTC_OK, N.cuda_lib().ptc_last_error(r._h).decode()
            assert np.array_equal(thit.cpu().numpy(), ref["hit"])
            assert np.array_equal(tst.cpu().numpy().view(np.uint32), ref["st"].view(np.uint32))
            assert np.array_equal(td2.cpu().numpy().view(np.uint32), ref["dist2"].view(np.uint32))


def test_static_scratch_symbol_and_constants():
    """CPU: the library exports the entry point and the Python constants are the header's."""
    import os
    import re
    lib = N.cuda_lib()
    assert hasattr(lib, "ptc_trace_rays_static")
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "pathtracer_cuda.h")).read()
    assert int(re.search(r"#define PTC_STATIC_MAX_RAYS (\d+)u", hdr).group(1)) == N.STATIC_MAX_RAYS
    m = re.search(r"#define PTC_STATIC_SCRATCH_BYTES \((\d+)u << (\d+)\)", hdr)
    assert int(m.group(1)) << int(m.group(2)) == N.STATIC_SCRATCH_BYTES
