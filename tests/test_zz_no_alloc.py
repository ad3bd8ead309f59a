"""mdi_interp_scale_map_static: the allocation-free variant of mdi_interp_scale_map (its owner-map
scratch is a static device array inside the library; no counterpart in the reference).
Called through the C ABI (ctypes) and compared bit for bit with the
existing entry point, on two calls in a row whose sizes differ -- the second, smaller view must see
nothing of the first call's scratch contents.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import mod

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _triangulation(H, W, n, seed):
    """Host-side inputs of the kernel: a scipy Delaunay over n pixel positions + the four corners."""
    from scipy.spatial import Delaunay
    g = np.random.default_rng(seed)
    pts = np.stack([g.integers(0, W, n), g.integers(0, H, n)], 1).astype(np.float64)
    pts = np.vstack([pts, [[0, 0], [0, H - 1], [W - 1, 0], [W - 1, H - 1]]])
    dt = Delaunay(pts)
    values = 1.0 + 0.2 * g.standard_normal(pts.shape[0])
    dev = lambda a, t: torch.from_numpy(np.ascontiguousarray(a, dtype=t)).to(DEV)
    base = torch.from_numpy((4.0 + g.random((H, W))).astype(np.float32)).to(DEV)
    ss = torch.tensor([1.7, -0.3], dtype=torch.float32, device=DEV)
    return dict(transform=dev(dt.transform, np.float64), simplices=dev(dt.simplices, np.int32),
                points=dev(dt.points, np.float64), values=dev(values, np.float64),
                n_tri=int(dt.simplices.shape[0]), fill=float(np.median(values)), H=H, W=W, base=base, ss=ss)


def _run(lib, L, t, static: bool, with_base: bool):
    out = torch.full((t["H"], t["W"]), float("nan"), dtype=torch.float32, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    b = t["base"].data_ptr() if with_base else None
    s = t["ss"].data_ptr() if with_base else None
    head = (t["transform"].data_ptr(), t["simplices"].data_ptr(), t["points"].data_ptr(), t["values"].data_ptr(),
            t["n_tri"], t["fill"], t["H"], t["W"], b, s)
    if static:
        L.check(lib.mdi_interp_scale_map_static(*head, out.data_ptr(), st), "mdi_interp_scale_map_static")
    else:
        ws = torch.empty(int(lib.mdi_interp_workspace_bytes(t["H"], t["W"])), dtype=torch.uint8, device=DEV)
        L.check(lib.mdi_interp_scale_map(*head, ws.data_ptr(), out.data_ptr(), st), "mdi_interp_scale_map")
    torch.cuda.synchronize()
    return out


def test_static_scratch_entry_point_equals_workspace_entry_point_twice_in_a_row():
    L = mod("_lib")
    lib = L.load()
    torch.cuda.set_device(0)
    big = _triangulation(480, 640, 3000, 1)      # first call fills 307 200 owner slots
    small = _triangulation(97, 151, 40, 2)       # second call: a smaller view over the same scratch
    refs = [_run(lib, L, big, False, True), _run(lib, L, small, False, False)]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    # two calls in a row through the allocation-free entry point (outputs preallocated by _run's caller side)
    got_big = _run(lib, L, big, True, True)
    got_small = _run(lib, L, small, True, False)
    assert torch.equal(got_big, refs[0]), "first call differs from mdi_interp_scale_map"
    assert torch.equal(got_small, refs[1]), "second call differs (stale scratch of the first call?)"
    assert not torch.isnan(got_big).any() and not torch.isnan(got_small).any()
    # and once more in the other order: the big view after the small one
    assert torch.equal(_run(lib, L, big, True, True), refs[0])
    # the library allocated nothing through torch's allocator (it cannot: it does not know torch); the
    # only growth is the two output tensors of this test that are still alive
    assert torch.cuda.memory_allocated() - before <= (got_big.numel() + got_small.numel()) * 4 + (1 << 20)


def test_static_scratch_entry_point_rejects_views_beyond_its_limit():
    L = mod("_lib")
    lib = L.load()
    H, W = 1200, 1800                               # 2 160 000 > 2 097 152 pixels
    assert H * W > L.INTERP_STATIC_MAX_PX
    out = torch.zeros(4, dtype=torch.float32, device=DEV)   # never written: the size check comes first
    rc = lib.mdi_interp_scale_map_static(None, None, None, None, 0, 1.0, H, W, None, None, out.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.mdi_last_error().decode()
    assert "exceed the static scratch" in msg and "mdi_interp_scale_map" in msg
    # the largest view it takes: 1920 x 1080 (the C3 shape) fits
    assert 1920 * 1080 <= L.INTERP_STATIC_MAX_PX
