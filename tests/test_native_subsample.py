"""a13/a14: minimal Gaussian extents and the k-d merge subsampling.

CPU tests check the C restatement (oracle/native_oracle.c) against hand-derived answers of the
reference algorithm (impl.cpp:70-126, :313-426); tests/test_native_ref_pin.py pins it bit for bit
against the reference's own module compiled into oracle/_ref.  GPU tests compare libmdi_b200.so with
the restatement bit for bit (counts, order, fp32 values)."""
import numpy as np
import pytest
import torch

from helpers import mod

import native_oracle as NO


def _cams(n_cams, seed, W=160, H=120):
    rng = np.random.default_rng(seed)
    K = np.zeros((n_cams, 3, 3), np.float32)
    K[:, 0, 0] = rng.uniform(90, 140, n_cams); K[:, 1, 1] = rng.uniform(90, 140, n_cams)
    K[:, 0, 2] = W / 2; K[:, 1, 2] = H / 2; K[:, 2, 2] = 1
    P = np.zeros((n_cams, 3, 4), np.float32)
    for c in range(n_cams):
        a = rng.uniform(-0.6, 0.6)
        R = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]], np.float32)
        t = rng.uniform(-0.5, 0.5, 3).astype(np.float32) + np.array([0, 0, 4], np.float32)
        P[c] = K[c] @ np.hstack([R, t[:, None]])
    sizes = np.tile(np.array([[W, H]], np.int32), (n_cams, 1))
    return K, P.astype(np.float32), sizes


def _cloud(kind, n, seed):
    rng = np.random.default_rng(seed)
    if kind == "gauss":
        pts = rng.normal(size=(n, 3)).astype(np.float32)
    elif kind == "surface":  # dense sheet: many merges with several points per leaf
        uv = rng.random((n, 2)).astype(np.float32) * 2 - 1
        pts = np.stack([uv[:, 0], uv[:, 1], 0.2 * np.sin(3 * uv[:, 0]) + 0.001 * rng.normal(size=n)], 1).astype(np.float32)
    elif kind == "clusters":  # duplicates, coplanar groups (inf / NaN aspect ratios), exact zeros
        base = rng.integers(-3, 4, size=(n // 8 + 1, 3)).astype(np.float32) * 0.25
        pts = base[rng.integers(0, base.shape[0], n)]
        jitter = rng.normal(size=(n, 3)).astype(np.float32) * 1e-3
        jitter[rng.random(n) < 0.5] = 0
        jitter[:, 2][rng.random(n) < 0.5] = 0
        pts = (pts + jitter).astype(np.float32)
    else:
        raise ValueError(kind)
    rgb = rng.random((n, 3)).astype(np.float32)
    return pts, rgb


# ---- CPU: the restatement against hand-derived answers ---------------------------------------
def test_oracle_extents_hand_case():
    K = np.array([[[100, 0, 50], [0, 80, 40], [0, 0, 1]]], np.float32)
    P = (K[0] @ np.hstack([np.eye(3, dtype=np.float32), np.zeros((3, 1), np.float32)]))[None]
    sizes = np.array([[100, 80]], np.int32)
    pts = np.array([[0, 0, 2], [0, 0, -1], [10, 0, 2], [0.1, -0.2, 4]], np.float32)
    ext = NO.min_extents(pts, K, P, sizes)
    # visible: depth 2 -> 2*2/min(100,80) ; behind camera -> -1 ; u = 550 outside -> -1 ; depth 4
    assert ext.tolist() == [np.float32(2 * (np.float32(2) / np.float32(80))), -1.0, -1.0,
                            np.float32(2 * (np.float32(4) / np.float32(80)))]


def test_oracle_merge_hand_cases():
    rgb = lambda n: np.arange(3 * n, dtype=np.float32).reshape(n, 3) / 10
    # 1 point: emitted as is
    p, c = NO.merge_subsample([[1, 2, 3]], rgb(1), [0.5])
    assert p.tolist() == [[1, 2, 3]]
    # 2 points on a diagonal (tight box is a cube, aspect 1) closer than their mean extent: merged to the mean
    pts = np.array([[0, 0, 0], [0.1, 0.1, 0.1]], np.float32)
    p, c = NO.merge_subsample(pts, rgb(2), [0.2, 0.2])
    assert p.shape == (1, 3) and np.array_equal(p[0], (pts[0] + pts[1]) / np.float32(2))
    assert np.array_equal(c[0], (rgb(2)[0] + rgb(2)[1]) / np.float32(2))
    # same two points, extent too small: both emitted in index order
    p, c = NO.merge_subsample(pts, rgb(2), [0.01, 0.01])
    assert np.array_equal(p, pts)
    # axis-aligned pair: tight aspect is x/0 = inf, root cube aspect is 1 -> min is 1 -> merge
    pts = np.array([[0, 0, 0], [0.1, 0, 0]], np.float32)
    p, _ = NO.merge_subsample(pts, rgb(2), [0.2, 0.2])
    assert p.shape == (1, 3)
    # 4 points, nothing merges: first split is on Y (prev axis X), the right half (y >= mid) is popped
    # first; a frame of two points is emitted in index order without another split.
    pts = np.array([[0, 0, 0], [0, 1, 0], [0, 1, 1], [1, 0, 1]], np.float32)
    p, _ = NO.merge_subsample(pts, rgb(4), [1e-3] * 4)
    assert p.tolist() == [[0, 1, 0], [0, 1, 1], [0, 0, 0], [1, 0, 1]]
    # 6 points: Y split, then the 3-point right frame splits on Z and emits z >= mid first
    pts = np.array([[0, 0, 0], [0, 1, 0], [0, 1, 1], [1, 0, 1], [1, 1, 0.25], [1, 0, 0]], np.float32)
    p, _ = NO.merge_subsample(pts, rgb(6), [1e-3] * 6)
    assert p.tolist() == [[0, 1, 1], [0, 1, 0], [1, 1, 0.25], [1, 0, 1], [0, 0, 0], [1, 0, 0]]
    # three coincident unobserved points never merge and can never be separated
    with pytest.raises(RuntimeError):
        NO.merge_subsample(np.zeros((3, 3), np.float32), rgb(3), [-1, -1, -1], max_frames=10000)


def test_oracle_merge_properties():
    pts, rgb = _cloud("surface", 20000, 1)
    ext = np.full(pts.shape[0], 0.02, np.float32)
    p, c = NO.merge_subsample(pts, rgb, ext)
    assert 0 < p.shape[0] < pts.shape[0]
    assert np.all(p.min(0) >= pts.min(0)) and np.all(p.max(0) <= pts.max(0))
    # a huge extent merges everything into the sequential fp32 mean
    p1, _ = NO.merge_subsample(pts, rgb, np.full(pts.shape[0], 100, np.float32), max_bbox_aspect_ratio=1e9)
    acc = np.zeros(3, np.float32)
    for row in pts:
        acc += row
    assert p1.shape == (1, 3) and np.array_equal(p1[0], acc / np.float32(pts.shape[0]))
    # a negative extent disables merging: the output is a permutation of the input
    p2, c2 = NO.merge_subsample(pts, rgb, np.full(pts.shape[0], -1, np.float32))
    assert p2.shape == pts.shape
    assert np.array_equal(np.sort(p2.view([("x", "f4"), ("y", "f4"), ("z", "f4")]).ravel()),
                          np.sort(pts.view([("x", "f4"), ("y", "f4"), ("z", "f4")]).ravel()))


# ---- GPU: libmdi_b200.so against the restatement ---------------------------------------------
def _gpu_merge(pts, rgb, ext, ratio=1.1, mult=1.0, max_levels=0):
    import ctypes as C
    L = mod("_lib")
    lib = L.load()
    dev = "cuda:0"
    n = pts.shape[0]
    tp, tc, te = (torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev) for a in (pts, rgb, ext))
    ws = torch.empty(int(lib.mdi_merge_workspace_bytes(n)), dtype=torch.uint8, device=dev)
    op = torch.full((max(n, 1), 3), float("nan"), device=dev)
    oc = torch.full((max(n, 1), 3), float("nan"), device=dev)
    n_out, n_lev = C.c_int64(-1), C.c_int32(-1)
    st = torch.cuda.current_stream().cuda_stream
    L.check(lib.mdi_merge_subsample(L.ptr(tp), L.ptr(tc), L.ptr(te), n, ratio, mult, max_levels, L.ptr(ws),
                                    L.ptr(op), L.ptr(oc), C.byref(n_out), C.byref(n_lev), st), "mdi_merge_subsample")
    m = n_out.value
    return op[:m].cpu().numpy(), oc[:m].cpu().numpy(), n_lev.value


@pytest.mark.gpu
@pytest.mark.parametrize("n,n_cams", [(1, 1), (5000, 7), (70000, 300)])
def test_min_extents_bit_exact(n, n_cams):
    L = mod("_lib")
    lib = L.load()
    K, P, sizes = _cams(n_cams, 5)
    pts, _ = _cloud("gauss", n, 6)
    pts[:, 2] *= 0.5
    want = NO.min_extents(pts, K, P, sizes)
    dev = "cuda:0"
    tp, tK, tP, ts = (torch.from_numpy(a).to(dev) for a in (pts, K, P, sizes))  # kept alive over the call
    ext = torch.empty(n, device=dev)
    L.check(lib.mdi_min_extents(L.ptr(tp), n, L.ptr(tK), L.ptr(tP), L.ptr(ts), n_cams, L.ptr(ext),
                                torch.cuda.current_stream().cuda_stream), "mdi_min_extents")
    got = ext.cpu().numpy()
    assert np.array_equal(got, want)
    if n > 1:
        assert (want > 0).sum() > n // 4 and (want < 0).sum() > 0


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n,ext_scale,ratio", [
    ("gauss", 2, 1.0, 1.1), ("gauss", 3, 0.01, 1.1), ("gauss", 1000, 0.3, 1.1), ("gauss", 50000, 0.08, 1.5),
    ("surface", 120000, 0.02, 1.1), ("surface", 300000, 0.01, 3.0), ("clusters", 40000, 0.004, 1.1),
    ("clusters", 40000, 0.004, float("inf")), ("gauss", 30000, 1000.0, 1e9)])
def test_merge_subsample_bit_exact(kind, n, ext_scale, ratio):
    pts, rgb = _cloud(kind, n, 11)
    rng = np.random.default_rng(12)
    ext = (ext_scale * rng.uniform(0.5, 1.5, n)).astype(np.float32)
    ext[rng.random(n) < 0.02] = -1.0  # points no camera sees
    if kind == "clusters":
        ext = np.abs(ext)  # coincident points must be able to merge, else the reference never terminates
    want_p, want_c = NO.merge_subsample(pts, rgb, ext, ratio, 1.0)
    got_p, got_c, levels = _gpu_merge(pts, rgb, ext, ratio, 1.0)
    assert got_p.shape == want_p.shape, (got_p.shape, want_p.shape, levels)
    assert np.array_equal(got_p.view(np.uint32), want_p.view(np.uint32))
    assert np.array_equal(got_c.view(np.uint32), want_c.view(np.uint32))


@pytest.mark.gpu
def test_merge_subsample_edge_cases():
    L = mod("_lib")
    z = np.zeros((0, 3), np.float32)
    p, c, _ = _gpu_merge(z, z, np.zeros(0, np.float32))
    assert p.shape == (0, 3)
    one = np.array([[1, 2, 3]], np.float32)
    p, c, _ = _gpu_merge(one, one / 10, np.array([0.5], np.float32))
    assert np.array_equal(p, one) and np.array_equal(c, one / 10)
    # the input on which the reference loops forever is reported, not spun on
    with pytest.raises(L.MdiError, match="still open"):
        _gpu_merge(np.zeros((3, 3), np.float32), np.zeros((3, 3), np.float32), -np.ones(3, np.float32), max_levels=600)


@pytest.mark.gpu
def test_subsample_pointcloud_api_matches_restatement():
    sub = mod("point_cloud_postprocess.subsampling")
    cfg = mod("point_cloud_postprocess.config")
    K, P, sizes = _cams(12, 21)
    pts, rgb = _cloud("surface", 60000, 22)
    params = cfg.PointCloudSubsamplingParams(max_bbox_aspect_ratio=2.0, min_extent_multiplier=1.5)
    p, c, e, dbg_p, dbg_c = sub.subsample_pointcloud(pts, rgb, list(K), list(P), sizes, params, device="cuda:0")
    wp, wc, we = NO.subsample_pointcloud(pts, rgb, K, P, sizes, 2.0, 1.5)
    assert np.array_equal(e, we)
    assert p.shape == wp.shape and 0 < p.shape[0] < pts.shape[0]
    assert np.array_equal(p, wp) and np.array_equal(c, wc)
    with pytest.raises(ValueError, match="shape"):
        sub.subsample_pointcloud(pts[:, :2], rgb, list(K), list(P), sizes, params, device="cuda:0")
    with pytest.raises(ValueError, match="must match"):
        sub.subsample_pointcloud(pts, rgb[:-1], list(K), list(P), sizes, params, device="cuda:0")
