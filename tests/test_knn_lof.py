"""a12 (LOF outlier removal) and the kNN of SURVEY 8f-1 against scikit-learn itself -- the
reference calls sklearn (postprocess.py:16-22, runner_utils.py:143-147), so sklearn run on the host
is the oracle here; tolerances are stated per test."""
import numpy as np
import pytest
import torch

from helpers import mod


def _cloud(n, seed, outliers=0.01):
    rng = np.random.default_rng(seed)
    uv = rng.random((n, 2)).astype(np.float32) * 4 - 2
    pts = np.stack([uv[:, 0], uv[:, 1], 0.3 * np.sin(2 * uv[:, 0]) * np.cos(uv[:, 1])], 1)
    pts += rng.normal(size=(n, 3)) * 2e-3
    m = rng.random(n) < outliers
    pts[m] += rng.normal(size=(int(m.sum()), 3)) * 0.8  # floaters off the surface
    return pts.astype(np.float32)


@pytest.mark.gpu
@pytest.mark.parametrize("n,k", [(5, 4), (300, 4), (20000, 4), (20000, 40), (150000, 16)])
def test_knn_matches_sklearn(n, k):
    from sklearn.neighbors import NearestNeighbors
    K = mod("knn")
    pts = _cloud(n, 3)
    want_d, want_i = NearestNeighbors(n_neighbors=k, metric="euclidean").fit(pts).kneighbors(pts)
    x = torch.from_numpy(pts).cuda()
    got = K.knn(x, k)
    assert got.dtype == torch.float32 and got.shape == (n, k) and got.device == x.device
    # same neighbour distances (fp32 differences, fp64 accumulation on both sides)
    np.testing.assert_allclose(got.cpu().numpy(), want_d.astype(np.float32), rtol=1e-6, atol=1e-9)
    idx, dist, kdist = K.knn_search(x, k, exclude_self=False)
    idx = idx.cpu().numpy()
    assert np.array_equal(idx[:, 0], np.arange(n))  # no duplicates in this cloud: self first
    same = (idx == want_i).mean()
    assert same > 0.999, same  # index sets differ only where two neighbours are equidistant
    assert np.array_equal(kdist.cpu().numpy(), dist[:, -1].cpu().numpy())


@pytest.mark.gpu
def test_knn_duplicates_far_points_and_errors():
    from sklearn.neighbors import NearestNeighbors
    K = mod("knn")
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(4000, 3)).astype(np.float32)
    pts[100:180] = pts[100]            # 80 coincident points
    pts[0] = [500, -300, 900]          # isolated points: the octree walk must stay exact and bounded
    pts[1] = [-1e4, 0, 0]
    want_d, _ = NearestNeighbors(n_neighbors=8).fit(pts).kneighbors(pts)
    got = K.knn(torch.from_numpy(pts).cuda(), 8).cpu().numpy()
    np.testing.assert_allclose(got, want_d.astype(np.float32), rtol=1e-6, atol=1e-9)
    with pytest.raises(ValueError, match="Expected n_neighbors <= n_samples_fit"):
        K.knn(torch.zeros((3, 3)).cuda(), 4)
    with pytest.raises(mod("_lib").MdiError):
        K.knn(torch.zeros((10, 3)), 4)  # CPU tensor: no fallback


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["floaters_everywhere", "one_far_cluster", "flat_axis"])
def test_knn_robust_grid_with_many_points_outside_the_cube(case):
    """The grid covers the 0.1 % .. 99.9 % coordinate range only; everything outside is clamped into
    the (open-ended) border cells.  The search must stay exact when far more than 0.1 % of the points
    lie outside -- far floaters on all sides, a distant cluster on one side, a degenerate (flat) axis."""
    from sklearn.neighbors import NearestNeighbors
    K = mod("knn")
    rng = np.random.default_rng(7)
    pts = _cloud(30000, 5)
    if case == "floaters_everywhere":
        m = rng.random(len(pts)) < 0.02
        pts[m] = (rng.normal(size=(int(m.sum()), 3)) * 300).astype(np.float32)
    elif case == "one_far_cluster":
        pts[:900] = (np.array([4000.0, -2500.0, 800.0]) + rng.normal(size=(900, 3)) * 0.05).astype(np.float32)
    else:
        pts[:, 2] = 1.25
        pts[:60, 2] = rng.normal(size=60).astype(np.float32) * 50
    for k in (4, 40):
        want_d, want_i = NearestNeighbors(n_neighbors=k, metric="euclidean").fit(pts).kneighbors(pts)
        idx, dist, _ = K.knn_search(torch.from_numpy(pts).cuda(), k, exclude_self=False)
        np.testing.assert_allclose(dist.cpu().numpy(), want_d.astype(np.float32), rtol=1e-6, atol=1e-9)
        assert (idx.cpu().numpy() == want_i).mean() > 0.999


@pytest.mark.gpu
@pytest.mark.parametrize("n,k", [(3000, 40), (60000, 40), (60000, 10)])
def test_lof_matches_sklearn(n, k):
    from sklearn.neighbors import LocalOutlierFactor
    K = mod("knn")
    pts = _cloud(n, 7, outliers=0.02)
    clf = LocalOutlierFactor(n_neighbors=k, n_jobs=-1)
    pred = clf.fit_predict(pts)
    want_lof = -clf.negative_outlier_factor_
    out, lof = K.local_outlier_factor(torch.from_numpy(pts).cuda(), k)
    lof = lof.cpu().numpy(); out = out.cpu().numpy()
    np.testing.assert_allclose(lof, want_lof, rtol=2e-6)
    assert (lof == want_lof).mean() > 0.99  # same fp32 arithmetic: bit-equal except around distance ties
    # SURVEY 8 a12 bar: identical outlier set except points with |LOF - 1.5| < 1e-4
    differ = out != (pred == -1)
    assert not np.any(differ & (np.abs(want_lof - 1.5) >= 1e-4))
    assert differ.sum() <= 2
    assert 0 < out.sum() < n // 10


@pytest.mark.gpu
def test_postprocess_lof_then_merge():
    pp = mod("point_cloud_postprocess.postprocess")
    cfg = mod("point_cloud_postprocess.config")
    from sklearn.neighbors import LocalOutlierFactor
    import native_oracle as NO
    from test_native_subsample import _cams
    pts = _cloud(30000, 9, outliers=0.02)
    rgb = np.random.default_rng(1).random((30000, 3)).astype(np.float32)
    K, P, sizes = _cams(6, 2)
    c = cfg.PointCloudPostprocessConfig(outlier_removal=cfg.OutlierRemovalMethod.lof, lof_num_neighbors=20,
                                        subsample=True)
    p, r = pp.postprocess_point_cloud(torch.from_numpy(pts).cuda(), torch.from_numpy(rgb), list(K), list(P), sizes,
                                      c, "cuda:0")
    keep = LocalOutlierFactor(n_neighbors=20).fit_predict(pts) == 1
    wp, wc, _ = NO.subsample_pointcloud(pts[keep], rgb[keep], K, P, sizes)
    assert np.array_equal(p.cpu().numpy(), wp) and np.array_equal(r.cpu().numpy(), wc)
