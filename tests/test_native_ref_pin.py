"""Pins oracle/native_oracle.c (the C restatement that the GPU tests of a13 / a14 compare against) to
the reference's OWN native module: native_modules/subsampling/src/{impl.cpp,pointcloud_subsampling.cpp}
compiled unmodified by oracle/build_ref_native.py into oracle/_ref/ (Eigen replaced by the stand-in
headers of oracle/eigen_standin/).  Count, emission order and every fp32 bit must agree on the
clouds the GPU tests use, including duplicate / coplanar groups (inf and NaN aspect ratios)."""
import contextlib
import os
from types import SimpleNamespace

import numpy as np
import pytest

import build_ref_native
import native_oracle as NO
from test_native_subsample import _cams, _cloud

pytestmark = pytest.mark.skipif(not (build_ref_native.available() or build_ref_native.built()),
                                reason="reference native sources not present and oracle/_ref not built")


@contextlib.contextmanager
def _quiet_in(tmp_path):
    """The reference prints its statistics and writes *_stats.txt into the CWD (impl.cpp:121-123,421-423)."""
    cwd = os.getcwd()
    os.chdir(tmp_path)
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        yield
    finally:
        os.dup2(saved, 1)
        os.close(devnull); os.close(saved)
        os.chdir(cwd)


def _ref(pts, rgb, K, P, sizes, ratio, mult, tmp_path):
    m = build_ref_native.load()
    params = SimpleNamespace(max_bbox_aspect_ratio=ratio, min_extent_multiplier=mult)   # read by attribute, .cpp:53-54
    with _quiet_in(tmp_path):
        p, c, e, dbg_p, dbg_c = m.subsample_pointcloud(pts, rgb, list(K), list(P), sizes, params)
    return p, c, e, dbg_p, dbg_c


@pytest.mark.parametrize("kind,n,n_cams,ratio,mult", [
    ("gauss", 1, 1, 1.1, 1.0), ("gauss", 2, 3, 1.1, 1.0), ("gauss", 3, 3, 1.1, 1.0), ("gauss", 1000, 7, 1.1, 1.0),
    ("gauss", 50000, 20, 1.5, 1.0), ("surface", 120000, 12, 1.1, 1.0), ("surface", 60000, 12, 2.0, 1.5),
    ("surface", 100000, 5, 3.0, 4.0), ("clusters", 40000, 9, 1.1, 1.0), ("clusters", 40000, 9, float("inf"), 2.0),
    ("gauss", 30000, 4, 1e9, 1000.0)])
def test_restatement_equals_reference_module(kind, n, n_cams, ratio, mult, tmp_path):
    K, P, sizes = _cams(n_cams, 31)
    pts, rgb = _cloud(kind, n, 32)
    if kind != "clusters":
        pts[:, 2] *= 0.5   # inside the cameras' frusta: most points are observed
    rp, rc, re, _, _ = _ref(pts, rgb, K, P, sizes, ratio, mult, tmp_path)
    oe = NO.min_extents(pts, K, P, sizes)
    assert np.array_equal(re.view(np.uint32), oe.view(np.uint32)), "a13: extents differ from impl.cpp:70-126"
    if n > 100:
        assert (oe > 0).sum() > n // 10
    op, oc = NO.merge_subsample(pts, rgb, oe, ratio, mult)
    assert rp.shape == op.shape, (rp.shape, op.shape)
    assert np.array_equal(rp.view(np.uint32), op.view(np.uint32)), "a14: positions / order differ from impl.cpp:313-426"
    assert np.array_equal(rc.view(np.uint32), oc.view(np.uint32))


def test_reference_module_rejects_bad_shapes_like_the_binding(tmp_path):
    """pointcloud_subsampling.cpp:31-51 -> ValueError; the GPU API raises the same messages."""
    K, P, sizes = _cams(2, 1)
    pts, rgb = _cloud("gauss", 10, 2)
    m = build_ref_native.load()
    params = SimpleNamespace(max_bbox_aspect_ratio=1.1, min_extent_multiplier=1.0)
    with pytest.raises(ValueError, match=r"shape \(N, 3\)"):
        m.subsample_pointcloud(pts[:, :2], rgb, list(K), list(P), sizes, params)
    with pytest.raises(ValueError, match="must match"):
        m.subsample_pointcloud(pts, rgb[:-1], list(K), list(P), sizes, params)
    with pytest.raises(ValueError, match="must match"):
        m.subsample_pointcloud(pts, rgb, list(K), list(P[:1]), sizes, params)
