"""View ranges on part streams (Engine.run, RunOptions.streams): a batch queued as 2 or 3 contiguous
view ranges on separate CUDA streams must give bit for bit what the one-stream run gives -- fits,
RANSAC iterations (the device sampler draws per batch view, MdiRansacConfig.view_base), quantiles,
keep mask and the cloud -- on the one-pass statistics path, on the exact path (small views), with
caller-supplied RANSAC samples, and when a view takes the exact-statistics retry."""
import numpy as np
import pytest
import torch

from helpers import mod

DEV = torch.device("cuda", 0) if torch.cuda.is_available() else None


def _cfg(aligner, subsample):
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum(aligner)
    cfg.mdi.subsample_factor = subsample
    return cfg


def _run(cfg, sc, streams, seed=7, sample_idx=None):
    mdi, L = mod("monocular_depth_init"), mod("_lib")
    torch.manual_seed(seed)   # the RANSAC device-sampler seed is drawn from torch's CPU generator
    r = mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points,
                              sc.point_indices, DEV, L.IMAGE_U8, sample_idx=sample_idx, streams=streams)
    torch.cuda.synchronize()
    return r


def _same(a, b):
    assert a.total == b.total > 0
    for f in ("status", "n_valid", "scale_shift", "ransac_iters", "ransac_inliers", "lo_hi", "quantiles",
              "patch_blocked", "keep", "view_counts", "view_offsets", "xyz", "rgb", "P", "d"):
        x, y = getattr(a, f), getattr(b, f)
        assert (x is None) == (y is None), f
        if x is not None:
            assert x.shape == y.shape and torch.equal(x.view(torch.uint8) if x.is_floating_point() else x,
                                                      y.view(torch.uint8) if y.is_floating_point() else y), f


@pytest.mark.gpu
@pytest.mark.parametrize("aligner,subsample,H,W", [("ransac", "adaptive", 300, 400), ("msac", "adaptive", 120, 160),
                                                   ("lstsqrs", 7, 300, 400), ("ransac", 4, 120, 160)])
def test_view_ranges_on_streams_equal_one_stream(aligner, subsample, H, W):
    syn = mod("synthetic")
    eng = mod("engine").get_engine(DEV)
    sc = syn.make_scene(n_views=50, H=H, W=W, n_sfm=1500, seed=31, device=str(DEV), low_conf_every=9)
    cfg = _cfg(aligner, subsample)
    ref = _run(cfg, sc, 1)
    assert int((ref.status != 0).sum()) > 0        # skipped (low-confidence) views sit inside the ranges
    for streams in (2, 3):
        got = _run(cfg, sc, streams)
        _same(got, ref)
        assert len(getattr(eng, "_part_streams", [])) >= streams - 1   # the ranges really ran on part streams
    _same(_run(cfg, sc, 1), ref)                   # and the one-stream path is unchanged afterwards


@pytest.mark.gpu
def test_view_ranges_with_given_samples_and_retry_view():
    syn = mod("synthetic")
    sc = syn.make_scene(n_views=40, H=300, W=400, n_sfm=1200, seed=32, device=str(DEV), low_conf_every=50)
    d21 = sc.depth[21]                             # two-valued map: its one-pass brackets cannot be certified
    sc.depth[21] = torch.where(d21 > d21.median(), torch.full_like(d21, 2.0), torch.full_like(d21, 1.0))
    cfg = _cfg("ransac", "adaptive")
    cfg.mdi.alignment.ransac.max_iters = 300
    g = torch.Generator().manual_seed(5)
    idx = torch.stack([torch.stack([torch.randperm(1000, generator=g)[:4] for _ in range(300)]) for _ in range(40)]).to(torch.int32)
    ref = _run(cfg, sc, 1, sample_idx=idx)
    assert ref.extras.get("stats_retry")
    got = _run(cfg, sc, 2, sample_idx=idx)
    assert got.extras.get("stats_retry")
    _same(got, ref)
    ref2, got2 = _run(cfg, sc, 1), _run(cfg, sc, 2)    # device sampler
    _same(got2, ref2)
    assert not torch.equal(ref2.scale_shift, ref.scale_shift)


def _range_run(depth, mask, ss, raw, exact=False):
    """a7 + keep mask of one view with a given scale / shift, statistics on the raw depth or behind the fit"""
    L = mod("_lib")
    single = mod("_single")
    eng = mod("engine").get_engine(DEV)
    vin = single.single_view_inputs(depth.to(DEV), mask.to(DEV), DEV)
    opt = mod("engine").RunOptions(aligner=None, scale_shift=torch.tensor([ss], dtype=torch.float32, device=DEV),
                                   subsample="adaptive", use_patch_mask=False,
                                   flags=L.KEEP_PRED_MASK | L.KEEP_SENTINEL | L.KEEP_SUBSAMPLE, want_xyz=False, want_rgb=False,
                                   stop_after="mask", stats_exact=exact, raw_stats=raw)
    res = eng.run(vin, opt)
    torch.cuda.synchronize()
    return res


@pytest.mark.gpu
def test_raw_depth_statistics_equal_statistics_behind_the_fit():
    """MDI_STATS_RAW + mdi_depth_range_map: the one pass over the pixels on the RAW depth (queued beside the
    alignment kernels) and the order statistics mapped through the fit afterwards give bit for bit the
    lo / hi, quantiles and keep mask of the pass that aligns every pixel -- for ordinary fits, fits that
    overflow, and, through the exact-path retry, for a decreasing fit and for views in which a pixel of the
    population aligns to exactly -42.0 (the sentinel the population excludes)."""
    g = torch.Generator().manual_seed(11)
    H, W = 480, 640
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    depth = 4 + 1.5 * torch.sin(xx / 97.0) + torch.cos(yy / 71.0) + 0.02 * torch.randn(H, W, generator=g)
    mask = torch.rand(H, W, generator=g) > 0.1
    hit = depth.clone()
    hit[100, 200] = 2.0                      # with (s, t) = (3, -48): 2 * 3 - 48 = -42 exactly, inside the mask
    mask[100, 200] = True
    cases = [("plain", depth, (1.3, 0.25), False), ("shrink", depth, (1e-3, -7.0), False), ("tiny", depth, (1e-30, 0.0), None),
             ("huge", depth, (3e38, 1.0), None), ("negative", depth, (-0.7, 9.0), True), ("zero", depth, (0.0, 2.0), True),
             ("sentinel_hit", hit, (3.0, -48.0), True), ("sentinel_near", depth, (3.0, -48.0), None)]
    for name, d, ss, expect_retry in cases:
        ref = _range_run(d, mask, ss, raw=False)
        got = _range_run(d, mask, ss, raw=True)
        assert torch.equal(got.lo_hi.view(torch.int32), ref.lo_hi.view(torch.int32)), (name, got.lo_hi, ref.lo_hi)
        assert torch.equal(got.quantiles.view(torch.int32), ref.quantiles.view(torch.int32)), name
        assert torch.equal(got.keep, ref.keep) and got.total == ref.total, name
        assert torch.equal(got.status, ref.status), name
        if expect_retry is not None:
            assert bool(got.extras.get("stats_retry", False)) == expect_retry, name
    # and through the whole pipeline: overlap on / off, the same cloud
    emod = mod("engine")
    sc = mod("synthetic").make_scene(n_views=12, H=300, W=400, n_sfm=1500, seed=33, device=str(DEV), low_conf_every=5)
    cfg = _cfg("ransac", "adaptive")
    saved = emod.OVERLAP_STATS
    try:
        emod.OVERLAP_STATS = False
        ref = _run(cfg, sc, 1)
        emod.OVERLAP_STATS = True
        got = _run(cfg, sc, 1)
    finally:
        emod.OVERLAP_STATS = saved
    assert hasattr(emod.get_engine(DEV), "_align_stream")
    _same(got, ref)
