"""GPU parity tests proper: the CUDA path, called through the drop-in API / C ABI, against the
golden vectors of the unmodified reference and against the oracle on the same seeded inputs.

Bars (north_star): keep masks, counts, indices, SfM filter: bit-exact.  RANSAC: identical
iteration count and inlier set for identical sample indices.  scale/shift and xyz/rgb: 1e-5
relative (FP32) -- REL below; xyz errors are taken relative to the extent of the cloud (xyz_err)."""
import numpy as np
import pytest
import torch

import mdi_oracle as O
from helpers import (BIG_CASES, FULL_CASES, SMALL_CASES, OneViewParser, case_config, load_case, mod, oracle_kwargs,
                     rgb_matches)

pytestmark = pytest.mark.gpu
REL = 1e-5
DEV = "cuda:0"


def rel_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30))) if a.size else 0.0


def xyz_err(a, b):
    """max_i |a_i - b_i| / max_j |b_j|: error relative to the magnitude of the cloud.  (Points
    of these scenes pass through the world origin, where a per-point ratio is meaningless: the
    fp32 roundings are relative to the camera translation |C| ~ 4, not to |p| ~ 0.)"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if not a.size:
        return 0.0
    return float(np.max(np.linalg.norm(a - b, axis=1)) / max(np.max(np.linalg.norm(b, axis=1)), 1e-30))


def fit_err(h, h_ref, d):
    """1e-5 bar on the fitted line: max_i |(s d_i + t) - (s_ref d_i + t_ref)| / max_i |s_ref d_i + t_ref|
    over the SfM sample depths.  (Parameter-wise ratios are not meaningful at 1e-5: the reference
    solves the normal equations with an fp32 SVD whose own error vs the exact fp64 solution is
    up to ~1e-5 on the smaller parameter -- tests/test_host_math.py shows ours is the closer one.
    They are still bounded at 1e-4 below.)"""
    d = np.asarray(d, np.float64)
    a = float(h[0]) * d + float(h[1])
    b = float(h_ref[0]) * d + float(h_ref[1])
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


def exact_fit(c):
    """fp64 least squares on the point set the reference's returned fit was computed from (all
    points for lstsqrs; the sample-inliers of the winning hypothesis for RANSAC / MSAC, found by
    replaying the oracle, which is bit-identical to the reference)."""
    d = c.depth[c.ref_uv[1], c.ref_uv[0]]
    z = torch.from_numpy(c.ref_z)
    sel = torch.ones_like(z, dtype=torch.bool)
    if c.meta["aligner"] != "lstsqrs":
        torch.manual_seed(c.meta["seed"])
        tr = O.ransac_fit(d, z, loss_kind=c.meta["aligner"])
        D = O.stack_ones(d)
        for idx in tr.sample_idx:
            idx = torch.as_tensor(idx)
            inl = O.residuals_sq(O.lstsq_fit(D[:, idx], z[idx]), d, z) < 0.01
            h = O.lstsq_fit(D[:, inl], z[inl])
            if float(h[0]) == float(tr.scale) and float(h[1]) == float(tr.shift):
                sel = inl
                break
    dd, gg = d[sel].double().numpy(), z[sel].double().numpy()
    return np.linalg.lstsq(np.stack([dd, np.ones_like(dd)], 1), gg, rcond=None)[0]


def run_case(c, with_samples=True):
    pfd = mod("depth_prediction.points_from_depth")
    dpi = mod("depth_prediction.predictors.depth_predictor_interface")
    types = mod("types")
    cfg = case_config(c)
    pred = dpi.PredictedDepth(c.depth.to(DEV), c.mask.to(DEV))
    image = types.InputImage(data=c.image.float() / 255.0, name=c.name, cam2world=c.c2w, K=c.K)
    sample_idx = None
    if with_samples and c.meta["aligner"] in ("ransac", "msac") and not c.raised:
        sample_idx = torch.from_numpy(O.draw_sample_idx(int(c.ref_uv.shape[1]), 2500, 4, c.meta["seed"]))
    return pfd.get_pts_from_depth(pred, image, OneViewParser(c), cfg, DEV, sample_idx=sample_idx, return_result=True)


@pytest.mark.parametrize("name", SMALL_CASES + FULL_CASES + BIG_CASES)
def test_view_matches_reference_golden(name):
    c = load_case(name)
    if c.raised:
        exc = mod("depth_alignment.exceptions").LowDepthAlignmentConfidenceError
        with pytest.raises(exc):
            run_case(c)
        return
    pts, keep, P, res = run_case(c)
    assert pts.is_cuda and not keep.is_cuda and keep.dtype == torch.bool and P.is_cuda
    # a1 / a2: projection matrix, SfM filter: bit-exact
    assert np.array_equal(P.cpu().numpy(), c.ref_P)
    n = int(res.n_valid[0])
    assert n == c.ref_uv.shape[1]
    assert np.array_equal(res.uv[:, :n].cpu().numpy(), c.ref_uv)
    assert np.array_equal(res.z[:n].cpu().numpy(), c.ref_z)
    # a3 / a4: scale, shift within 1e-5 relative; RANSAC trace identical
    ss = res.scale_shift[0].cpu().numpy()
    d = c.depth.numpy()[c.ref_uv[1], c.ref_uv[0]]
    e_ref = fit_err(ss, (c.ref_scale, c.ref_shift), d)
    if e_ref > REL:
        # ill-conditioned inlier set: the reference's fp32 SVD is itself further than 1e-5 from the
        # exact solution of its own normal equations; then ours must be the closer one, within 1e-5
        exact = exact_fit(c)
        e_ours, e_theirs = fit_err(ss, exact, d), fit_err((c.ref_scale, c.ref_shift), exact, d)
        assert e_ours <= REL and e_ours < e_theirs, (ss, c.ref_scale, c.ref_shift, exact)
    assert rel_err(ss[0], c.ref_scale) <= 10 * REL and rel_err(ss[1], c.ref_shift) <= 10 * REL
    if c.meta["aligner"] in ("ransac", "msac"):
        assert int(res.ransac_iters[0]) == int(c.ref_ransac_iters)
        assert int(res.ransac_inliers[0]) == int(c.ref_ransac_inliers)
        # north_star: the selected inlier SET is identical for identical sample indices (ransacs.py:150-151)
        got_inl = res.inlier_mask[:n].bool().cpu().numpy()
        assert np.array_equal(got_inl, c.ref_ransac_inlier_mask), \
            f"{int((got_inl != c.ref_ransac_inlier_mask).sum())} of {n} inlier flags differ"
    # a5-a10: keep mask bit-exact, count, order
    assert np.array_equal(keep.numpy(), c.ref_keep), f"{int((keep.numpy() != c.ref_keep).sum())} mask bits differ"
    assert pts.shape[0] == c.ref_pts.shape[0] == int(c.ref_keep.sum())
    # where the reference's own fit is off by more than 1e-5 (tie-break above) its error carries over
    # linearly into the depths it unprojects; the stage-wise test pins xyz bit-exactly for equal fits
    tol = REL if e_ref <= REL else REL + 2 * e_ref
    assert xyz_err(pts.cpu().numpy(), c.ref_pts) <= tol
    if "ref_lo" in c.__dict__:
        lh = res.lo_hi[0].cpu().numpy()
        span = max(abs(float(c.ref_lo)), abs(float(c.ref_hi)))
        assert abs(lh[0] - c.ref_lo) <= tol * span and abs(lh[1] - c.ref_hi) <= tol * span


@pytest.mark.parametrize("name", ["small_lstsq_static4", "small_adaptive_range_3_9", "c1_lstsq_static10",
                                  "c2_ransac_adaptive"] + BIG_CASES)
def test_stagewise_bit_exact_given_reference_fit(name):
    """L1 parity: with the reference's (scale, shift) fed in, keep mask, xyz, rgb and the quantile
    statistics must be bit-identical -- at every BASELINE shape, including n ~ 8.3 M valid pixels at
    3840x2160 where 0.75 * (n - 1) is not representable in fp32."""
    c = load_case(name)
    eng_m = mod("engine")
    single = mod("_single")
    L = mod("_lib")
    cfg = case_config(c)
    opt = single.options_from_config(cfg, aligner=None, want_rgb=True,
                                     scale_shift=torch.tensor([[float(c.ref_scale), float(c.ref_shift)]], device=DEV))
    sfm = c.sfm.to(DEV).float().contiguous()
    vin = single.single_view_inputs(c.depth.to(DEV), c.mask.to(DEV), DEV, K=c.K, c2w=c.c2w,
                                    image=(c.image.to(DEV)), image_kind=L.IMAGE_U8,
                                    sfm_xyz=sfm, sfm_counts=[len(sfm)])
    res = eng_m.get_engine(DEV).run(vin, opt)
    assert np.array_equal(res.keep.bool().cpu().numpy(), c.ref_keep)
    assert np.array_equal(res.xyz.cpu().numpy(), c.ref_pts)
    assert rgb_matches(c, res.rgb.cpu().numpy())
    if "ref_lo" in c.__dict__:
        assert np.float32(res.lo_hi[0, 0].item()) == c.ref_lo and np.float32(res.lo_hi[0, 1].item()) == c.ref_hi
        assert np.float32(res.quantiles[0, 0].item()) == c.ref_q1 and np.float32(res.quantiles[0, 1].item()) == c.ref_q3


def test_batched_scene_equals_per_view_oracle():
    """Four views (one skipped for low confidence, one with an invalid rectangle) in ONE batch:
    view-major, row-major order; identical counts; rgb gather from uint8 and float images."""
    syn = mod("synthetic")
    mdi = mod("monocular_depth_init")
    L = mod("_lib")
    sc = syn.make_scene(n_views=4, H=120, W=160, n_sfm=2500, seed=11, low_conf_every=4)
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = "adaptive"
    pts_o, rgb_o, traces, kept = O.scene_points(sc, aligner="lstsqrs", subsample="adaptive", include_sfm_points=False)
    assert kept == [0, 1, 2]
    for kind, imgs in [(L.IMAGE_U8, [im.to(DEV) for im in sc.image]),
                       (L.IMAGE_F32_255, [im.float().pin_memory() for im in sc.image])]:
        res = mdi.points_from_views(cfg, [d.to(DEV) for d in sc.depth], [m.to(DEV) for m in sc.mask], imgs,
                                    sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, DEV, kind)
        torch.cuda.synchronize()
        assert res.status.cpu().tolist() == [0, 0, 0, 1]
        counts = res.view_counts.cpu().tolist()
        assert counts == [int(t.keep.sum()) for t in traces[:3]] + [0]
        keep = res.keep.bool().cpu().view(4, -1)
        for i in range(3):
            assert torch.equal(keep[i], traces[i].keep)
        assert not keep[3].any()
        assert xyz_err(res.xyz.cpu().numpy(), pts_o.numpy()) <= REL
        assert np.array_equal(res.rgb.cpu().numpy(), rgb_o.numpy())


def test_subsampler_and_mask_apis_vs_oracle():
    c = load_case("tall_lstsq_adaptive")
    aligned = (c.depth * torch.tensor(c.ref_scale) + torch.tensor(c.ref_shift))
    a_dev, m_dev = aligned.to(DEV), c.mask.to(DEV)
    rgb = torch.zeros(*c.depth.shape, 3)
    st = mod("depth_subsampling.static_subsampler").StaticDepthSubsampler(7)
    assert torch.equal(st.get_mask(rgb, a_dev, m_dev).cpu(), O.static_mask(c.mask, 7))
    ad_cfg = mod("depth_subsampling.config").AdaptiveSubsamplingConfig(4, 11)
    ad = mod("depth_subsampling.adaptive_subsampling").AdaptiveDepthSubsampler(ad_cfg)
    assert torch.equal(ad.get_mask(rgb, a_dev, m_dev).cpu(), O.adaptive_mask(aligned, c.mask, 4, 11))
    # more than 32 strides: the flat lattice walk with the exact per-point stride arithmetic
    ad_wide = mod("depth_subsampling.adaptive_subsampling").AdaptiveDepthSubsampler(
        mod("depth_subsampling.config").AdaptiveSubsamplingConfig(2, 40))
    assert torch.equal(ad_wide.get_mask(rgb, a_dev, m_dev).cpu(), O.adaptive_mask(aligned, c.mask, 2, 40))
    nm = mod("depth_subsampling.num_sfm_points_mask")
    uv = torch.from_numpy(c.ref_uv)
    for n_small, thr in [(10, 8), (20, 2), (5, 30)]:
        pc = mod("depth_subsampling.config").NumSfMPointsMaskConfig(n_small, thr)
        got = nm.num_sfm_points_mask(uv.to(DEV), tuple(c.depth.shape), pc)
        assert torch.equal(got, O.sfm_density_mask_loop(uv, tuple(c.depth.shape), n_small, thr))
    gm = mod("depth_prediction.points_from_depth").depth_gradient_mask(a_dev, 0.03)
    assert torch.equal(gm.cpu(), O.gradient_mask(aligned, 0.03))


def test_alignment_strategy_api():
    c = load_case("small_lstsq_static4")
    dpi = mod("depth_prediction.predictors.depth_predictor_interface")
    pred = dpi.PredictedDepth(c.depth.to(DEV), c.mask.to(DEV))
    uv, z = torch.from_numpy(c.ref_uv).to(DEV), torch.from_numpy(c.ref_z).to(DEV)
    cfg = case_config(c)
    r = mod("depth_alignment.pipeline").DepthAlignmentPipeline.from_config(cfg).align(None, pred, uv, z, cfg, None)
    ref_aligned = c.depth * torch.tensor(c.ref_scale) + torch.tensor(c.ref_shift)
    assert rel_err(r.aligned_depth.cpu().numpy(), ref_aligned.numpy()) <= 4 * REL
    assert torch.equal(r.mask.cpu(), c.mask & (r.aligned_depth.cpu() != -42.0))
    d = c.depth[c.ref_uv[1], c.ref_uv[0]].to(DEV)
    s, t = mod("depth_alignment.alignment.lstsqrs").align_depth_least_squares(torch.vstack([d, torch.ones_like(d)]), z)
    assert fit_err((float(s), float(t)), (c.ref_scale, c.ref_shift), d.cpu().numpy()) <= REL
    # rank-deficient sample: pinv's minimum-norm answer
    d4 = torch.tensor([2.0, 2, 2, 2], device=DEV)
    s, t = mod("depth_alignment.alignment.lstsqrs").align_depth_least_squares(
        torch.vstack([d4, torch.ones_like(d4)]), torch.tensor([3.0, 3.1, 2.9, 3.0], device=DEV))
    assert abs(float(s) - 1.2) < 1e-5 and abs(float(t) - 0.6) < 1e-5


def test_ransac_device_sampler_recovers_truth():
    """Throughput mode (device-generated samples): the fit must land on the generating
    scale/shift of the synthetic views and be reproducible under torch.manual_seed."""
    syn = mod("synthetic")
    mdi = mod("monocular_depth_init")
    L = mod("_lib")
    sc = syn.make_scene(n_views=6, H=240, W=320, n_sfm=3000, seed=5, low_conf_every=50)
    cfg = mod("config").Config()
    outs = []
    for seed in (1234, 1234, 1, 2):
        torch.manual_seed(seed)
        res = mdi.points_from_views(cfg, sc.depth.to(DEV), sc.mask.to(DEV), sc.image.to(DEV), sc.K, sc.c2w,
                                    sc.names, sc.points, sc.point_indices, DEV, L.IMAGE_U8)
        outs.append((res.scale_shift.cpu().numpy().copy(), res.xyz.cpu().numpy().copy()))
        # A view either lands on the truth or left the loop at iteration 0 through the reference's own
        # stopping rule: a first sample with a gross outlier leaves a handful of inliers, ratio**4
        # vanishes in float32, log(1 - 0) = 0 and _required_samples returns 0 (ransacs.py:88-91)
        close = np.isclose(outs[-1][0], sc.true_scale_shift, rtol=0.02, atol=0.02).all(axis=1)
        ratio = (res.ransac_inliers.cpu().numpy() / res.n_valid.cpu().numpy()).astype(np.float32)
        quirk = (res.ransac_iters.cpu().numpy() == 0) & (np.float32(1.0) - ratio ** 4 == np.float32(1.0))
        assert (close | quirk).all(), (seed, outs[-1][0], sc.true_scale_shift)
        assert close.sum() >= 4
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert not np.array_equal(outs[0][0], outs[2][0])


def _range_via_engine(aligned, mask, exact):
    L = mod("_lib")
    single = mod("_single")
    eng = mod("engine").get_engine(DEV)
    vin = single.single_view_inputs(aligned.to(DEV), mask.to(DEV), DEV)
    opt = mod("engine").RunOptions(aligner=None, subsample="adaptive", use_patch_mask=False,
                                   flags=L.KEEP_PRED_MASK | L.KEEP_SUBSAMPLE, want_xyz=False, want_rgb=False,
                                   stop_after="mask", stats_exact=exact)
    res = eng.run(vin, opt)
    return res


def test_quantile_one_pass_path_equals_exact_path_and_torch():
    """a7: the one-pass (sample bracket + certified select) statistics, the three-pass radix
    select and torch.quantile on the CPU agree bit for bit; degenerate inputs that cannot be
    bracketed (constant / few-valued depth) go through the retry path and still agree."""
    g = torch.Generator().manual_seed(7)
    H, W = 600, 800
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    smooth = 4 + 1.5 * torch.sin(xx / 123.0) + torch.cos(yy / 107.0) + 0.01 * torch.randn(H, W, generator=g)
    mask_all = torch.ones(H, W, dtype=torch.bool)
    mask_rect = mask_all.clone(); mask_rect[100:300, 200:700] = False
    heavy = smooth.clone(); heavy[: H // 2] = 300.0          # half the image clamped to one value
    const = torch.full((H, W), 2.5)
    few = torch.randint(0, 3, (H, W), generator=g).float() + 1.0
    neg = -smooth
    cases = [("smooth", smooth, mask_all, False), ("rect", smooth, mask_rect, False), ("neg", neg, mask_all, False),
             ("heavy", heavy, mask_all, None), ("const", const, mask_all, True), ("few", few, mask_rect, True)]
    for name, a, m, expect_retry in cases:
        lo, hi = O.adaptive_range(a, m)[:2]
        keep_ref = O.adaptive_mask(a, m, 5, 15)
        fast = _range_via_engine(a, m, exact=False)
        exact = _range_via_engine(a, m, exact=True)
        for tag, r in (("fast", fast), ("exact", exact)):
            lh = r.lo_hi[0].cpu()
            assert float(lh[0]) == float(lo) and float(lh[1]) == float(hi), (name, tag, lh, lo, hi)
            assert torch.equal(r.keep.view(-1).bool().cpu(), keep_ref), (name, tag)
        assert torch.equal(fast.quantiles.cpu(), exact.quantiles.cpu()), name
        if expect_retry is not None:
            assert bool(fast.extras.get("stats_retry", False)) == expect_retry, name
        assert not exact.extras.get("stats_retry", False)


@pytest.mark.gpu
def test_quantile_depth_domain_loop_with_scale_shift_and_odd_populations():
    """a7 with a scale/shift applied on the fly: the one-pass kernel tests raw depths against
    depth-domain brackets (monotone alignment, stats_bracket_kernel).  Positive scales take that
    loop, a negative scale the key-domain loop; NaN / inf in the population and pixels that hit
    the -42.0 sentinel after alignment must be noticed and sent through the exact path.  All
    bit-identical to torch.quantile on the aligned values."""
    L = mod("_lib")
    single = mod("_single")
    E = mod("engine")
    g = torch.Generator().manual_seed(11)
    H, W = 600, 800
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    d = 4 + 1.5 * torch.sin(xx / 123.0) + torch.cos(yy / 107.0) + 0.01 * torch.randn(H, W, generator=g)
    mask = torch.ones(H, W, dtype=torch.bool); mask[50:200, 100:500] = False
    odd_mask = mask.to(torch.uint8) * 7  # any non-zero byte counts
    d_nan = d.clone(); d_nan[300, 400] = float("nan")
    d_nan_out = d.clone(); d_nan_out[100, 200] = float("nan")   # outside the population: harmless
    d_inf = d.clone(); d_inf[10, 10] = float("inf")
    s_sent, t_sent = 2.0, -50.0
    d_sent = d.clone(); d_sent[400:420, 300:340] = 4.0            # 4 * 2 - 50 = -42 exactly
    d_far = d.clone(); d_far[77, 99] = -42.0
    cases = [("plain", d, mask, 2.37, -0.81, False, False),
             ("cross_zero", d, mask, 1.7, -7.0, False, False),
             ("tiny_scale", d, mask, 1e-30, 0.0, False, None),
             ("neg_scale", d, mask, -1.3, 9.0, False, False),
             ("nan_in", d_nan, mask, 1.5, 0.25, False, True),
             ("nan_out", d_nan_out, mask, 1.5, 0.25, False, False),
             ("inf_in", d_inf, mask, 1.5, 0.25, False, True),
             ("sentinel_hit", d_sent, mask, s_sent, t_sent, True, False),   # in the sample's range: tested per pixel
             ("sentinel_far_hit", d_far, mask, 1.0, 0.0, True, True),        # outside it: noticed, exact path
             ("sentinel_flag_no_hit", d, mask, 2.37, -0.81, True, False)]
    for name, dd, m, sc, sh, sentinel, expect_retry in cases:
        a = dd * torch.tensor(sc, dtype=torch.float32) + torch.tensor(sh, dtype=torch.float32)
        pop = m & (a != -42.0) if sentinel else m
        lo, hi = O.adaptive_range(a, pop)[:2]
        flags = L.KEEP_PRED_MASK | L.KEEP_SUBSAMPLE | (L.KEEP_SENTINEL if sentinel else 0)
        out = {}
        for exact in (False, True):
            eng = E.get_engine(DEV)
            vin = single.single_view_inputs(dd.to(DEV), m.to(DEV), DEV)
            opt = E.RunOptions(aligner=None, scale_shift=torch.tensor([[sc, sh]], dtype=torch.float32, device=DEV),
                               subsample="adaptive", use_patch_mask=False, flags=flags, want_xyz=False,
                               want_rgb=False, stop_after="mask", stats_exact=exact)
            out[exact] = eng.run(vin, opt)
            lh = out[exact].lo_hi[0].cpu()
            ref = torch.tensor([float(lo), float(hi)])
            assert torch.equal(lh.isnan(), ref.isnan()) and torch.equal(lh.nan_to_num(0.0), ref.nan_to_num(0.0)), \
                (name, exact, lh, lo, hi)
        qf, qe = out[False].quantiles.cpu(), out[True].quantiles.cpu()
        assert torch.equal(qf.isnan(), qe.isnan()) and torch.equal(qf.nan_to_num(0.0), qe.nan_to_num(0.0)), name
        assert torch.equal(out[False].keep.cpu(), out[True].keep.cpu()), name
        if expect_retry is not None:
            assert bool(out[False].extras.get("stats_retry", False)) == expect_retry, name
    # a mask whose non-zero bytes are not 1
    eng = E.get_engine(DEV)
    base = dict(aligner=None, scale_shift=torch.tensor([[2.37, -0.81]], dtype=torch.float32, device=DEV),
                subsample="adaptive", use_patch_mask=False, flags=L.KEEP_PRED_MASK | L.KEEP_SUBSAMPLE,
                want_xyz=False, want_rgb=False, stop_after="mask")
    r1 = eng.run(single.single_view_inputs(d.to(DEV), mask.to(DEV), DEV), E.RunOptions(**base))
    r7 = eng.run(single.single_view_inputs(d.to(DEV), odd_mask.to(DEV), DEV), E.RunOptions(**base))
    assert torch.equal(r1.quantiles.cpu(), r7.quantiles.cpu()) and torch.equal(r1.lo_hi.cpu(), r7.lo_hi.cpu())


@pytest.mark.gpu
@pytest.mark.parametrize("chunks", [1, 3, 16])
def test_host_views_end_to_end_path_equals_resident_path(chunks):
    """points_from_host_views (pinned host depth / mask / image, chunked copy-compute overlap, rgb
    back on the host) gives the cloud of the resident path, whatever the chunking."""
    syn = mod("synthetic")
    mdi = mod("monocular_depth_init")
    L = mod("_lib")
    sc = syn.make_scene(n_views=7, H=120, W=160, n_sfm=2500, seed=13, low_conf_every=3)
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = "adaptive"
    ref = mdi.points_from_views(cfg, sc.depth.to(DEV), sc.mask.to(DEV), sc.image.to(DEV), sc.K, sc.c2w, sc.names,
                                sc.points, sc.point_indices, DEV, L.IMAGE_U8)
    for _ in range(2):  # twice: the staging buffers are reused
        res, rgb_h, h2d, d2h = mdi.points_from_host_views(
            cfg, sc.depth.pin_memory(), sc.mask.pin_memory(), sc.image.pin_memory(), sc.K, sc.c2w, sc.names,
            sc.points, sc.point_indices, DEV, L.IMAGE_U8, chunks=chunks)
        assert res.extras["chunks"] == min(chunks, 7)
        assert res.total == ref.total and torch.equal(res.xyz, ref.xyz)
        assert torch.equal(rgb_h, ref.rgb.cpu())
        assert torch.equal(res.view_counts, ref.view_counts) and torch.equal(res.status, ref.status)
        assert torch.equal(res.extras["counts_host"], ref.view_counts.cpu())
        assert h2d > sc.depth.numel() * 4 and d2h >= res.total * 12


@pytest.mark.gpu
def test_tma_staged_statistics_kernel_is_bit_identical():
    """The opt-in TMA-staged statistics pass (MDI_STATS_MAIN=tma, chosen once per process) must pass
    the same quantile and golden-view checks as the default kernel: run them in a child process."""
    import os
    import subprocess
    import sys
    r = subprocess.run([sys.executable, "-m", "pytest", __file__, "-q", "-m", "gpu", "-x", "-k",
                        "quantile_one_pass or golden or batched_scene"],
                       capture_output=True, text=True, timeout=900, env={**os.environ, "MDI_STATS_MAIN": "tma"})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.gpu
@pytest.mark.parametrize("subsample", ["adaptive", 7])
def test_mixed_size_views_in_one_batch(subsample):
    """Views of different sizes passed as lists in ONE batch (the lattice grid is sized for the
    largest view, patch geometry is per view) equal the same views run scene by scene."""
    syn = mod("synthetic")
    mdi = mod("monocular_depth_init")
    L = mod("_lib")
    a = syn.make_scene(n_views=2, H=120, W=160, n_sfm=2500, seed=21, low_conf_every=50)
    b = syn.make_scene(n_views=3, H=208, W=96, n_sfm=2000, seed=22, low_conf_every=50)
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = subsample
    run = lambda sc: mdi.points_from_views(cfg, [d.to(DEV) for d in sc.depth], [m.to(DEV) for m in sc.mask],
                                           [im.to(DEV) for im in sc.image], sc.K, sc.c2w, sc.names, sc.points,
                                           sc.point_indices, DEV, L.IMAGE_U8)
    ra, rb = run(a), run(b)
    names = [f"a_{n}" for n in a.names] + [f"b_{n}" for n in b.names]
    points = np.concatenate([a.points, b.points])
    pidx = {f"a_{n}": a.point_indices[n] for n in a.names}
    pidx.update({f"b_{n}": np.asarray(b.point_indices[n]) + a.points.shape[0] for n in b.names})
    both = mdi.points_from_views(cfg, [d.to(DEV) for d in list(a.depth) + list(b.depth)],
                                 [m.to(DEV) for m in list(a.mask) + list(b.mask)],
                                 [im.to(DEV) for im in list(a.image) + list(b.image)],
                                 torch.cat([a.K, b.K]), torch.cat([a.c2w, b.c2w]), names, points, pidx, DEV, L.IMAGE_U8)
    assert both.view_counts.cpu().tolist() == ra.view_counts.cpu().tolist() + rb.view_counts.cpu().tolist()
    assert torch.equal(both.xyz, torch.cat([ra.xyz, rb.xyz])) and torch.equal(both.rgb, torch.cat([ra.rgb, rb.rgb]))
    assert torch.equal(both.keep, torch.cat([ra.keep, rb.keep]))
    assert both.total == ra.total + rb.total > 0


@pytest.mark.gpu
def test_speculative_emit_equals_plain_emit():
    """The emit kernel launched ahead of the size read-back (buffers sized from the previous run)
    returns the same cloud as the plain path when the guess covers the output, when it is far too
    large, and when it is too small (the kernel must then write nothing and the engine re-emit)."""
    syn = mod("synthetic")
    mdi = mod("monocular_depth_init")
    L = mod("_lib")
    eng = mod("engine").get_engine(DEV)
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = 2
    sc = syn.make_scene(n_views=5, H=120, W=160, n_sfm=2500, seed=21, device=str(DEV), low_conf_every=4)

    def run():
        return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points,
                                     sc.point_indices, DEV, L.IMAGE_U8)

    eng._emit_hint = 0
    ref = run()                                   # no hint yet: plain path
    assert not ref.extras.get("speculative_emit", False) and ref.total > 10000
    for hint, expect in ((ref.total, True), (ref.total * 50, True), (ref.total // 3, False), (1, False)):
        assert (ref.total <= hint + (hint >> 3) + 4096) == expect   # the engine's buffer rule
        eng._emit_hint = hint
        res = run()
        assert bool(res.extras.get("speculative_emit", False)) == expect, hint
        assert res.total == ref.total and eng._emit_hint == ref.total
        assert res.xyz.shape == ref.xyz.shape and torch.equal(res.xyz, ref.xyz) and torch.equal(res.rgb, ref.rgb)
        assert res.xyz.is_contiguous() and res.rgb.is_contiguous()


@pytest.mark.gpu
def test_cached_descriptor_upload_follows_camera_changes():
    """A repeated call on the same resident tensors reuses the uploaded descriptor table; the reuse
    must notice a camera edited in place (same tensors, new values) and give what a cold engine gives."""
    syn = mod("synthetic")
    mdi = mod("monocular_depth_init")
    L = mod("_lib")
    eng = mod("engine").get_engine(DEV)
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = 4
    sc = syn.make_scene(n_views=4, H=96, W=128, n_sfm=2000, seed=5, device=str(DEV))

    def run():
        return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points,
                                     sc.point_indices, DEV, L.IMAGE_U8)

    eng._plan = None
    cold = run()
    assert eng._plan is not None
    warm = run()                                   # served from the cached table
    assert torch.equal(warm.xyz, cold.xyz) and torch.equal(warm.rgb, cold.rgb) and torch.equal(warm.P, cold.P)
    sc.c2w[:, :3, 3] += 0.25                       # move every camera, in place
    moved = run()
    eng._plan = None
    moved_cold = run()
    assert not torch.equal(moved.xyz, cold.xyz)
    assert torch.equal(moved.xyz, moved_cold.xyz) and torch.equal(moved.P, moved_cold.P)
    assert torch.equal(moved.view_counts, moved_cold.view_counts)
