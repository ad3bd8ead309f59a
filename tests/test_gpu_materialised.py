"""GPU parity of the alignments that are not one scale / shift per view: the
`interp` aligner (interp.py:75-110,161-201,204-232,281-361) and region-wise alignment
(pipeline.py:196-288), against fixtures written by the unmodified reference and against the oracle.

Bars: aligned depth maps 1e-5 relative (fp32), masks / keep masks bit-exact whenever the
scale-outlier set agrees with sklearn's (integer pixel coordinates make kNN distance ties common and
their order is implementation defined -- the disagreement is counted and bounded)."""
import numpy as np
import pytest
import torch

import mdi_oracle as O
from helpers import MATERIALISED_CASES, OneViewParser, case_config, case_segmentation, load_case, mod

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REL = 1e-5


def map_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-30))


def _pred(c):
    dpi = mod("depth_prediction.predictors.depth_predictor_interface")
    return dpi.PredictedDepth(c.depth.to(DEV), c.mask.to(DEV))


def test_interp_scale_map_kernel_equals_scipy():
    """mdi_interp_scale_map against scipy's LinearNDInterpolator (through the oracle), including
    duplicate points, points on the border and a view with very few points."""
    interp = mod("depth_alignment.alignment.interp")
    cfg = mod("depth_alignment.config").InterpConfig()
    g = torch.Generator().manual_seed(3)
    for H, W, n in [(120, 160, 2000), (97, 151, 50), (64, 48, 3), (300, 400, 20000)]:
        coords = torch.stack([torch.randint(0, W, (n,), generator=g), torch.randint(0, H, (n,), generator=g)])
        coords[:, : n // 10] = coords[:, n // 10: 2 * (n // 10)][:, : n // 10]   # duplicates
        values = (1.0 + 0.2 * torch.randn(n, generator=g)).float()
        want = O.interp_linear_map(coords, values, W, H)
        got = interp.linear_interpolation(coords.to(DEV), values.to(DEV), cfg, DEV, W, H)
        assert got.shape == (H, W) and got.dtype == torch.float32
        assert map_err(got.cpu().numpy(), want.numpy()) <= 1e-6, (H, W, n)
        base = (4 + torch.rand(H, W, generator=g)).float()
        ss = torch.tensor([1.7, -0.3])
        fused = interp.linear_interpolation(coords.to(DEV), values.to(DEV), cfg, DEV, W, H, base=base.to(DEV),
                                            scale_shift=ss.to(DEV))
        assert torch.equal(fused.cpu(), got.cpu() * (base * ss[0] + ss[1]))     # interp.py:361, separately rounded


@pytest.mark.parametrize("name", [n for n in MATERIALISED_CASES if "interp" in n])
def test_interp_aligner_matches_reference_golden(name):
    c = load_case(name)
    interp = mod("depth_alignment.alignment.interp")
    cfg = case_config(c)
    uv, z = torch.from_numpy(c.ref_uv), torch.from_numpy(c.ref_z)
    n = uv.shape[1]
    init = c.meta.get("init", "ransac")
    sample_idx = torch.from_numpy(O.draw_sample_idx(n, 2500, 4, c.meta["seed"])) if init == "ransac" else None
    # the reference's outlier set (sklearn, through the oracle) fed in: every pixel within 1e-5
    torch.manual_seed(c.meta["seed"])
    _, removed_ref, h_ref = O.interp_align(c.depth, uv, z, init=init)
    r = interp.DepthAlignmentInterpolate.align(_pred(c), uv.to(DEV), z.to(DEV), cfg, None, sample_idx=sample_idx,
                                               outlier_mask=removed_ref)
    assert map_err(r.aligned_depth.cpu().numpy(), c.ref_aligned) <= 2 * REL
    assert torch.equal(r.mask.cpu(), c.mask)
    # own outlier removal (mdi_knn / mdi_lof): the sets may differ by distance ties only
    d = c.depth[uv[1], uv[0]]
    factors = (z / (d * h_ref[0] + h_ref[1])).to(DEV)
    cls = interp.scale_factor_outlier_removal(uv.T.to(DEV), factors)
    ours = cls.scale_only_outliers.cpu()
    sc_o, both_o, pos_o, reg_o = O.interp_outlier_classes(uv.T, factors.cpu())
    n_diff = int((ours != sc_o).sum())
    assert n_diff <= 0.005 * n, f"{n_diff} of {n} scale-outlier flags differ from sklearn's"
    pos_diff = int(((cls.position_only_outliers.cpu() | cls.both_outliers.cpu()) != (pos_o | both_o)).sum())
    assert pos_diff <= 0.005 * n, f"{pos_diff} of {n} LOF position-outlier flags differ from sklearn's (k-distance ties)"
    r2 = interp.DepthAlignmentInterpolate.align(_pred(c), uv.to(DEV), z.to(DEV), cfg, None, sample_idx=sample_idx)
    a, b = r2.aligned_depth.cpu().numpy(), c.ref_aligned
    bad = np.abs(a - b) > 2 * REL * np.abs(b).max()
    assert bad.mean() <= (0.0 if n_diff == 0 else 0.03), f"{bad.mean():.4f} of the pixels differ ({n_diff} outlier ties)"


@pytest.mark.parametrize("name", MATERIALISED_CASES)
def test_view_through_get_pts_from_depth_matches_reference_golden(name):
    """The whole per-view path with aligner = interp / a registered segmenter: keep mask, count, xyz."""
    c = load_case(name)
    pfd = mod("depth_prediction.points_from_depth")
    types = mod("types")
    segm = mod("depth_alignment.segmentation")
    cfg = case_config(c)
    seg = case_segmentation(c)
    uv, z = torch.from_numpy(c.ref_uv), torch.from_numpy(c.ref_z)
    sample_idx = None
    if seg is not None:
        cfg.mdi.alignment.segmenter = mod("depth_alignment.config").DepthSegmentationStrategyEnum("slic")
        segm.register_segmenter("slic", lambda pd, ckpt, scfg: seg.to(pd.depth.device))
        if c.meta["aligner"] == "ransac":
            torch.manual_seed(c.meta["seed"])
            fits = O.align_regions(c.depth, c.mask, uv, z, seg, "ransac")[3]
            rows = []
            for r in sorted(fits):
                used = fits[r][2]
                pad = np.full((2500, 4), -1, dtype=np.int64)
                pad[: used.shape[0]] = used
                rows.append(torch.from_numpy(pad))
            sample_idx = torch.stack(rows)
    elif c.meta.get("init", "ransac") == "ransac":
        sample_idx = torch.from_numpy(O.draw_sample_idx(uv.shape[1], 2500, 4, c.meta["seed"]))
    try:
        image = types.InputImage(data=c.image.float() / 255.0, name=c.name, cam2world=c.c2w, K=c.K)
        pts, keep, P, res = pfd.get_pts_from_depth(_pred(c), image, OneViewParser(c), cfg, DEV, sample_idx=sample_idx,
                                                   return_result=True)
    finally:
        segm.register_segmenter("slic", None)
    aligned = res.extras["aligned_depth"][0].cpu().numpy()
    sent = c.ref_aligned == -42.0
    assert np.array_equal(aligned == -42.0, sent)                       # regions without a fit keep the sentinel
    flips = int((keep.numpy() != c.ref_keep).sum())
    if seg is not None:
        assert map_err(np.where(sent, 0, aligned), np.where(sent, 0, c.ref_aligned)) <= 2 * REL
        assert flips == 0 and pts.shape[0] == c.ref_pts.shape[0]
        err = np.max(np.linalg.norm(pts.cpu().numpy() - c.ref_pts, axis=1)) / np.max(np.linalg.norm(c.ref_pts, axis=1))
        assert err <= 2 * REL
    else:   # interp: outlier ties can move the scale map locally (see the aligner test)
        assert flips <= 0.01 * c.ref_keep.sum(), f"{flips} keep bits differ"
        if flips == 0:
            err = np.max(np.linalg.norm(pts.cpu().numpy() - c.ref_pts, axis=1)) / np.max(np.linalg.norm(c.ref_pts, axis=1))
            assert err <= 0.02


def test_region_pipeline_api_and_empty_regions():
    """DepthAlignmentPipeline.align with a segmenter: batched per-region lstsq fits equal the oracle;
    a region without SfM points stays at the sentinel and is masked out; propagate_mask narrows the
    predicted mask by the dead zone (pipeline.py:219-236,264-288)."""
    c = load_case("small_regions_lstsq")
    pipe_m = mod("depth_alignment.pipeline")
    segm = mod("depth_alignment.segmentation")
    cfg = case_config(c)
    cfg.mdi.alignment.segmenter = mod("depth_alignment.config").DepthSegmentationStrategyEnum("sam")
    uv, z = torch.from_numpy(c.ref_uv), torch.from_numpy(c.ref_z)
    H, W = c.depth.shape
    seg = case_segmentation(c).clone()
    seg[:6, :8] = 3                                   # a fourth region ...
    inside = (uv[0] < 8) & (uv[1] < 6)
    uv2, z2 = uv[:, ~inside], z[~inside]              # ... that holds no SfM point
    with pytest.raises(NotImplementedError, match="register_segmenter"):
        pipe_m.DepthAlignmentPipeline.from_config(cfg)
    segm.register_segmenter("sam", lambda pd, ckpt, scfg: seg.to(pd.depth.device))
    try:
        for propagate in (False, True):
            cfg.mdi.alignment.segmentation.propagate_mask = propagate
            pred = _pred(c)
            r = pipe_m.DepthAlignmentPipeline.from_config(cfg).align(None, pred, uv2.to(DEV), z2.to(DEV), cfg, None)
            want, wmask, wpred, fits = O.align_regions(c.depth, c.mask, uv2, z2, seg, "lstsqrs", propagate_mask=propagate)
            got = r.aligned_depth.cpu().numpy()
            assert sorted(fits) == [0, 1, 2]
            assert np.array_equal(got == -42.0, want.numpy() == -42.0) and (got[:6, :8] == -42.0).all()
            assert map_err(np.where(got == -42.0, 0, got), np.where(got == -42.0, 0, want.numpy())) <= 2 * REL
            assert torch.equal(r.mask.cpu(), wmask) and torch.equal(pred.mask.cpu(), wpred)
    finally:
        segm.register_segmenter("sam", None)


def test_batched_scene_with_interp_equals_per_view_runs():
    """points_from_views(aligner = interp): the batched init fit + per-view scale maps give the clouds of
    the per-view calls, in view order; a low-confidence view is skipped."""
    syn, mdi, L = mod("synthetic"), mod("monocular_depth_init"), mod("_lib")
    pfd, types = mod("depth_prediction.points_from_depth"), mod("types")
    dpi = mod("depth_prediction.predictors.depth_predictor_interface")
    sc = syn.make_scene(n_views=4, H=120, W=160, n_sfm=2500, seed=11, low_conf_every=4)
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("interp")
    cfg.mdi.alignment.interp.init = "lstsqrs"
    cfg.mdi.subsample_factor = "adaptive"
    res = mdi.points_from_views(cfg, sc.depth.to(DEV), sc.mask.to(DEV), sc.image.to(DEV), sc.K, sc.c2w, sc.names,
                                sc.points, sc.point_indices, DEV, L.IMAGE_U8)
    assert res.status.cpu().tolist() == [0, 0, 0, 1]
    parts = []
    parser = syn.SyntheticParser(sc)
    for i in range(3):
        image = types.InputImage(data=sc.image[i].float() / 255.0, name=sc.names[i], cam2world=sc.c2w[i], K=sc.K[i])
        pts, keep, P = pfd.get_pts_from_depth(dpi.PredictedDepth(sc.depth[i].to(DEV), sc.mask[i].to(DEV)), image, parser,
                                              cfg, DEV)
        parts.append(pts)
        assert torch.equal(res.keep.view(4, -1)[i].bool().cpu(), keep)
    assert torch.equal(res.xyz, torch.cat(parts))
    pts_o, rgb_o, traces, kept = O.scene_points(sc, aligner="interp", subsample="adaptive", include_sfm_points=False,
                                                interp_init="lstsqrs")
    assert kept == [0, 1, 2] and abs(res.xyz.shape[0] - pts_o.shape[0]) <= 0.01 * pts_o.shape[0]
