"""CPU: the oracle restatement against the golden vectors produced by the unmodified reference
(bit-exact), and against the reference itself when /root/reference is present."""
import json

import numpy as np
import pytest
import torch

import mdi_oracle as O
import ref_import
from helpers import (BIG_CASES, FULL_CASES, GOLDEN, MATERIALISED_CASES, SMALL_CASES, case_segmentation, load_case,
                     oracle_kwargs, rgb_matches)


@pytest.mark.parametrize("name", SMALL_CASES + FULL_CASES + BIG_CASES + MATERIALISED_CASES)
def test_oracle_matches_golden(name):
    c = load_case(name)
    torch.manual_seed(c.meta["seed"])
    if c.raised:
        with pytest.raises(O.LowConfidence):
            O.points_from_depth(c.depth, c.mask, c.K, c.c2w, c.sfm, **oracle_kwargs(c))
        return
    nt = torch.get_num_threads()
    torch.set_num_threads(c.meta.get("torch_threads", nt))  # large views: see make_golden.py (MKL sgemm kernel choice)
    try:
        tr = O.points_from_depth(c.depth, c.mask, c.K, c.c2w, c.sfm, **oracle_kwargs(c))
    finally:
        torch.set_num_threads(nt)
    assert np.array_equal(tr.P.numpy(), c.ref_P)
    assert np.array_equal(tr.uv.numpy(), c.ref_uv) and np.array_equal(tr.z.numpy(), c.ref_z)
    if hasattr(c, "ref_scale"):
        assert np.float32(tr.scale) == c.ref_scale and np.float32(tr.shift) == c.ref_shift
    assert np.array_equal(tr.keep.numpy(), c.ref_keep)
    assert np.array_equal(tr.pts_world.numpy(), c.ref_pts)
    rgb = (c.image.float() / 255.0).view(-1, 3)[tr.keep].numpy()
    assert rgb_matches(c, rgb)
    if tr.ransac is not None:
        assert tr.ransac.last_iter == int(c.ref_ransac_iters)
        assert tr.ransac.n_inliers == int(c.ref_ransac_inliers)
        assert np.array_equal(tr.ransac.inliers.numpy(), c.ref_ransac_inlier_mask)
    if tr.lo_hi is not None and hasattr(c, "ref_lo"):
        assert np.float32(tr.lo_hi[0]) == c.ref_lo and np.float32(tr.lo_hi[1]) == c.ref_hi


def test_sample_idx_replays_rng():
    """Feeding draw_sample_idx() reproduces the run that consumed the global RNG."""
    c = load_case("small_ransac_adaptive")
    d = c.depth[c.ref_uv[1], c.ref_uv[0]]
    z = torch.from_numpy(c.ref_z)
    torch.manual_seed(c.meta["seed"])
    a = O.ransac_fit(d, z)
    idx = O.draw_sample_idx(len(z), 2500, 4, c.meta["seed"])
    b = O.ransac_fit(d, z, sample_idx=idx)
    assert a.last_iter == b.last_iter and float(a.scale) == float(b.scale) and float(a.shift) == float(b.shift)
    assert torch.equal(a.inliers, b.inliers)


def test_patch_mask_histogram_equals_loop():
    c = load_case("small_lstsq_static4")
    uv = torch.from_numpy(c.ref_uv)
    for n_small, thr in [(20, 15), (10, 3), (7, 0)]:
        a = O.sfm_density_mask(uv, tuple(c.depth.shape), n_small, thr)
        b = O.sfm_density_mask_loop(uv, tuple(c.depth.shape), n_small, thr)
        assert torch.equal(a, b)


def test_known_answers():
    ka = json.loads((GOLDEN / "known_answers.json").read_text())
    for n, tot, want in ka["required_samples"]:
        assert float(O.required_samples(torch.tensor(n), tot, 4, 0.999)) == want
    for h, w, n, ph, pw, gh, gw in ka["patch_sizes"]:
        assert O.patch_sizes((h, w), n) == ((ph, pw), (gh, gw))
    for d, g, s, t in ka["lstsq"]:
        dd = torch.tensor(d, dtype=torch.float32)
        hs = O.lstsq_fit(O.stack_ones(dd), torch.tensor(g, dtype=torch.float32))
        assert float(hs[0]) == s and float(hs[1]) == t


@pytest.mark.skipif(not ref_import.available(), reason="reference tree not present (GPU box)")
def test_oracle_vs_reference_live():
    import contextlib, io
    R = ref_import.load()
    c = load_case("tall_msac_static5")
    cfg = R.Config()
    cfg.mdi.alignment.aligner = R.DepthAlignmentStrategyEnum("ransac")
    cfg.mdi.subsample_factor = "adaptive"

    class P:
        points = c.sfm.numpy()
        point_indices = {"v": np.arange(len(c.sfm))}

    torch.manual_seed(99)
    with contextlib.redirect_stdout(io.StringIO()):
        pts, keep, Pm = R.points_from_depth.get_pts_from_depth(
            R.PredictedDepth(c.depth.clone(), c.mask.clone()),
            R.InputImage(data=c.image.float() / 255.0, name="v", cam2world=c.c2w, K=c.K), P, cfg, "cpu")
    torch.manual_seed(99)
    tr = O.points_from_depth(c.depth, c.mask, c.K, c.c2w, c.sfm, aligner="ransac", subsample="adaptive")
    assert torch.equal(keep, tr.keep) and torch.equal(pts, tr.pts_world) and torch.equal(Pm, tr.P)


@pytest.mark.parametrize("name", MATERIALISED_CASES)
def test_oracle_aligned_map_matches_golden(name):
    """interp.py:281-361 / pipeline.py:196-288: the aligned depth MAP of the reference, bit for bit."""
    c = load_case(name)
    uv, z = torch.from_numpy(c.ref_uv), torch.from_numpy(c.ref_z)
    torch.manual_seed(c.meta["seed"])
    seg = case_segmentation(c)
    if seg is not None:
        out, mask, _, fits = O.align_regions(c.depth, c.mask, uv, z, seg, c.meta["aligner"])
        assert len(fits) == 3
    else:
        out, removed, _ = O.interp_align(c.depth, uv, z, init=c.meta.get("init", "ransac"))
        mask = (out != -42.0) & c.mask
        assert 0 < int(removed.sum()) < 60        # ~1 % scale-only outliers are dropped
    assert np.array_equal(out.numpy(), c.ref_aligned) and np.array_equal(mask.numpy(), c.ref_aligned_mask)
