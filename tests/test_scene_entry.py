"""Scene-level entry on the GPU, called the way runner.py:77-78 calls it --
`pts_and_rgb_from_monocular_depth(config, parser, device)` imported from `gs_init_compare`, depth
read from a predicted-depth cache directory -- against the oracle's scene loop
(monocular_depth_init.py:120-224)."""
import importlib
import sys

import numpy as np
import pytest
import torch

import mdi_oracle as O
from helpers import ROOT, mod

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REL = 1e-5


def entry_module():
    """`gs_init_compare.monocular_depth_init` through the alias package at the repo root (falls back
    to the implementation module if this process already imported the real reference under that name)."""
    ref = sys.modules.get("gs_init_compare")
    if ref is not None and "/root/reference" in str(getattr(ref, "__file__", "")):
        return mod("monocular_depth_init")
    if str(ROOT) not in sys.path:
        sys.path.insert(0, str(ROOT))
    return importlib.import_module("gs_init_compare.monocular_depth_init")


def xyz_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.linalg.norm(a - b, axis=1)) / max(np.max(np.linalg.norm(b, axis=1)), 1e-30))


@pytest.fixture()
def scene(tmp_path):
    syn = mod("synthetic")
    dpi = mod("depth_prediction.predictors.depth_predictor_interface")
    dc = mod("depth_prediction.depth_cache")
    sc = syn.make_scene(n_views=5, H=120, W=160, n_sfm=2500, seed=11, low_conf_every=4)   # view 3 is skipped
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("lstsqrs")
    cfg.mdi.subsample_factor = "adaptive"
    cfg.mdi.cache_dir = str(tmp_path / "__mono_depth_cache__")
    for i, name in enumerate(sc.names):   # what the reference's predict_depth_or_get_cached_depth leaves behind
        p = dc.cache_path(cfg, sc.dataset_name, name)
        p.parent.mkdir(parents=True, exist_ok=True)
        torch.save(dpi.PredictedDepth(sc.depth[i].to(DEV), sc.mask[i].to(DEV)), p)
    return sc, syn.SyntheticParser(sc), cfg


def test_entry_as_runner_calls_it_matches_oracle_scene(scene):
    sc, parser, cfg = scene
    f = entry_module().pts_and_rgb_from_monocular_depth
    pts, rgbs, scales = f(cfg, parser, DEV)                      # runner.py:77-78
    pts_o, rgb_o, traces, kept = O.scene_points(sc, aligner="lstsqrs", subsample="adaptive", include_sfm_points=True)
    assert kept == [0, 1, 2, 4] and traces[3] is None           # the low-confidence view contributes nothing
    assert pts.is_cuda and not rgbs.is_cuda and scales is None   # pts on device, rgbs on CPU (:168,:181)
    assert pts.dtype == torch.float32 and rgbs.dtype == torch.float32
    assert pts.shape == pts_o.shape and rgbs.shape == rgb_o.shape
    n_sfm = sc.points.shape[0]
    assert torch.equal(pts[-n_sfm:].cpu(), torch.from_numpy(sc.points).float())   # include_sfm_points rows last
    assert xyz_err(pts.cpu().numpy(), pts_o.numpy()) <= REL
    assert np.array_equal(rgbs.numpy(), rgb_o.numpy())            # rgbs bit-exact
    cfg.mdi.include_sfm_points = False
    pts2, rgbs2, _ = f(cfg, parser, DEV)
    assert pts2.shape[0] == pts.shape[0] - n_sfm and torch.equal(pts2, pts[:-n_sfm]) and torch.equal(rgbs2, rgbs[:-n_sfm])


def test_entry_cache_miss_and_ignore_cache_raise(scene):
    sc, parser, cfg = scene
    dc = mod("depth_prediction.depth_cache")
    f = entry_module().pts_and_rgb_from_monocular_depth
    dc.cache_path(cfg, sc.dataset_name, sc.names[2]).unlink()
    with pytest.raises(dc.DepthCacheMiss, match=sc.names[2]):
        f(cfg, parser, DEV)
    cfg.mdi.ignore_cache = True
    with pytest.raises(dc.DepthCacheMiss, match="ignore_cache"):
        f(cfg, parser, DEV)


def test_entry_limit_init_scale_matches_sklearn_knn(scene):
    """monocular_depth_init.py:215-223 with the reference's knn (sklearn NearestNeighbors on the CPU,
    runner_utils.py:143-147) evaluated on the returned cloud."""
    from sklearn.neighbors import NearestNeighbors
    sc, parser, cfg = scene
    cfg.mdi.limit_init_scale = True
    cfg.init_scale = 0.7
    pts, rgbs, scales = entry_module().pts_and_rgb_from_monocular_depth(cfg, parser, DEV)
    x = pts.cpu().numpy()
    d, _ = NearestNeighbors(n_neighbors=4, metric="euclidean").fit(x).kneighbors(x)
    dist_avg = torch.sqrt((torch.from_numpy(d).to(pts)[:, 1:] ** 2).mean(dim=-1).cpu())
    dist_avg = torch.clamp(dist_avg, max=torch.quantile(dist_avg, cfg.mdi.init_scale_clamp_quantile))
    want = torch.log(dist_avg * cfg.init_scale).unsqueeze(-1).repeat(1, 3)
    assert scales.shape == (pts.shape[0], 3) and scales.is_cuda
    assert torch.allclose(scales.cpu(), want, rtol=1e-5, atol=1e-6)


def test_entry_postprocess_subsample_hands_compacted_views_to_the_merge(scene):
    """postprocess.subsample=True: K / P / image_sizes rows exist for the kept views only (the
    skipped view has none) and the merge equals the C restatement run on the same cloud."""
    import native_oracle as NO
    sc, parser, cfg = scene
    f = entry_module().pts_and_rgb_from_monocular_depth
    base_p, base_c, _ = f(cfg, parser, DEV)
    cfg.mdi.postprocess.subsample = True
    got_p, got_c, _ = f(cfg, parser, DEV)
    keep = [0, 1, 2, 4]
    P = torch.stack([O.projection_matrix(sc.K[i], sc.c2w[i]) for i in keep]).numpy()
    sizes = np.array([[160, 120]] * len(keep), dtype=np.int32)   # (W, H) rows (monocular_depth_init.py:175)
    want_p, want_c, _ = NO.subsample_pointcloud(base_p.cpu().numpy(), base_c.numpy(), sc.K[keep].numpy(), P, sizes)
    assert 0 < got_p.shape[0] < base_p.shape[0]
    assert np.array_equal(got_p.cpu().numpy(), want_p) and np.array_equal(got_c.cpu().numpy(), want_c)


def test_entry_lof_postprocess_matches_sklearn(scene):
    from sklearn.neighbors import LocalOutlierFactor
    sc, parser, cfg = scene
    f = entry_module().pts_and_rgb_from_monocular_depth
    base_p, base_c, _ = f(cfg, parser, DEV)
    cfg.mdi.postprocess.outlier_removal = mod("point_cloud_postprocess.config").OutlierRemovalMethod.lof
    got_p, got_c, _ = f(cfg, parser, DEV)
    out = LocalOutlierFactor(n_neighbors=cfg.mdi.postprocess.lof_num_neighbors).fit_predict(base_p.cpu().numpy()) == -1
    assert torch.equal(got_p.cpu(), base_p.cpu()[~out]) and torch.equal(got_c, base_c[~out])


def test_entry_noise_pts_only_and_ply_export(scene, tmp_path):
    sc, parser, cfg = scene
    pe = mod("utils.point_cloud_export")
    f = entry_module().pts_and_rgb_from_monocular_depth
    base_p, base_c, _ = f(cfg, parser, DEV)
    cfg.mdi.noise_std_scene_frac = 0.01                          # monocular_depth_init.py:163-166
    torch.manual_seed(3)
    noisy, _, _ = f(cfg, parser, DEV)
    n_sfm = sc.points.shape[0]
    delta = (noisy - base_p)[:-n_sfm]
    assert torch.equal(noisy[-n_sfm:], base_p[-n_sfm:])          # SfM rows are appended after the noise
    assert 0.5 < float(delta.std()) / (parser.scene_scale * 0.01) < 1.5
    cfg.mdi.noise_std_scene_frac = None
    cfg.mdi.pts_output_dir = str(tmp_path / "out")
    cfg.mdi.pts_output_per_image = True
    cfg.mdi.pts_only = True
    with pytest.raises(SystemExit) as e:                          # :212-213
        f(cfg, parser, DEV)
    assert e.value.code == 0
    d = tmp_path / "out" / sc.dataset_name
    files = sorted(p.name for p in d.iterdir())
    assert files == ["Metric3d_vitl_adaptive_None_lstsqrs.ply", "sfm.ply"], files   # :201
    cloud = d / files[0]
    p_xyz, p_rgb = pe.read_ply(cloud)
    assert np.array_equal(p_xyz, base_p.cpu().numpy().astype(np.float64))
    assert np.array_equal(p_rgb, np.round(np.clip(base_c.numpy().astype(np.float64), 0, 1) * 255).astype(np.uint8))
    s_xyz, _ = pe.read_ply(d / "sfm.ply")
    assert np.array_equal(s_xyz, sc.points.astype(np.float64))
    per = tmp_path / "out" / sc.names[0]
    assert sorted(p.name for p in per.iterdir()) == ["camera_plane.ply", "my_points_world.ply",
                                                     "sfm_points_plane_proj.ply", "sfm_points_world.ply"]
    assert not (tmp_path / "out" / sc.names[3]).exists()          # skipped view: no export
    mine, _ = pe.read_ply(per / "my_points_world.ply")
    assert mine.shape[0] > 0 and np.array_equal(mine, base_p.cpu().numpy()[: mine.shape[0]].astype(np.float64))


def test_entry_default_ransac_recovers_the_generating_fit(scene):
    """Default aligner (RANSAC, device-side sampler): no sample indices can be passed through the
    reference signature, so the bar is the generating scale/shift of the synthetic views."""
    sc, parser, cfg = scene
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
    torch.manual_seed(5)
    pts, rgbs, _ = entry_module().pts_and_rgb_from_monocular_depth(cfg, parser, DEV)
    torch.manual_seed(5)
    pts_o, rgb_o, traces, kept = O.scene_points(sc, aligner="ransac", subsample="adaptive", include_sfm_points=True)
    assert kept == [0, 1, 2, 4]
    assert abs(pts.shape[0] - pts_o.shape[0]) <= 0.02 * pts_o.shape[0]
