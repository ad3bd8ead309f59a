"""Shared test helpers: golden-case loading and scene regeneration."""
import importlib
import json
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"
PKG = "3dgs_monocular_depth_init_b200"


def mod(name=""):
    return importlib.import_module(PKG + ("." + name if name else ""))


SMALL_CASES = ["small_lstsq_static4", "small_ransac_adaptive", "small_msac_static3_grad", "small_lowconf",
               "small_ransac_nopatch", "small_adaptive_range_3_9", "tall_lstsq_adaptive", "tall_msac_static5"]
FULL_CASES = ["c1_lstsq_static10", "c2_ransac_adaptive"]
# one view of each larger BASELINE shape (C3 1920x1080, C4 3840x2160, C5 1600x1200), RANSAC + adaptive
BIG_CASES = ["c3_view_1920x1080", "c4_view_3840x2160", "c5_view_1600x1200"]
# 8f-3 / 8f-4: interp aligner and region-wise alignment (the fixtures also hold the aligned depth map)
MATERIALISED_CASES = ["small_interp_ransac_init", "small_interp_lstsq_init", "small_regions_ransac", "small_regions_lstsq"]


def case_segmentation(c):
    """Region map of the region-wise fixtures (oracle/make_golden.py::test_segmentation), or None."""
    if not c.meta.get("segmented"):
        return None
    H, W = c.depth.shape
    seg = torch.zeros((H, W), dtype=torch.int64)
    seg[:, W // 2:] = 1
    seg[H // 2:, : W // 3] = 2
    return seg

_scene_cache = {}


def load_case(name):
    """Returns a namespace: meta dict, inputs (depth, mask, image u8, K, c2w, sfm) as CPU tensors,
    reference outputs as numpy arrays."""
    z = np.load(GOLDEN / f"{name}.npz")
    meta = json.loads(bytes(z["meta"]).decode())
    c = SimpleNamespace(name=name, meta=meta, raised=int(z["raised"]))
    if "in_depth" in z:
        c.depth = torch.from_numpy(z["in_depth"])
        c.mask = torch.from_numpy(z["in_mask"])
        c.image = torch.from_numpy(z["in_image"])
        c.K = torch.from_numpy(z["in_K"])
        c.c2w = torch.from_numpy(z["in_c2w"])
        c.sfm = torch.from_numpy(z["in_sfm"])
    else:
        key = json.dumps(meta["gen"], sort_keys=True)
        if key not in _scene_cache:
            _scene_cache[key] = mod("synthetic").make_scene(**meta["gen"])
        sc = _scene_cache[key]
        i = meta["view"]
        chk = np.array([float(sc.depth[i].double().sum()), float(sc.mask[i].sum()),
                        float(sc.image[i].double().sum()), float(sc.points.astype(np.float64).sum())])
        assert np.array_equal(chk, z["in_checksum"]), "synthetic generator drifted from the golden fixture"
        c.depth, c.mask, c.image, c.K, c.c2w = sc.depth[i], sc.mask[i], sc.image[i], sc.K[i], sc.c2w[i]
        c.sfm = torch.from_numpy(sc.points[sc.point_indices[sc.names[i]]])
    for k in z.files:
        if not k.startswith("in_") and k not in ("meta", "raised"):
            setattr(c, "ref_" + k, z[k])
    if not c.raised:
        H, W = c.depth.shape
        c.ref_keep = np.unpackbits(c.ref_keep)[: H * W].astype(bool)
        if hasattr(c, "ref_aligned_mask"):
            c.ref_aligned_mask = np.unpackbits(c.ref_aligned_mask)[: H * W].astype(bool).reshape(H, W)
        if hasattr(c, "ref_ransac_inlier_mask"):  # final inlier set of the reference loop (ransacs.py:150-151)
            c.ref_ransac_inlier_mask = np.unpackbits(c.ref_ransac_inlier_mask)[: c.ref_uv.shape[1]].astype(bool)
    return c


def rgb_matches(c, rgb) -> bool:
    """Bit-exact comparison of a gathered rgb array [Nk,3] float32 with the reference's: against the
    stored array, or (large views) against the SHA-256 of the reference's bytes."""
    import hashlib
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    if hasattr(c, "ref_rgb"):
        return np.array_equal(rgb, c.ref_rgb)
    return rgb.shape[0] == int(c.ref_rgb_rows) and \
        hashlib.sha256(rgb.tobytes()).digest() == bytes(c.ref_rgb_sha256.tobytes())


def case_config(c):
    """Config (this package's mirror) for a golden case."""
    cfgm = mod("config")
    dac = mod("depth_alignment.config")
    m = c.meta
    cfg = cfgm.Config()
    cfg.mdi.alignment.aligner = dac.DepthAlignmentStrategyEnum(m["aligner"])
    cfg.mdi.subsample_factor = m["subsample"]
    cfg.mdi.depth_grad_mask_thresh = m.get("grad")
    cfg.mdi.use_num_sfm_points_mask = m.get("use_patch", True)
    cfg.mdi.adaptive_subsampling.factor_range_min = m.get("fmin", 5)
    cfg.mdi.adaptive_subsampling.factor_range_max = m.get("fmax", 15)
    cfg.mdi.num_sfm_points_mask.num_patches_small_axis = m.get("n_small", 20)
    cfg.mdi.num_sfm_points_mask.threshold = m.get("thr", 15)
    cfg.mdi.alignment.interp.init = m.get("init", "ransac")
    return cfg


def oracle_kwargs(c):
    m = c.meta
    return dict(aligner=m["aligner"], subsample=m["subsample"], use_patch_mask=m.get("use_patch", True),
                n_small=m.get("n_small", 20), patch_threshold=m.get("thr", 15), grad_thresh=m.get("grad"),
                fmin=m.get("fmin", 5), fmax=m.get("fmax", 15), interp_init=m.get("init", "ransac"),
                segmentation=case_segmentation(c))


class OneViewParser:
    def __init__(self, c):
        self.points = c.sfm.numpy()
        self.point_indices = {c.name: np.arange(len(self.points))}
