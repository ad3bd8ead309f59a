"""Per-view geometry on the B200 (reference: depth_prediction/points_from_depth.py:111-329):
SfM projection + filter, alignment, mask composition, order-preserving compaction and
unprojection run as one batched kernel sequence (engine.Engine.run) with V = 1."""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import numpy as np
import torch

from .. import _lib as L
from .._single import options_from_config, ransac_seed, single_view_inputs
from ..depth_alignment.exceptions import LowDepthAlignmentConfidenceError
from ..depth_subsampling.adaptive_subsampling import AdaptiveDepthSubsampler
from ..depth_subsampling.static_subsampler import StaticDepthSubsampler
from ..engine import RunOptions, get_engine
from ..types import InputImage
from .predictors.depth_predictor_interface import PredictedDepth


def get_subsampler(cfg):
    """points_from_depth.py:183-189."""
    if cfg.mdi.subsample_factor == "adaptive":
        return AdaptiveDepthSubsampler(cfg.mdi.adaptive_subsampling)
    elif isinstance(cfg.mdi.subsample_factor, int):
        return StaticDepthSubsampler(cfg.mdi.subsample_factor)
    raise ValueError(f"Unsupported subsampling factor: {cfg.mdi.subsample_factor}")


def _raise_low_confidence(n_inb: int, m: int):
    raise LowDepthAlignmentConfidenceError(
        "Less than 1/4 of SFM points", f" ({n_inb} / {m})", " reprojected into image bounds.")


def project_and_filter_sfm_pts(image, sfm_points: torch.Tensor, P: torch.Tensor, imsize,
                               predicted_depth: PredictedDepth, debug_export_dir=None):
    """points_from_depth.py:145-180: returns (coords int64 [2,M'], depths float32 [M']).
    imsize = (W, H).  P is used as given (no camera recomputation)."""
    dev = sfm_points.device
    eng = get_engine(dev)
    M = int(sfm_points.shape[0])
    vin = single_view_inputs(predicted_depth.depth, predicted_depth.mask, dev,
                             sfm_xyz=sfm_points.float().contiguous(), sfm_counts=[M],
                             P_override=P.detach().cpu().float()[None])
    res = eng.run(vin, RunOptions(stop_after="project", use_patch_mask=False))
    n_inb, n_valid, status = (int(x) for x in torch.stack([res.n_inbounds[0], res.n_valid[0], res.status[0]]).tolist())
    if status & L.STATUS_LOW_CONFIDENCE:
        _raise_low_confidence(n_inb, M)
    return res.uv[:, :n_valid].long(), res.z[:n_valid].clone()


def depth_gradient_mask(depth: torch.Tensor, gradient_threshold: float) -> torch.Tensor:
    """points_from_depth.py:192-212."""
    dev = depth.device
    vin = single_view_inputs(depth, None, dev)
    opt = RunOptions(aligner=None, use_patch_mask=False, flags=0, subsample=1, grad_thresh=float(gradient_threshold),
                     want_xyz=False, want_rgb=False, stop_after="mask")
    res = get_engine(dev).run(vin, opt)
    return res.keep.view(depth.shape).bool()


def get_pts_from_depth(predicted_depth: PredictedDepth, image: InputImage, parser, config, device: str,
                       debug_export_dir: Optional[Path] = None, *, sample_idx: Optional[torch.Tensor] = None,
                       return_result: bool = False):
    """Returns (pts_world float32 [N,3] on device, keep mask bool [H*W] on CPU, P float32 [3,4] on
    device) -- points_from_depth.py:215-329.  Raises LowDepthAlignmentConfidenceError like the
    reference.  `sample_idx` ([max_iters, sample] int) feeds explicit RANSAC samples (parity)."""
    eng = get_engine(device)
    dev = eng.device
    sfm = torch.from_numpy(np.ascontiguousarray(parser.points[parser.point_indices[image.name]])).to(dev).float()
    M = int(sfm.shape[0])
    opt = options_from_config(config, want_rgb=False)
    if opt.aligner in ("ransac", "msac"):
        if sample_idx is not None and config.mdi.alignment.segmenter is None:
            opt.sample_idx = sample_idx.to(device=dev, dtype=torch.int32).contiguous()[None]
        elif sample_idx is None:
            opt.ransac_seed = ransac_seed()
        opt.want_inliers = bool(return_result)   # RunResult.inlier_mask: the final inlier set (ransacs.py:150-151)
    vin = single_view_inputs(predicted_depth.depth, predicted_depth.mask, dev, K=image.K, c2w=image.cam2world,
                             sfm_xyz=sfm.contiguous(), sfm_counts=[M])
    if config.mdi.alignment.segmenter is not None:
        from ..depth_alignment.pipeline import run_views_materialised
        res = run_views_materialised(eng, vin, config, opt, None if sample_idx is None else [sample_idx])
    elif opt.aligner == "interp":
        from ..depth_alignment.alignment.interp import run_views_with_interp
        res = run_views_with_interp(eng, vin, config, opt, None if sample_idx is None else sample_idx[None])
    else:
        res = eng.run(vin, opt)
    status, n_inb = int(res.status[0].item()), int(res.n_inbounds[0].item())
    if status & L.STATUS_LOW_CONFIDENCE:
        _raise_low_confidence(n_inb, M)
    if status & L.STATUS_EMPTY_STATS:
        raise RuntimeError("quantile() input tensor must be non-empty")
    out = (res.xyz, res.keep.bool().cpu(), res.P[0])
    return out + (res,) if return_result else out
