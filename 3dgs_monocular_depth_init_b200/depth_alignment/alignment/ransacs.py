"""RANSAC / MSAC scale-shift alignment on the B200 (reference: depth_alignment/alignment/ransacs.py).

The reference draws torch.randperm(N)[:4] from the global CPU generator once per iteration; here
the hypotheses are scored in parallel and sampled by a counter-based device generator seeded
from that same global generator (so torch.manual_seed still makes a run reproducible).  Parity
runs pass explicit sample indices through the engine (RunOptions.sample_idx)."""
from pathlib import Path
from typing import Optional

import torch

from ..._single import options_from_config, ransac_seed
from ...depth_prediction.predictors.depth_predictor_interface import PredictedDepth
from ..interface import DepthAlignmentResult, DepthAlignmentStrategy
from .lstsqrs import aligned_depth, fit_view

LAST_STATS = {}  # iterations / inliers of the last call (the reference prints them, ransacs.py:182)


def _align(kind, predicted_depth, coords, gt, config, sample_idx=None):
    opt = options_from_config(config, aligner=kind)
    if sample_idx is not None:   # parity runs: explicit samples [max_iters, sample_size]
        opt.sample_idx = sample_idx.to(device=predicted_depth.depth.device, dtype=torch.int32).contiguous()[None]
    else:
        opt.ransac_seed = ransac_seed()
    res = fit_view(predicted_depth, coords, gt, opt)
    LAST_STATS.update(iterations=res.ransac_iters, inliers=res.ransac_inliers, scale_shift=res.scale_shift)
    return DepthAlignmentResult(aligned_depth=aligned_depth(predicted_depth.depth, res.scale_shift),
                                mask=predicted_depth.mask)


class DepthAlignmentRansac(DepthAlignmentStrategy):
    @classmethod
    def align(cls, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, config, debug_export_dir: Optional[Path] = None,
              *args, **kwargs) -> DepthAlignmentResult:
        return _align("ransac", predicted_depth, sfm_points_camera_coords, sfm_points_depth, config,
                      kwargs.get("sample_idx"))


class DepthAlignmentMsac(DepthAlignmentStrategy):
    @classmethod
    def align(cls, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, config, debug_export_dir: Optional[Path] = None,
              *args, **kwargs) -> DepthAlignmentResult:
        return _align("msac", predicted_depth, sfm_points_camera_coords, sfm_points_depth, config,
                      kwargs.get("sample_idx"))
