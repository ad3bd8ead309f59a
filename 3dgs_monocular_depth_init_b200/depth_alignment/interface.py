"""Reference: gs_init_compare/depth_alignment/interface.py:14-40."""
import abc
from pathlib import Path
from typing import NamedTuple, Optional

import torch

from ..depth_prediction.predictors.depth_predictor_interface import PredictedDepth


class DepthAlignmentResult(NamedTuple):
    aligned_depth: torch.Tensor
    mask: torch.Tensor


class DepthAlignmentStrategy(abc.ABC):
    @classmethod
    @abc.abstractmethod
    def align(cls, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, config, debug_export_dir: Optional[Path]) -> DepthAlignmentResult:
        """predicted_depth.depth [H,W]; sfm_points_camera_coords int [2,M'] (row 0 = x, row 1 = y);
        sfm_points_depth float32 [M']."""
