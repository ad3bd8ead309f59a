"""Alignment pipeline (reference: depth_alignment/pipeline.py:170-293, no-segmenter branch)."""
from dataclasses import dataclass
from pathlib import Path
from typing import Optional

import torch

from ..depth_prediction.predictors.depth_predictor_interface import PredictedDepth
from .interface import DepthAlignmentResult, DepthAlignmentStrategy

INVALID_DEPTH_VAL = -42.0


@dataclass
class DepthAlignmentPipeline:
    config: object
    segmentation: Optional[object]
    alignment: DepthAlignmentStrategy

    @staticmethod
    def from_config(config):
        if config.mdi.alignment.segmenter is not None:
            raise NotImplementedError("depth segmentation is outside the B200 path (SURVEY.md 8f)")
        return DepthAlignmentPipeline(config, None, config.mdi.alignment.aligner.get_implementation())

    def align(self, image, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, config, debug_export_dir: Optional[Path]):
        # one region holding every point (pipeline.py:248-251); the stitched map equals the
        # aligner's output and the sentinel test (pipeline.py:285-288) is applied to it
        r = self.alignment.align(predicted_depth, sfm_points_camera_coords, sfm_points_depth, config, debug_export_dir)
        return DepthAlignmentResult(aligned_depth=r.aligned_depth,
                                    mask=(r.aligned_depth != INVALID_DEPTH_VAL) & predicted_depth.mask)
