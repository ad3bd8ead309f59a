"""Output contract of the depth predictors (reference:
depth_prediction/predictors/depth_predictor_interface.py:8-71).  The predictors themselves are
out of scope; synthetic or cached depth is fed in."""
from abc import ABCMeta, abstractmethod
from dataclasses import dataclass
from typing import NamedTuple, Optional

import torch


@dataclass
class PredictedDepth:
    depth: torch.Tensor
    mask: torch.Tensor
    depth_confidence: Optional[torch.Tensor] = None
    normal: Optional[torch.Tensor] = None
    normal_confidence: Optional[torch.Tensor] = None


# torch.load(weights_only=True) of a cached PredictedDepth (reference :42-43).  Cache files written
# by the reference name the class by ITS module path; both spellings resolve to this dataclass.
REFERENCE_QUALNAME = "gs_init_compare.depth_prediction.predictors.depth_predictor_interface.PredictedDepth"
torch.serialization.add_safe_globals([PredictedDepth, (PredictedDepth, REFERENCE_QUALNAME)])


class CameraIntrinsics(NamedTuple):
    K: torch.Tensor

    @property
    def fx(self):
        return self.K[0, 0].item()

    @property
    def fy(self):
        return self.K[1, 1].item()

    @property
    def cx(self):
        return self.K[0, 2].item()

    @property
    def cy(self):
        return self.K[1, 2].item()


class DepthPredictor(metaclass=ABCMeta):
    @abstractmethod
    def __init__(self, config, device):
        pass

    @property
    @abstractmethod
    def name(self) -> str:
        """Name of the predictor."""

    def predict_depth(self, img: torch.Tensor, intrinsics: CameraIntrinsics) -> PredictedDepth:
        raise NotImplementedError
