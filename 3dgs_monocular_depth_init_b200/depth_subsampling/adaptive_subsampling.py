"""Reference: gs_init_compare/depth_subsampling/adaptive_subsampling.py:82-122."""
from dataclasses import dataclass

from ._run import subsample_mask
from .config import AdaptiveSubsamplingConfig
from .interface import DepthSubsampler


@dataclass
class AdaptiveDepthSubsampler(DepthSubsampler):
    config: AdaptiveSubsamplingConfig

    def get_mask(self, rgb, depth, mask):
        return subsample_mask(depth, mask, subsample="adaptive", factor_min=self.config.factor_range_min,
                              factor_max=self.config.factor_range_max)
