"""Reference: gs_init_compare/depth_subsampling/static_subsampler.py:11-22."""
from dataclasses import dataclass

from ._run import subsample_mask
from .interface import DepthSubsampler


@dataclass
class StaticDepthSubsampler(DepthSubsampler):
    subsample_factor: int

    def get_mask(self, rgb, depth, mask):
        return subsample_mask(depth, mask, subsample=int(self.subsample_factor))
