"""Reference: gs_init_compare/depth_subsampling/config.py:4-26 (same fields and defaults)."""
from dataclasses import dataclass


@dataclass
class AdaptiveSubsamplingConfig:
    factor_range_min: int = 5
    factor_range_max: int = 15


@dataclass
class NumSfMPointsMaskConfig:
    num_patches_small_axis: int = 20
    threshold: int = 15
