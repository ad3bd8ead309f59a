"""Reference: gs_init_compare/point_cloud_postprocess/config.py:8-18 and
native_modules/subsampling/.../subsampling_params.py:4-17 (same fields and defaults)."""
from dataclasses import dataclass, field
from enum import Enum


@dataclass
class PointCloudSubsamplingParams:
    max_bbox_aspect_ratio: float = 1.1
    min_extent_multiplier: float = 1.0


class OutlierRemovalMethod(str, Enum):
    off = "none"
    lof = "lof"


@dataclass
class PointCloudPostprocessConfig:
    outlier_removal: OutlierRemovalMethod = OutlierRemovalMethod.off
    lof_num_neighbors: int = 40
    subsample: bool = False
    subsample_params: PointCloudSubsamplingParams = field(default_factory=PointCloudSubsamplingParams)
