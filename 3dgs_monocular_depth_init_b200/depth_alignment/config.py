"""Alignment configuration: same field names, types and defaults as the reference
(gs_init_compare/depth_alignment/config.py:6-142), with default_factory so the dataclasses
import on Python >= 3.11.  The SAM / SLIC segmenters themselves are outside this path (SURVEY
section 8f): their enum members are kept so CLI strings still parse, and region-wise alignment takes
a caller-supplied integer segmentation instead (depth_alignment/pipeline.py)."""
from dataclasses import dataclass, field
from enum import Enum
from typing import Literal, Optional


class DepthAlignmentStrategyEnum(str, Enum):
    lstsqrs = "lstsqrs"
    ransac = "ransac"
    msac = "msac"
    interp = "interp"

    def get_implementation(self):
        if self == self.lstsqrs:
            from .alignment.lstsqrs import DepthAlignmentLstSqrs

            return DepthAlignmentLstSqrs
        if self == self.ransac:
            from .alignment.ransacs import DepthAlignmentRansac

            return DepthAlignmentRansac
        if self == self.msac:
            from .alignment.ransacs import DepthAlignmentMsac

            return DepthAlignmentMsac
        if self == self.interp:
            from .alignment.interp import DepthAlignmentInterpolate

            return DepthAlignmentInterpolate
        raise NotImplementedError(f"Unknown depth alignment strategy: {self}")


class DepthSegmentationStrategyEnum(str, Enum):
    slic = "slic"
    sam = "sam"

    def get_implementation(self):
        raise NotImplementedError(f"Depth segmentation is out of scope of the B200 path: {self}")


@dataclass
class SAMSegmentationconfig:
    use_normals: bool = True
    degenerate_mask_thresh: float = 0.9
    expansion_radius: int = 4
    tiny_region_area_fraction: float = 1e-4


@dataclass
class SLICSegmentationConfig:
    compactness = 0.01
    num_regions = 40


@dataclass
class DepthSegmentationConfig:
    region_margin: int = 10
    propagate_mask: bool = False
    min_border_grad_threshold: float = 0.0005
    min_sfm_pts_in_region: int = 5
    sam: SAMSegmentationconfig = field(default_factory=SAMSegmentationconfig)
    slic: SLICSegmentationConfig = field(default_factory=SLICSegmentationConfig)


@dataclass
class RansacConfig:
    inlier_threshold: float = 0.01
    max_iters: int = 2500
    confidence: float = 0.999
    sample_size: int = 4
    min_iters: int = 0


@dataclass
class InterpConfig:
    method: Literal["rbf", "linear"] = "linear"
    init: Optional[Literal["lstsqrs", "ransac"]] = "ransac"
    scale_outlier_removal: bool = True
    smoothing: float = 0.001
    kernel: str = "thin_plate_spline"
    max_rbf_points: int = 5000


@dataclass
class DepthAlignmentConfig:
    segmenter: Optional[DepthSegmentationStrategyEnum] = None
    aligner: DepthAlignmentStrategyEnum = DepthAlignmentStrategyEnum.ransac
    segmentation: DepthSegmentationConfig = field(default_factory=DepthSegmentationConfig)
    ransac: RansacConfig = field(default_factory=RansacConfig)
    interp: InterpConfig = field(default_factory=InterpConfig)
