"""Config surface of the path (reference: gs_init_compare/config.py:20-66 for `mdi.*`).

`Config` carries only what this path reads (`mdi`, `init_scale`); the reference's own Config
object is accepted everywhere by duck typing, so runner.py can pass its instance unchanged."""
from dataclasses import dataclass, field
from typing import Literal, Optional, Union

from .depth_alignment.config import DepthAlignmentConfig
from .depth_prediction.configs import DepthAnythingV2Config, Metric3dV2Config, MogeConfig, UnidepthConfig
from .depth_subsampling.config import AdaptiveSubsamplingConfig, NumSfMPointsMaskConfig
from .point_cloud_postprocess.config import PointCloudPostprocessConfig


@dataclass
class MonocularDepthInitConfig:
    predictor: Optional[Literal["metric3d", "moge", "unidepth", "depth_anything_v2", "depth_pro"]] = "metric3d"
    # predictor sub-configs: only their `name` matters here (component of the depth-cache path)
    metric3d: Metric3dV2Config = field(default_factory=Metric3dV2Config)
    unidepth: UnidepthConfig = field(default_factory=UnidepthConfig)
    depthanything: DepthAnythingV2Config = field(default_factory=DepthAnythingV2Config)
    moge: MogeConfig = field(default_factory=MogeConfig)
    alignment: DepthAlignmentConfig = field(default_factory=DepthAlignmentConfig)
    depth_grad_mask_thresh: Optional[float] = None
    include_sfm_points: bool = True
    subsample_factor: Union[int, Literal["adaptive"]] = 10
    adaptive_subsampling: AdaptiveSubsamplingConfig = field(default_factory=AdaptiveSubsamplingConfig)
    use_num_sfm_points_mask: bool = True
    num_sfm_points_mask: NumSfMPointsMaskConfig = field(default_factory=NumSfMPointsMaskConfig)
    postprocess: PointCloudPostprocessConfig = field(default_factory=PointCloudPostprocessConfig)
    limit_init_scale: bool = False
    init_scale_clamp_quantile: float = 0.75
    pts_output_dir: Optional[str] = None
    pts_output_per_image: bool = False
    pts_only: bool = False
    noise_std_scene_frac: Optional[float] = None
    ignore_cache: bool = False
    cache_dir: str = "__mono_depth_cache__"


@dataclass
class Config:
    init_type: str = "monocular_depth"
    init_scale: float = 1.0
    mdi: MonocularDepthInitConfig = field(default_factory=MonocularDepthInitConfig)
