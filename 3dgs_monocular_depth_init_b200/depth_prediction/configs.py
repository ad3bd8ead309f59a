"""Predictor configuration schema (reference: gs_init_compare/depth_prediction/configs.py:6-75).

The depth predictors themselves are outside the B200 path; these dataclasses are kept because
(a) callers address them by dotted path (`--mdi.metric3d.backbone vitg`) and (b) the predictor's
`name` -- which is built from them -- is a component of the predicted-depth cache path
(monocular_depth_init.py:68) that `depth_cache.py` reads."""
from dataclasses import dataclass
from enum import Enum
from typing import Literal, Optional


class DepthAnythingV2Backbone(str, Enum):
    vits = "vits"
    vitb = "vitb"
    vitl = "vitl"


@dataclass
class DepthAnythingV2Config:
    backbone: DepthAnythingV2Backbone = DepthAnythingV2Backbone.vitl
    metric: bool = True
    metric_model_type: Optional[Literal["indoor", "outdoor"]] = "indoor"


class Metric3dBackbone(str, Enum):
    vits = "vits"
    vitl = "vitl"
    vitg = "vitg"


@dataclass
class Metric3dV2Config:
    backbone: Metric3dBackbone = Metric3dBackbone.vitl


class MogeBackbone(str, Enum):
    vits = "vits"
    vitl = "vitl"
    vitg = "vitg"


@dataclass
class MogeConfig:
    backbone: MogeBackbone = MogeBackbone.vitl


class UnidepthBackbone(str, Enum):
    vits = "vits"
    vitb = "vitb"
    vitl = "vitl"


@dataclass
class UnidepthConfig:
    backbone: UnidepthBackbone = UnidepthBackbone.vitl
