"""Where a view's predicted depth comes from (reference: monocular_depth_init.py:32-87).

The reference instantiates a depth predictor (`pick_model(config)(config, device)`) and wraps it in
a per-image disk cache: `<cache_dir>/<model.name>/<dataset_name>/<image_name>.pth`, a
`torch.save`d `PredictedDepth`.  The predictor networks are outside the B200 path, so the default
depth source here is the cache alone: a hit is loaded exactly like the reference does, a miss (or
`ignore_cache=True`) raises `DepthCacheMiss` naming the file -- unless a predictor has been plugged
in with `register_predictor`, in which case the reference's predict-then-save behaviour is
reproduced (monocular_depth_init.py:80-86)."""
from __future__ import annotations

import logging
from pathlib import Path
from typing import Callable, Dict, Optional

import torch

from .predictors.depth_predictor_interface import CameraIntrinsics, DepthPredictor, PredictedDepth

_LOGGER = logging.getLogger(__name__)

_PREDICTORS = ("metric3d", "depth_pro", "moge", "unidepth", "depth_anything_v2")
# predictor key (config.mdi.predictor) -> factory(config, device) -> DepthPredictor
_FACTORIES: Dict[str, Callable] = {}


class DepthCacheMiss(FileNotFoundError):
    """No cached depth for an image and no predictor registered to compute it."""


def register_predictor(key: str, factory: Optional[Callable]) -> None:
    """Plug a DepthPredictor implementation in for `config.mdi.predictor == key` (None removes it).
    `factory(config, device)` must return an object with `.name` and `.predict_depth(img, intrinsics)`
    (depth_predictor_interface.py:46-71)."""
    if factory is None:
        _FACTORIES.pop(key, None)
    else:
        _FACTORIES[key] = factory


def _value(x):
    # the reference formats `str, Enum` members with f-strings under Python 3.10, which yields the value
    return getattr(x, "value", x)


def model_name(config) -> str:
    """`model.name` of the predictor `pick_model(config)` would build, without building it:
    metric3d.py:26, moge.py:15, unidepth.py:38, depth_anything_v2.py:64,80, apple_depth_pro.py:54.
    Raises like pick_model (monocular_depth_init.py:33-57)."""
    mdi = config.mdi
    key = mdi.predictor
    if key is None:
        raise ValueError("No depth predictor model specified in config.")
    if key == "metric3d":
        return f"Metric3d_{_value(mdi.metric3d.backbone)}"
    if key == "depth_pro":
        return "AppleDepthPro"
    if key == "moge":
        return f"MoGe_{_value(mdi.moge.backbone)}"
    if key == "unidepth":
        return f"UniDepth_{_value(mdi.unidepth.backbone)}"
    if key == "depth_anything_v2":
        enc = _value(mdi.depthanything.backbone)
        if mdi.depthanything.metric:
            kind = mdi.depthanything.metric_model_type
            if kind not in ("indoor", "outdoor"):
                raise ValueError(f"Unsupported metric model type '{kind}' for DepthAnythingV2")
            return f"DepthAnythingV2_{enc}_metric_{kind}"
        return f"DepthAnythingV2_{enc}_relative"
    raise ValueError(f"Unsupported monodepth model: {key}")


def cache_path(config, dataset_name: str, image_name: str, name: Optional[str] = None) -> Path:
    """monocular_depth_init.py:68-71."""
    return Path(config.mdi.cache_dir) / (name or model_name(config)) / dataset_name / f"{image_name}.pth"


def load_cached_depth(path: Path, device) -> PredictedDepth:
    """`torch.load` of a cached PredictedDepth (monocular_depth_init.py:76), moved to `device` (the
    reference asserts the depth already lives there, :140; a cache written on another GPU index is
    accepted here).  Files written by the reference pickle the class under its `gs_init_compare...`
    path: the alias package at the repo root resolves it to this package's dataclass."""
    pred = torch.load(path, map_location=device, weights_only=True)
    if not hasattr(pred, "depth") or not hasattr(pred, "mask"):
        raise TypeError(f"{path}: expected a PredictedDepth, got {type(pred).__name__}")
    return pred


class CachedDepthSource:
    """Default `depth_source` of pts_and_rgb_from_monocular_depth: callable
    (image: InputImage holding the dataset's 0..255 image, data: dataset dict, device) -> PredictedDepth that mirrors
    predict_depth_or_get_cached_depth (monocular_depth_init.py:60-87)."""

    def __init__(self, config, dataset_name: str, device):
        self.config, self.dataset_name, self.device = config, dataset_name, device
        self.name = model_name(config)   # raises for a missing / unknown predictor, like pick_model
        self._model: Optional[DepthPredictor] = None
        _LOGGER.info("Using depth predictor model: %s", self.name)

    def _predictor(self) -> Optional[DepthPredictor]:
        fac = _FACTORIES.get(self.config.mdi.predictor)
        if fac is not None and self._model is None:
            self._model = fac(self.config, self.device)
            if self._model.name != self.name:  # the cache directory is keyed by the name
                self.name = self._model.name
        return self._model

    def __call__(self, image, data: dict, device) -> PredictedDepth:
        cfg = self.config
        model = self._predictor()
        path = cache_path(cfg, self.dataset_name, image.name, self.name)
        depth = None
        if not cfg.mdi.ignore_cache and path.exists():
            try:
                depth = load_cached_depth(path, device)
            except Exception as e:  # noqa: BLE001 - the reference logs and falls through to the predictor
                if model is None:
                    raise
                _LOGGER.warning("Failed to load cached depth for image %s: %s", image.name, e)
        if depth is None:
            if model is None:
                why = "mdi.ignore_cache is set" if cfg.mdi.ignore_cache else "no cached depth"
                raise DepthCacheMiss(
                    f"{path}: {why} and no depth predictor is registered for '{cfg.mdi.predictor}'. The "
                    "predictor networks are outside the B200 path: fill the cache with the reference "
                    "(torch.save(PredictedDepth) per image), or plug one in with "
                    "depth_prediction.depth_cache.register_predictor().")
            path.parent.mkdir(exist_ok=True, parents=True)
            # the reference hands the predictor image / 255 (monocular_depth_init.py:128,133); the
            # depth_source contract carries the dataset's 0..255 image
            depth = model.predict_depth(image.data / 255.0, CameraIntrinsics(image.K))
            try:
                torch.save(depth, path)
            except KeyboardInterrupt:
                path.unlink(missing_ok=True)
                raise
        return depth
