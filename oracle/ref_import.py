"""TEST INFRASTRUCTURE -- not product code.  Imports the UNMODIFIED reference hot path from
/root/reference on CPU (only possible in the build container; the GPU box has no
/root/reference).  Used by oracle/make_golden.py to generate tests/golden/*.npz and by the
CPU tests (when the reference is present) to validate oracle/mdi_oracle.py.

Two shims, both outside the reference tree:
  1. Python 3.12 rejects the reference's dataclass-instance defaults -> dataclasses.dataclass
     is temporarily wrapped to add unsafe_hash=True while the reference modules import.
  2. sys.modules stubs for packages absent here but imported at module top level.
"""
from __future__ import annotations

import dataclasses
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("MDI_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "gs_init_compare"))


_ref = None


def load():
    """Returns a namespace with the reference's hot-path symbols."""
    global _ref
    if _ref is not None:
        return _ref
    if not available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT}")
    import numpy, scipy, sklearn.neighbors, torch, tyro  # noqa: F401,E401  real deps first

    # The repo root holds an ALIAS package of the same name (gs_init_compare/ -> the B200 implementation).
    # If anything imported it in this process, drop it and its finder: the oracle must never be checked
    # against the product by accident.
    for k in [k for k in sys.modules if k == "gs_init_compare" or k.startswith("gs_init_compare.")]:
        del sys.modules[k]
    sys.meta_path[:] = [f for f in sys.meta_path if type(f).__name__ != "_AliasFinder"]
    sys.path.insert(0, REFERENCE_ROOT)
    orig = dataclasses.dataclass

    def compat(cls=None, /, **kw):
        kw.setdefault("unsafe_hash", True)
        return (lambda c: orig(c, **kw)) if cls is None else orig(cls, **kw)

    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _S:
        def __init__(self, *a, **k):
            pass

    dataclasses.dataclass = compat
    try:
        stub("gsplat")
        stub("gsplat.strategy", DefaultStrategy=_S, MCMCStrategy=_S)

        @compat
        class PointCloudSubsamplingParams:
            max_bbox_aspect_ratio: float = 1.1
            min_extent_multiplier: float = 1.0

        stub("pointcloud_subsampling", subsample_pointcloud=None)
        stub("pointcloud_subsampling.subsampling_params",
             PointCloudSubsamplingParams=PointCloudSubsamplingParams)
        mpl = stub("matplotlib")
        mpl.pyplot = stub("matplotlib.pyplot")
        stub("matplotlib.colors", ListedColormap=None)
        stub("open3d"); stub("skimage"); stub("pycolmap"); stub("imageio"); stub("imageio.v2")
        stub("segment_anything", SamAutomaticMaskGenerator=None, sam_model_registry=None)
        stub("nerfbaselines", Method=object, MethodInfo=dict, ModelInfo=dict, Dataset=dict,
             OptimizeEmbeddingOutput=dict)
        stub("nerfbaselines.utils", pad_poses=None, image_to_srgb=None, convert_image_dtype=None)
        stub("torchrbf", RBFInterpolator=None)          # interp.py:23 (rbf method only)
        stub("skimage.segmentation"); stub("skimage.measure"); stub("skimage.graph")

        from gs_init_compare.config import Config
        from gs_init_compare.depth_alignment.alignment import lstsqrs, ransacs
        from gs_init_compare.depth_alignment.alignment import interp
        from gs_init_compare.depth_alignment import pipeline
        from gs_init_compare.depth_alignment.segmentation import region_margin
        from gs_init_compare.depth_alignment.config import DepthAlignmentStrategyEnum
        from gs_init_compare.depth_alignment.exceptions import LowDepthAlignmentConfidenceError
        from gs_init_compare.depth_alignment.pipeline import DepthAlignmentPipeline
        from gs_init_compare.depth_prediction import points_from_depth
        from gs_init_compare.depth_prediction.predictors.depth_predictor_interface import PredictedDepth
        from gs_init_compare.depth_subsampling import adaptive_subsampling, num_sfm_points_mask, static_subsampler
        from gs_init_compare.types import InputImage
    finally:
        dataclasses.dataclass = orig
    import gs_init_compare
    where = os.path.realpath(os.path.dirname(gs_init_compare.__file__))
    assert where.startswith(os.path.realpath(REFERENCE_ROOT)), f"gs_init_compare resolved to {where}, not the reference"
    assert "3dgs_monocular_depth_init_b200" not in points_from_depth.get_pts_from_depth.__module__
    ns = types.SimpleNamespace(
        Config=Config, lstsqrs=lstsqrs, ransacs=ransacs, interp=interp, pipeline=pipeline, region_margin=region_margin,
        DepthAlignmentStrategyEnum=DepthAlignmentStrategyEnum,
        LowDepthAlignmentConfidenceError=LowDepthAlignmentConfidenceError,
        DepthAlignmentPipeline=DepthAlignmentPipeline, points_from_depth=points_from_depth,
        PredictedDepth=PredictedDepth, adaptive_subsampling=adaptive_subsampling,
        num_sfm_points_mask=num_sfm_points_mask, static_subsampler=static_subsampler,
        InputImage=InputImage,
    )
    _ref = ns
    return ns
