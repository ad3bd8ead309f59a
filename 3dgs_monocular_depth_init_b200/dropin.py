"""Overlay of an installed reference package with the B200 hot path.

`gs_init_compare/` at the repo root REPLACES the reference's package for the hot-path modules.  When
the reference itself is installed (its trainer, datasets and nerfbaselines glue are needed) the same
effect is had by patching its modules in place: `install()` rebinds every hot-path entry point of
SURVEY.md 8(b) inside the reference's own modules, so `runner.py:77-78`, `method.py` and user code
keep their imports and call sites.  Call it before `gs_init_compare.runner` is imported, or after --
already imported callers that did `from ... import name` are rebound too."""
from __future__ import annotations

import importlib
import sys

_IMPL = __name__.rsplit(".", 1)[0]

# reference module (relative to the package) -> names rebound to this package's objects
ENTRY_POINTS = {
    "monocular_depth_init": ["pts_and_rgb_from_monocular_depth", "add_noise_to_point_cloud"],
    "depth_prediction.points_from_depth": ["get_pts_from_depth", "project_and_filter_sfm_pts", "depth_gradient_mask",
                                           "get_subsampler"],
    "depth_alignment.pipeline": ["DepthAlignmentPipeline"],
    "depth_alignment.alignment.lstsqrs": ["align_depth_least_squares", "DepthAlignmentLstSqrs"],
    "depth_alignment.alignment.ransacs": ["DepthAlignmentRansac", "DepthAlignmentMsac"],
    "depth_subsampling.static_subsampler": ["StaticDepthSubsampler"],
    "depth_subsampling.adaptive_subsampling": ["AdaptiveDepthSubsampler"],
    "depth_subsampling.num_sfm_points_mask": ["num_sfm_points_mask"],
    "point_cloud_postprocess.postprocess": ["postprocess_point_cloud", "lof_outlier_removal"],
    "utils.runner_utils": ["knn"],
}
# modules that import those names with `from ... import` at their top level (runner.py:42,
# monocular_depth_init.py:19-25, points_from_depth.py:13-21)
CALLERS = ["runner", "monocular_depth_init", "depth_prediction.points_from_depth", "depth_alignment.pipeline"]


def install(reference_package: str = "gs_init_compare", strict: bool = False) -> list:
    """Rebinds the hot-path entry points inside the modules of `reference_package`.  Modules that
    cannot be imported (missing optional dependencies of the reference) are skipped unless `strict`.
    Returns the list of `module.name` strings that were rebound."""
    done = []
    ours = {}
    for rel, names in ENTRY_POINTS.items():
        impl = importlib.import_module(f"{_IMPL}.{rel}")
        try:
            ref = importlib.import_module(f"{reference_package}.{rel}")
        except ImportError:
            if strict:
                raise
            continue
        if ref is impl:   # the alias package is in use: nothing to patch
            continue
        for n in names:
            obj = getattr(impl, n)
            ours[n] = obj
            setattr(ref, n, obj)
            done.append(f"{reference_package}.{rel}.{n}")
    for rel in CALLERS:   # names already copied into caller namespaces
        mod = sys.modules.get(f"{reference_package}.{rel}")
        if mod is None:
            continue
        for n, obj in ours.items():
            if n in mod.__dict__ and mod.__dict__[n] is not obj:
                setattr(mod, n, obj)
                done.append(f"{reference_package}.{rel}.{n}")
    return done
