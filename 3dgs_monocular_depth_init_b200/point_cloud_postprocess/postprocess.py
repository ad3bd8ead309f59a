"""Post-processing of the merged cloud (reference: point_cloud_postprocess/postprocess.py:32-77)."""
from __future__ import annotations

import numpy as np
import torch

from .config import OutlierRemovalMethod, PointCloudPostprocessConfig


def postprocess_point_cloud(pts, rgbs, intrinsic_matrices, proj_matrices, image_sizes,
                            config: PointCloudPostprocessConfig, device):
    method = getattr(config.outlier_removal, "value", config.outlier_removal)
    if method == OutlierRemovalMethod.lof.value:
        raise NotImplementedError("LOF outlier removal (SURVEY.md a12) is not built yet")
    elif method != OutlierRemovalMethod.off.value:
        raise ValueError(f"Unsupported outlier removal method: {config.outlier_removal}")
    if config.subsample:
        raise NotImplementedError("native merge subsampling (SURVEY.md a13/a14) is not built yet")
    return pts, rgbs
