"""Post-processing of the merged cloud (reference: point_cloud_postprocess/postprocess.py:32-77)."""
from __future__ import annotations

import numpy as np
import torch

from .config import OutlierRemovalMethod, PointCloudPostprocessConfig
from .subsampling import subsample_pointcloud_device


def postprocess_point_cloud(pts, rgbs, intrinsic_matrices, proj_matrices, image_sizes,
                            config: PointCloudPostprocessConfig, device):
    method = getattr(config.outlier_removal, "value", config.outlier_removal)
    if method == OutlierRemovalMethod.lof.value:
        raise NotImplementedError("LOF outlier removal (SURVEY.md a12) is not built yet")
    elif method != OutlierRemovalMethod.off.value:
        raise ValueError(f"Unsupported outlier removal method: {config.outlier_removal}")
    if config.subsample:
        # postprocess.py:60-75; the cloud stays on the device instead of a numpy round trip
        dev = torch.device(device)
        K = torch.as_tensor(np.asarray(intrinsic_matrices, dtype=np.float32).reshape(-1, 3, 3))
        P = torch.as_tensor(np.asarray(proj_matrices, dtype=np.float32).reshape(-1, 3, 4))
        sizes = torch.as_tensor(np.asarray(image_sizes, dtype=np.int32).reshape(-1, 2))
        pts, rgbs, _ = subsample_pointcloud_device(pts.to(dev), rgbs.to(dev), K, P, sizes,
                                                   config.subsample_params)
    return pts, rgbs
