"""Post-processing of the merged cloud (reference: point_cloud_postprocess/postprocess.py:32-77)."""
from __future__ import annotations

import numpy as np
import torch

from .config import OutlierRemovalMethod, PointCloudPostprocessConfig
from .subsampling import subsample_pointcloud_device
from ..knn import local_outlier_factor, local_outlier_factor_sharded


def lof_outlier_removal(pts: torch.Tensor, config: PointCloudPostprocessConfig, group=None) -> torch.Tensor:
    """postprocess.py:16-22: bool [N], True where LocalOutlierFactor(n_neighbors).fit_predict == -1.
    `group`: ranks that all hold this cloud split the queries (knn.local_outlier_factor_sharded)."""
    if group is not None:
        return local_outlier_factor_sharded(pts, config.lof_num_neighbors, group)[0]
    outliers, _ = local_outlier_factor(pts, config.lof_num_neighbors)
    return outliers


def postprocess_point_cloud(pts, rgbs, intrinsic_matrices, proj_matrices, image_sizes,
                            config: PointCloudPostprocessConfig, device, group=None):
    """`group` (beyond the reference's signature): a torch.distributed group whose ranks all hold the
    same cloud -- the LOF queries and the extent pass are then split over them."""
    method = getattr(config.outlier_removal, "value", config.outlier_removal)
    if method == OutlierRemovalMethod.lof.value:
        # postprocess.py:16-22,41-58 (the filtered_* / outliers_* PLY exports are not written)
        outliers = lof_outlier_removal(pts, config, group)
        keep = ~outliers
        pts = pts[keep]
        rgbs = rgbs[keep.to(rgbs.device)]
    elif method != OutlierRemovalMethod.off.value:
        raise ValueError(f"Unknown outlier removal method: {config.outlier_removal}")
    if config.subsample:
        # postprocess.py:60-75; the cloud stays on the device instead of a numpy round trip
        dev = torch.device(device)
        K = torch.as_tensor(np.asarray(intrinsic_matrices, dtype=np.float32).reshape(-1, 3, 3))
        P = torch.as_tensor(np.asarray(proj_matrices, dtype=np.float32).reshape(-1, 3, 4))
        sizes = torch.as_tensor(np.asarray(image_sizes, dtype=np.int32).reshape(-1, 2))
        pts, rgbs, _ = subsample_pointcloud_device(pts.to(dev), rgbs.to(dev), K, P, sizes,
                                                   config.subsample_params, group=group)
    return pts, rgbs
