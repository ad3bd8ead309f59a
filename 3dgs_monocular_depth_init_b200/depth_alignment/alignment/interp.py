"""`interp` aligner on the B200 (reference: depth_alignment/alignment/interp.py).

Per-SfM-point scale factors gt / pre-aligned depth (:329-334) -> scale-factor outlier removal
(:161-201) -> piecewise-linear interpolation over a Delaunay triangulation of the points plus the
four image corners (:75-110) -> dense scale map times the pre-aligned depth (:361).

What runs where:
  * init stage (:204-232): the lstsq / RANSAC kernels of the engine (mdi_gather_fit / mdi_ransac);
  * outlier removal: exact kNN and LOF on the GPU (mdi_knn / mdi_lof) over the pixel coordinates;
  * triangulation: scipy.spatial.Delaunay (Qhull) on the host, as the reference does -- a sequential
    incremental algorithm over ~5 k points per view;
  * evaluation at all H x W pixels and the product with the aligned depth: mdi_interp_scale_map
    (csrc/k8_interp.cu), with scipy's own barycentric transforms so the values are LinearNDInterpolator's.
`method="rbf"` needs the third-party torchrbf package (absent here) and raises.

Tie note: pixel coordinates are integers, so equal neighbour distances are common; sklearn's KDTree
returns tied neighbours in tree order, the GPU search in (distance, index) order.  The two can pick
different 5th neighbours, i.e. a different median for a handful of points."""
from __future__ import annotations

import logging
from enum import IntEnum
from pathlib import Path
from typing import NamedTuple, Optional

import numpy as np
import torch

from ... import _lib as L
from ..._single import options_from_config, ransac_seed
from ...depth_prediction.predictors.depth_predictor_interface import PredictedDepth
from ...engine import RunOptions, get_engine
from ...knn import knn_search, local_outlier_factor
from ..interface import DepthAlignmentResult, DepthAlignmentStrategy
from .lstsqrs import fit_view

LOGGER = logging.getLogger(__name__)


def rbf_interpolation(coords, values, config, device, W: int, H: int):
    """interp.py:30-72 (torchrbf RBFInterpolator on a 256-wide grid + bilinear upsampling)."""
    raise NotImplementedError("interp.method='rbf' needs the torchrbf package, which is not part of the B200 path; "
                              "use method='linear'")


def _triangulate(coords_np: np.ndarray, values_np: np.ndarray, W: int, H: int):
    """Host part of interp.py:83-109: corners appended, Delaunay, corner values by inverse-distance
    weighting over their Delaunay neighbours.  -> (scipy Delaunay, values float32 [N+4], fill value)."""
    from scipy.spatial import Delaunay
    corner_coords = np.array([[0, 0], [0, H - 1], [W - 1, 0], [W - 1, H - 1]])
    corner_indices = np.arange(coords_np.shape[0], coords_np.shape[0] + 4)
    coords_np = np.vstack((coords_np, corner_coords))
    values_np = np.hstack((values_np, np.empty(4, dtype=values_np.dtype)))
    dt = Delaunay(coords_np)
    for corner_ix in corner_indices:
        indptr, indices = dt.vertex_neighbor_vertices
        neighbors = indices[indptr[corner_ix]: indptr[corner_ix + 1]]
        neighbors = np.setdiff1d(neighbors, corner_indices)   # exclude other corners
        distances = np.linalg.norm(coords_np[neighbors] - coords_np[corner_ix], axis=1)
        weights = 1.0 / (distances + 1e-8)
        weights /= np.sum(weights)
        corner_value = np.sum(values_np[neighbors] * weights)
        if np.isnan(corner_value):
            corner_value = np.median(values_np[neighbors])
        values_np[corner_ix] = corner_value
    return dt, values_np, float(np.median(values_np))


def linear_interpolation(coords: torch.Tensor, values: torch.Tensor, config, device, W: int, H: int,
                         base: Optional[torch.Tensor] = None, scale_shift: Optional[torch.Tensor] = None):
    """interp.py:75-110: coords int [2,N] (row 0 = x), values float32 [N] -> scale map float32 [H,W]
    on the device.  With `base` (predicted depth [H,W]) and `scale_shift` ([2], device) the product
    scale_map * (base * s + t) of interp.py:361 is returned instead (fused into the same kernel)."""
    eng = get_engine(device)
    dev = eng.device
    dt, values_np, fill = _triangulate(coords.T.cpu().numpy(), values.detach().cpu().numpy(), W, H)
    to_dev = lambda a, dtp: torch.from_numpy(np.ascontiguousarray(a, dtype=dtp)).to(dev)
    transform = to_dev(dt.transform, np.float64)       # [T,3,2] scipy's barycentric transforms
    simplices = to_dev(dt.simplices, np.int32)         # [T,3]
    points = to_dev(dt.points, np.float64)             # [P,2]
    vals = to_dev(values_np, np.float64)               # LinearNDInterpolator evaluates in float64
    out = torch.empty((H, W), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        ws = torch.empty(int(eng.lib.mdi_interp_workspace_bytes(H, W)), dtype=torch.uint8, device=dev)
        b = None if base is None else base.to(device=dev, dtype=torch.float32).contiguous()
        ss = None if scale_shift is None else scale_shift.to(device=dev, dtype=torch.float32).reshape(-1)[:2].contiguous()
        L.check(eng.lib.mdi_interp_scale_map(transform.data_ptr(), simplices.data_ptr(), points.data_ptr(), vals.data_ptr(),
                                             int(simplices.shape[0]), fill, int(H), int(W), L.ptr(b), L.ptr(ss),
                                             ws.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream),
                "mdi_interp_scale_map")
    return out


def interpolate_scale(coords, values, config, device, W: int, H: int, **kw):
    """interp.py:113-127."""
    if config.method == "rbf":
        return rbf_interpolation(coords, values, config, device, W, H)
    elif config.method == "linear":
        return linear_interpolation(coords, values, config, device, W, H, **kw)
    raise ValueError(f"Unknown interpolation method: {config.method}")


class OutlierType(IntEnum):
    REGULAR = 0
    SCALE_ONLY = 1
    POSITION_ONLY = 2
    BOTH = 3


class OutlierClassification(NamedTuple):
    scale_only_outliers: torch.Tensor
    both_outliers: torch.Tensor
    position_only_outliers: torch.Tensor
    regular: torch.Tensor


def scale_factor_outlier_removal(coords: torch.Tensor, scales: torch.Tensor, debug_export_dir: Optional[Path] = None):
    """interp.py:161-201.  coords int [N,2] (x, y), scales float32 [N], both on the device."""
    K_lof = 10
    K_scale_knn = 5
    num_pts = coords.shape[0]
    dev = scales.device
    if num_pts < min(K_lof + 1, K_scale_knn + 1):
        z = torch.zeros(num_pts, dtype=torch.bool, device=dev)
        return OutlierClassification(z, z.clone(), z.clone(), torch.ones(num_pts, dtype=torch.bool, device=dev))
    # pixel coordinates as points of the z = 0 plane (exact in float32 below 2^24)
    xyz = torch.zeros((num_pts, 3), dtype=torch.float32, device=dev)
    xyz[:, :2] = coords.to(device=dev, dtype=torch.float32)
    position_outliers, _ = local_outlier_factor(xyz, K_lof)
    idx, _, _ = knn_search(xyz, K_scale_knn + 1, exclude_self=False)
    knn_indices = idx[:, 1:].long()   # first column dropped like the reference (the point itself / a duplicate of it)
    knn_median_scale = torch.median(scales[knn_indices], dim=1).values
    scale_diff = torch.abs(scales - knn_median_scale)
    scale_diff_threshold = torch.quantile(scale_diff, 0.99)
    scale_outliers = scale_diff > scale_diff_threshold
    return OutlierClassification(
        scale_only_outliers=scale_outliers & ~position_outliers,
        both_outliers=scale_outliers & position_outliers,
        position_only_outliers=position_outliers & ~scale_outliers,
        regular=~(scale_outliers | position_outliers),
    )


def initial_fit(predicted_depth: PredictedDepth, coords: torch.Tensor, gt_depth: torch.Tensor, config,
                sample_idx: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """Scale / shift of the init stage (interp.py:204-232) as a device tensor [2], None for init=None."""
    init = config.mdi.alignment.interp.init
    if init is None:
        return None
    if init not in ("lstsqrs", "ransac"):
        raise ValueError(f"Unknown interp alignment init method: {init}")
    opt = options_from_config(config, aligner=init)
    if init == "ransac":
        if sample_idx is not None:
            opt.sample_idx = sample_idx.to(device=predicted_depth.depth.device, dtype=torch.int32).contiguous()[None]
        else:
            opt.ransac_seed = ransac_seed()
    return fit_view(predicted_depth, coords, gt_depth, opt).scale_shift[0]


def _interp_from_fit(depth_map: torch.Tensor, coords: torch.Tensor, gt: torch.Tensor, ss: Optional[torch.Tensor],
                     interp_config, outlier_mask: Optional[torch.Tensor] = None, debug_export_dir=None) -> torch.Tensor:
    """interp.py:329-361 given the init fit `ss` ([2] device tensor or None): -> aligned depth [H,W]."""
    H, W = depth_map.shape
    dev = depth_map.device
    if interp_config.method == "rbf":
        rbf_interpolation(None, None, interp_config, dev, W, H)
    d = depth_map[coords[1].long(), coords[0].long()].float()
    pre = d if ss is None else d * ss[0] + ss[1]          # two roundings, like depth * s + t
    scale_factors = gt / pre
    if outlier_mask is None and interp_config.scale_outlier_removal:
        outlier_mask = scale_factor_outlier_removal(coords.T, scale_factors, debug_export_dir).scale_only_outliers
    if outlier_mask is not None:
        outlier_mask = outlier_mask.to(dev)
        n_out = int(outlier_mask.sum())
        if n_out > 0:
            LOGGER.info("Removed %d/%d scale outlier points.", n_out, coords.shape[1])
        scale_factors = scale_factors[~outlier_mask]
        coords = coords[:, ~outlier_mask]
    try:
        return interpolate_scale(coords, scale_factors, interp_config, dev, W, H, base=depth_map, scale_shift=ss)
    except NotImplementedError:
        raise
    except Exception as e:  # noqa: BLE001 - the reference falls back to the median scale (:349-359)
        LOGGER.warning("Scale factor interpolation failed (%s); using median scale instead of interpolation.", e)
        base = depth_map.float()
        pre_map = base if ss is None else base * ss[0] + ss[1]
        return scale_factors.median() * pre_map


def align_depth_interpolate(predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
                            gt_depth: torch.Tensor, config, debug_export_dir: Optional[Path] = None, *,
                            sample_idx: Optional[torch.Tensor] = None, outlier_mask: Optional[torch.Tensor] = None):
    """interp.py:281-361.  `sample_idx` feeds explicit RANSAC samples to the init stage and
    `outlier_mask` (bool [N]) replaces the outlier-removal step (both: parity runs)."""
    dev = predicted_depth.depth.device
    coords = sfm_points_camera_coords.to(dev)
    gt = gt_depth.to(device=dev, dtype=torch.float32)
    ss = initial_fit(predicted_depth, coords, gt, config, sample_idx)
    aligned = _interp_from_fit(predicted_depth.depth, coords, gt, ss, config.mdi.alignment.interp, outlier_mask,
                               debug_export_dir)
    return DepthAlignmentResult(aligned, predicted_depth.mask)


def run_views_with_interp(eng, vin, config, opt2: RunOptions, sample_idx: Optional[torch.Tensor] = None):
    """The batched pipeline with `aligner = interp`: (1) project + init fit of every view in one
    engine run, (2) per view: scale factors, outlier removal, host triangulation, dense scale map
    times the pre-aligned depth -> a materialised aligned depth map, (3) the mask / compaction /
    unprojection stages on those maps (aligner None: the maps are used as they are).
    `opt2`: options of the second run (subsampling, masks, outputs); its aligner is overridden."""
    import dataclasses
    V = int(vin.K.shape[0])
    init = config.mdi.alignment.interp.init
    if init not in (None, "lstsqrs", "ransac"):
        raise ValueError(f"Unknown interp alignment init method: {init}")
    opt1 = options_from_config(config, aligner=init, stop_after="fit", use_patch_mask=False)
    if init == "ransac":
        if sample_idx is not None:
            opt1.sample_idx = sample_idx.to(device=eng.device, dtype=torch.int32).contiguous()
        else:
            opt1.ransac_seed = ransac_seed()
    if V == 0:
        opt2.aligner, opt2.scale_shift = None, None
        return eng.run(vin, opt2)
    r1 = eng.run(vin, opt1)
    n_valid = r1.n_valid.cpu().tolist()
    status = r1.status.cpu().tolist()
    aligned = []
    for v in range(V):
        depth_v = vin.depth[v]
        if status[v] != L.STATUS_OK:
            aligned.append(depth_v)          # skipped view: never read again
            continue
        b, n = int(r1.sfm_begin[v]), int(n_valid[v])
        ss = None if r1.scale_shift is None else r1.scale_shift[v]
        aligned.append(_interp_from_fit(depth_v, r1.uv[:, b:b + n], r1.z[b:b + n], ss, config.mdi.alignment.interp))
    vin2 = dataclasses.replace(vin, depth=aligned)
    opt2.aligner, opt2.scale_shift = None, None
    res = eng.run(vin2, opt2)
    res.scale_shift = r1.scale_shift
    res.ransac_iters, res.ransac_inliers = r1.ransac_iters, r1.ransac_inliers
    res.extras["aligned_depth"] = aligned
    return res


class DepthAlignmentInterpolate(DepthAlignmentStrategy):
    @classmethod
    def align(cls, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, config, debug_export_dir: Optional[Path] = None,
              *args, **kwargs) -> DepthAlignmentResult:
        return align_depth_interpolate(predicted_depth, sfm_points_camera_coords, sfm_points_depth, config,
                                       debug_export_dir, **kwargs)
