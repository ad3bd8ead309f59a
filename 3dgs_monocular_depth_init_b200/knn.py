"""Exact k nearest neighbours and the Local Outlier Factor on the GPU.

Reference: utils/runner_utils.py:143-147 (`knn`, sklearn NearestNeighbors on the CPU) and
point_cloud_postprocess/postprocess.py:16-22 (`lof_outlier_removal`, sklearn LocalOutlierFactor).
Same signatures; the search runs in libmdi_b200.so (mdi_knn / mdi_lof) -- there is no CPU path.
"""
from __future__ import annotations

import warnings

import torch

from . import _lib as L


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def knn_search(x: torch.Tensor, k: int, exclude_self: bool):
    """-> (idx int32 [N,k], dist float32 [N,k], kdist float32 [N]) in ascending (distance, index) order."""
    lib = L.load()
    if not x.is_cuda:
        raise L.MdiError("knn needs a CUDA tensor; there is no CPU fallback")
    if x.ndim != 2 or x.shape[1] != 3:
        raise ValueError("knn: expected points of shape (N, 3)")
    pts = x.to(torch.float32).contiguous()
    n = pts.shape[0]
    if n and not bool(torch.isfinite(pts).all()):
        # sklearn's check_array (NearestNeighbors / LocalOutlierFactor .fit) rejects such input
        raise ValueError("Input X contains NaN or infinity.")
    with torch.cuda.device(pts.device):
        ws = torch.empty(int(lib.mdi_knn_workspace_bytes(n)), dtype=torch.uint8, device=pts.device)
        idx = torch.empty((n, k), dtype=torch.int32, device=pts.device)
        dist = torch.empty((n, k), dtype=torch.float32, device=pts.device)
        kdist = torch.empty(n, dtype=torch.float32, device=pts.device)
        rc = lib.mdi_knn(L.ptr(pts), n, k, int(exclude_self), L.ptr(ws), L.ptr(idx), L.ptr(dist), L.ptr(kdist),
                         _stream())
        if rc != 0:
            msg = lib.mdi_last_error().decode(errors="replace")
            if "Expected n_neighbors" in msg:
                raise ValueError(msg.split("mdi_knn: ")[-1])  # sklearn raises ValueError here
            L.check(rc, "mdi_knn")
    return idx, dist, kdist


def knn(x: torch.Tensor, K: int = 4) -> torch.Tensor:
    """runner_utils.py:143-147: distances [N,K] to the K nearest points of x, the point itself
    included (column 0 is 0), in x's dtype and on x's device."""
    _, dist, _ = knn_search(x, K, exclude_self=False)
    return dist.to(x)


def local_outlier_factor(x: torch.Tensor, n_neighbors: int = 40):
    """sklearn LocalOutlierFactor(n_neighbors).fit(x): -> (outlier bool [N], lof float32 [N]);
    outlier = fit_predict(x) == -1, lof = -negative_outlier_factor_."""
    lib = L.load()
    n = x.shape[0]
    k = n_neighbors
    if n < 2:
        raise ValueError("local_outlier_factor needs at least 2 points")
    if k >= n:  # sklearn _lof.py: clamps with this warning
        warnings.warn(f"n_neighbors ({k}) is greater than the total number of samples ({n}). "
                      f"n_neighbors will be set to (n_samples - 1) for estimation.")
        k = max(1, n - 1)
    pts = x.to(torch.float32).contiguous()
    idx, dist, kdist = knn_search(pts, k, exclude_self=True)
    with torch.cuda.device(pts.device):
        lrd = torch.empty(n, dtype=torch.float32, device=pts.device)
        lof = torch.empty(n, dtype=torch.float32, device=pts.device)
        out = torch.empty(n, dtype=torch.uint8, device=pts.device)
        L.check(lib.mdi_lof(n, k, L.ptr(idx), L.ptr(dist), L.ptr(kdist), L.ptr(lrd), L.ptr(lof), L.ptr(out), _stream()),
                "mdi_lof")
    return out.bool(), lof
