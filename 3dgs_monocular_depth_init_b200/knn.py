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


# ---- query split over ranks (LOF needs the whole cloud; every rank has it after the exchange) ----
def query_range(n: int, world: int, rank: int):
    """Contiguous block of the n (Morton-sorted) query positions owned by `rank`."""
    per = -(-n // world) if world else n
    return min(rank * per, n), min((rank + 1) * per, n)


def gather_rows_by_index(rows: torch.Tensor, self_idx: torch.Tensor, n: int, group=None) -> torch.Tensor:
    """Every rank contributes `rows` ([m, ...]) for the points `self_idx` ([m] original indices, the
    ranks' index sets partition 0 .. n-1); returns the full [n, ...] array in original order on every
    rank.  One all-gather of the rows padded to the largest block plus one of the 4-byte indices."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    m = int(rows.shape[0])
    biggest = torch.tensor([m], dtype=torch.int64, device=rows.device)
    dist.all_reduce(biggest, op=dist.ReduceOp.MAX, group=group)   # blocks are cell-aligned, not exactly n / world
    per = max(int(biggest), 1)
    pad_rows = torch.zeros((per,) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
    pad_rows[:m] = rows
    pad_idx = torch.full((per,), -1, dtype=torch.int64, device=rows.device)
    pad_idx[:m] = self_idx.to(torch.int64)
    all_rows = torch.empty((world * per,) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
    all_idx = torch.empty(world * per, dtype=torch.int64, device=rows.device)
    dist.all_gather_into_tensor(all_rows, pad_rows, group=group)
    dist.all_gather_into_tensor(all_idx, pad_idx, group=group)
    ok = all_idx >= 0
    out = torch.empty((n,) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
    out[all_idx[ok]] = all_rows[ok]
    return out


def local_outlier_factor_sharded(x: torch.Tensor, n_neighbors: int = 40, group=None):
    """`local_outlier_factor` with the queries split over the ranks of `group`: every rank holds the
    same cloud x, answers about n / G of the kNN queries (mdi_knn_build + mdi_knn_split + mdi_knn_query) and the per-point k-distance,
    lrd and outlier flag are exchanged (4 + 4 + 1 bytes per point).  Bit-identical to the single-GPU
    result on every rank.  -> (outlier bool [N], lof float32 [N])."""
    import torch.distributed as dist
    lib = L.load()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = int(x.shape[0])
    k = n_neighbors
    if n < 2:
        raise ValueError("local_outlier_factor needs at least 2 points")
    if k >= n:
        warnings.warn(f"n_neighbors ({k}) is greater than the total number of samples ({n}). "
                      f"n_neighbors will be set to (n_samples - 1) for estimation.")
        k = max(1, n - 1)
    if not x.is_cuda:
        raise L.MdiError("knn needs a CUDA tensor; there is no CPU fallback")
    pts = x.to(torch.float32).contiguous()
    if not bool(torch.isfinite(pts).all()):
        raise ValueError("Input X contains NaN or infinity.")
    dev = pts.device
    with torch.cuda.device(dev):
        ws = torch.empty(int(lib.mdi_knn_workspace_bytes(n)), dtype=torch.uint8, device=dev)
        L.check(lib.mdi_knn_build(L.ptr(pts), n, L.ptr(ws), _stream()), "mdi_knn_build")
        # cell-aligned rank boundaries, identical on every rank (the grid is a function of the cloud)
        bounds = torch.empty(world + 1, dtype=torch.int64, device=dev)
        L.check(lib.mdi_knn_split(L.ptr(ws), n, world, L.ptr(bounds), _stream()), "mdi_knn_split")
        a, b = (int(v) for v in bounds[rank:rank + 2].tolist())
        m = b - a
        self_idx = torch.empty(max(m, 1), dtype=torch.int32, device=dev)[:m]
        idx = torch.empty((m, k), dtype=torch.int32, device=dev)
        dist_rows = torch.empty((m, k), dtype=torch.float32, device=dev)
        kdist_rows = torch.empty(m, dtype=torch.float32, device=dev)
        L.check(lib.mdi_knn_query(L.ptr(ws), n, k, 1, a, b, L.ptr(self_idx), L.ptr(idx), L.ptr(dist_rows),
                                  L.ptr(kdist_rows), _stream()), "mdi_knn_query")
        del ws
        kdist = gather_rows_by_index(kdist_rows, self_idx, n, group)
        lrd_rows = torch.empty(m, dtype=torch.float32, device=dev)
        L.check(lib.mdi_lof_lrd(m, k, L.ptr(idx), L.ptr(dist_rows), L.ptr(kdist), L.ptr(lrd_rows), _stream()), "mdi_lof_lrd")
        lrd = gather_rows_by_index(lrd_rows, self_idx, n, group)
        lof_rows = torch.empty(m, dtype=torch.float32, device=dev)
        out_rows = torch.empty(m, dtype=torch.uint8, device=dev)
        L.check(lib.mdi_lof_score(m, k, L.ptr(idx), L.ptr(lrd), L.ptr(lrd_rows), L.ptr(lof_rows), L.ptr(out_rows), _stream()),
                "mdi_lof_score")
        packed = gather_rows_by_index(torch.stack([lof_rows, out_rows.float()], 1), self_idx, n, group)
    return packed[:, 1] != 0, packed[:, 0].contiguous()
