"""Merge subsampling of the final cloud on the GPU -- the reference's only native module.

Reference: native_modules/subsampling/src/pointcloud_subsampling.cpp:22-67 (binding and argument
checks), impl.cpp:70-126 (minimal Gaussian extents), impl.cpp:313-426 (the k-d merge).  Same
argument meaning and error behaviour; the arithmetic runs in libmdi_b200.so (mdi_min_extents,
mdi_merge_subsample) -- there is no CPU path.  The reference's debug cloud (4th/5th return value,
only ever exported to `merged.ply`) is returned empty, and no *_stats.txt files are written.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib as L


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def subsample_pointcloud_device(points: torch.Tensor, rgbs: torch.Tensor, intrinsic_matrices: torch.Tensor,
                                proj_matrices: torch.Tensor, image_sizes: torch.Tensor, params,
                                max_levels: int = 0, group=None):
    """Device tensors in, device tensors out: (pts[N',3], rgbs[N',3], extents[N]).
    `group` (torch.distributed group whose ranks all hold this same cloud): the O(N x cameras) extent
    pass is split over the ranks; the k-d merge itself depends on the global order and runs on every
    rank redundantly ("replicas only")."""
    lib = L.load()
    if points.ndim != 2 or points.shape[1] != 3:
        raise ValueError("Input points array must have shape (N, 3)")
    if rgbs.ndim != 2 or rgbs.shape[1] != 3:
        raise ValueError("Input rgbs array must have shape (N, 3)")
    if image_sizes.ndim != 2 or image_sizes.shape[1] != 2:
        raise ValueError("Input image_sizes array must have shape (N, 2)")
    if rgbs.shape[0] != points.shape[0]:
        raise ValueError("Number of points must match number of rgbs.")
    n_cams = intrinsic_matrices.shape[0]
    if n_cams != proj_matrices.shape[0] or n_cams != image_sizes.shape[0]:
        raise ValueError("Number of intrinsic_matrices must match number of camera_2_world_matrices and image_sizes.")
    if not points.is_cuda:
        raise L.MdiError("subsample_pointcloud_device needs CUDA tensors; there is no CPU fallback")
    dev = points.device
    pts = points.to(torch.float32).contiguous()
    col = rgbs.to(device=dev, dtype=torch.float32).contiguous()
    K = intrinsic_matrices.to(device=dev, dtype=torch.float32).contiguous()
    P = proj_matrices.to(device=dev, dtype=torch.float32).contiguous()
    sizes = image_sizes.to(device=dev, dtype=torch.int32).contiguous()
    n = pts.shape[0]
    with torch.cuda.device(dev):
        if group is not None:
            # a13 is independent per point: every rank computes the extents of its block of the
            # (replicated) cloud and the 4-byte results are all-gathered
            import torch.distributed as dist
            world, rank = dist.get_world_size(group), dist.get_rank(group)
            per = -(-n // world)
            a, b = min(rank * per, n), min((rank + 1) * per, n)
            part = torch.zeros(per, dtype=torch.float32, device=dev)
            if b > a:
                L.check(lib.mdi_min_extents(pts.data_ptr() + 12 * a, b - a, L.ptr(K), L.ptr(P), L.ptr(sizes), n_cams,
                                            L.ptr(part), _stream()), "mdi_min_extents")
            full = torch.empty(world * per, dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(full, part, group=group)
            ext = full[:n].contiguous()
        else:
            ext = torch.empty(n, dtype=torch.float32, device=dev)
            L.check(lib.mdi_min_extents(L.ptr(pts), n, L.ptr(K), L.ptr(P), L.ptr(sizes), n_cams, L.ptr(ext), _stream()),
                    "mdi_min_extents")
        ws = torch.empty(int(lib.mdi_merge_workspace_bytes(n)), dtype=torch.uint8, device=dev)
        out_p = torch.empty((n, 3), dtype=torch.float32, device=dev)
        out_c = torch.empty((n, 3), dtype=torch.float32, device=dev)
        n_out, n_levels = C.c_int64(0), C.c_int32(0)
        L.check(lib.mdi_merge_subsample(L.ptr(pts), L.ptr(col), L.ptr(ext), n,
                                        float(params.max_bbox_aspect_ratio), float(params.min_extent_multiplier),
                                        int(max_levels), L.ptr(ws), L.ptr(out_p), L.ptr(out_c), C.byref(n_out),
                                        C.byref(n_levels), _stream()), "mdi_merge_subsample")
    m = int(n_out.value)
    return out_p[:m], out_c[:m], ext


def subsample_pointcloud(points, rgbs, intrinsic_matrices, proj_matrices, image_sizes, params, device="cuda"):
    """Signature of the reference binding (numpy in, numpy out):
    -> (points', rgbs', min_gaussian_extents, debug_points, debug_rgbs)."""
    dev = torch.device(device)
    as_t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(np.asarray(a)), dtype=dt).to(dev)
    K = as_t(np.asarray(intrinsic_matrices, dtype=np.float32).reshape(-1, 3, 3), torch.float32)
    P = as_t(np.asarray(proj_matrices, dtype=np.float32).reshape(-1, 3, 4), torch.float32)
    p, c, e = subsample_pointcloud_device(as_t(points, torch.float32), as_t(rgbs, torch.float32), K, P,
                                          as_t(image_sizes, torch.int32), params)
    empty = np.zeros((0, 3), np.float32)
    return p.cpu().numpy(), c.cpu().numpy(), e.cpu().numpy(), empty, empty.copy()
