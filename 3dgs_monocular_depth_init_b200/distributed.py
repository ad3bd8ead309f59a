"""Multi-GPU: views are block-partitioned over ranks (one process per GPU); each rank runs the
whole per-view pipeline on its own views with no data-path collective, then the per-view point
clouds are concatenated in rank (= view) order by one NCCL all-gather exchange, so the result is
bit-identical to the single-GPU order.  gloo works for the CPU tests."""
from __future__ import annotations

import torch
import torch.distributed as dist


def view_partition(n_views: int, world_size: int, rank: int) -> range:
    """Contiguous block of views owned by `rank` (dataset order)."""
    per = -(-n_views // world_size)
    return range(min(rank * per, n_views), min((rank + 1) * per, n_views))


def allgather_points(xyz: torch.Tensor, rgb: torch.Tensor, group=None):
    """Variable-length all-gather of [n_r, 3] clouds -> ([sum n_r, 3], [sum n_r, 3], counts[G]).
    Header exchange (counts), then one padded all_gather_into_tensor of xyz|rgb rows and a
    rank-ordered compaction."""
    world = dist.get_world_size(group)
    if world == 1:
        return xyz, rgb, torch.tensor([xyz.shape[0]], dtype=torch.int64)
    n_local = torch.tensor([xyz.shape[0]], dtype=torch.int64, device=xyz.device)
    counts = torch.empty(world, dtype=torch.int64, device=xyz.device)
    dist.all_gather_into_tensor(counts, n_local, group=group)
    counts_h = counts.cpu()
    n_max = int(counts_h.max())
    send = torch.zeros((n_max, 6), dtype=torch.float32, device=xyz.device)
    send[: xyz.shape[0], :3] = xyz
    send[: xyz.shape[0], 3:] = rgb
    recv = torch.empty((world, n_max, 6), dtype=torch.float32, device=xyz.device)
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
    parts = [recv[r, : int(counts_h[r])] for r in range(world)]
    out = torch.cat(parts, dim=0)
    return out[:, :3].contiguous(), out[:, 3:].contiguous(), counts_h
