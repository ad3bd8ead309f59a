"""Multi-GPU: views are block-partitioned over ranks (one process per GPU); each rank runs the
whole per-view pipeline on its own views with no data-path collective, then the per-view point
clouds are concatenated in rank (= view) order by one NCCL all-gather exchange, so the result is
bit-identical to the single-GPU order.  gloo works for the CPU tests."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def view_partition(n_views: int, world_size: int, rank: int) -> range:
    """Contiguous block of views owned by `rank` (dataset order)."""
    per = -(-n_views // world_size)
    return range(min(rank * per, n_views), min((rank + 1) * per, n_views))


def allgather_points(xyz: torch.Tensor, rgb: torch.Tensor, group=None):
    """Variable-length all-gather of [n_r, 3] clouds -> ([sum n_r, 3], [sum n_r, 3], counts[G]) in
    rank (= view) order.  Header exchange (counts), then one all_gather_into_tensor per array padded
    to the largest count (padding rows are never initialised nor read) and a rank-ordered
    compaction.  NCCL's grouped send/recv of the exact sizes was measured slower on NVLink (fewer
    channels than the collective), so the padded collective stays."""
    world = dist.get_world_size(group)
    if world == 1:
        return xyz, rgb, torch.tensor([xyz.shape[0]], dtype=torch.int64)
    n_local = torch.tensor([xyz.shape[0]], dtype=torch.int64, device=xyz.device)
    counts = torch.empty(world, dtype=torch.int64, device=xyz.device)
    dist.all_gather_into_tensor(counts, n_local, group=group)
    counts_h = counts.cpu()
    cnt = [int(c) for c in counts_h]
    n_max = max(cnt)
    outs = []
    for src in (xyz, rgb):
        if n_max == 0:
            outs.append(torch.empty((0, 3), dtype=src.dtype, device=src.device))
            continue
        if src.shape[0] == n_max and src.is_contiguous():
            send = src
        else:
            send = torch.empty((n_max, 3), dtype=src.dtype, device=src.device)
            send[: src.shape[0]].copy_(src)
        recv = torch.empty((world, n_max, 3), dtype=src.dtype, device=src.device)
        dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
        outs.append(torch.cat([recv[r, : cnt[r]] for r in range(world)], dim=0))
    return outs[0], outs[1], counts_h


class CloudExchange:
    """The exchange step of SURVEY 8e fused into the emit kernel: every rank owns a symmetric
    (peer-mapped over NVLink / NVSwitch) pair of output clouds, and each rank's emit kernel writes
    its points into ALL of them at the rank's global row offset (mdi_emit_points_multi).  After a
    barrier every rank holds the whole scene in view order: no NCCL data collective, no padding,
    no compaction pass.  Only the per-rank counts travel through NCCL (one 8-byte all-gather).

    The returned clouds alias the exchange buffers: they are valid until the next `begin` on any
    rank (the count exchange of the next scene doubles as the "everybody is done reading" barrier
    for work enqueued on the current stream).  Construction and growth are collective.
    Raises if torch's symmetric memory is unavailable; callers then use `allgather_points`."""

    def __init__(self, device, group=None, capacity_rows: int = 1 << 20):
        import torch.distributed._symmetric_memory as symm_mem
        self._symm = symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.device = torch.device(device)
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > 16:
            raise ValueError("CloudExchange: at most 16 ranks (MDI_MAX_EMIT_DST)")
        self.capacity = 0
        self._alloc(int(capacity_rows))

    def _alloc(self, rows: int) -> None:
        self.xyz = self._symm.empty(rows * 3, dtype=torch.float32, device=self.device)
        self.rgb = self._symm.empty(rows * 3, dtype=torch.float32, device=self.device)
        self._hx = self._symm.rendezvous(self.xyz, group=self.group)
        self._hr = self._symm.rendezvous(self.rgb, group=self.group)
        self.xyz_ptrs = [int(p) for p in self._hx.buffer_ptrs]
        self.rgb_ptrs = [int(p) for p in self._hr.buffer_ptrs]
        # NVSwitch multicast addresses of the two clouds (0 when the fabric has no NVLS)
        self.xyz_mc = int(getattr(self._hx, "multicast_ptr", 0) or 0)
        self.rgb_mc = int(getattr(self._hr, "multicast_ptr", 0) or 0)
        # opt-in: with scalar multimem stores the multicast emit measured slower than the unicast one
        # on 2 GPUs (0.36 vs 0.18 ms); it needs 16-byte vector stores to pay off
        self.use_multicast = bool(self.xyz_mc and self.rgb_mc) and os.environ.get("MDI_MULTICAST") == "1"
        self.capacity = rows

    def begin(self, n_local: int):
        """Collective.  -> (row_offset of this rank, total rows, counts list)."""
        n = torch.tensor([int(n_local)], dtype=torch.int64, device=self.device)
        counts = torch.empty(self.world, dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(counts, n, group=self.group)
        cnt = [int(c) for c in counts.cpu()]
        total = sum(cnt)
        if total > self.capacity:  # same decision on every rank
            rows = self.capacity
            while rows < total:
                rows *= 2
            self._alloc(rows)
        self.counts = cnt
        self.total = total
        return sum(cnt[: self.rank]), total, cnt

    def finish(self):
        """Collective, stream ordered: returns once every rank's emit kernel has completed, i.e.
        this rank's clouds are whole.  -> (xyz [total,3], rgb [total,3]) views of the buffers."""
        self._hx.barrier()
        return self.xyz[: 3 * self.total].view(-1, 3), self.rgb[: 3 * self.total].view(-1, 3)
