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
    its points into ALL of them at the rank's global row offset (mdi_emit_points_exchange, 16-byte
    stores; one multimem store per 16 bytes through the NVSwitch multicast address when the fabric
    offers one).  After a device barrier every rank holds the whole scene in view order: no NCCL data
    collective, no padding, no compaction pass.

    The per-rank row counts travel on the device too (mdi_exchange_counts: each rank stores its count
    into every rank's peer-mapped header and spins until all counts of the step have arrived), so the
    emit kernel is launched without a host synchronisation or an NCCL call in front of it; the host
    reads the counts back once, after everything has been queued.

    The returned clouds alias the exchange buffers: they are valid until the next step on any rank
    (the count exchange of the next step doubles as the "everybody is done reading" barrier for work
    enqueued on the current stream).  Construction and growth are collective.
    Raises if torch's symmetric memory is unavailable; callers then use `allgather_points`."""

    MAX_RANKS = 16   # MDI_MAX_EMIT_DST

    def __init__(self, device, group=None, capacity_rows: int = 1 << 20, byte_rgb: bool = False):
        """byte_rgb (the same on every rank; uint8 images only): peers receive the gathered rgb as
        3 bytes per point and expand them locally (u8 / 255, bit-identical) -- 15 instead of 24 bytes
        of NVLink ingress per remote point."""
        import torch.distributed._symmetric_memory as symm_mem
        from . import _lib as L
        self._symm, self._L = symm_mem, L
        self.group = group if group is not None else dist.group.WORLD
        self.device = torch.device(device)
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > self.MAX_RANKS:
            raise ValueError("CloudExchange: at most 16 ranks (MDI_MAX_EMIT_DST)")
        self.capacity = 0
        self.byte_rgb = bool(byte_rgb)
        self._alloc(int(capacity_rows))
        # count-exchange headers: 2 slot sets x world int64 per rank, zero before anybody publishes
        self._hdr = symm_mem.empty(2 * self.MAX_RANKS, dtype=torch.int64, device=self.device)
        self._hdr.zero_()
        self._hh = symm_mem.rendezvous(self._hdr, group=self.group)
        self.hdr_ptrs = [int(p) for p in self._hh.buffer_ptrs]
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)   # every header is zero before the first publish
        self.epoch = 0
        self.record = torch.zeros(3 + self.MAX_RANKS, dtype=torch.int64, device=self.device)
        self._record_pin = torch.zeros(3 + self.MAX_RANKS + 2, dtype=torch.int64, pin_memory=True)
        self._zero2 = torch.zeros(2, dtype=torch.int64, device=self.device)
        self.counts, self.total = [0] * self.world, 0

    def _alloc(self, rows: int) -> None:
        self.xyz = self._symm.empty(rows * 3, dtype=torch.float32, device=self.device)
        self.rgb = self._symm.empty(rows * 3, dtype=torch.float32, device=self.device)
        self._hx = self._symm.rendezvous(self.xyz, group=self.group)
        self._hr = self._symm.rendezvous(self.rgb, group=self.group)
        self.xyz_ptrs = [int(p) for p in self._hx.buffer_ptrs]
        self.rgb_ptrs = [int(p) for p in self._hr.buffer_ptrs]
        self.rgb8_ptrs = None
        if self.byte_rgb:
            self.rgb8 = self._symm.empty(rows * 3 + 16, dtype=torch.uint8, device=self.device)
            self._h8 = self._symm.rendezvous(self.rgb8, group=self.group)
            self.rgb8_ptrs = [int(p) for p in self._h8.buffer_ptrs]
        # NVSwitch multicast addresses of the two clouds (0 when the fabric has no NVLS)
        self.xyz_mc = int(getattr(self._hx, "multicast_ptr", 0) or 0)
        self.rgb_mc = int(getattr(self._hr, "multicast_ptr", 0) or 0)
        # MDI_MULTICAST=1 / 0 forces the choice; default: unicast peer stores
        env = os.environ.get("MDI_MULTICAST")
        self.use_multicast = bool(self.xyz_mc and self.rgb_mc) and env == "1"
        self.capacity = rows

    # ---- device-side protocol --------------------------------------------------------------------
    def publish_counts(self, local_ptr, stream: int) -> None:
        """Collective, stream ordered.  local_ptr: device int64 [2] = (rows of this rank, retry flag),
        None for a rank without views.  Fills self.record (see mdi_exchange_counts)."""
        import ctypes as C
        L = self._L
        self.epoch += 1
        hp = (C.c_void_p * self.world)(*self.hdr_ptrs)
        L.check(L.load().mdi_exchange_counts(hp, self.rank, self.world, self.epoch,
                                             local_ptr if local_ptr is not None else self._zero2.data_ptr(),
                                             self.record.data_ptr(), stream), "mdi_exchange_counts")

    def expand_bytes(self, stream: int) -> None:
        """After the step's barrier (byte_rgb): fill the other ranks' rows of this rank's float rgb
        cloud from the bytes they sent.  Local, stream ordered."""
        L = self._L
        L.check(L.load().mdi_expand_rgb8(self.rgb8.data_ptr(), self.rgb.data_ptr(), self.record.data_ptr(), self.rank,
                                         int(self.capacity), stream), "mdi_expand_rgb8")

    def barrier(self) -> None:
        """Collective, stream ordered: returns (on the stream) once every rank's emit kernel of this
        step has completed, i.e. this rank's clouds are whole."""
        self._hx.barrier()

    def read_record(self, extra: torch.Tensor = None):
        """One pinned read-back + stream synchronisation: -> (row_offset, total rows of all ranks,
        any-rank retry flag, counts list, extra values).  `extra`: up to 2 more device int64 to fetch."""
        n = 3 + self.world
        self._record_pin[:n].copy_(self.record[:n], non_blocking=True)
        if extra is not None:
            self._record_pin[n:n + extra.numel()].copy_(extra, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        vals = self._record_pin.tolist()
        if vals[1] < 0:
            raise self._L.MdiError("CloudExchange: a peer rank did not publish its row count (mdi_exchange_counts timed out)")
        self.counts = [int(c) for c in vals[3:n]]
        self.total = int(vals[1])
        return int(vals[0]), self.total, bool(vals[2]), self.counts, vals[n:n + (extra.numel() if extra is not None else 0)]

    def grow(self, rows_needed: int) -> None:
        """Collective (every rank sees the same total): reallocate clouds that hold rows_needed rows."""
        rows = max(self.capacity, 1)
        while rows < rows_needed:
            rows *= 2
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)   # nobody still writes into the old clouds
        self._alloc(rows)

    def clouds(self):
        """-> (xyz [total,3], rgb [total,3]) views of the exchange buffers."""
        return self.xyz[: 3 * self.total].view(-1, 3), self.rgb[: 3 * self.total].view(-1, 3)


def bind_process_to_gpu_node(device_index: int) -> dict:
    """Pins this process (and with it the first-touch placement of the pinned staging buffers it
    allocates afterwards) to the CPUs of the NUMA node its GPU hangs off.  One process per GPU: without
    this every rank's host buffers land on the node the launcher happened to start on, and the
    host->device copies of the far GPUs cross the socket interconnect.  Reads sysfs only; returns what
    it found ({"numa_node": n | None, "cpus": k, "bound": bool}) and does nothing when the topology is
    not exposed (containers often report node -1)."""
    info = {"numa_node": None, "cpus": 0, "bound": False}
    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return info
        info["numa_node"] = node
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
            info.update(cpus=len(allowed), bound=True)
    except (OSError, ValueError, AttributeError):
        pass
    return info
