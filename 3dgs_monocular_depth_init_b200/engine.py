"""Batched host driver of the kernels: builds the per-view descriptor table, owns the device
workspaces (torch tensors: plumbing only) and sequences the C-ABI calls on the current CUDA
stream.  Both the per-view drop-in functions and the scene-level entry point go through
`Engine.run`; nothing here computes on the CPU except inv(K) of a1 (3x3, batched).

Reference path restated: points_from_depth.py:215-329 (per view) and
monocular_depth_init.py:120-184 (scene loop), batched over views.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib as L

TensorOrList = Union[torch.Tensor, Sequence[torch.Tensor]]

# view ranges of one run queued on separate CUDA streams (RunOptions.streams overrides)
DEFAULT_STREAMS = max(1, int(os.environ.get("MDI_STREAMS", "1") or 1))
MIN_VIEWS_PER_RANGE = 16
# a7 statistics on the raw depth beside the alignment kernels, which then run on a high-priority side
# stream (mdi_depth_range with MDI_STATS_RAW + mdi_depth_range_map).  Bit-identical, but off by default
# (MDI_RAW_STATS=1 turns it on): on C2 the two chains share the SMs poorly (1.10 instead of 1.14 ms per
# step when no view needs the fallback), and a view whose fit is not increasing or reaches the sentinel
# -- the garbage fits the reference's iteration-0 exit leaves in 1-3 % of the views -- sends the batch
# through the exact-path retry (1.75 ms)
OVERLAP_STATS = os.environ.get("MDI_RAW_STATS", "0") == "1"


def patch_geometry(H: int, W: int, n_small: int):
    """(patch_h, patch_w, grid_h, grid_w) -- num_sfm_points_mask.py:7-35."""
    small_is_h = H <= W
    small, large = (H, W) if small_is_h else (W, H)
    ps_small = small // n_small
    if ps_small <= 0:
        raise ZeroDivisionError("image smaller than num_patches_small_axis")
    n_large = -(-large // ps_small)
    ps_large = large // n_large
    if small_is_h:
        return ps_small, ps_large, n_small, n_large
    return ps_large, ps_small, n_large, n_small


@dataclass
class ViewInputs:
    """Inputs of a batch of views.  depth / mask / image are either one stacked tensor
    ([V,H,W], [V,H,W], [V,H,W,3]) or per-view lists (sizes may differ)."""

    depth: TensorOrList                       # float32 cuda
    mask: Optional[TensorOrList]              # bool/uint8 cuda; None (or None entries) = all valid
    image: Optional[TensorOrList]             # float32 / uint8, cuda or pinned host; None = no rgb
    image_kind: int                           # L.IMAGE_*
    K: torch.Tensor                           # [V,3,3] cpu
    c2w: torch.Tensor                         # [V,4,4] cpu
    sfm_xyz: Optional[torch.Tensor] = None    # [Np,3] float32 cuda
    sfm_index: Optional[torch.Tensor] = None  # int32 cuda, concatenated per view (padded slots)
    sfm_counts: Optional[np.ndarray] = None   # [V] points per view
    # alternative to projection (per-view align API): filtered coords given, slot layout
    uv: Optional[torch.Tensor] = None         # [2,S] int32 cuda
    z: Optional[torch.Tensor] = None          # [S] float32 cuda
    n_valid: Optional[torch.Tensor] = None    # [V] int32 cuda
    P_override: Optional[torch.Tensor] = None  # [V,3,4] float32 cpu: use instead of computing P


@dataclass
class RunOptions:
    aligner: Optional[str] = "lstsqrs"       # "lstsqrs" | "ransac" | "msac" | None (scale_shift given)
    scale_shift: Optional[torch.Tensor] = None  # [V,2] float32 cuda when aligner is None
    ransac_threshold: float = 0.01
    ransac_max_iters: int = 2500
    ransac_confidence: float = 0.999
    ransac_sample_size: int = 4
    ransac_min_iters: int = 0
    ransac_seed: int = 0
    sample_idx: Optional[torch.Tensor] = None   # [V,max_iters,sample] int32 cuda (parity runs)
    subsample: object = 10                       # int or "adaptive"
    factor_min: int = 5
    factor_max: int = 15
    use_patch_mask: bool = True
    num_patches_small_axis: int = 20
    patch_threshold: int = 15
    grad_thresh: Optional[float] = None
    flags: int = L.KEEP_PRED_MASK | L.KEEP_SENTINEL | L.KEEP_NONNEG | L.KEEP_SUBSAMPLE
    want_keep_mask: bool = True
    want_xyz: bool = True
    want_rgb: bool = True
    want_inliers: bool = False
    stop_after: Optional[str] = None             # "project" | "fit" | "mask": partial pipelines
    stats_exact: bool = False                    # force the three-pass exact quantile select
    exchange: object = None                      # distributed.CloudExchange: emit into every rank's cloud
    speculative_emit: bool = True                # launch a10/a11 before the output size reaches the host
    defer_emit: bool = False                     # stop after the mask; Engine.emit_deferred() writes the cloud later
    streams: Optional[int] = None                # view ranges queued on separate streams (None: MDI_STREAMS, default 1)
    raw_stats: Optional[bool] = None             # a7 on the raw depth beside the alignment (None: MDI_RAW_STATS, default off)


@dataclass
class RunResult:
    n_views: int
    px_begin: np.ndarray                # [V+1] int64
    sfm_begin: np.ndarray               # [V+1] int64 (padded slot starts)
    status: Optional[torch.Tensor] = None        # [V] int32 cuda
    n_inbounds: Optional[torch.Tensor] = None
    n_valid: Optional[torch.Tensor] = None
    uv: Optional[torch.Tensor] = None   # [2,S] int32 cuda
    z: Optional[torch.Tensor] = None
    d: Optional[torch.Tensor] = None
    P: Optional[torch.Tensor] = None    # [V,3,4] float32 cuda
    sums: Optional[torch.Tensor] = None
    scale_shift: Optional[torch.Tensor] = None   # [V,2] float32 cuda
    ransac_iters: Optional[torch.Tensor] = None
    ransac_inliers: Optional[torch.Tensor] = None
    inlier_mask: Optional[torch.Tensor] = None
    lo_hi: Optional[torch.Tensor] = None
    quantiles: Optional[torch.Tensor] = None
    grad_range: Optional[torch.Tensor] = None
    patch_blocked: Optional[torch.Tensor] = None
    keep: Optional[torch.Tensor] = None  # [sum HW] uint8 cuda
    xyz: Optional[torch.Tensor] = None   # [N,3] float32 cuda
    rgb: Optional[torch.Tensor] = None   # [N,3] float32 cuda
    view_counts: Optional[torch.Tensor] = None   # [V] int64 cuda
    view_offsets: Optional[torch.Tensor] = None  # [V] int64 cuda
    total: int = 0
    extras: dict = field(default_factory=dict)


def _stream() -> int:
    """Raw handle of the current CUDA stream (the Stream-object route costs ~25 us per call)."""
    try:
        return int(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))
    except AttributeError:
        return torch.cuda.current_stream().cuda_stream


def _ptrs(x: Optional[TensorOrList], V: int, esize_check: Optional[int] = None):
    """(pointer array uint64 [V], shapes) for a stacked tensor or a list (None entries -> 0)."""
    if x is None:
        return np.zeros(V, dtype=np.uint64), None
    if isinstance(x, torch.Tensor):
        assert x.is_contiguous() and x.shape[0] == V
        stride = x[0].numel() * x.element_size() if V else 0
        p = x.data_ptr() + np.arange(V, dtype=np.uint64) * np.uint64(stride)
        return p.astype(np.uint64), [tuple(x.shape[1:])] * V
    p = np.zeros(V, dtype=np.uint64)
    shapes = []
    for i, t in enumerate(x):
        if t is None:
            shapes.append(None)
            continue
        assert t.is_contiguous()
        p[i] = t.data_ptr()
        shapes.append(tuple(t.shape))
    return p, shapes


# kernels launched by each C-ABI entry point (for the bench's `gpu_launches` claim)
KERNELS_PER_CALL = {
    "mdi_setup_cameras": 1, "mdi_upload_table": 1, "mdi_project_filter": 1, "mdi_setup_project_filter": 1, "mdi_gather_fit": 1, "mdi_ransac": 1,
    "mdi_apply_scale_shift": 1, "mdi_depth_range": 5, "mdi_depth_range_map": 1, "mdi_grad_range": 3, "mdi_factor_table": 1, "mdi_keep_mask": 4,
    "mdi_scan_tiles": 2, "mdi_emit_points": 1, "mdi_emit_points_capped": 1, "mdi_emit_points_multi": 1, "mdi_emit_points_multicast": 1,
    "mdi_emit_points_exchange": 1, "mdi_exchange_counts": 1, "mdi_expand_rgb8": 1,
    "mdi_min_extents": 1,
}


class _CountingLib:
    """Proxy over the ctypes library: counts kernel launches and, when the engine's `profile`
    flag is set, brackets every launching call with CUDA events on the current stream."""

    def __init__(self, lib, engine):
        self._lib, self._engine = lib, engine

    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        n = KERNELS_PER_CALL.get(name)
        if n is None:
            self.__dict__[name] = fn
            return fn
        eng = self._engine

        def call(*args):
            if eng.profile:
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                rc = fn(*args)
                e1.record()
                eng.events.append((name, e0, e1))
            else:
                rc = fn(*args)
            eng.launches += n
            return rc

        self.__dict__[name] = call  # later look-ups bypass __getattr__
        return call


class Engine:
    """Holds cached workspaces per device.  One instance per process / GPU."""

    def __init__(self, device: "torch.device | str" = "cuda"):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.MdiError("the B200 path runs on a CUDA device only (no CPU fallback)")
        self.lib = _CountingLib(L.load(), self)
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        L.check(self.lib.mdi_check_device(idx), "mdi_check_device")
        self._ws: dict[str, torch.Tensor] = {}
        self._pin: dict[str, torch.Tensor] = {}
        self._pin_ev: dict[str, torch.cuda.Event] = {}   # last device read of each pinned staging buffer
        self._plan = None   # (key, descriptor table + its uploaded image) of the previous run
        self.launches = 0   # kernels launched through this engine (bench `gpu_launches`)
        self.profile = False
        self.events: list = []  # (entry point, start event, end event) when profile is on

    def kernel_times_ms(self) -> dict:
        """Sums the recorded per-call device times by entry point and clears the record
        (call after a synchronize)."""
        out: dict = {}
        for name, e0, e1 in self.events:
            t = out.setdefault(name, [0.0, 0])
            t[0] += e0.elapsed_time(e1)
            t[1] += 1
        self.events = []
        return {k: {"ms": v[0], "calls": v[1]} for k, v in out.items()}

    # cached, grow-only workspace buffers (never shrink: results that alias them are only
    # valid until the next run() unless the caller asked for owned outputs)
    def _buf(self, name: str, n: int, dtype: torch.dtype) -> torch.Tensor:
        t = self._ws.get(name)
        if t is None or t.numel() < n or t.dtype != dtype:
            t = torch.empty(max(int(n * 1.25), 16), dtype=dtype, device=self.device)
            self._ws[name] = t
        return t[:n]

    # ---- view ranges on part streams ---------------------------------------------------------------
    def _part_stream(self, i: int) -> "torch.cuda.Stream":
        ps = self.__dict__.setdefault("_part_streams", [])
        while len(ps) <= i:
            ps.append(torch.cuda.Stream(self.device))
        return ps[i]

    @staticmethod
    def _view_ranges(vin: ViewInputs, opt: RunOptions, V: int, desc, adaptive: bool, have_proj: bool):
        """Contiguous view ranges of a run: one (the whole batch) unless RunOptions.streams / MDI_STREAMS
        asks for more and the run is the plain full pipeline (every kernel of it indexes its per-view
        arguments by the view's position in the range, so a range is the same call with offset
        pointers; the per-pixel mask kernel is the exception and keeps such runs in one range)."""
        n = opt.streams if opt.streams is not None else DEFAULT_STREAMS
        lattice = bool(opt.flags & L.KEEP_SUBSAMPLE) and (
            (adaptive and opt.factor_min >= 1 and opt.factor_max >= opt.factor_min and opt.factor_max - opt.factor_min < 4096)
            or (not adaptive and opt.subsample not in (0, -2**31)))
        if (n <= 1 or V < MIN_VIEWS_PER_RANGE * 2 or not have_proj or opt.stop_after not in (None, "mask")
                or opt.aligner not in ("lstsqrs", "ransac", "msac") or opt.grad_thresh is not None or opt.defer_emit
                or opt.stats_exact or not lattice or int(desc["H"].max()) >= 65536 or int(desc["W"].max()) >= 65536):
            return [(0, V)]
        n = min(n, V // MIN_VIEWS_PER_RANGE)
        cuts = [(V * i) // n for i in range(n + 1)]
        return [(cuts[i], cuts[i + 1]) for i in range(n)]

    # ---- descriptor table ---------------------------------------------------------------------
    def build_views(self, vin: ViewInputs, opt: RunOptions):
        V = vin.K.shape[0]
        dptr, dshapes = _ptrs(vin.depth, V)
        mptr, _ = _ptrs(vin.mask, V)
        iptr, _ = _ptrs(vin.image, V)
        if isinstance(vin.depth, torch.Tensor):
            assert vin.depth.dtype == torch.float32 and vin.depth.is_cuda
        else:
            for t in vin.depth:
                assert t.dtype == torch.float32 and t.is_cuda
        counts = (np.asarray(vin.sfm_counts, dtype=np.int64) if vin.sfm_counts is not None
                  else np.zeros(V, dtype=np.int64))
        # everything but the three pointer columns depends only on the shapes, the SfM counts and the
        # patch option: kept from the previous call when those repeat (same scene again, equal chunks)
        key = (V, tuple(dshapes) if not isinstance(vin.depth, torch.Tensor) else ("stacked",) + tuple(vin.depth.shape),
               counts.tobytes(), bool(opt.use_patch_mask), int(opt.num_patches_small_axis))
        cached = getattr(self, "_geom_cache", None)
        if cached is not None and cached[0] == key:
            desc, px_begin, tile_begin, sfm_begin = (a.copy() for a in cached[1])
        else:
            desc, px_begin, tile_begin, sfm_begin = self._build_geometry(V, dshapes, counts, opt)
            self._geom_cache = (key, tuple(a.copy() for a in (desc, px_begin, tile_begin, sfm_begin)))
        desc["depth"], desc["mask"], desc["image"] = dptr, mptr, iptr
        return desc, px_begin, tile_begin, sfm_begin

    @staticmethod
    def _build_geometry(V, dshapes, counts, opt):
        Hs = np.array([s[0] for s in dshapes], dtype=np.int64)
        Ws = np.array([s[1] for s in dshapes], dtype=np.int64)
        HW = Hs * Ws
        desc = np.zeros(V, dtype=L.VIEW_DTYPE)
        desc["H"], desc["W"] = Hs, Ws
        padded = (counts + 3) & ~3  # 16-byte aligned slot ranges (bulk copies in mdi_ransac)
        sfm_begin = np.zeros(V + 1, dtype=np.int64)
        np.cumsum(padded, out=sfm_begin[1:])
        desc["sfm_begin"], desc["sfm_count"] = sfm_begin[:-1], counts
        px_begin = np.zeros(V + 1, dtype=np.int64)
        np.cumsum(HW, out=px_begin[1:])
        n_tiles = -(-HW // L.TILE_PX)
        tile_begin = np.zeros(V + 1, dtype=np.int64)
        np.cumsum(n_tiles, out=tile_begin[1:])
        desc["px_begin"], desc["tile_begin"], desc["n_tiles"] = px_begin[:-1], tile_begin[:-1], n_tiles
        if opt.use_patch_mask and V and (Hs == Hs[0]).all() and (Ws == Ws[0]).all():
            g = patch_geometry(int(Hs[0]), int(Ws[0]), opt.num_patches_small_axis)
            if g[2] * g[3] > L.PATCH_STRIDE:
                raise L.MdiError(f"patch grid {g[2]}x{g[3]} exceeds {L.PATCH_STRIDE} patches")
            desc["patch_h"], desc["patch_w"], desc["grid_h"], desc["grid_w"] = g
        elif opt.use_patch_mask and V:
            geo = {}
            pg = np.empty((V, 4), dtype=np.int32)
            for i, key in enumerate(zip(Hs.tolist(), Ws.tolist())):
                g = geo.get(key)
                if g is None:
                    g = geo[key] = patch_geometry(key[0], key[1], opt.num_patches_small_axis)
                    if g[2] * g[3] > L.PATCH_STRIDE:
                        raise L.MdiError(f"patch grid {g[2]}x{g[3]} exceeds {L.PATCH_STRIDE} patches")
                pg[i] = g
            desc["patch_h"], desc["patch_w"], desc["grid_h"], desc["grid_w"] = pg.T
        else:
            desc["patch_h"] = desc["patch_w"] = 1
        return desc, px_begin, tile_begin, sfm_begin

    @staticmethod
    def _plan_key(vin: ViewInputs, opt: RunOptions, V: int, kn: np.ndarray, c2wn: np.ndarray):
        """Identity of everything the uploaded descriptor table depends on, or None when the inputs
        are per-view lists / carry an externally supplied P (those always take the full path)."""
        if vin.P_override is not None:
            return None
        parts = []
        for x in (vin.depth, vin.mask, vin.image):
            if x is None:
                parts.append(None)
            elif isinstance(x, torch.Tensor) and x.is_contiguous() and x.shape[0] == V:
                parts.append((x.data_ptr(), tuple(x.shape), x.dtype, x.is_cuda))
            else:
                return None
        counts = None if vin.sfm_counts is None else np.asarray(vin.sfm_counts, dtype=np.int64).tobytes()
        return (V, tuple(parts), counts, kn.tobytes(), c2wn.tobytes(), bool(opt.use_patch_mask),
                int(opt.num_patches_small_axis))

    def _upload(self, name: str, parts, owned: bool = False) -> torch.Tensor:
        """One H2D copy of the concatenated byte images of `parts` (numpy arrays) through a
        persistent pinned staging buffer.  The buffer is rewritten only after the copy that last read
        it has completed (partial pipelines -- stop_after="project" / "fit" -- return without a
        stream synchronisation)."""
        sizes = [int(a.nbytes) for a in parts]
        n = sum(sizes)
        ev = self._pin_ev.pop(name, None)
        if ev is not None:
            ev.synchronize()
        pin = self._pin.get(name)
        if pin is None or pin.numel() < n:
            pin = self._pin[name] = torch.empty(max(int(n * 1.25), 256), dtype=torch.uint8, pin_memory=True)
        host = pin.numpy()
        o = 0
        for a, sz in zip(parts, sizes):
            host[o:o + sz] = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
            o += sz
        if owned and n % 4 == 0:
            # a block of a chunked host run: bulk copies of later blocks are already queued on the copy
            # engine, and a memcpy would wait behind all of them -- the SMs fetch the table instead
            dev = torch.empty(n, dtype=torch.uint8, device=self.device)
            L.check(self.lib.mdi_upload_table(dev.data_ptr(), pin.data_ptr(), n, _stream()), "mdi_upload_table")
        else:
            dev = torch.empty(n, dtype=torch.uint8, device=self.device) if owned else self._buf(name, n, torch.uint8)
            dev.copy_(pin[:n], non_blocking=True)
        done = torch.cuda.Event()
        done.record()
        self._pin_ev[name] = done
        return dev

    # ---- the pipeline -------------------------------------------------------------------------
    def run(self, vin: ViewInputs, opt: RunOptions) -> RunResult:
        lib, st = self.lib, _stream()
        run_start = torch.cuda.Event()
        run_start.record()
        V = int(vin.K.shape[0])
        kn = np.ascontiguousarray(vin.K.detach().cpu().numpy(), dtype=np.float32).reshape(V, 3, 3)
        c2wn = np.ascontiguousarray(vin.c2w.detach().cpu().numpy(), dtype=np.float32).reshape(V, 4, 4)
        # The same resident views again (every step of a training-set sweep, every bench step): the
        # descriptor table, K and cam2world already sit on the device from the previous call, and
        # rebuilding + re-uploading them is ~0.1 ms of host time in front of the first launch.
        # (a deferred-emit run owns its table until Engine.emit_deferred: it neither takes nor leaves a plan)
        key = self._plan_key(vin, opt, V, kn, c2wn) if V and not opt.defer_emit else None
        plan = self._plan[1] if key is not None and self._plan is not None and self._plan[0] == key else None
        if plan is not None:
            desc, px_begin, tile_begin, sfm_begin, blob, pinhole = plan
            px_begin, sfm_begin = px_begin.copy(), sfm_begin.copy()
        else:
            self._plan = None
            desc, px_begin, tile_begin, sfm_begin = self.build_views(vin, opt)
        total_tiles = int(tile_begin[-1])
        max_tiles = int(desc["n_tiles"].max()) if V else 0
        slots = int(sfm_begin[-1]) + 4
        res = RunResult(n_views=V, px_begin=px_begin, sfm_begin=sfm_begin)
        if V == 0:
            if opt.exchange is not None and opt.stop_after is None and (opt.want_xyz or opt.want_rgb):
                # a rank without views still takes part in the collective exchange
                self._exchange_rounds(opt.exchange, res, None, None, None)
            return res
        if plan is None:
            # a1: cameras.  One upload: [descriptors][K][cam2world][inv(K) unless pinhole]
            pinhole = bool(((kn[:, 0, 1] == 0) & (kn[:, 1, 0] == 0) & (kn[:, 2, 0] == 0) & (kn[:, 2, 1] == 0)
                            & (kn[:, 2, 2] == 1)).all())
            # zero-skew pinhole intrinsics: inv(K) in closed form on the device (bit-identical to
            # torch.linalg.inv); anything else: batched LU inverse on the host like the reference
            if vin.P_override is not None:  # stage-wise parity / project_and_filter_sfm_pts: P as given
                desc["P"] = vin.P_override.reshape(V, 12).numpy()
            blob_parts = [desc.view(np.uint8).reshape(-1), kn, c2wn]
            if not pinhole:
                blob_parts.append(torch.linalg.inv(torch.from_numpy(kn)).numpy())
            blob = self._upload("views", blob_parts, owned=opt.defer_emit)
            if key is not None:
                self._plan = (key, (desc, px_begin.copy(), tile_begin, sfm_begin.copy(), blob, pinhole))
        vp = blob.data_ptr()
        kp = vp + desc.nbytes
        c2wp = kp + 4 * 9 * V
        kinvp = None if pinhole else c2wp + 4 * 16 * V
        res.P = torch.empty((V, 3, 4), dtype=torch.float32, device=self.device)
        i32, f32 = torch.int32, torch.float32
        res.status = torch.zeros(V, dtype=i32, device=self.device)
        VS = desc.itemsize
        adaptive = opt.subsample == "adaptive"
        if not adaptive and not isinstance(opt.subsample, int) and opt.stop_after not in ("project", "fit"):
            raise ValueError(f"Unsupported subsampling factor: {opt.subsample}")
        have_proj = vin.uv is None and vin.sfm_xyz is not None
        ranges = self._view_ranges(vin, opt, V, desc, adaptive, have_proj)
        parts = len(ranges)
        keep_P = 0 if vin.P_override is None else 1
        # a1 rides in the projection launch when the whole batch is projected by one (mdi_setup_project_filter)
        fuse_cam = have_proj and parts == 1
        if not fuse_cam:
            L.check(lib.mdi_setup_cameras(vp, V, kp, c2wp, kinvp, res.P.data_ptr(), keep_P, st), "mdi_setup_cameras")

        def at(t, a, per_view=1):
            """address of view a's row in a per-view array (None stays NULL)"""
            return None if t is None else t.data_ptr() + a * per_view * t.element_size()

        use_stats = bool(adaptive and (opt.flags & L.KEEP_SUBSAMPLE) and opt.stop_after not in ("project", "fit"))
        # views under 64K pixels give the one-pass path too small a sample to bracket: exact select
        first_mode = L.STATS_EXACT if (opt.stats_exact or max_tiles <= 16) else 0
        pop_flags = opt.flags & (L.KEEP_PRED_MASK | L.KEEP_SENTINEL)
        want_raw = opt.raw_stats if opt.raw_stats is not None else (
            OVERLAP_STATS and have_proj and opt.aligner in ("lstsqrs", "ransac", "msac"))
        raw_stats = bool(use_stats and want_raw and first_mode == 0 and parts == 1 and not opt.defer_emit
                         and opt.stop_after in (None, "mask"))

        # ---- a2: outputs and, when the batch is one range on this stream, the launch itself -- the device
        # gets its first kernel before the host allocates everything else
        if vin.uv is not None:
            res.uv, res.z, res.n_valid = vin.uv, vin.z, vin.n_valid
            assert res.uv.shape[1] >= slots - 4 and res.uv.dtype == i32
            slots = int(res.uv.shape[1])
        elif have_proj:
            res.uv = torch.empty((2, slots), dtype=i32, device=self.device)
            res.z = torch.empty(slots, dtype=f32, device=self.device)
            res.n_inbounds = torch.empty(V, dtype=i32, device=self.device)
            res.n_valid = torch.empty(V, dtype=i32, device=self.device)

        def launch_project(a: int, b: int, s: int):
            """a2 of views a..b-1: projection + filtering"""
            if fuse_cam:
                L.check(lib.mdi_setup_project_filter(vp, V, kp, c2wp, kinvp, res.P.data_ptr(), keep_P,
                                                     vin.sfm_xyz.data_ptr(), L.ptr(vin.sfm_index), slots,
                                                     res.uv.data_ptr(), res.z.data_ptr(), res.n_inbounds.data_ptr(),
                                                     res.n_valid.data_ptr(), res.status.data_ptr(), s),
                        "mdi_setup_project_filter")
                return
            L.check(lib.mdi_project_filter(vp + a * VS, b - a, vin.sfm_xyz.data_ptr(), L.ptr(vin.sfm_index), slots,
                                           res.uv.data_ptr(), res.z.data_ptr(), at(res.n_inbounds, a),
                                           at(res.n_valid, a), at(res.status, a), s),
                    "mdi_project_filter")

        projected = False
        if parts == 1 and have_proj and not raw_stats:
            launch_project(0, V, st)
            projected = True
        if opt.stop_after == "project" and parts == 1 and not raw_stats:
            return res

        # ---- a7 workspaces: with an increasing fit the order statistics of the aligned depth are the
        # aligned order statistics of the RAW depth, so the one pass over the pixels does not have to wait
        # for the fit (raw_stats: queued beside projection / fit / RANSAC)
        # [view counts V][view offsets V][total][stats retry flag]
        meta = torch.zeros(2 * V + 2, dtype=torch.int64, device=self.device)
        ws_part = [0] * parts
        ws = None
        if use_stats:
            wsb = int(lib.mdi_depth_range_workspace_bytes(V, max_tiles))
            if parts > 1:   # one workspace per view range, 256-byte aligned; the whole-batch retry reuses the buffer
                o = 0
                for i, (a, b) in enumerate(ranges):
                    ws_part[i] = o
                    o += (int(lib.mdi_depth_range_workspace_bytes(b - a, max_tiles)) + 255) & ~255
                wsb = max(wsb, o)
            ws = self._buf("stats_ws", wsb, torch.uint8)
            res.lo_hi = torch.empty((V, 2), dtype=f32, device=self.device)
            res.quantiles = torch.empty((V, 4), dtype=f32, device=self.device)

        # ---- a3 / a4 outputs (all allocated on the calling stream before anything is queued on a part
        # stream, so that the allocator never sees those streams)
        have_pts = res.uv is not None
        use_patch = bool(opt.use_patch_mask and have_pts)
        want_fit = opt.stop_after != "project"
        write_fit = opt.aligner == "lstsqrs"
        cfg = None
        if have_pts and want_fit:
            res.d = torch.zeros(slots, dtype=f32, device=self.device)
            res.sums = torch.empty((V, 5), dtype=torch.float64, device=self.device)
            if write_fit:
                res.scale_shift = torch.empty((V, 2), dtype=f32, device=self.device)
            if use_patch:
                res.patch_blocked = torch.empty((V, L.PATCH_STRIDE), dtype=torch.uint8, device=self.device)
        if opt.aligner in ("ransac", "msac") and want_fit:
            assert have_pts, "RANSAC needs SfM points"
            cfg = L.RansacConfigC(float(opt.ransac_threshold), int(opt.ransac_max_iters),
                                  float(opt.ransac_confidence), int(opt.ransac_sample_size),
                                  int(opt.ransac_min_iters),
                                  L.LOSS_RANSAC if opt.aligner == "ransac" else L.LOSS_MSAC, 0,
                                  int(opt.ransac_seed) & (2**64 - 1))
            res.scale_shift = torch.empty((V, 2), dtype=f32, device=self.device)
            res.ransac_iters = torch.empty(V, dtype=i32, device=self.device)
            res.ransac_inliers = torch.empty(V, dtype=i32, device=self.device)
            if opt.want_inliers:
                res.inlier_mask = torch.zeros(slots, dtype=torch.uint8, device=self.device)
            if opt.sample_idx is not None:
                assert opt.sample_idx.dtype == i32 and opt.sample_idx.is_cuda and opt.sample_idx.is_contiguous()
                assert tuple(opt.sample_idx.shape) == (V, opt.ransac_max_iters, opt.ransac_sample_size)
            max_pts = int(desc["sfm_count"].max())
        elif opt.aligner is None:
            res.scale_shift = opt.scale_shift
            if res.scale_shift is not None:
                assert res.scale_shift.dtype == f32 and tuple(res.scale_shift.shape) == (V, 2)
        elif opt.aligner not in ("lstsqrs", "ransac", "msac"):
            raise NotImplementedError(f"aligner {opt.aligner!r} is not on the B200 path")

        def launch_fit(a: int, b: int, s: int):
            """a3 + a8 (gather, normal equations, patch histogram) and a4 (RANSAC / MSAC) of views a..b-1"""
            if have_pts:
                L.check(lib.mdi_gather_fit(vp + a * VS, b - a, slots, res.uv.data_ptr(), res.z.data_ptr(), at(res.n_valid, a),
                                           res.d.data_ptr(), at(res.sums, a, 5), at(res.scale_shift, a, 2) if write_fit else None,
                                           1 if write_fit else 0, int(opt.patch_threshold),
                                           at(res.patch_blocked, a, L.PATCH_STRIDE), s), "mdi_gather_fit")
            if cfg is not None:
                cfg.view_base = a   # the device sampler draws per batch view, whatever the launch ranges are
                L.check(lib.mdi_ransac(vp + a * VS, b - a, res.d.data_ptr(), res.z.data_ptr(), at(res.n_valid, a),
                                       at(res.status, a), at(opt.sample_idx, a, opt.ransac_max_iters * opt.ransac_sample_size),
                                       C.byref(cfg), max_pts, at(res.scale_shift, a, 2), L.ptr(res.inlier_mask),
                                       at(res.ransac_iters, a), at(res.ransac_inliers, a), s),
                        "mdi_ransac")

        fit_done = None
        if parts == 1 and raw_stats:
            # the alignment kernels (one CTA per view, latency-bound) go to a HIGH-priority side stream and
            # the statistics pass over the raw depth stays here: the SMs take the alignment CTAs as soon as
            # they are ready and fill up with the streaming pass in between
            if not hasattr(self, "_align_stream"):
                hi_prio = torch.cuda.Stream.priority_range()[1] if hasattr(torch.cuda.Stream, "priority_range") else -1
                self._align_stream = torch.cuda.Stream(self.device, priority=hi_prio)
            sa = self._align_stream
            cams = torch.cuda.Event()
            cams.record()
            sa.wait_event(cams)
            with torch.cuda.stream(sa):
                s_a = _stream()
                if have_proj:
                    launch_project(0, V, s_a)
                launch_fit(0, V, s_a)
                fit_done = torch.cuda.Event()
                fit_done.record(sa)
            L.check(lib.mdi_depth_range(vp, V, max_tiles, None, None, pop_flags | L.STATS_RAW, ws.data_ptr(),
                                        None, None, None, st), "mdi_depth_range")
        elif parts == 1:
            if have_proj and not projected:
                launch_project(0, V, st)
            if opt.stop_after == "project":
                return res
            launch_fit(0, V, st)
            if opt.stop_after == "fit":
                return res

        # a7 / a9: per-view statistics
        ftab = None
        flags = opt.flags
        if use_patch:
            flags |= L.KEEP_PATCH
        ssp = L.ptr(res.scale_shift)
        res.view_counts, res.view_offsets = meta[:V], meta[V:2 * V]
        if use_stats:
            ftab = self._buf("factor_table", V * L.FACTOR_TABLE_STRIDE, f32)
        if opt.grad_thresh is not None:
            flags |= L.KEEP_GRAD
            res.grad_range = torch.empty((V, 2), dtype=f32, device=self.device)
            L.check(lib.mdi_grad_range(vp, V, max_tiles, ssp, res.status.data_ptr(),
                                       res.grad_range.data_ptr(), st), "mdi_grad_range")
        scfg = L.SubsampleConfigC(L.SUBSAMPLE_ADAPTIVE if adaptive else L.SUBSAMPLE_STATIC,
                                  0 if adaptive else int(opt.subsample), int(opt.factor_min),
                                  int(opt.factor_max), int(flags),
                                  float(opt.grad_thresh) if opt.grad_thresh is not None else 0.0)
        if opt.want_keep_mask:
            res.keep = torch.empty(int(px_begin[-1]), dtype=torch.uint8, device=self.device)
        if opt.defer_emit:  # the emit of this run comes after later runs: its inputs must not be recycled
            bits = torch.empty(total_tiles * L.TILE_WORDS, dtype=i32, device=self.device)
            tile_offsets = torch.empty(total_tiles, dtype=torch.int64, device=self.device)
        else:
            bits = self._buf("bits", total_tiles * L.TILE_WORDS, i32)
            tile_offsets = self._buf("tile_offsets", total_tiles, torch.int64)
        tile_counts = self._buf("tile_counts", total_tiles, i32)
        # zero-fill of the mask outputs on a side stream: it only has to wait for what was queued before
        # this call, so it runs under the alignment / statistics kernels instead of after them
        main = torch.cuda.current_stream(self.device)
        if not hasattr(self, "_zero_stream"):
            self._zero_stream = torch.cuda.Stream(self.device)
        zs = self._zero_stream
        zs.wait_event(run_start)
        with torch.cuda.stream(zs):
            if res.keep is not None:
                res.keep.record_stream(zs)
                res.keep.zero_()
            bits.zero_()
            tile_counts.zero_()
            zero_done = torch.cuda.Event()
            zero_done.record(zs)
        prezeroed = [True]
        raw_pending = [fit_done]

        want_xyz = bool(opt.want_xyz)
        want_rgb = bool(opt.want_rgb and vin.image is not None)
        # The output size is data dependent.  With the size of the previous run as a hint, a10/a11 are
        # launched into buffers of hint + 12.5 % rows right behind the scan, and the 16-byte size
        # read-back goes through a side stream, so the device never waits for the host; the kernel
        # writes nothing when the real size exceeds the buffers, and the plain path below re-emits.
        hint = getattr(self, "_emit_hint", 0)
        may_speculate = bool(opt.speculative_emit and opt.exchange is None and opt.stop_after != "mask"
                             and not opt.defer_emit and (want_xyz or want_rgb) and hint > 0)
        Hmax, Wmax = int(desc["H"].max()), int(desc["W"].max())

        def launch_mask_range(a: int, b: int, stats_mode: int, s: int, ws_ptr: int, zeroed: bool):
            """a7 statistics -> a5-a10 keep mask of views a..b-1 (launches only)"""
            if use_stats and raw_pending[0] is not None and stats_mode == 0:
                # the pass over the pixels ran on the raw depth beside the alignment: map it through the fit
                main.wait_event(raw_pending[0])
                raw_pending[0] = None
                L.check(lib.mdi_depth_range_map(V, max_tiles, ssp, res.status.data_ptr(), pop_flags, ws_ptr,
                                                res.lo_hi.data_ptr(), res.quantiles.data_ptr(),
                                                meta[2 * V + 1:].data_ptr(), s), "mdi_depth_range_map")
            elif use_stats:
                L.check(lib.mdi_depth_range(vp + a * VS, b - a, max_tiles, at(res.scale_shift, a, 2), at(res.status, a),
                                            pop_flags | stats_mode,
                                            ws_ptr, at(res.lo_hi, a, 2), at(res.quantiles, a, 4),
                                            meta[2 * V + 1:].data_ptr(), s), "mdi_depth_range")
            if use_stats:
                L.check(lib.mdi_factor_table(b - a, at(res.lo_hi, a, 2), int(opt.factor_min), int(opt.factor_max),
                                             at(ftab, a, L.FACTOR_TABLE_STRIDE), s), "mdi_factor_table")
            scfg.flags = int(flags) | (L.KEEP_PREZEROED if zeroed else 0)
            L.check(lib.mdi_keep_mask(vp + a * VS, b - a, total_tiles, int(px_begin[-1]), Hmax, Wmax, at(res.scale_shift, a, 2),
                                      at(res.status, a), at(res.lo_hi, a, 2), at(res.grad_range, a, 2),
                                      at(res.patch_blocked, a, L.PATCH_STRIDE), at(ftab, a, L.FACTOR_TABLE_STRIDE),
                                      C.byref(scfg), L.ptr(res.keep), bits.data_ptr(), tile_counts.data_ptr(), s),
                    "mdi_keep_mask")

        def launch_scan():
            L.check(lib.mdi_scan_tiles(vp, V, total_tiles, tile_counts.data_ptr(), tile_offsets.data_ptr(),
                                       res.view_counts.data_ptr(), res.view_offsets.data_ptr(),
                                       meta[2 * V:].data_ptr(), st), "mdi_scan_tiles")

        def launch_mask(stats_mode: int):
            """a7 statistics -> a5-a10 keep mask -> compaction offsets (launches only).  The first call of a
            run with several view ranges also carries their alignment stages: range i goes to part stream
            i (range 0 stays on the calling stream), so that the latency-bound kernels of one range
            (projection, fit, RANSAC, brackets, in-bin select) share the SMs with the HBM-bound ones of
            another; the ranges meet again in front of the scan."""
            if prezeroed[0] and parts > 1:
                prezeroed[0] = False
                forked = torch.cuda.Event()
                forked.record(main)
                joined = []
                for i, (a, b) in enumerate(ranges):
                    ps = main if i == 0 else self._part_stream(i - 1)
                    if i:
                        ps.wait_event(forked)
                    with torch.cuda.stream(ps):
                        s = _stream()
                        launch_project(a, b, s)
                        launch_fit(a, b, s)
                        ps.wait_event(zero_done)
                        launch_mask_range(a, b, stats_mode, s, ws.data_ptr() + ws_part[i] if use_stats else 0, True)
                        if i:
                            e = torch.cuda.Event()
                            e.record(ps)
                            joined.append(e)
                for e in joined:
                    main.wait_event(e)
                launch_scan()
                return
            zeroed = prezeroed[0]
            if zeroed:
                main.wait_event(zero_done)
                prezeroed[0] = False
            # (not zeroed: the exact-statistics retry fills again, in stream order)
            launch_mask_range(0, V, stats_mode, st, ws.data_ptr() if use_stats else 0, zeroed)
            launch_scan()

        def mask_stage(stats_mode: int, speculate: bool = False):
            """launch_mask + the read-back of (total, retry); returns (total, retry, spec)."""
            launch_mask(stats_mode)
            if not hasattr(self, "_rb_stream"):
                self._rb_stream = torch.cuda.Stream(self.device)
                self._tail_pin = torch.empty(2, dtype=torch.int64, pin_memory=True)
            if not speculate:
                # one 16-byte read-back (the only sync of the pipeline), into pinned memory: a pageable
                # destination makes the driver stage the copy and wait on the bulk copies in flight
                self._tail_pin.copy_(meta[2 * V:], non_blocking=True)
                done = torch.cuda.Event()
                done.record(main)
                done.synchronize()
                return int(self._tail_pin[0]), bool(self._tail_pin[1]), None
            scanned = torch.cuda.Event()
            scanned.record(main)
            cap = hint + (hint >> 3) + 4096
            xyz = torch.empty((cap, 3), dtype=f32, device=self.device) if want_xyz else None
            rgb = torch.empty((cap, 3), dtype=f32, device=self.device) if want_rgb else None
            L.check(lib.mdi_emit_points_capped(vp, V, total_tiles, ssp, bits.data_ptr(), tile_offsets.data_ptr(),
                                               res.view_offsets.data_ptr(), int(vin.image_kind), L.ptr(xyz), L.ptr(rgb),
                                               meta[2 * V:].data_ptr(), cap, st), "mdi_emit_points_capped")
            rb = self._rb_stream
            rb.wait_event(scanned)
            with torch.cuda.stream(rb):
                meta.record_stream(rb)
                self._tail_pin.copy_(meta[2 * V:], non_blocking=True)
                copied = torch.cuda.Event()
                copied.record(rb)
            copied.synchronize()
            return int(self._tail_pin[0]), bool(self._tail_pin[1]), (cap, xyz, rgb)

        if (opt.exchange is not None and opt.stop_after is None and not opt.defer_emit
                and (opt.want_xyz or opt.want_rgb)):
            # multi-GPU: counts travel on the device, the emit kernel writes this rank's points into every
            # rank's cloud; the host synchronises once, after everything has been queued
            def launch_emit():
                ex = opt.exchange
                xp = rp = None
                if ex.byte_rgb and want_rgb and int(vin.image_kind) != L.IMAGE_U8:
                    raise L.MdiError("CloudExchange(byte_rgb=True) needs uint8 images on every rank")
                n_dst = 0 if (ex.use_multicast and not ex.byte_rgb) else ex.world
                r8 = (C.c_void_p * n_dst)(*ex.rgb8_ptrs) if (ex.byte_rgb and want_rgb and n_dst) else None
                if n_dst:
                    xp = (C.c_void_p * n_dst)(*ex.xyz_ptrs) if want_xyz else None
                    rp = (C.c_void_p * n_dst)(*ex.rgb_ptrs) if want_rgb else None
                L.check(lib.mdi_emit_points_exchange(vp, V, total_tiles, ssp, bits.data_ptr(), tile_offsets.data_ptr(),
                                                     res.view_offsets.data_ptr(), int(vin.image_kind), n_dst, xp, rp,
                                                     ex.xyz_mc if (not n_dst and want_xyz) else None,
                                                     ex.rgb_mc if (not n_dst and want_rgb) else None,
                                                     ex.record.data_ptr(), int(ex.capacity), r8, int(ex.rank), st),
                        "mdi_emit_points_exchange")

            def remask(mode):
                if mode == L.STATS_RETRY_FLAGGED:
                    meta[2 * V + 1] = 0
                    res.extras["stats_retry"] = True
                launch_mask(mode)

            self._exchange_rounds(opt.exchange, res, lambda: remask(first_mode), lambda: remask(L.STATS_RETRY_FLAGGED),
                                  launch_emit, meta[2 * V:])
            return res
        res.total, retry, spec = mask_stage(first_mode, may_speculate)
        if retry:
            # a view's one-pass quantile brackets could not be certified (e.g. a constant depth
            # map): exact three-pass select for the flagged views, then mask + scan again
            meta[2 * V + 1] = 0
            res.extras["stats_retry"] = True
            spec = None
            res.total, retry, _ = mask_stage(L.STATS_RETRY_FLAGGED)
            assert not retry
        if opt.stop_after == "mask" or not (opt.want_xyz or opt.want_rgb):
            return res
        N = res.total
        if opt.defer_emit:
            res.extras["emit_state"] = dict(blob=blob, n_views=V, total_tiles=total_tiles, scale_shift=res.scale_shift,
                                            bits=bits, tile_offsets=tile_offsets, view_offsets=res.view_offsets,
                                            image_kind=int(vin.image_kind), has_image=vin.image is not None, total=N)
            return res
        self._emit_hint = N
        if spec is not None and N <= spec[0]:
            # the speculative launch covered it; a much too large buffer is not kept alive by the result
            shrink = N < spec[0] // 2
            res.xyz = None if spec[1] is None else (spec[1][:N].clone() if shrink else spec[1][:N])
            res.rgb = None if spec[2] is None else (spec[2][:N].clone() if shrink else spec[2][:N])
            res.extras["speculative_emit"] = True
            return res
        if want_xyz:
            res.xyz = torch.empty((N, 3), dtype=f32, device=self.device)
        if want_rgb:
            res.rgb = torch.empty((N, 3), dtype=f32, device=self.device)
        if N:
            L.check(lib.mdi_emit_points(vp, V, total_tiles, ssp, bits.data_ptr(), tile_offsets.data_ptr(),
                                        res.view_offsets.data_ptr(), int(vin.image_kind), L.ptr(res.xyz), L.ptr(res.rgb), st),
                    "mdi_emit_points")
        return res

    def _exchange_rounds(self, ex, res: RunResult, first_mask, retry_mask, launch_emit, local=None):
        """The collective part of a multi-GPU step.  Every rank runs the same sequence of rounds:
        [mask stage] -> publish counts -> emit into all clouds -> barrier -> one read-back.  A rank
        whose one-pass statistics need the exact retry raises its flag in the count exchange, and
        then EVERY rank goes through a second round (ranks without flagged views just emit again),
        so the collective calls stay matched; clouds that turn out too small are grown collectively
        and the emit repeated (the kernel writes nothing when the scene exceeds the capacity)."""
        st = _stream()
        if first_mask is not None:
            first_mask()
        while True:
            ex.publish_counts(None if local is None else local.data_ptr(), st)
            if launch_emit is not None:
                launch_emit()
            ex.barrier()
            if ex.byte_rgb:
                ex.expand_bytes(st)
            row_offset, total_all, retry_any, counts, mine = ex.read_record(local)
            if retry_any:
                # ranks whose own flag is clear keep their masks and simply emit again next round
                if local is not None and retry_mask is not None and mine[1]:
                    retry_mask()
                continue
            if total_all > ex.capacity:
                ex.grow(total_all)
                if launch_emit is not None:
                    launch_emit()
                ex.barrier()
                if ex.byte_rgb:
                    ex.expand_bytes(st)
                torch.cuda.current_stream(self.device).synchronize()
            break
        res.total = counts[ex.rank]
        res.xyz, res.rgb = ex.clouds()
        res.extras["rank_counts"] = counts
        res.extras["row_offset"] = row_offset

    def emit_deferred(self, states, want_xyz: bool = True, want_rgb: bool = True):
        """a10/a11 of runs made with RunOptions.defer_emit, in order, into ONE cloud: run i's points go
        to rows sum(total[:i]) .. (the row_offset form of the emit kernel).  Returns (xyz, rgb)."""
        lib, st = self.lib, _stream()
        total = sum(s["total"] for s in states)
        f32 = torch.float32
        want_rgb = bool(want_rgb and all(s["has_image"] for s in states))
        xyz = torch.empty((total, 3), dtype=f32, device=self.device) if want_xyz else None
        rgb = torch.empty((total, 3), dtype=f32, device=self.device) if want_rgb else None
        row = 0
        for s in states:
            if s["total"]:
                xp = (C.c_void_p * 1)(xyz.data_ptr()) if xyz is not None else None
                rp = (C.c_void_p * 1)(rgb.data_ptr()) if rgb is not None else None
                L.check(lib.mdi_emit_points_multi(s["blob"].data_ptr(), s["n_views"], s["total_tiles"], L.ptr(s["scale_shift"]),
                                                  s["bits"].data_ptr(), s["tile_offsets"].data_ptr(),
                                                  s["view_offsets"].data_ptr(), s["image_kind"], 1, xp, rp, row, st),
                        "mdi_emit_points_multi")
            row += s["total"]
        return xyz, rgb


_engines: dict[int, Engine] = {}


def get_engine(device: "torch.device | str" = "cuda") -> Engine:
    dev = torch.device(device)
    if dev.type != "cuda":
        raise L.MdiError(f"device {device!r}: the B200 path has no CPU fallback")
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    eng = _engines.get(idx)
    if eng is None:
        with torch.cuda.device(idx):
            eng = _engines[idx] = Engine(torch.device("cuda", idx))
    return eng
