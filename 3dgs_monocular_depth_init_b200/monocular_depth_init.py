"""Scene-level entry point on the B200 (reference: gs_init_compare/monocular_depth_init.py:95-224).

All views of the scene go through ONE batched kernel sequence (engine.Engine.run); views that the
reference would skip with LowDepthAlignmentConfidenceError contribute no points and no K/P/size
rows.  Depth predictors are out of scope: depth comes from `predict_fn(image, data) ->
PredictedDepth` (cached depth, a predictor, or synthetic depth)."""
from __future__ import annotations

import logging
import sys
from pathlib import Path
from typing import Callable, List, Optional

import numpy as np
import torch

from . import _lib as L
from ._single import options_from_config, ransac_seed
from .engine import RunResult, ViewInputs, get_engine
from .knn import knn
from .point_cloud_postprocess.postprocess import postprocess_point_cloud
from .types import InputImage

_LOGGER = logging.getLogger(__name__)
DEPTH_SOURCE: Optional[Callable] = None  # set by the caller: (InputImage, data dict, device) -> PredictedDepth


def add_noise_to_point_cloud(pts: torch.Tensor, noise_std: float):
    """monocular_depth_init.py:90-92."""
    return pts + torch.randn_like(pts) * noise_std


def gather_scene(parser, device, depth_source: Callable):
    """Loads the train split into per-view lists (monocular_depth_init.py:110-139)."""
    dataset = type(parser).DatasetCls(parser, split="train")
    depths, masks, images, Ks, c2ws, names = [], [], [], [], [], []
    for data in dataset:
        assert data["image"].max() > 1  # image must be 0..255 (monocular_depth_init.py:122)
        image = InputImage(name=data["image_name"], cam2world=data["camtoworld"], K=data["K"], data=data["image"])
        pred = depth_source(image, data, device)
        assert pred.depth.device == torch.device(device) or str(pred.depth.device).startswith(str(device))
        depths.append(pred.depth.float().contiguous())
        masks.append(pred.mask.contiguous())
        images.append(data["image"])
        Ks.append(data["K"]); c2ws.append(data["camtoworld"]); names.append(data["image_name"])
    return depths, masks, images, torch.stack(Ks), torch.stack(c2ws), names


class _PinnedPool:
    """Grow-only pinned staging buffers reused across calls (a fresh pinned allocation is a
    cudaHostAlloc: tens of milliseconds).  A buffer is handed out again only after the copy that
    last used it has completed (event recorded by `mark`)."""

    def __init__(self):
        self._buf, self._ev = {}, {}

    def get(self, name: str, shape, dtype) -> torch.Tensor:
        n = int(np.prod(shape)) if len(shape) else 1
        ev = self._ev.pop(name, None)
        if ev is not None:
            ev.synchronize()
        t = self._buf.get(name)
        if t is None or t.dtype != dtype or t.numel() < n:
            # 25 % headroom: result sizes vary from call to call (RANSAC seeds), and every growth is a
            # cudaHostAlloc of tens of milliseconds
            t = self._buf[name] = torch.empty(max(n + (n >> 2), 1), dtype=dtype, pin_memory=True)
        return t[:n].view(*shape)

    def mark(self, name: str) -> None:
        ev = torch.cuda.Event()
        ev.record()
        self._ev[name] = ev


_PINNED = _PinnedPool()


class SfmTable:
    """Device-resident SfM inputs of a scene: parser.points and the per-view index lists
    (parser.point_indices) concatenated into 16-byte aligned slot ranges."""

    def __init__(self, points: np.ndarray, point_indices: dict, names, device):
        lists = [np.asarray(point_indices[n]) for n in names]
        counts = np.fromiter((len(a) for a in lists), dtype=np.int64, count=len(lists))
        padded = (counts + 3) & ~3
        begin = np.zeros(len(lists) + 1, dtype=np.int64)
        np.cumsum(padded, out=begin[1:])
        # staged in reused pinned memory so both uploads are asynchronous
        use_pin = torch.device(device).type == "cuda"
        n_idx = int(begin[-1]) + 4
        pts_np = np.asarray(points)
        if use_pin:
            idx_t = _PINNED.get("sfm_idx", (n_idx,), torch.int32)
            idx_t.zero_()
            xyz_t = _PINNED.get("sfm_xyz", (int(pts_np.shape[0]), 3), torch.float32)
        else:
            idx_t = torch.zeros(n_idx, dtype=torch.int32)
            xyz_t = torch.empty((int(pts_np.shape[0]), 3), dtype=torch.float32)
        idx = idx_t.numpy()
        for b, a in zip(begin[:-1].tolist(), lists):  # one slice copy per view
            idx[b:b + len(a)] = a
        np.copyto(xyz_t.numpy(), pts_np.reshape(-1, 3), casting="same_kind")
        self.counts = counts
        self.begin = begin
        self.host_bytes = int(pts_np.shape[0]) * 12 + idx.nbytes
        self.xyz = xyz_t.to(device, non_blocking=True)
        self.index = idx_t.to(device, non_blocking=True)
        if use_pin:
            with torch.cuda.device(self.xyz.device):
                _PINNED.mark("sfm_idx"); _PINNED.mark("sfm_xyz")
        self._staging = (xyz_t, idx_t)

    def slice(self, a: int, b: int) -> "SfmTable":
        """The table of views a..b-1 (shares the uploaded arrays)."""
        t = object.__new__(SfmTable)
        t.counts = self.counts[a:b]
        t.begin = self.begin[a:b + 1] - self.begin[a]
        t.host_bytes = 0
        t.xyz = self.xyz
        t.index = self.index[int(self.begin[a]):]
        t._staging = self._staging
        return t


@torch.no_grad()
def points_from_views(config, depths, masks, images, K, c2w, names, points: np.ndarray, point_indices: dict,
                      device, image_kind: int = L.IMAGE_F32_255, sample_idx: Optional[torch.Tensor] = None,
                      sfm: Optional[SfmTable] = None, exchange=None, defer_emit: bool = False,
                      streams: Optional[int] = None):
    """The batched hot path.  depths / masks / images: stacked tensors or per-view lists; images
    are [H,W,3] float32 0..255 or uint8, on the device or in pinned host memory.  `sfm` is an
    already uploaded SfmTable of (points, point_indices, names); built here when omitted.
    `exchange` (distributed.CloudExchange, one per rank): this rank's views are emitted into every
    rank's cloud, and xyz / rgb of the result are the whole scene (see distributed.py).
    `streams`: number of contiguous view ranges queued on separate CUDA streams (engine.py; None = the
    MDI_STREAMS default); the result does not depend on it.
    Returns the engine RunResult (xyz / rgb on the device)."""
    eng = get_engine(device)
    dev = eng.device
    if sfm is None:
        sfm = SfmTable(points, point_indices, names, dev)
    opt = options_from_config(config)
    if opt.aligner in ("ransac", "msac"):
        if sample_idx is not None and config.mdi.alignment.segmenter is None:
            opt.sample_idx = sample_idx.to(device=dev, dtype=torch.int32).contiguous()
        elif sample_idx is None:
            opt.ransac_seed = ransac_seed()
    opt.exchange = exchange
    opt.defer_emit = bool(defer_emit)
    opt.streams = streams
    vin = ViewInputs(depth=depths, mask=masks, image=images, image_kind=image_kind, K=K, c2w=c2w,
                     sfm_xyz=sfm.xyz, sfm_index=sfm.index, sfm_counts=sfm.counts)
    if config.mdi.alignment.segmenter is not None:   # region-wise alignment: aligned map materialised per view
        from .depth_alignment.pipeline import run_views_materialised
        return run_views_materialised(eng, vin, config, opt, sample_idx)
    if opt.aligner == "interp":   # per-pixel scale map: aligned depth materialised per view, then the mask stages
        from .depth_alignment.alignment.interp import run_views_with_interp
        return run_views_with_interp(eng, vin, config, opt, sample_idx)
    return eng.run(vin, opt)


def _export_per_image(out_dir, parser, names, images, K_all, c2w_all, res, counts, status):
    """`mdi.pts_output_per_image`: the four clouds the reference dumps per view
    (points_from_depth.py:47-108 `debug_export_point_clouds`): the image plane in world space, the
    view's SfM points projected onto that plane, the SfM points, and the view's unprojected points.
    (The matplotlib depth / alignment plots of :33-45 and pipeline.py:30-167 are not produced.)"""
    from .utils.point_cloud_export import export_point_cloud_to_ply
    P_all = res.P.cpu()
    xyz, rgb = res.xyz.cpu().numpy(), (res.rgb.cpu().numpy() if res.rgb is not None else None)
    offs = np.concatenate([[0], np.cumsum(counts)])
    for i, name in enumerate(names):
        if status[i] & L.STATUS_LOW_CONFIDENCE:   # the reference raises before its export
            continue
        d = Path(out_dir) / name
        d.mkdir(exist_ok=True, parents=True)
        im = images[i].cpu()
        H, W = int(im.shape[0]), int(im.shape[1])
        K, c2w = K_all[i].float(), c2w_all[i].float()

        def to_world(pc):   # points_from_depth.py:302-312
            cam = torch.linalg.inv(K) @ pc.reshape(-1, 3).T
            return (c2w @ torch.vstack([cam, torch.ones(cam.shape[1])]))[:3].T

        plane = torch.dstack([torch.from_numpy(np.mgrid[0:W, 0:H].T).float(), torch.ones((W, H)).T]).reshape(-1, 3)
        export_point_cloud_to_ply(to_world(plane).numpy(), (im.float() / 255.0).reshape(-1, 3).numpy(), d, "camera_plane")
        idx = np.asarray(parser.point_indices[name])
        sfm = torch.from_numpy(np.ascontiguousarray(parser.points[idx])).float()
        sfm_rgb = parser.points_rgb[idx] / 255.0
        proj = P_all[i] @ torch.vstack([sfm.T, torch.ones(sfm.shape[0])])
        proj = proj / proj[2]
        export_point_cloud_to_ply(to_world(proj.T).numpy(), sfm_rgb, d, "sfm_points_plane_proj")
        export_point_cloud_to_ply(sfm.numpy(), sfm_rgb, d, "sfm_points_world")
        a, b = int(offs[i]), int(offs[i + 1])
        export_point_cloud_to_ply(xyz[a:b], None if rgb is None else rgb[a:b], d, "my_points_world")


@torch.no_grad()
def pts_and_rgb_from_monocular_depth(config, parser, device: str = "cuda", depth_source: Optional[Callable] = None):
    """-> (pts float32 [N,3] on device, rgbs float32 [N,3] on CPU, scales [N,3] | None);
    monocular_depth_init.py:95-224, same signature (runner.py:77-78 calls it unchanged).

    Depth per view comes from `depth_source(image, data, device) -> PredictedDepth` (image.data is
    the dataset's 0..255 image); the default is the reference's predicted-depth cache
    (`depth_prediction.depth_cache.CachedDepthSource`:
    `<mdi.cache_dir>/<model.name>/<parser.dataset_name>/<image_name>.pth`), which raises
    `DepthCacheMiss` for an image that is not cached unless a predictor has been registered."""
    eng = get_engine(device)
    dev = eng.device
    depth_source = depth_source or DEPTH_SOURCE
    model_label = getattr(depth_source, "name", None)
    if depth_source is None:
        from .depth_prediction.depth_cache import CachedDepthSource
        depth_source = CachedDepthSource(config, getattr(parser, "dataset_name", "dataset"), dev)
        model_label = depth_source.name
    if config.mdi.pts_output_dir is not None and model_label is None:
        from .depth_prediction.depth_cache import model_name
        model_label = model_name(config)
    depths, masks, images, K, c2w, names = gather_scene(parser, dev, depth_source)
    # images stay where the dataset put them (CPU in the reference): pin so the gather kernel
    # reads only the kept pixels over PCIe instead of copying 12 B/px
    imgs = [im if im.is_cuda else im.contiguous().pin_memory() for im in images]
    kind = L.IMAGE_U8 if imgs and imgs[0].dtype == torch.uint8 else L.IMAGE_F32_255
    res = points_from_views(config, depths, masks, imgs, K, c2w, names, parser.points, parser.point_indices, dev, kind)
    status = res.status.cpu().numpy() if res.status is not None else np.zeros(0, np.int32)
    counts = res.view_counts.cpu().numpy() if res.view_counts is not None else np.zeros(0, np.int64)
    for i in np.nonzero(status & L.STATUS_LOW_CONFIDENCE)[0]:
        _LOGGER.warning("Low depth alignment confidence for image %s", names[i])
    if (status & L.STATUS_EMPTY_STATS).any():
        raise RuntimeError("quantile() input tensor must be non-empty")
    pts = res.xyz if res.xyz is not None else torch.zeros((0, 3), dtype=torch.float32, device=dev)
    rgbs = res.rgb if res.rgb is not None else torch.zeros((0, 3), dtype=torch.float32, device=dev)
    if config.mdi.pts_output_dir and config.mdi.pts_output_per_image:   # monocular_depth_init.py:142-146
        _export_per_image(config.mdi.pts_output_dir, parser, names, images, K, c2w, res, counts, status)
    if config.mdi.noise_std_scene_frac is not None:
        pts = add_noise_to_point_cloud(pts, parser.scene_scale * config.mdi.noise_std_scene_frac)
    rgbs = rgbs.cpu()
    ok = (status & L.STATUS_LOW_CONFIDENCE) == 0
    Kn = K.cpu().numpy()
    Pn = res.P.cpu().numpy() if res.P is not None else np.zeros((0, 3, 4), np.float32)
    intrinsic_matrices = [Kn[i] for i in range(len(names)) if ok[i]]
    proj_matrices = [Pn[i] for i in range(len(names)) if ok[i]]
    # deviation from the reference: image_sizes is compacted to the kept views
    image_sizes = np.array([[images[i].shape[1], images[i].shape[0]] for i in range(len(names)) if ok[i]],
                           dtype=np.int32).reshape(-1, 2)
    if config.mdi.include_sfm_points:
        pts = torch.cat([pts, torch.from_numpy(parser.points).float().to(dev)], dim=0)
        rgbs = torch.cat([rgbs, torch.from_numpy(parser.points_rgb / 255.0).float()], dim=0)
    _LOGGER.info("Num points before postprocess: %d", pts.shape[0])
    pts, rgbs = postprocess_point_cloud(pts, rgbs, intrinsic_matrices, proj_matrices, image_sizes,
                                        config.mdi.postprocess, dev)
    _LOGGER.info("Num points after postprocess: %d", pts.shape[0])
    if config.mdi.pts_output_dir is not None:   # monocular_depth_init.py:198-211
        from .utils.point_cloud_export import export_point_cloud_to_ply
        output_dir = Path(config.mdi.pts_output_dir) / getattr(parser, "dataset_name", "dataset")
        output_dir.mkdir(exist_ok=True, parents=True)
        # the reference formats the `str, Enum` members under Python 3.10, i.e. by value
        seg, alg = config.mdi.alignment.segmenter, config.mdi.alignment.aligner
        filename = (f"{model_label}_{config.mdi.subsample_factor}_{getattr(seg, 'value', seg)}_"
                    f"{getattr(alg, 'value', alg)}")
        export_point_cloud_to_ply(pts.cpu().numpy(), rgbs.cpu().numpy(), output_dir, filename, outlier_std_dev=None)
        export_point_cloud_to_ply(parser.points, parser.points_rgb / 255.0, output_dir, "sfm")
    if config.mdi.pts_only:                     # :212-213
        sys.exit(0)
    scales = None
    if config.mdi.limit_init_scale:  # monocular_depth_init.py:215-223
        dist2_avg = (knn(pts, 4)[:, 1:] ** 2).mean(dim=-1)
        dist_avg = torch.sqrt(dist2_avg)
        quantile = torch.quantile(dist_avg, config.mdi.init_scale_clamp_quantile)
        dist_avg = torch.clamp(dist_avg, max=quantile)
        scales = torch.log(dist_avg * config.init_scale).unsqueeze(-1).repeat(1, 3)
    return pts, rgbs, scales


_TRACE: Optional[list] = None   # debugging: a list here collects (label, timing event) marks of the host path


def _mark(label: str):
    if _TRACE is not None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        _TRACE.append((label, ev))


@torch.no_grad()
def points_from_host_views(config, depth_h, mask_h, image_h, K, c2w, names, points, point_indices, device,
                           image_kind: int = L.IMAGE_U8, sample_idx: Optional[torch.Tensor] = None, exchange=None,
                           chunks: int = 1):
    """End-to-end form of the hot path for HOST inputs (what a dataset loader hands over):
    depth [V,H,W] float32 and mask [V,H,W] bool in pinned host memory are copied to the device,
    the image [V,H,W,3] stays in pinned host memory and only its kept pixels are read over PCIe by
    the gather kernel; rgb comes back to the host (the reference returns rgbs on the CPU,
    monocular_depth_init.py:168), xyz stays on the device.  rgb_host is a view of a reused pinned
    buffer: copy it if it has to outlive the next call.
    With chunks > 1 the views go through in blocks: all host->device copies are queued on a copy
    stream up front, block c is aligned / masked while blocks c+1.. are still in flight, and the
    clouds of all blocks are written at the end (Engine.emit_deferred) -- the in-place image reads of
    the emit kernel use the same PCIe direction as the bulk copies and must not overlap them.
    Measured on a B200 (200-view C2 scene) this still does NOT pay: 23.4 ms at 1 block, 24.1 at 2,
    25.3 at 4 -- although the same 4-block step takes 23.2 ms under the torch profiler, whose time
    line shows the intended overlap (copies back to back, each block's kernels inside the next
    block's copy).  The cause of the un-profiled slowdown is not understood; the default stays one
    block.  Always one block when `exchange` is given.
    Returns (RunResult, rgb_host, h2d_bytes, d2h_bytes)."""
    eng = get_engine(device)
    dev = eng.device
    assert depth_h.is_pinned() and mask_h.is_pinned() and image_h.is_pinned()
    V = int(depth_h.shape[0])
    depth_d = eng._buf("e2e_depth", depth_h.numel(), torch.float32).view(depth_h.shape)
    mask_d = eng._buf("e2e_mask", mask_h.numel(), torch.uint8).view(mask_h.shape)
    mask_u8 = mask_h.view(torch.uint8)
    n_chunks = 1 if (exchange is not None or V == 0) else max(1, min(int(chunks), V))
    bounds = [(c * V) // n_chunks for c in range(n_chunks + 1)]
    main = torch.cuda.current_stream(dev)
    if not hasattr(eng, "_copy_stream"):
        eng._copy_stream = torch.cuda.Stream(dev)
    cs = eng._copy_stream
    _mark("start")
    cs.wait_stream(main)  # earlier work may still be reading the staging buffers
    arrived = []

    def queue_block(c):
        a, b = bounds[c], bounds[c + 1]
        with torch.cuda.stream(cs):
            depth_d[a:b].copy_(depth_h[a:b], non_blocking=True)
            mask_d[a:b].copy_(mask_u8[a:b], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
            arrived.append(ev)

    # the copy engine serves host->device copies in the order they are handed over, whatever their
    # stream: the SfM table goes right behind the first block (its host-side build hides under that
    # copy), and nothing else of a block's processing uses the copy engine (mdi_upload_table)
    queue_block(0)
    sfm = SfmTable(points, point_indices, names, dev)
    for c in range(1, n_chunks):
        queue_block(c)
    parts = []
    for c in range(n_chunks):
        a, b = bounds[c], bounds[c + 1]
        main.wait_event(arrived[c])
        parts.append(points_from_views(config, depth_d[a:b], mask_d[a:b], image_h[a:b], K[a:b], c2w[a:b], names[a:b],
                                       points, point_indices, dev, image_kind,
                                       sample_idx[a:b] if sample_idx is not None else None,
                                       sfm=sfm if n_chunks == 1 else sfm.slice(a, b), exchange=exchange,
                                       defer_emit=n_chunks > 1))
    _mark("masks done")
    if n_chunks == 1:
        res = parts[0]
    else:
        # every block's cloud is written after the last copy has landed: the image reads of the emit
        # kernel and the bulk copies use the same PCIe direction and must not overlap
        res = RunResult(n_views=V, px_begin=None, sfm_begin=None)
        res.xyz, res.rgb = eng.emit_deferred([p.extras["emit_state"] for p in parts])
        res.view_counts = torch.cat([p.view_counts for p in parts])
        res.status = torch.cat([p.status for p in parts])
        res.scale_shift = torch.cat([p.scale_shift for p in parts])
        res.total = sum(p.total for p in parts)
    _mark("clouds written")
    if exchange is not None:  # res.rgb is the whole scene: this rank's rows go back to its host
        r0 = res.extras["row_offset"]
        rgb_local = res.rgb[r0:r0 + res.total]
    else:
        rgb_local = res.rgb
    # reused pinned result buffers: rgb_h / counts_host are valid until the next call
    rgb_h = _PINNED.get("e2e_rgb", tuple(rgb_local.shape), torch.float32)
    rgb_h.copy_(rgb_local, non_blocking=True)
    counts_h = _PINNED.get("e2e_counts", tuple(res.view_counts.shape), torch.int64)
    counts_h.copy_(res.view_counts, non_blocking=True)
    _mark("results on host")
    main.synchronize()
    h2d = depth_h.numel() * 4 + mask_h.numel() + res.total * 3 * image_h.element_size() \
        + sfm.host_bytes + len(names) * (L.VIEW_DTYPE.itemsize + 25 * 4)
    d2h = res.total * 12 + counts_h.numel() * 8 + 8 * n_chunks
    res.extras = dict(res.extras or {}, counts_host=counts_h, chunks=n_chunks)
    return res, rgb_h, int(h2d), int(d2h)
