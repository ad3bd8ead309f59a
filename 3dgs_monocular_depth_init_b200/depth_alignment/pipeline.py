"""Alignment pipeline (reference: depth_alignment/pipeline.py:170-293).

No segmenter (default): one region holding every SfM point -- the aligner's output is the result.
With a segmenter (region-wise alignment, pipeline.py:196-283): the same aligner per region.  For
the scale/shift aligners all regions of the view are fitted in ONE batched engine run (each region
is a "view" sharing the depth map, with its own SfM subset: mdi_gather_fit / mdi_ransac) and
mdi_apply_region_scale_shift stitches the aligned map; `interp` runs per region."""
from dataclasses import dataclass
from pathlib import Path
from typing import Callable, Optional

import numpy as np
import torch

from .. import _lib as L
from ..depth_prediction.predictors.depth_predictor_interface import PredictedDepth
from .interface import DepthAlignmentResult, DepthAlignmentStrategy

INVALID_DEPTH_VAL = -42.0


def fit_regions(predicted_depth: PredictedDepth, coords: torch.Tensor, gt: torch.Tensor, region_of_point: torch.Tensor,
                n_regions: int, config, sample_idx: Optional[torch.Tensor] = None):
    """Scale / shift of every region in one batched run.  region_of_point int64 [M] in -1 .. R-1 (-1:
    the point takes part in no region).  -> (table float32 [R,2], valid bool [R], RunResult)."""
    from .._single import options_from_config, ransac_seed
    from ..engine import ViewInputs, get_engine
    dev = predicted_depth.depth.device
    eng = get_engine(dev)
    order = torch.argsort(region_of_point, stable=True)            # region-major, original order inside
    reg_sorted = region_of_point[order]
    counts = torch.bincount(reg_sorted[reg_sorted >= 0], minlength=n_regions).cpu().numpy().astype(np.int64)
    keep = np.nonzero(counts)[0]                                   # regions with at least one point
    table = torch.zeros((n_regions, 2), dtype=torch.float32, device=dev)
    valid = torch.zeros(n_regions, dtype=torch.bool, device=dev)
    if len(keep) == 0:
        return table, valid, None
    first = int((reg_sorted < 0).sum())
    starts = first + np.concatenate([[0], np.cumsum(counts)[:-1]])
    padded = (counts[keep] + 3) & ~3
    begin = np.zeros(len(keep) + 1, dtype=np.int64)
    np.cumsum(padded, out=begin[1:])
    slots = int(begin[-1]) + 4
    uv_s = torch.zeros((2, slots), dtype=torch.int32, device=dev)
    z_s = torch.zeros(slots, dtype=torch.float32, device=dev)
    for b, r in zip(begin[:-1].tolist(), keep.tolist()):
        idx = order[int(starts[r]): int(starts[r]) + int(counts[r])]
        uv_s[:, b:b + len(idx)] = coords[:, idx].to(torch.int32)
        z_s[b:b + len(idx)] = gt[idx].float()
    R = len(keep)
    eye3, eye4 = torch.eye(3)[None].repeat(R, 1, 1), torch.eye(4)[None].repeat(R, 1, 1)
    depth = predicted_depth.depth.float().contiguous()
    vin = ViewInputs(depth=[depth] * R, mask=None, image=None, image_kind=L.IMAGE_F32_UNIT, K=eye3, c2w=eye4,
                     uv=uv_s, z=z_s, n_valid=torch.from_numpy(counts[keep]).to(device=dev, dtype=torch.int32),
                     sfm_counts=counts[keep])
    opt = options_from_config(config, stop_after="fit", use_patch_mask=False)
    if opt.aligner in ("ransac", "msac"):
        if sample_idx is not None:
            opt.sample_idx = sample_idx[torch.from_numpy(keep)].to(device=dev, dtype=torch.int32).contiguous()
        else:
            opt.ransac_seed = ransac_seed()
    res = eng.run(vin, opt)
    kept = torch.from_numpy(keep).to(dev)
    table[kept] = res.scale_shift
    valid[kept] = True
    return table, valid, res


@dataclass
class DepthAlignmentPipeline:
    config: object
    segmentation: Optional[Callable]
    alignment: DepthAlignmentStrategy

    @staticmethod
    def from_config(config):
        segmentation = None
        if config.mdi.alignment.segmenter is not None:
            from .segmentation import get_segmenter
            segmentation = get_segmenter(config.mdi.alignment.segmenter)
        return DepthAlignmentPipeline(config, segmentation, config.mdi.alignment.aligner.get_implementation())

    def align(self, image, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, config, debug_export_dir: Optional[Path] = None, *,
              sample_idx: Optional[torch.Tensor] = None):
        """`sample_idx` (parity runs): RANSAC samples, [max_iters, s] without a segmenter, [R, max_iters, s]
        (one set per region, in region order) with one."""
        if self.segmentation is None:
            # one region holding every point (pipeline.py:248-251); the stitched map equals the
            # aligner's output and the sentinel test (pipeline.py:285-288) is applied to it
            kw = {} if sample_idx is None else {"sample_idx": sample_idx}
            r = self.alignment.align(predicted_depth, sfm_points_camera_coords, sfm_points_depth, config,
                                     debug_export_dir, **kw)
            return DepthAlignmentResult(aligned_depth=r.aligned_depth,
                                        mask=(r.aligned_depth != INVALID_DEPTH_VAL) & predicted_depth.mask)
        return self._align_regions(predicted_depth, sfm_points_camera_coords, sfm_points_depth, config,
                                   debug_export_dir, sample_idx)

    def _align_regions(self, predicted_depth, coords, gt, config, debug_export_dir, sample_idx):
        from .segmentation.region_margin import calculate_region_margin_mask
        from ..engine import get_engine
        dev = predicted_depth.depth.device
        seg_cfg = config.mdi.alignment.segmentation
        segmentation = self.segmentation(predicted_depth, Path(config.mdi.cache_dir) / "checkpoints", seg_cfg)
        segmentation = torch.as_tensor(segmentation).to(dev)
        deadzone = calculate_region_margin_mask(segmentation, seg_cfg.region_margin)       # pipeline.py:219-221
        region_ids = torch.unique(segmentation[predicted_depth.mask])                      # :232
        if seg_cfg.propagate_mask:                                                         # :235-236
            predicted_depth.mask = predicted_depth.mask & deadzone
        coords = coords.to(dev)
        u, v = coords[0].long(), coords[1].long()
        sfm_regions = segmentation[v, u]                                                   # :238-243
        sfm_ok = deadzone[v, u]
        R = int(region_ids.numel())
        # position of every point's / pixel's region in region_ids (-1: not an aligned region).  The
        # reference indexes its per-region lists by region id (:259,:262), i.e. assumes ids 0 .. R-1.
        def index_of(x):
            if R == 0:
                return torch.full_like(x, -1, dtype=torch.int64)
            pos = torch.searchsorted(region_ids, x.contiguous()).clamp_(max=R - 1)
            return torch.where(region_ids[pos] == x, pos, torch.full_like(pos, -1))
        point_region = torch.where(sfm_ok, index_of(sfm_regions), torch.full_like(u, -1))
        H, W = predicted_depth.depth.shape
        aligner = getattr(config.mdi.alignment.aligner, "value", config.mdi.alignment.aligner)
        if aligner == "interp":
            out = torch.full((H, W), INVALID_DEPTH_VAL, dtype=torch.float32, device=dev)
            for r in range(R):
                sel = point_region == r
                if not bool(sel.any()):
                    continue   # "No SfM points found in region": stays at the sentinel (:264-270)
                kw = {} if sample_idx is None else {"sample_idx": sample_idx[r]}
                res = self.alignment.align(predicted_depth, coords[:, sel], gt[sel], config, debug_export_dir, **kw)
                region_mask = segmentation == region_ids[r]
                out[region_mask] = res.aligned_depth[region_mask]
        else:
            table, valid, _ = fit_regions(predicted_depth, coords, gt.to(dev), point_region, R, config, sample_idx)
            pix_region = index_of(segmentation.reshape(-1)).to(torch.int32).contiguous()
            depth = predicted_depth.depth.float().contiguous()
            out = torch.empty((H, W), dtype=torch.float32, device=dev)
            valid_u8 = valid.to(torch.uint8).contiguous()
            eng = get_engine(dev)
            L.check(eng.lib.mdi_apply_region_scale_shift(depth.data_ptr(), pix_region.data_ptr(), depth.numel(),
                                                         table.data_ptr(), valid_u8.data_ptr(), R, out.data_ptr(),
                                                         torch.cuda.current_stream().cuda_stream),
                    "mdi_apply_region_scale_shift")
        return DepthAlignmentResult(aligned_depth=out, mask=(out != INVALID_DEPTH_VAL) & predicted_depth.mask)


def run_views_materialised(eng, vin, config, opt2, sample_idx=None):
    """The batched pipeline when the alignment is not one scale / shift per view (a segmenter is
    configured): (1) project + filter the SfM points of every view in one engine run, (2) per view,
    DepthAlignmentPipeline.align -> a materialised aligned depth map (sentinel -42 where no region
    was aligned) and the possibly narrowed mask, (3) the mask / compaction / unprojection stages on
    those maps with the alignment switched off.  `opt2`: options of the last run."""
    import dataclasses
    from .._single import options_from_config  # noqa: F401  (config validation happens in the callers)
    from ..engine import RunOptions
    V = int(vin.K.shape[0])
    opt2.aligner, opt2.scale_shift = None, None
    if V == 0:
        return eng.run(vin, opt2)
    r1 = eng.run(vin, RunOptions(stop_after="project", use_patch_mask=False))
    n_valid = r1.n_valid.cpu().tolist()
    status = r1.status.cpu().tolist()
    pipe = DepthAlignmentPipeline.from_config(config)
    depths, masks = [], []
    for v in range(V):
        depth_v = vin.depth[v]
        mask_v = None if vin.mask is None else vin.mask[v]
        if status[v] != L.STATUS_OK:
            depths.append(depth_v); masks.append(mask_v)
            continue
        b, n = int(r1.sfm_begin[v]), int(n_valid[v])
        pred = PredictedDepth(depth_v, torch.ones_like(depth_v, dtype=torch.bool) if mask_v is None else mask_v.bool())
        si = None if sample_idx is None else sample_idx[v]
        res = pipe.align(None, pred, r1.uv[:, b:b + n].long(), r1.z[b:b + n], config, None, sample_idx=si)
        depths.append(res.aligned_depth.float().contiguous())
        masks.append(pred.mask.contiguous())   # propagate_mask may have narrowed it (pipeline.py:235-236)
    res = eng.run(dataclasses.replace(vin, depth=depths, mask=masks), opt2)
    res.extras["aligned_depth"] = depths
    return res
