"""Helpers turning the reference's per-view arguments into one-view engine batches."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import _lib as L
from .engine import Engine, RunOptions, RunResult, ViewInputs, get_engine


def _dev_f32(t: torch.Tensor, device) -> torch.Tensor:
    return t.to(device=device, dtype=torch.float32).contiguous()


def coords_to_slots(uv: torch.Tensor, z: torch.Tensor, device):
    """[2,M'] integer coords + [M'] depths -> padded slot-layout int32/float32 device buffers."""
    m = int(uv.shape[1])
    slots = ((m + 3) & ~3) + 4
    uv_s = torch.zeros((2, slots), dtype=torch.int32, device=device)
    z_s = torch.zeros(slots, dtype=torch.float32, device=device)
    uv_s[:, :m] = uv.to(device=device, dtype=torch.int32)
    z_s[:m] = z.to(device=device, dtype=torch.float32)
    n_valid = torch.tensor([m], dtype=torch.int32, device=device)
    return uv_s, z_s, n_valid, m


def identity_cameras():
    return torch.eye(3)[None], torch.eye(4)[None]


def single_view_inputs(depth: torch.Tensor, mask: Optional[torch.Tensor], device, K=None, c2w=None,
                       image=None, image_kind=L.IMAGE_F32_UNIT, **kw) -> ViewInputs:
    d = _dev_f32(depth, device)
    m = None
    if mask is not None:
        m = mask.to(device=device).contiguous()
        if m.dtype != torch.bool and m.dtype != torch.uint8:
            m = m != 0
    if K is None:
        K, c2w = identity_cameras()
    else:
        K, c2w = K[None], c2w[None]
    return ViewInputs(depth=d[None], mask=None if m is None else m[None],
                      image=None if image is None else image[None], image_kind=image_kind,
                      K=K, c2w=c2w, **kw)


def options_from_config(config, **over) -> RunOptions:
    """RunOptions from the reference's Config duck type (config.mdi.*)."""
    mdi = config.mdi
    aligner = mdi.alignment.aligner
    aligner = getattr(aligner, "value", aligner)
    rc = mdi.alignment.ransac
    opt = RunOptions(
        aligner=None if aligner is None else str(aligner),
        ransac_threshold=rc.inlier_threshold, ransac_max_iters=rc.max_iters,
        ransac_confidence=rc.confidence, ransac_sample_size=rc.sample_size,
        ransac_min_iters=rc.min_iters,
        subsample=mdi.subsample_factor,
        factor_min=mdi.adaptive_subsampling.factor_range_min,
        factor_max=mdi.adaptive_subsampling.factor_range_max,
        use_patch_mask=bool(mdi.use_num_sfm_points_mask),
        num_patches_small_axis=mdi.num_sfm_points_mask.num_patches_small_axis,
        patch_threshold=mdi.num_sfm_points_mask.threshold,
        grad_thresh=mdi.depth_grad_mask_thresh,
    )
    if opt.aligner not in ("lstsqrs", "ransac", "msac", "interp"):
        raise NotImplementedError(f"Unknown depth alignment strategy: {opt.aligner}")
    if opt.subsample != "adaptive" and not isinstance(opt.subsample, int):
        raise ValueError(f"Unsupported subsampling factor: {opt.subsample}")
    for k, v in over.items():
        setattr(opt, k, v)
    return opt


def ransac_seed() -> int:
    """Device-sampler seed tied to torch's global CPU generator, which the reference consumes
    (ransacs.py:131): torch.manual_seed(s) makes runs reproducible, and each call advances it."""
    return int(torch.randint(0, 2**62, (1,)).item())
