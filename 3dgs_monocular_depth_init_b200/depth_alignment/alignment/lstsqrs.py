"""Least-squares scale/shift alignment on the B200 (reference: depth_alignment/alignment/lstsqrs.py)."""
import torch

from ... import _lib as L
from ..._single import coords_to_slots, single_view_inputs
from ...depth_prediction.predictors.depth_predictor_interface import PredictedDepth
from ...engine import RunOptions, get_engine
from ..interface import DepthAlignmentResult, DepthAlignmentStrategy


def align_depth_least_squares(depth: torch.Tensor, gt_depth: torch.Tensor):
    """depth [2,N] = (predicted depth ; ones), gt_depth [N] -> (scale, shift) 0-dim tensors on the
    input device (lstsqrs.py:9-26).  The second row is ignored (it is ones by contract): the
    normal equations are accumulated in fp64 on the GPU and solved with pinv's rank rule."""
    dev = depth.device
    n = int(gt_depth.numel())
    eng = get_engine(dev)
    # a 1 x n "image" whose pixel i holds d_i, with coords (u=i, v=0): reuses mdi_gather_fit
    d = depth[0].reshape(1, n).float().contiguous()
    uv = torch.stack([torch.arange(n, device=dev), torch.zeros(n, dtype=torch.long, device=dev)])
    uv_s, z_s, n_valid, m = coords_to_slots(uv, gt_depth, dev)
    if n == 0:
        z = torch.zeros((), device=dev)
        return z, z.clone()
    vin = single_view_inputs(d, None, dev, uv=uv_s, z=z_s, n_valid=n_valid, sfm_counts=[m])
    res = eng.run(vin, RunOptions(aligner="lstsqrs", use_patch_mask=False, stop_after="fit"))
    return res.scale_shift[0, 0], res.scale_shift[0, 1]


def aligned_depth(depth: torch.Tensor, scale_shift: torch.Tensor) -> torch.Tensor:
    """depth * scale + shift, each op rounded separately (lstsqrs.py:52)."""
    eng = get_engine(depth.device)
    d = depth.float().contiguous()
    out = torch.empty_like(d)
    ss = scale_shift.reshape(-1)[:2].contiguous()
    L.check(eng.lib.mdi_apply_scale_shift(d.data_ptr(), d.numel(), ss.data_ptr(), out.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream), "mdi_apply_scale_shift")
    return out


def fit_view(predicted_depth: PredictedDepth, coords: torch.Tensor, gt: torch.Tensor, opt: RunOptions):
    dev = predicted_depth.depth.device
    eng = get_engine(dev)
    uv_s, z_s, n_valid, m = coords_to_slots(coords, gt, dev)
    vin = single_view_inputs(predicted_depth.depth, None, dev, uv=uv_s, z=z_s, n_valid=n_valid, sfm_counts=[m])
    opt.use_patch_mask = False
    opt.stop_after = "fit"
    return eng.run(vin, opt)


class DepthAlignmentLstSqrs(DepthAlignmentStrategy):
    @classmethod
    def align(cls, predicted_depth: PredictedDepth, sfm_points_camera_coords: torch.Tensor,
              sfm_points_depth: torch.Tensor, *args, **kwargs) -> DepthAlignmentResult:
        res = fit_view(predicted_depth, sfm_points_camera_coords, sfm_points_depth, RunOptions(aligner="lstsqrs"))
        return DepthAlignmentResult(aligned_depth=aligned_depth(predicted_depth.depth, res.scale_shift),
                                    mask=predicted_depth.mask)
