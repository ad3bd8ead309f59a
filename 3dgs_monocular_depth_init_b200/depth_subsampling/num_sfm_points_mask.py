"""SfM-density patch mask (reference: gs_init_compare/depth_subsampling/num_sfm_points_mask.py)."""
import torch

from .. import _lib as L
from .._single import coords_to_slots, single_view_inputs
from ..engine import RunOptions, get_engine, patch_geometry
from .config import NumSfMPointsMaskConfig


def calculate_patch_sizes(image_shape, num_patches_small_axis):
    """((patch_h, patch_w), (grid_h, grid_w)) -- num_sfm_points_mask.py:7-35."""
    ph, pw, gh, gw = patch_geometry(int(image_shape[0]), int(image_shape[1]), int(num_patches_small_axis))
    return (ph, pw), (gh, gw)


def num_sfm_points_mask(sfm_points_camera: torch.Tensor, imsize, sfm_pts_mask_config: NumSfMPointsMaskConfig,
                        device=None) -> torch.Tensor:
    """bool [H,W]: False in patches holding more than `threshold` SfM points.  The reference
    returns a CPU tensor (num_sfm_points_mask.py:43); so does this (computed on the GPU)."""
    dev = torch.device(device) if device is not None else (
        sfm_points_camera.device if sfm_points_camera.is_cuda else torch.device("cuda"))
    H, W = int(imsize[0]), int(imsize[1])
    eng = get_engine(dev)
    uv_s, z_s, n_valid, m = coords_to_slots(sfm_points_camera, torch.zeros(sfm_points_camera.shape[1]), dev)
    dummy = torch.zeros((H, W), dtype=torch.float32, device=dev)
    vin = single_view_inputs(dummy, None, dev, uv=uv_s, z=z_s, n_valid=n_valid, sfm_counts=[m])
    opt = RunOptions(aligner=None, use_patch_mask=True, flags=0, subsample=1,
                     num_patches_small_axis=sfm_pts_mask_config.num_patches_small_axis,
                     patch_threshold=sfm_pts_mask_config.threshold, want_xyz=False, want_rgb=False,
                     stop_after="mask")
    res = eng.run(vin, opt)
    return res.keep.view(H, W).bool().cpu()
