"""Dead zone around region boundaries (reference: depth_alignment/segmentation/region_margin.py).

The reference runs these torch ops on the depth map's device; they are restated with the same ops
(replicate padding + F.conv2d box filter, utils/image_filtering.py:24-56) so the snapped comparison
sees the same float32 values."""
import torch
import torch.nn.functional as F


def snap_to_int_if_close(x: torch.Tensor):
    """region_margin.py:6-13."""
    nearest = x.round()
    return torch.where(torch.isclose(x, nearest), nearest, x)


def get_actual_margin_size(image_shape, region_margin) -> int:
    """region_margin.py:16-18."""
    KERNEL_REFERENCE_IMSIZE = 1297
    return int(region_margin * max(image_shape) / KERNEL_REFERENCE_IMSIZE)


def box_blur2d(x: torch.Tensor, ksize: int) -> torch.Tensor:
    """image_filtering.py box_blur_kernel + filter2d: [B,C,H,W] float, replicate padding."""
    if ksize % 2 == 0:
        raise ValueError("Box blur kernel size must be odd.")
    kernel = (torch.ones((ksize, ksize)) / (ksize * ksize)).to(x.device)
    pad = ksize // 2
    padded = F.pad(x, (pad, pad, pad, pad), mode="replicate")
    kernel = kernel.flip(0, 1)[None, None].repeat(x.size(1), 1, 1, 1)
    return F.conv2d(padded, kernel, stride=1, groups=x.size(1))


def calculate_region_margin_mask(region_map: torch.Tensor, region_margin: int) -> torch.Tensor:
    """region_margin.py:21-33: True where the box-blurred region map still equals the region id."""
    if region_margin == 0:
        return torch.ones_like(region_map, dtype=torch.bool)
    kernel_size = 2 * get_actual_margin_size(region_map.shape, region_margin) + 1
    blurred = box_blur2d(region_map[None, None].float(), ksize=kernel_size)[0, 0]
    return snap_to_int_if_close(blurred) == region_map
