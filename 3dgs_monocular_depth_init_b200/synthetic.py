"""Deterministic synthetic scenes of the shapes named in BASELINE.json: smooth
ground-truth depth, predicted depth with a known scale/shift and noise, look-at poses on a
ring, SfM points sampled from the ground truth with gross outliers, foreign points and
clustered blobs (so some SfM-density patches exceed the threshold).

Generated with torch on the requested device (CPU for parity fixtures, CUDA for benchmarks);
a view is a pure function of (seed, view index, shape), so every rank generates identical data.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch


@dataclass
class SyntheticScene:
    depth: torch.Tensor        # [V,H,W] float32
    mask: torch.Tensor         # [V,H,W] bool
    image: torch.Tensor        # [V,H,W,3] uint8 (0..255)
    K: torch.Tensor            # [V,3,3] float32 cpu
    c2w: torch.Tensor          # [V,4,4] float32 cpu
    points: np.ndarray         # [Np,3] float32 (parser.points)
    points_rgb: np.ndarray     # [Np,3] uint8
    point_indices: dict        # image name -> int64 index array
    names: list
    true_scale_shift: np.ndarray  # [V,2]
    scene_scale: float = 4.0
    dataset_name: str = "synthetic"


def _look_at(eye: np.ndarray, target: np.ndarray) -> np.ndarray:
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, 0.0, 1.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, eye
    return c2w


def gt_depth(H: int, W: int, device, phase: float = 0.0) -> torch.Tensor:
    y = torch.arange(H, device=device, dtype=torch.float32)[:, None]
    x = torch.arange(W, device=device, dtype=torch.float32)[None, :]
    return (4.0 + 1.5 * torch.sin(x / (W / 6.5) + phase) + torch.cos(y / (H / 5.6))
            + 0.25 * torch.sin((x + y) / 37.0))


def make_scene(n_views: int, H: int, W: int, n_sfm: int = 5000, seed: int = 20261104,
               device: str = "cpu", low_conf_every: int = 50, image_u8: bool = True,
               outlier_frac: Optional[float] = None, keep_views: Optional[range] = None) -> SyntheticScene:
    """`keep_views` (a contiguous range): generate the scene of n_views views -- SfM table, foreign
    indices and all -- but materialise depth / mask / image, K, cam2world and names of those views only
    (a rank's block of a sharded scene; identical to slicing the full scene)."""
    dev = torch.device(device)
    rng = np.random.default_rng(seed)
    kept = range(n_views) if keep_views is None else keep_views
    k0, n_keep = (kept[0] if len(kept) else 0), len(kept)
    depth = torch.empty((n_keep, H, W), dtype=torch.float32, device=dev)
    mask = torch.ones((n_keep, H, W), dtype=torch.bool, device=dev)
    image = torch.empty((n_keep, H, W, 3), dtype=torch.uint8, device=dev)
    Ks, c2ws, names, true_h = [], [], [], []
    pts_all, idx_all = [], []
    n_pts = 0
    for v in range(n_views):
        g = torch.Generator(device=dev)
        g.manual_seed(seed + 1000 * v + 7)
        f = 1.2 * max(W, H) * rng.uniform(0.9, 1.1)
        K = np.array([[f, 0, W / 2], [0, f * rng.uniform(0.995, 1.005), H / 2], [0, 0, 1]])
        ang = 2 * math.pi * v / max(n_views, 1) + rng.uniform(-0.05, 0.05)
        eye = np.array([4 * math.cos(ang), 4 * math.sin(ang), rng.uniform(-0.5, 0.5)])
        c2w = _look_at(eye, np.array([0.0, 0.0, 0.0]) + rng.uniform(-0.2, 0.2, 3))
        G = gt_depth(H, W, dev, phase=0.37 * v)
        s_true, t_true = rng.uniform(0.5, 3.0), rng.uniform(-0.3, 0.3)
        here = v in kept
        if here:
            D = (G - t_true) / s_true * (1 + 0.005 * torch.randn((H, W), generator=g, device=dev))
            # 0.1 % of pixels: zero or negative-after-alignment depth
            bad = torch.rand((H, W), generator=g, device=dev) < 0.001
            D = torch.where(bad, -D * 0.5, D)
            depth[v - k0] = D
        if v % 2 == 1:  # MoGe-like mask: one invalid rectangle of ~5 % area
            rh, rw = int(H * 0.25), int(W * 0.2)
            r0, c0 = int(rng.integers(0, H - rh)), int(rng.integers(0, W - rw))
            if here:
                mask[v - k0, r0:r0 + rh, c0:c0 + rw] = False
        if here:
            image[v - k0] = torch.randint(0, 256, (H, W, 3), generator=g, device=dev, dtype=torch.uint8)
            image[v - k0, 0, 0, 0] = 255  # the reference asserts image.max() > 1
        # SfM points: half uniform, half in 8 blobs
        n_own = n_sfm if v % max(low_conf_every, 1) != max(low_conf_every, 1) - 1 else n_sfm // 8
        n_uni = n_own // 2
        px = rng.uniform(0, W - 1, n_uni)
        py = rng.uniform(0, H - 1, n_uni)
        centers = np.stack([rng.uniform(0.1 * W, 0.9 * W, 8), rng.uniform(0.1 * H, 0.9 * H, 8)], 1)
        which = rng.integers(0, 8, n_own - n_uni)
        bx = np.clip(centers[which, 0] + rng.normal(0, 0.03 * W, n_own - n_uni), 0, W - 1)
        by = np.clip(centers[which, 1] + rng.normal(0, 0.03 * H, n_own - n_uni), 0, H - 1)
        px, py = np.concatenate([px, bx]), np.concatenate([py, by])
        order = rng.permutation(n_own)
        px, py = px[order], py[order]
        Gc = G.cpu().numpy()
        z = Gc[np.clip(np.rint(py).astype(int), 0, H - 1), np.clip(np.rint(px).astype(int), 0, W - 1)]
        z = z * (1 + 0.002 * rng.normal(size=n_own))
        frac = outlier_frac if outlier_frac is not None else rng.uniform(0.1, 0.4)
        out = rng.uniform(size=n_own) < frac
        z = np.where(out, z * rng.uniform(1.3, 2.0, n_own), z)
        cam = np.linalg.inv(K) @ np.stack([px * z, py * z, z])
        world = (c2w[:3, :3] @ cam + c2w[:3, 3:4]).T
        if n_own < n_sfm:  # low-confidence view: the rest of its points lie behind the camera
            n_beh = n_sfm - n_own
            cam_b = np.stack([rng.uniform(-1, 1, n_beh), rng.uniform(-1, 1, n_beh), -rng.uniform(1, 5, n_beh)])
            world = np.concatenate([world, (c2w[:3, :3] @ cam_b + c2w[:3, 3:4]).T], 0)
            n_own = n_sfm
        pts_all.append(world.astype(np.float32))
        own_idx = np.arange(n_pts, n_pts + n_own)
        n_pts += n_own
        idx_all.append(own_idx)
        Ks.append(K.astype(np.float32))
        c2ws.append(c2w.astype(np.float32))
        names.append(f"view_{v:05d}")
        true_h.append((s_true, t_true))
    points = np.concatenate(pts_all, 0) if pts_all else np.zeros((0, 3), np.float32)
    # foreign points: 20 % extra indices drawn from all views' points
    point_indices = {}
    for v in range(n_views):
        own = idx_all[v]
        n_for = n_sfm // 5
        foreign = rng.integers(0, max(n_pts, 1), n_for)
        idx = np.concatenate([own, foreign])
        point_indices[names[v]] = idx[rng.permutation(len(idx))].astype(np.int64)
    points_rgb = rng.integers(0, 256, (n_pts, 3)).astype(np.uint8)
    sel = slice(k0, k0 + n_keep)
    return SyntheticScene(
        depth=depth, mask=mask, image=image,
        K=torch.from_numpy(np.stack(Ks[sel])) if n_keep else torch.zeros((0, 3, 3)),
        c2w=torch.from_numpy(np.stack(c2ws[sel])) if n_keep else torch.zeros((0, 4, 4)),
        points=points, points_rgb=points_rgb, point_indices=point_indices, names=names[sel],
        true_scale_shift=np.array(true_h[sel], dtype=np.float64).reshape(-1, 2),
    )


class SyntheticDataset:
    """Iterates dicts like the reference's Dataset (datasets/colmap.py:353-415)."""

    def __init__(self, parser: "SyntheticParser", split: str = "train"):
        self.parser = parser

    def __len__(self):
        return len(self.parser.scene.names)

    def __getitem__(self, i):
        s = self.parser.scene
        if i >= len(self):
            raise IndexError(i)
        return {
            "K": s.K[i],
            "camtoworld": s.c2w[i],
            "image": s.image[i].cpu().float(),
            "image_name": s.names[i],
        }


class SyntheticParser:
    """Duck type of the COLMAP Parser the path consumes (monocular_depth_init.py:103-181)."""

    DatasetCls = SyntheticDataset

    def __init__(self, scene: SyntheticScene):
        self.scene = scene
        self.points = scene.points
        self.points_rgb = scene.points_rgb
        self.point_indices = scene.point_indices
        self.scene_scale = scene.scene_scale
        self.dataset_name = scene.dataset_name
