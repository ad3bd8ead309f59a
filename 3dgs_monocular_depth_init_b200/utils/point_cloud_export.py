"""PLY export of point clouds (reference: gs_init_compare/utils/point_cloud_export.py:9-100).

The reference writes through open3d (`write_point_cloud(..., write_ascii=False)`); open3d is not a
dependency here, so the same file layout is written directly: binary little-endian PLY, vertices as
`double x y z`, colours as `uchar red green blue` with open3d's colour conversion
`round(clamp(c, 0, 1) * 255)`.  Host-side I/O, not on the kernel path."""
from __future__ import annotations

import logging
from pathlib import Path
from typing import Optional

import numpy as np


def write_ply(path: Path, pts: np.ndarray, rgbs: Optional[np.ndarray]) -> None:
    pts = np.asarray(pts, dtype=np.float64).reshape(-1, 3)
    n = pts.shape[0]
    fields = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")]
    header = ["ply", "format binary_little_endian 1.0", "comment Created by 3dgs_monocular_depth_init_b200",
              f"element vertex {n}", "property double x", "property double y", "property double z"]
    if rgbs is not None:
        rgbs = np.asarray(rgbs, dtype=np.float64).reshape(-1, 3)
        if rgbs.shape[0] != n:
            raise ValueError(f"write_ply: {n} points but {rgbs.shape[0]} colours")
        fields += [("red", "u1"), ("green", "u1"), ("blue", "u1")]
        header += ["property uchar red", "property uchar green", "property uchar blue"]
    header.append("end_header")
    rec = np.empty(n, dtype=np.dtype(fields))
    rec["x"], rec["y"], rec["z"] = pts[:, 0], pts[:, 1], pts[:, 2]
    if rgbs is not None:
        c = np.round(np.clip(rgbs, 0.0, 1.0) * 255.0).astype(np.uint8)
        rec["red"], rec["green"], rec["blue"] = c[:, 0], c[:, 1], c[:, 2]
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(rec.tobytes())


def read_ply(path: Path):
    """Reader of the files written above -> (pts float64 [N,3], rgbs uint8 [N,3] | None)."""
    with open(path, "rb") as f:
        n, has_rgb = 0, False
        while True:
            line = f.readline().decode("ascii").strip()
            if line.startswith("element vertex"):
                n = int(line.split()[-1])
            elif line == "property uchar red":
                has_rgb = True
            elif line == "end_header":
                break
        fields = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")] + ([("red", "u1"), ("green", "u1"), ("blue", "u1")] if has_rgb else [])
        rec = np.frombuffer(f.read(), dtype=np.dtype(fields), count=n)
    pts = np.stack([rec["x"], rec["y"], rec["z"]], 1)
    return pts, (np.stack([rec["red"], rec["green"], rec["blue"]], 1) if has_rgb else None)


def export_point_cloud_to_ply(pts: np.ndarray, rgbs: Optional[np.ndarray], output_dir: Path, depth_pts_filename: str,
                              outlier_std_dev: Optional[float] = None, sphere_sizes: Optional[np.ndarray] = None):
    """Same arguments and files as the reference (point_cloud_export.py:9-100): `<name>.ply`, plus
    `<name>_size_color.ply` when sphere_sizes is given.  The `<name>_spheres.ply` triangle mesh
    (open3d sphere primitives; no caller in the reference passes sphere_sizes) is not written."""
    output_dir = Path(output_dir)
    pts = np.asarray(pts)
    if outlier_std_dev is not None:   # :27-35
        mean = np.mean(pts, axis=0)
        std_dev = np.std(pts, axis=0)
        mask = np.all(np.abs((pts - mean) / std_dev) < outlier_std_dev, axis=1)
        pts = pts[mask]
        if rgbs is not None:
            rgbs = np.asarray(rgbs)[mask]
        if sphere_sizes is not None:
            sphere_sizes = np.asarray(sphere_sizes)[mask]
    if sphere_sizes is not None:      # :43-66
        sphere_sizes = np.asarray(sphere_sizes)
        min_size = np.percentile(sphere_sizes, 1)
        max_size = np.percentile(sphere_sizes, 80)
        norm = (sphere_sizes - min_size) / (max_size - min_size + 1e-8)
        write_ply(output_dir / f"{depth_pts_filename}_size_color.ply", pts,
                  np.stack([norm, np.zeros_like(norm), 1.0 - norm], axis=1))
        logging.info("Saved point cloud with sphere_size-based color to %s",
                     output_dir / f"{depth_pts_filename}_size_color.ply")
        logging.warning("export_point_cloud_to_ply: the %s_spheres.ply mesh needs open3d and is not written",
                        depth_pts_filename)
    write_ply(output_dir / f"{depth_pts_filename}.ply", pts, rgbs)
    logging.info("Saved point cloud to %s", output_dir / f"{depth_pts_filename}.ply")
