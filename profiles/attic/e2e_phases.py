"""Breaks the end-to-end step (host buffers in, rgb out) into its phases with CUDA events."""
import sys, time, importlib
sys.path.insert(0, "/root/repo")
import torch
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); mdi = mod("monocular_depth_init"); syn = mod("synthetic")
dev = torch.device("cuda", 0)
cfg = mod("config").Config()
cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
cfg.mdi.subsample_factor = "adaptive"
sc = syn.make_scene(n_views=200, H=840, W=1297, n_sfm=5000, seed=1, device=str(dev), low_conf_every=50)
depth_h = sc.depth.cpu().pin_memory(); mask_h = sc.mask.cpu().pin_memory(); image_h = sc.image.cpu().pin_memory()
depth_d = torch.empty_like(sc.depth); mask_d = torch.empty(sc.mask.shape, dtype=torch.uint8, device=dev)
image_d = torch.empty_like(sc.image)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
def run(image_src, copy_image):
    e = [ev() for _ in range(5)]
    e[0].record()
    depth_d.copy_(depth_h, non_blocking=True); mask_d.copy_(mask_h.view(torch.uint8), non_blocking=True)
    e[1].record()
    if copy_image: image_d.copy_(image_h, non_blocking=True)
    e[2].record()
    res = mdi.points_from_views(cfg, depth_d, mask_d, image_d if copy_image else image_src, sc.K, sc.c2w, sc.names,
                                sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
    e[3].record()
    rgb_h = torch.empty(res.rgb.shape, dtype=torch.float32, pin_memory=True); rgb_h.copy_(res.rgb, non_blocking=True)
    e[4].record(); torch.cuda.synchronize()
    return [e[i].elapsed_time(e[i + 1]) for i in range(4)], res.total
for name, ci in (("image read in place over PCIe", False), ("image copied H2D", True)):
    for _ in range(2): run(image_h, ci)
    t0 = time.perf_counter(); ph, tot = run(image_h, ci); wall = (time.perf_counter() - t0) * 1e3
    print(f"{name}: h2d depth+mask {ph[0]:.2f} ms, h2d image {ph[1]:.2f} ms, pipeline {ph[2]:.2f} ms, d2h rgb {ph[3]:.2f} ms, wall {wall:.2f} ms, kept {tot}")
gb = (depth_h.numel() * 4 + mask_h.numel()) / 1e9
print(f"depth+mask {gb:.3f} GB")
