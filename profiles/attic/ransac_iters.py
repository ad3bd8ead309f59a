import sys, importlib
sys.path.insert(0, "/root/repo")
import torch, numpy as np
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); mdi = mod("monocular_depth_init"); syn = mod("synthetic")
dev = torch.device("cuda", 0)
cfg = mod("config").Config()
cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
cfg.mdi.subsample_factor = "adaptive"
sc = syn.make_scene(n_views=200, H=840, W=1297, n_sfm=5000, seed=1, device=str(dev), low_conf_every=50)
res = mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8)
it = res.ransac_iters.cpu().numpy()
print("n_iter percentiles", None if it is None else np.percentile(it, [0, 10, 50, 90, 100]))
nv = res.n_valid.cpu().numpy(); ni = res.ransac_inliers.cpu().numpy()
print("n_valid percentiles", np.percentile(nv, [0, 10, 50, 90, 100]))
print("inlier ratio percentiles", np.percentile(ni / np.maximum(nv, 1), [0, 10, 50, 90, 100]))
print("hist of iters", np.bincount(np.minimum(it // 16, 20)))
a = cfg.mdi.alignment
print({k: getattr(a.ransac, k) for k in dir(a.ransac) if not k.startswith("_")} if hasattr(a, "ransac") else dir(a))
