import sys
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
from helpers import mod
DEV = torch.device("cuda", 0)
syn = mod("synthetic"); mdi = mod("monocular_depth_init"); L = mod("_lib")
sc = syn.make_scene(n_views=6, H=240, W=320, n_sfm=3000, seed=5, low_conf_every=50)
cfg = mod("config").Config()
for seed in (1234, 1, 2, 3):
    torch.manual_seed(seed)
    res = mdi.points_from_views(cfg, sc.depth.to(DEV), sc.mask.to(DEV), sc.image.to(DEV), sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, DEV, L.IMAGE_U8)
    ss = res.scale_shift.cpu().numpy()
    print(seed, "iters", res.ransac_iters.cpu().tolist(), "inl", res.ransac_inliers.cpu().tolist(), "nvalid", res.n_valid.cpu().tolist())
    print("   err", np.abs(ss - sc.true_scale_shift).max(axis=1).round(4).tolist())
