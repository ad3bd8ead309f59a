import sys, time, importlib, cProfile, pstats, io
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
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
def step():
    return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
for _ in range(3): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): step()
torch.cuda.synchronize()
print("ms/step", (time.perf_counter() - t0) / 20 * 1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(20): step()
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(35); print(s.getvalue())
