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
depth_h = sc.depth.cpu().pin_memory(); mask_h = sc.mask.cpu().pin_memory(); image_h = sc.image.cpu().pin_memory()
def run(ch):
    return mdi.points_from_host_views(cfg, depth_h, mask_h, image_h, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, chunks=ch)
for ch in (1, 4, 8):
    for _ in range(2): run(ch)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): run(ch)
    torch.cuda.synchronize(); print(f"chunks={ch}: {(time.perf_counter()-t0)/5*1e3:.2f} ms/step")
pr = cProfile.Profile(); pr.enable()
for _ in range(5): run(4)
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(14); print(s.getvalue()[:3500])
