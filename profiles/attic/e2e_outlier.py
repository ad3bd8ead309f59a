import sys, time, importlib, gc
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
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
for _ in range(5):
    mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
torch.cuda.synchronize()
def st():
    s = torch.cuda.memory_stats()
    return s.get("num_device_alloc", 0), s.get("num_device_free", 0), s.get("num_alloc_retries", 0)
for i in range(20):
    a = st(); g0 = gc.get_count(); t0 = time.perf_counter()
    r = mdi.points_from_host_views(cfg, depth_h, mask_h, image_h, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8)
    dt = (time.perf_counter() - t0) * 1e3
    b = st()
    print(f"step {i}: {dt:7.2f} ms  device allocs +{b[0]-a[0]} frees +{b[1]-a[1]} retries +{b[2]-a[2]} gc {g0}->{gc.get_count()} reserved {torch.cuda.memory_reserved()/1e9:.2f} GB")
