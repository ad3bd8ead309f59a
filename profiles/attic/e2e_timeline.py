# device time line of one chunked end-to-end step (copies + kernels, with stream ids)
import sys, time, importlib
sys.path.insert(0, "/root/repo")
import torch, numpy as np
from torch.profiler import profile, ProfilerActivity
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); mdi = mod("monocular_depth_init"); syn = mod("synthetic")
dev = torch.device("cuda", 0)
cfg = mod("config").Config()
cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
cfg.mdi.subsample_factor = "adaptive"
sc = syn.make_scene(n_views=200, H=840, W=1297, n_sfm=5000, seed=1, device="cpu", low_conf_every=50)
dh, mh, ih = sc.depth.pin_memory(), sc.mask.pin_memory(), sc.image.pin_memory()
chunks = int(sys.argv[1]) if len(sys.argv) > 1 else 4
def step():
    return mdi.points_from_host_views(cfg, dh, mh, ih, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev,
                                      L.IMAGE_U8, chunks=chunks)
for _ in range(3): step()
torch.cuda.synchronize()
mdi._TRACE = []
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    torch.cuda.synchronize(); w0 = time.perf_counter(); step(); torch.cuda.synchronize(); wall = (time.perf_counter() - w0) * 1e3
tr = mdi._TRACE; mdi._TRACE = None
print("marks:", "  ".join(f"{l}: {tr[0][1].elapsed_time(e):.2f}" for l, e in tr[1:]), f" wall {wall:.2f}")
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
for e in evs:
    d = e.time_range.end - e.time_range.start
    if d > 30 or "Memcpy" in e.name or (e.time_range.start - t0) > 19000:
        print(f"{(e.time_range.start - t0) / 1e3:8.3f} ms  dur {d / 1e3:7.3f} ms  {e.name[:70]}")
