"""cProfile of the host side of a step (small views: the GPU work is negligible)."""
import sys, cProfile, pstats, io
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench, torch
dev = torch.device("cuda", 0)
L, mdi, emod = bench.mod("_lib"), bench.mod("monocular_depth_init"), bench.mod("engine")
cfg = bench.make_config("ransac", "adaptive")
V, H, W, n_sfm = 200, 264, 264, 300
sc = bench.mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev), low_conf_every=50)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
torch.manual_seed(1234)
def step():
    return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
for _ in range(5): step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(200): step()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28)
print(s.getvalue()[:6000])
