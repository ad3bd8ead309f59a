"""Host-side floor of a step: the C2 call sequence on views so small that the GPU work is negligible."""
import sys, json, time
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench, torch
dev = torch.device("cuda", 0)
L, mdi, emod = bench.mod("_lib"), bench.mod("monocular_depth_init"), bench.mod("engine")
cfg = bench.make_config("ransac", "adaptive")
for (V, H, W, n_sfm) in ((200, 264, 264, 300), (200, 840, 1297, 5000)):
    sc = bench.mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev), low_conf_every=50)
    sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
    torch.manual_seed(1234)
    def step():
        return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
    for _ in range(5): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(50): r = step()
    e1.record(); torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 50 * 1e3
    print(json.dumps({"V": V, "H": H, "W": W, "ms_per_step_events": round(e0.elapsed_time(e1) / 50, 4), "ms_per_step_wall": round(wall, 4), "launches": emod.get_engine(dev).launches}))
