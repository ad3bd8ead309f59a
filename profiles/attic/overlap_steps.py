import sys, json
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench, torch, numpy as np
dev = torch.device("cuda", 0)
L, mdi, emod = bench.mod("_lib"), bench.mod("monocular_depth_init"), bench.mod("engine")
V, H, W, n_sfm, aligner, subsample = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
cfg = bench.make_config(aligner, subsample)
sc = bench.mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev), low_conf_every=50)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
def step():
    return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
for ov in (False, True):
    emod.OVERLAP_STATS = ov
    torch.manual_seed(1234)
    for _ in range(4): step()
    out = []
    evs = []
    torch.cuda.synchronize()
    for i in range(24):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = step(); e1.record()
        evs.append((e0, e1, bool(r.extras.get("stats_retry", False))))
    torch.cuda.synchronize()
    ms = [(round(a.elapsed_time(b), 3), rt) for a, b, rt in evs]
    clean = [m for m, rt in ms if not rt]
    print(json.dumps({"overlap": ov, "clean_steps": len(clean), "median_clean_ms": float(np.median(clean)) if clean else None,
                      "median_retry_ms": float(np.median([m for m, rt in ms if rt])) if len(clean) < len(ms) else None, "all": ms}))
