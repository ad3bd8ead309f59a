#!/usr/bin/env python
"""C2 step time with / without the raw-depth statistics overlap (engine.OVERLAP_STATS), same process."""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench
import torch
dev = torch.device("cuda", 0)
L, mdi, emod = bench.mod("_lib"), bench.mod("monocular_depth_init"), bench.mod("engine")
wl = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
V, H, W, n_sfm, aligner, subsample = bench.WORKLOADS[wl]
cfg = bench.make_config(aligner, subsample)
sc = bench.mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev), low_conf_every=50)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
torch.manual_seed(1234)
def step():
    return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
for ov in (False, True, False, True):
    emod.OVERLAP_STATS = ov
    for _ in range(4): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): r = step()
    e1.record(); torch.cuda.synchronize()
    print(json.dumps({"workload": wl, "overlap": ov, "ms_per_step": round(e0.elapsed_time(e1) / 20, 4), "retry": bool(r.extras.get("stats_retry", False))}), flush=True)
