import sys
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
torch.manual_seed(1234)
eng = emod.get_engine(dev)
orig = eng.run
def run(vin, opt):
    # stop the retry from hiding the flags: run once to the mask with retry visible
    return orig(vin, opt)
for it in range(3):
    emod.OVERLAP_STATS = True
    r = mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
    ss = r.scale_shift.cpu().numpy(); st = r.status.cpu().numpy(); it_ = r.ransac_iters.cpu().numpy(); inl = r.ransac_inliers.cpu().numpy()
    bad = np.where(~((ss[:, 0] > 0) & np.isfinite(ss[:, 0]) & np.isfinite(ss[:, 1])))[0]
    print("retry", r.extras.get("stats_retry"), "non-increasing fits at views", bad.tolist(), "ss", ss[bad].tolist(), "iters", it_[bad].tolist(), "inl", inl[bad].tolist(), "status", st[bad].tolist())
    lo = r.lo_hi.cpu().numpy()
    # views whose aligned range includes -42
    print("  min aligned", float(r.quantiles[:, 2].min()), "views with min <= -42:", int((r.quantiles[:, 2] <= -42).sum()))
