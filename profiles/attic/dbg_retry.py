import sys, importlib, numpy as np, torch
sys.path.insert(0, "/root/repo")
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); mdi = mod("monocular_depth_init"); syn = mod("synthetic"); E = mod("engine")
dev = torch.device("cuda", 0)
cfg = mod("config").Config()
cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
cfg.mdi.subsample_factor = "adaptive"
sc = syn.make_scene(n_views=200, H=840, W=1297, n_sfm=5000, seed=20261104, device=str(dev), low_conf_every=50)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
eng = E.get_engine(dev)
snaps = []
class P:
    def __init__(s, lib): s._lib = lib
    def __getattr__(s, n):
        f = getattr(s._lib, n)
        if n != "mdi_depth_range": return f
        def g(*a):
            rc = f(*a); torch.cuda.synchronize()
            snaps.append(eng._ws["stats_ws"][:200 * 216].cpu().numpy().view(np.uint32).reshape(200, 54).copy()); return rc
        return g
eng.lib = P(eng.lib)
torch.manual_seed(1234)
names = ["n_sample", "n_cand", "min", "max", "count", "below0", "below1", "brlo0", "brlo1", "brhi0", "brhi1", "sh0", "sh1", "run_exact"]
for step in range(40):
    snaps.clear()
    res = mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
    h = snaps[0]
    fc = h[:, 32]
    bad = np.nonzero(fc)[0]
    if step == 0: print('fast views', int(h[:, 46].sum()), 'of', len(h), 'force_retry', int(h[:, 47].sum()))
    if len(bad) == 0: continue
    print(step, "retry", res.extras.get("stats_retry"), "bad views", bad.tolist(), [int(fc[i]) for i in bad])
    for i in bad[:3]:
        d = dict(zip(names, h[i, :14].tolist()))
        n = d["count"]
        print("    fast", int(h[i, 46]), "force", int(h[i, 47]), "dlo/dhi/ds", h[i, 48:54].view(np.float32).tolist())
        print("   ", d, "r1", 0.25 * (n - 1), "r3", 0.75 * (n - 1), "ss", res.scale_shift[i].tolist())
