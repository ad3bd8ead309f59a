# host-side time line of one step: when (relative to the call) each C-ABI launch is issued
import sys, time, importlib
sys.path.insert(0, "/root/repo")
import torch, numpy as np
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); mdi = mod("monocular_depth_init"); syn = mod("synthetic")
dev = torch.device("cuda", 0)
cfg = mod("config").Config()
cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum("ransac")
cfg.mdi.subsample_factor = "adaptive"
sc = syn.make_scene(n_views=200, H=840, W=1297, n_sfm=5000, seed=1, device=str(dev), low_conf_every=50)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
eng = mod("engine").get_engine(dev)
def step():
    return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
for _ in range(5): step()
torch.cuda.synchronize()
marks = []
names = ["mdi_setup_cameras", "mdi_project_filter", "mdi_gather_fit", "mdi_ransac", "mdi_depth_range", "mdi_keep_mask",
         "mdi_scan_tiles", "mdi_emit_points_capped"]
for nm in names:
    fn = getattr(eng.lib, nm)
    def wrap(*a, _fn=fn, _nm=nm):
        marks.append((_nm, time.perf_counter()))
        r = _fn(*a)
        marks.append((_nm + " ret", time.perf_counter()))
        return r
    eng.lib.__dict__[nm] = wrap
engm = mod("engine")
def wrap_obj(obj, nm):
    fn = getattr(obj, nm)
    def w(*a, **k):
        marks.append((nm, time.perf_counter()))
        r = fn(*a, **k)
        marks.append((nm + " ret", time.perf_counter()))
        return r
    setattr(obj, nm, w)
wrap_obj(eng, "build_views"); wrap_obj(eng, "_upload"); wrap_obj(mdi, "options_from_config"); wrap_obj(eng, "run")
rows = []
for _ in range(30):
    marks.clear()
    t0 = time.perf_counter()
    step()
    t1 = time.perf_counter()
    rows.append([(m[1] - t0) * 1e6 for m in marks] + [(t1 - t0) * 1e6])
    labels = [m[0] for m in marks] + ["return"]
torch.cuda.synchronize()
med = np.median(np.array(rows), axis=0)
for l, m in zip(labels, med): print(f"{m:8.1f} us  {l}")
