"""Phase timing of ransac_kernel (clock64 deltas of thread 0 summed per phase into a __device__ array and read
back through a debug entry point mdi_debug_ransac_prof).  It needs the instrumented build: the RP(slot) marks
and the debug entry point were patched into k2_ransac.cu for the measurement and are NOT in the tree (DESIGN section 6
quotes the result: 123 k cycles per view before the rewrite, 36 % scoring, 23 % barrier wait, 13 % refits,
12 % sample fits, 10 % replay)."""
import sys, ctypes as C
sys.path.insert(0, "/root/repo")
import bench, torch, numpy as np
dev = torch.device("cuda", 0)
L, mdi = bench.mod("_lib"), bench.mod("monocular_depth_init")
V, H, W, n_sfm, aligner, subsample = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
cfg = bench.make_config(aligner, subsample)
sc = bench.mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev), low_conf_every=50)
sfm = mdi.SfmTable(sc.points, sc.point_indices, sc.names, dev)
torch.manual_seed(1234)
lib = L.load()
out = (C.c_ulonglong * 16)()
def step():
    return mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices, dev, L.IMAGE_U8, sfm=sfm)
for _ in range(3): step()
lib.mdi_debug_ransac_prof(out, 1)
N = 10
for _ in range(N): r = step()
lib.mdi_debug_ransac_prof(out, 1)
o = np.array(list(out), dtype=np.float64)
views = o[8]
names = ["stage", "sample_fit", "scoring", "score_barrier", "refit_batch", "scan->refit barrier", "scan"]
print("views", views / N, "rounds/view", o[7] / views, "KH1/2/4 rounds per view", o[10:13] / views)
print("cycles per view total", o[9] / views, "batches/view", o[13] / views, "refits/view", o[14] / views)
for i, nme in enumerate(names):
    print(f"{nme:14s} {o[i] / views:10.0f} cycles/view  {100 * o[i] / o[9]:5.1f}%")
it = r.ransac_iters.cpu().numpy()
print("iters percentiles", np.percentile(it, [0, 10, 25, 50, 75, 90, 100]))
