import sys, importlib, numpy as np, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
PKG = "3dgs_monocular_depth_init_b200"
mod = lambda n: importlib.import_module(PKG + "." + n)
L = mod("_lib"); single = mod("_single"); E = mod("engine")
DEV = "cuda:0"
g = torch.Generator().manual_seed(7)
H, W = 600, 800
yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
a = 4 + 1.5 * torch.sin(xx / 123.0) + torch.cos(yy / 107.0) + 0.01 * torch.randn(H, W, generator=g)
m = torch.ones(H, W, dtype=torch.bool)
eng = E.get_engine(DEV)
vin = single.single_view_inputs(a.to(DEV), m.to(DEV), DEV)
opt = E.RunOptions(aligner=None, subsample="adaptive", use_patch_mask=False, flags=L.KEEP_PRED_MASK | L.KEEP_SUBSAMPLE,
                   want_xyz=False, want_rgb=False, stop_after="mask")
snaps = []
real = eng.lib._lib.mdi_depth_range
class P:
    def __init__(s, lib): s._lib = lib
    def __getattr__(s, n):
        f = getattr(s._lib, n)
        if n != "mdi_depth_range": return f
        def g(*a):
            rc = f(*a); torch.cuda.synchronize()
            snaps.append(eng._ws["stats_ws"][:48].cpu().numpy().view(np.uint32).copy()); return rc
        return g
eng.lib = P(eng.lib)
res = eng.run(vin, opt)
hdr = snaps[0]
names = ["n_sample", "n_cand", "min", "max", "count", "below0", "below1", "brlo0", "brlo1", "brhi0", "brhi1", "run_exact"]
print(dict(zip(names, hdr.tolist())), res.extras)
def k2f(k):
    u = (k & 0x7FFFFFFF) if (k & 0x80000000) else (~k & 0xFFFFFFFF)
    return np.array([u], dtype=np.uint32).view(np.float32)[0]
for n in ["min", "max", "brlo0", "brhi0", "brlo1", "brhi1"]:
    print(n, k2f(int(hdr[names.index(n)])))
v = np.sort(a.numpy().ravel())
n = v.size
print("true q1,q3", v[int(0.25 * (n - 1))], v[int(0.75 * (n - 1))], "count below brlo0", (v < k2f(int(hdr[7]))).sum(), "in br0", ((v >= k2f(int(hdr[7]))) & (v <= k2f(int(hdr[9])))).sum())
