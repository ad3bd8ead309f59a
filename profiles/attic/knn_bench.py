"""Times mdi_knn (k=4 incl. self) and LOF (k=40) on a synthetic dense surface cloud with floaters."""
import sys, time, importlib
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests")]
L = importlib.import_module("3dgs_monocular_depth_init_b200._lib")
K = importlib.import_module("3dgs_monocular_depth_init_b200.knn")
from test_knn_lof import _cloud
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
cpu_n = int(sys.argv[2]) if len(sys.argv) > 2 else 0
pts = _cloud(n, 5, outliers=0.01)
x = torch.from_numpy(pts).cuda()
lib = L.load()
for name, fn in (("knn k=4", lambda: K.knn(x, 4)), ("lof k=40", lambda: K.local_outlier_factor(x, 40))):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter(); r = fn(); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"{name}: n={n} wall {dt*1e3:.1f} ms  ({n/dt/1e6:.2f} Mpts/s)")
    lib.mdi_profile_enable(1); fn(); torch.cuda.synchronize(); lib.mdi_profile_enable(0)
    for k, v in L.profile_collect().items():
        print(f"    {k:24s} {v['ms']:9.3f} ms")
if cpu_n:
    from sklearn.neighbors import LocalOutlierFactor, NearestNeighbors
    s = pts[:cpu_n]
    t = time.perf_counter(); NearestNeighbors(n_neighbors=4).fit(s).kneighbors(s); print("sklearn knn k=4", cpu_n, f"{time.perf_counter()-t:.2f} s")
    t = time.perf_counter(); LocalOutlierFactor(n_neighbors=40, n_jobs=-1).fit_predict(s); print("sklearn lof k=40", cpu_n, f"{time.perf_counter()-t:.2f} s")
