"""Times mdi_min_extents + mdi_merge_subsample on a synthetic dense cloud; prints per-kernel ms."""
import sys, time, ctypes as C, importlib
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests")]
L = importlib.import_module("3dgs_monocular_depth_init_b200._lib")
from test_native_subsample import _cams, _cloud
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
n_cams = int(sys.argv[2]) if len(sys.argv) > 2 else 200
cpu_n = int(sys.argv[3]) if len(sys.argv) > 3 else 0
lib = L.load(); dev = "cuda:0"
K, P, sizes = _cams(n_cams, 5, W=1297, H=840)
pts, rgb = _cloud("surface", n, 3)
pts[:, :2] *= 1.5
tp, tc, tK, tP, ts = (torch.from_numpy(a).to(dev) for a in (pts, rgb, K, P, sizes))
ext = torch.empty(n, device=dev)
st = torch.cuda.current_stream().cuda_stream
ws = torch.empty(int(lib.mdi_merge_workspace_bytes(n)), dtype=torch.uint8, device=dev)
op = torch.empty((n, 3), device=dev); oc = torch.empty((n, 3), device=dev)
n_out, n_lev = C.c_int64(), C.c_int32()
def run():
    L.check(lib.mdi_min_extents(L.ptr(tp), n, L.ptr(tK), L.ptr(tP), L.ptr(ts), n_cams, L.ptr(ext), st))
    L.check(lib.mdi_merge_subsample(L.ptr(tp), L.ptr(tc), L.ptr(ext), n, 1.1, 1.0, 0, L.ptr(ws), L.ptr(op), L.ptr(oc),
                                    C.byref(n_out), C.byref(n_lev), st))
run(); torch.cuda.synchronize()
t = time.perf_counter(); run(); torch.cuda.synchronize(); dt = time.perf_counter() - t
print(f"n={n} cams={n_cams} -> {n_out.value} points, {n_lev.value} levels, wall {dt*1e3:.2f} ms, ws {ws.numel()/1e6:.0f} MB")
print("visible frac", float((ext > 0).float().mean()))
lib.mdi_profile_enable(1); run(); torch.cuda.synchronize(); lib.mdi_profile_enable(0)
for k, v in L.profile_collect().items():
    print(f"  {k:28s} {v['ms']:9.3f} ms  {v['calls']:5d} calls")
if cpu_n:
    import native_oracle as NO
    e = ext[:cpu_n].cpu().numpy()
    t = time.perf_counter(); NO.merge_subsample(pts[:cpu_n], rgb[:cpu_n], e); print("cpu merge", cpu_n, time.perf_counter() - t, "s")
    m = min(cpu_n, 20000)
    t = time.perf_counter(); NO.min_extents(pts[:m], K, P, sizes); print("cpu extents", m, time.perf_counter() - t, "s")
