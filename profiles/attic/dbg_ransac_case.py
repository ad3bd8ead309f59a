import sys, math
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np, torch
from helpers import load_case, mod, case_config, OneViewParser
import test_gpu_parity as T
import mdi_oracle as O
name = sys.argv[1] if len(sys.argv) > 1 else "small_ransac_nopatch"
c = load_case(name)
pts, keep, P, res = T.run_case(c)
n = int(res.n_valid[0])
d = res.d[:n].cpu(); g = res.z[:n].cpu()
sidx = O.draw_sample_idx(n, 2500, 4, c.meta["seed"])
# oracle event list
D = O.stack_ones(d)
best_lo = best_sample = float("inf"); best_n = 0; have = False
for it in range(80):
    idx = torch.as_tensor(sidx[it], dtype=torch.long)
    h = O.lstsq_fit(D[:, idx], g[idx]); r = O.residuals_sq(h, D[0], g); loss = int(torch.sum(r >= 0.01))
    ev = ""
    if loss < best_sample:
        inl = r < 0.01
        h_lo = O.lstsq_fit(D[:, inl], g[inl]); r_lo = O.residuals_sq(h_lo, D[0], g); loss_lo = int(torch.sum(r_lo >= 0.01))
        ev = f"refit loss_lo {loss_lo}"
        if loss_lo < best_lo:
            best_lo, best_sample, best_n, have = loss_lo, loss, int(torch.sum(r_lo < 0.01)), True
            ev += f" ACCEPT ninl {best_n} req {O.required_samples(torch.tensor(best_n), n, 4, 0.999):.2f}"
    if ev:
        print("it", it, "loss_s", loss, ev)
    if have and O.required_samples(torch.tensor(best_n), n, 4, 0.999) <= it:
        print("STOP at", it); break
eng = mod("engine").get_engine(T.DEV)
single = mod("_single")
for mi in (16, 17, 20, 24, 28, 32, 40, 48, 54, 64, 80, 2500):
    cfg = case_config(c)
    cfg.mdi.alignment.ransac.max_iters = mi
    pfd = mod("depth_prediction.points_from_depth"); dpi = mod("depth_prediction.predictors.depth_predictor_interface"); types = mod("types")
    pred = dpi.PredictedDepth(c.depth.to(T.DEV), c.mask.to(T.DEV))
    image = types.InputImage(data=c.image.float() / 255.0, name=c.name, cam2world=c.c2w, K=c.K)
    out = pfd.get_pts_from_depth(pred, image, OneViewParser(c), cfg, T.DEV, sample_idx=torch.from_numpy(sidx[:mi].copy()), return_result=True)
    r2 = out[3]
    tr = O.ransac_fit(d, g, "ransac", max_iters=mi, sample_idx=sidx)
    print("max_iters", mi, "gpu iters/inl", int(r2.ransac_iters[0]), int(r2.ransac_inliers[0]), "oracle", tr.last_iter, tr.n_inliers)
