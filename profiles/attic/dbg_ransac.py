import sys
sys.path[:0] = ["/root/repo", "/root/repo/oracle", "/root/repo/tests"]
import numpy as np, torch
import mdi_oracle as O
from helpers import load_case, case_config, OneViewParser, mod
DEV = "cuda:0"
for name in ["small_ransac_nopatch", "c2_ransac_adaptive", "c5_view_1600x1200", "tall_msac_static5"]:
    c = load_case(name)
    pfd = mod("depth_prediction.points_from_depth"); dpi = mod("depth_prediction.predictors.depth_predictor_interface"); types = mod("types")
    cfg = case_config(c)
    n = int(c.ref_uv.shape[1])
    sidx = O.draw_sample_idx(n, 2500, 4, c.meta["seed"])
    pred = dpi.PredictedDepth(c.depth.to(DEV), c.mask.to(DEV))
    image = types.InputImage(data=c.image.float() / 255.0, name=c.name, cam2world=c.c2w, K=c.K)
    pts, keep, P, res = pfd.get_pts_from_depth(pred, image, OneViewParser(c), cfg, DEV, sample_idx=torch.from_numpy(sidx), return_result=True)
    print(name, "iters", int(res.ransac_iters[0]), int(c.ref_ransac_iters), "inl", int(res.ransac_inliers[0]), int(c.ref_ransac_inliers),
          "ss", res.scale_shift[0].cpu().numpy(), c.ref_scale, c.ref_shift)
    # oracle trace of accepted models
    d = c.depth[c.ref_uv[1], c.ref_uv[0]]; z = torch.from_numpy(c.ref_z)
    D = O.stack_ones(d)
    best_s = best_lo = float("inf"); acc = []
    for it in range(int(c.ref_ransac_iters) + 1):
        idx = torch.as_tensor(sidx[it], dtype=torch.long)
        h = O.lstsq_fit(D[:, idx], z[idx]); r = O.residuals_sq(h, d, z); loss = int((r >= 0.01).sum())
        if loss < best_s:
            inl = r < 0.01
            hl = O.lstsq_fit(D[:, inl], z[inl]); rl = O.residuals_sq(hl, d, z); ll = int((rl >= 0.01).sum())
            ok = ll < best_lo
            acc.append((it, loss, ll, ok))
            if ok: best_s, best_lo = loss, ll
    print("   oracle refits (it, loss, loss_lo, accepted):", acc[:12])
