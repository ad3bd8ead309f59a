import sys, time
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np, torch
from helpers import load_case, mod
import test_gpu_parity as T
emod = mod("engine")
c = load_case("c2_ransac_adaptive")
for ov in (False, True, True):
    emod.OVERLAP_STATS = ov
    pts, keep, P, res = T.run_case(c)
    print("overlap", ov, "lo_hi", res.lo_hi[0].tolist(), "q", res.quantiles[0].tolist(), "ref", float(c.ref_lo), float(c.ref_hi), float(c.ref_q1), float(c.ref_q3),
          "ss", res.scale_shift[0].tolist(), "retry", res.extras.get("stats_retry"), "status", res.status.tolist(), "kept", int(keep.sum()))
