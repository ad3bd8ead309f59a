import sys
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np, torch
from helpers import load_case
import test_gpu_parity as T
c = load_case(sys.argv[1] if len(sys.argv) > 1 else "small_ransac_nopatch")
pts, keep, P, res = T.run_case(c)
print("iters", int(res.ransac_iters[0]), int(c.ref_ransac_iters), "inl", int(res.ransac_inliers[0]), int(c.ref_ransac_inliers))
