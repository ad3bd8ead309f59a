"""How often can the MSAC loss rounding change a decision?  
The reference sums float32 min(r, thr) with torch.sum (vectorised pairwise fp32); the kernel
accumulates in fp64 and rounds once to fp32.  For 10^4 random 4-point hypotheses on the C2 golden
view this counts (a) losses whose fp32 bits differ, (b) decisions `loss < best_so_far` of the
reference's running scan (ransacs.py:141) that come out differently, (c) the same for strict ties
between distinct hypotheses.  CPU only:  python profiles/msac_tie_study.py"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests")]
import mdi_oracle as O  # noqa: E402
from helpers import load_case  # noqa: E402

c = load_case("c2_ransac_adaptive")
d = c.depth[c.ref_uv[1], c.ref_uv[0]]
z = torch.from_numpy(c.ref_z)
D = O.stack_ones(d)
n = d.numel()
g = torch.Generator().manual_seed(1)
thr = torch.tensor(0.01)
N = 10000
l32 = np.empty(N, np.float32)
l64 = np.empty(N, np.float32)
for i in range(N):
    idx = torch.randperm(n, generator=g)[:4]
    h = O.lstsq_fit(D[:, idx], z[idx])
    r = O.residuals_sq(h, d, z)
    t = torch.minimum(r, torch.full_like(r, 0.01))
    l32[i] = float(torch.sum(t))                       # the reference (ransacs.py:64-65)
    l64[i] = np.float32(t.double().sum().item())       # fp64 accumulation, one rounding (k2_ransac.cu)
diff_bits = int((l32 != l64).sum())
rel = np.abs(l32.astype(np.float64) - l64) / np.maximum(np.abs(l64), 1e-30)


def decisions(l):
    best, out = np.inf, np.zeros(N, bool)
    for i in range(N):
        if l[i] < best:
            out[i] = True
            best = l[i]
    return out


flips = int((decisions(l32) != decisions(l64)).sum())
print(f"points {n}, hypotheses {N}")
print(f"(a) losses with different fp32 bits: {diff_bits} ({100 * diff_bits / N:.1f} %), max relative difference {rel.max():.2e}")
print(f"(b) running-best decisions that differ: {flips} of {N}  (improvements seen by the reference scan: {int(decisions(l32).sum())})")
gaps = np.diff(np.sort(l64.astype(np.float64)))
print(f"(c) smallest relative gap between two distinct hypotheses' losses: {np.min(gaps[gaps > 0] / np.sort(l64)[1:][gaps > 0]):.2e}"
      f"  (a flip needs a gap below the rounding difference of (a))")
