"""CPU: the scalar arithmetic shared by the kernels (csrc/mdi_math.cuh, compiled here as plain
C++) against torch / the oracle / the reference's golden vectors."""
import ctypes as C
import json
import subprocess
from pathlib import Path

import numpy as np
import pytest
import torch

import mdi_oracle as O
from helpers import GOLDEN, ROOT, load_case

SHIM_SRC = ROOT / "tests" / "host_math_shim.cpp"
SHIM_SO = ROOT / "tests" / "_host_math_shim.so"


@pytest.fixture(scope="module")
def shim():
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-mfma", "-shared", "-fPIC", "-x", "c++", str(SHIM_SRC),
                    "-o", str(SHIM_SO)], check=True)
    lib = C.CDLL(str(SHIM_SO))
    lib.t_required.restype = C.c_double
    lib.t_required.argtypes = [C.c_longlong, C.c_longlong, C.c_int, C.c_double]
    lib.t_solve.argtypes = [C.c_double] * 5 + [C.c_void_p]
    lib.t_lerp.restype = C.c_float
    lib.t_lerp.argtypes = [C.c_float] * 3
    lib.t_residual.restype = C.c_float
    lib.t_residual.argtypes = [C.c_float] * 4
    lib.t_quantile_rank.argtypes = [C.c_float, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.t_iqr.argtypes = [C.c_float] * 4 + [C.c_void_p] * 2
    lib.t_factor.argtypes = [C.c_void_p, C.c_longlong, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p]
    lib.t_patch.argtypes = [C.c_int] * 3 + [C.c_void_p]
    lib.t_keys.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
    lib.t_project.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]
    lib.t_unproject.argtypes = [C.c_void_p] * 5 + [C.c_longlong, C.c_void_p]
    lib.t_camera.argtypes = [C.c_void_p] * 3
    lib.t_pinhole_inv.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p]
    return lib


def solve(lib, d, g):
    d = np.asarray(d, np.float64); g = np.asarray(g, np.float64)
    h = np.zeros(2, np.float32)
    lib.t_solve((d * d).sum(), d.sum(), float(len(d)), (d * g).sum(), g.sum(), h.ctypes.data)
    return h


def quantile(lib, x, q):
    xs = np.sort(x)
    lo, hi, w = C.c_longlong(), C.c_longlong(), C.c_float()
    lib.t_quantile_rank(q, len(xs), C.byref(lo), C.byref(hi), C.byref(w))
    return lib.t_lerp(float(xs[lo.value]), float(xs[hi.value]), w.value)


def test_known_answers(shim):
    ka = json.loads((GOLDEN / "known_answers.json").read_text())
    for n, tot, want in ka["required_samples"]:
        assert shim.t_required(n, tot, 4, 0.999) == want, (n, tot)
    for h, w, n, ph, pw, gh, gw in ka["patch_sizes"]:
        out = np.zeros(4, np.int32)
        shim.t_patch(h, w, n, out.ctypes.data)
        assert out.tolist() == [ph, pw, gh, gw]
    for d, g, s, t in ka["lstsq"]:
        h = solve(shim, np.float32(d), np.float32(g))
        if np.allclose(h, [s, t], rtol=2e-5, atol=1e-6):
            continue
        # ill-conditioned but full rank: the reference's fp32 SVD is off by more than the
        # tolerance; ours must then be the one closer to the exact fp64 solution
        d64, g64 = np.float32(d).astype(np.float64), np.float32(g).astype(np.float64)
        exact = np.linalg.lstsq(np.stack([d64, np.ones(len(d64))], 1), g64, rcond=None)[0]
        assert np.abs(h - exact).max() <= np.abs(np.array([s, t]) - exact).max(), (d, g, h, s, t)


def test_solve_vs_pinv_random(shim):
    rng = np.random.default_rng(0)
    worst = 0.0
    for _ in range(300):
        n = int(rng.choice([4, 4, 4, 50, 3000]))
        d = rng.uniform(0.5, 5, n).astype(np.float32)
        g = (1.7 * d + 0.3 + rng.normal(0, 0.05, n)).astype(np.float32)
        ref = O.lstsq_fit(O.stack_ones(torch.from_numpy(d)), torch.from_numpy(g))
        h = solve(shim, d, g)
        exact = np.linalg.lstsq(np.stack([d.astype(np.float64), np.ones(n)], 1), g.astype(np.float64), rcond=None)[0]
        # ours must be at least as close to the exact fp64 answer as the reference's fp32 pinv
        e_ours = np.abs(h - exact).max(); e_ref = max(abs(float(ref[0]) - exact[0]), abs(float(ref[1]) - exact[1]))
        assert e_ours <= e_ref + 1e-6 * np.abs(exact).max()
        worst = max(worst, e_ours / np.abs(exact).max())
    assert worst < 1e-6


def test_quantile_vs_torch(shim):
    g = torch.Generator().manual_seed(1)
    for n in [1, 2, 3, 5, 6, 7, 100, 1001, 65537, 1 << 20, 8294400]:
        x = torch.randn(n, generator=g)
        for q in (0.25, 0.75):
            assert quantile(shim, x.numpy(), q) == float(torch.quantile(x, q)), (n, q)


def test_iqr_and_factor_vs_oracle(shim):
    for name in ["small_ransac_adaptive", "small_adaptive_range_3_9", "tall_lstsq_adaptive"]:
        c = load_case(name)
        aligned = c.depth * torch.tensor(c.ref_scale) + torch.tensor(c.ref_shift)
        mask = (aligned != -42.0) & c.mask
        lo_o, hi_o, q1, q3, mn, mx = O.adaptive_range(aligned, mask)
        lo, hi = C.c_float(), C.c_float()
        shim.t_iqr(float(q1), float(q3), float(mn), float(mx), C.byref(lo), C.byref(hi))
        assert lo.value == float(lo_o) == float(c.ref_lo) and hi.value == float(hi_o) == float(c.ref_hi)
        fmin, fmax = c.meta.get("fmin", 5), c.meta.get("fmax", 15)
        fac = O.adaptive_factor_map(aligned, mask, fmin, fmax).numpy()
        a = np.ascontiguousarray(aligned.numpy())
        out = np.zeros(a.size, np.int32)
        shim.t_factor(a.ctypes.data, a.size, lo.value, hi.value, fmin, fmax, out.ctypes.data)
        m = mask.numpy().reshape(-1)
        assert np.array_equal(out[m], fac.reshape(-1)[m])


def test_keys_monotone_roundtrip(shim):
    x = np.array([-np.inf, -3.5, -1e-30, -0.0, 0.0, 1e-38, 1.0, 1.0000001, 7e20, np.inf], np.float32)
    k = np.zeros(len(x), np.uint32); back = np.zeros(len(x), np.float32)
    shim.t_keys(x.ctypes.data, len(x), k.ctypes.data, back.ctypes.data)
    assert np.all(np.diff(k.astype(np.int64)) > 0)
    assert np.array_equal(back.view(np.uint32), x.view(np.uint32))


def test_projection_and_unprojection_bit_exact_vs_reference(shim):
    """FMA-chain dot products reproduce the reference's sgemm results bit for bit."""
    c = load_case("c1_lstsq_static10")
    P = np.ascontiguousarray(c.ref_P)
    X = np.ascontiguousarray(c.sfm.numpy())
    out = np.zeros((len(X), 3), np.float32)
    shim.t_project(P.ctypes.data, X.ctypes.data, len(X), out.ctypes.data)
    cam = torch.from_numpy(P) @ torch.vstack([torch.from_numpy(X).T, torch.ones(len(X))])
    assert np.array_equal(out.T, cam.numpy())
    H, W = c.depth.shape
    idx = np.nonzero(c.ref_keep)[0]
    row = (idx // W).astype(np.int32); col = (idx % W).astype(np.int32)
    z = (c.depth * torch.tensor(c.ref_scale) + torch.tensor(c.ref_shift)).reshape(-1)[torch.from_numpy(idx)].numpy()
    Ki = np.ascontiguousarray(torch.linalg.inv(c.K).numpy())
    M = np.ascontiguousarray(c.c2w.numpy())
    o = np.zeros((len(idx), 3), np.float32)
    shim.t_unproject(Ki.ctypes.data, M.ctypes.data, np.ascontiguousarray(z).ctypes.data, row.ctypes.data,
                     col.ctypes.data, len(idx), o.ctypes.data)
    assert np.array_equal(o, c.ref_pts)


def test_projection_matrix_bit_exact_on_every_golden_view(shim):
    from helpers import FULL_CASES, SMALL_CASES
    for name in SMALL_CASES + FULL_CASES:
        c = load_case(name)
        if c.raised:
            continue
        K = np.ascontiguousarray(c.K.numpy()); M = np.ascontiguousarray(c.c2w.numpy())
        P = np.zeros((3, 4), np.float32)
        shim.t_camera(K.ctypes.data, M.ctypes.data, P.ctypes.data)
        assert np.array_equal(P, c.ref_P), name
    # and on random poses against the oracle
    g = torch.Generator().manual_seed(0)
    for _ in range(200):
        Q, _r = torch.linalg.qr(torch.randn(3, 3, generator=g))
        c2w = torch.eye(4); c2w[:3, :3] = Q; c2w[:3, 3] = torch.randn(3, generator=g) * 5
        f = float(torch.rand(1, generator=g)) * 2000 + 300
        K = torch.tensor([[f, 0, 640.5], [0, f * 1.01, 360.25], [0, 0, 1]])
        P = np.zeros((3, 4), np.float32)
        shim.t_camera(np.ascontiguousarray(K.numpy()).ctypes.data, np.ascontiguousarray(c2w.numpy()).ctypes.data, P.ctypes.data)
        assert np.array_equal(P, O.projection_matrix(K, c2w).numpy())


def test_pinhole_inverse_equals_torch_inv(shim):
    rng = np.random.default_rng(0)
    n = 5000
    K = np.zeros((n, 3, 3), np.float32)
    K[:, 0, 0] = rng.uniform(300, 4000, n); K[:, 1, 1] = K[:, 0, 0] * rng.uniform(0.99, 1.01, n)
    K[:, 0, 2] = rng.uniform(100, 2000, n); K[:, 1, 2] = rng.uniform(100, 1200, n); K[:, 2, 2] = 1
    out = np.zeros_like(K)
    shim.t_pinhole_inv(K.ctypes.data, n, out.ctypes.data)
    assert np.array_equal(out, torch.linalg.inv(torch.from_numpy(K)).numpy())
