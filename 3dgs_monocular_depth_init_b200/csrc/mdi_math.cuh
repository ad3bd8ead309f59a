// mdi_math.cuh -- scalar arithmetic shared by the kernels, written so that it also compiles
// as plain C++ (tests build it with g++ -ffp-contract=off to pin the numerics on CPU).
//
// Everything here restates, operation by operation, the fp32 arithmetic of the reference
// (file:line in each comment).  The library is compiled with -fmad=false, and the places
// where the op order matters use explicit mul/add helpers so a host compiler cannot
// contract them either.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define MDI_HD __host__ __device__ __forceinline__
#else
#define MDI_HD inline
#endif

namespace mdi {

// ---- non-contracted fp32 primitives ------------------------------------------------------
MDI_HD float fmul(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fmul_rn(a, b);
#else
  volatile float r = a * b;
  return r;
#endif
}
MDI_HD float fadd(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fadd_rn(a, b);
#else
  volatile float r = a + b;
  return r;
#endif
}
MDI_HD float fsub(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fsub_rn(a, b);
#else
  volatile float r = a - b;
  return r;
#endif
}
MDI_HD float fdiv(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdiv_rn(a, b);
#else
  volatile float r = a / b;
  return r;
#endif
}

// a*x + b*y + c*z + d*1 the way an sgemm micro-kernel accumulates it: products folded in with
// FMA in k order (verified against torch CPU mm: bit-identical; plain mul/add is not).
MDI_HD float dot4_fma(float a, float b, float c, float d, float x, float y, float z) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(d, 1.0f, __fmaf_rn(c, z, __fmaf_rn(b, y, __fmul_rn(a, x))));
#else
  return fmaf(d, 1.0f, fmaf(c, z, fmaf(b, y, fmul(a, x))));
#endif
}
MDI_HD float dot3_fma(float a, float b, float c, float x, float y, float z) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(c, z, __fmaf_rn(b, y, __fmul_rn(a, x)));
#else
  return fmaf(c, z, fmaf(b, y, fmul(a, x)));
#endif
}

// inv(K) @ pc of points_from_depth.py:303.  inv() returns a column-major matrix and the
// right operand is a transposed view, which sends sgemm down a dot-product kernel: three
// separately rounded products summed as (p0 + p2) + p1.  (Measured against torch-CPU/MKL: exact
// for every N on one thread; with several threads MKL switches to an FMA kernel above
// N ~ 2-3e4 columns -- a thread-count dependent choice this build does not chase.)
MDI_HD float dot3_kinv(float a, float b, float c, float x, float y, float z) {
  return fadd(fadd(fmul(a, x), fmul(c, z)), fmul(b, y));
}

// P = K @ R @ [I | -C], R = c2w[:3,:3].T, C = c2w[:3,3] (points_from_depth.py:231-233).
// K row-major 3x3, c2w row-major 4x4, P row-major 3x4.  K @ R has a transposed view on the
// right: dot-product sgemm kernel (dot3_kinv); T @ [I | -C] is an FMA chain over k.
MDI_HD void projection_matrix(const float* k, const float* m, float* P) {
  float T[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)  // R[kk][j] = c2w[j][kk]
      T[i][j] = dot3_kinv(k[3 * i], k[3 * i + 1], k[3 * i + 2], m[4 * j], m[4 * j + 1], m[4 * j + 2]);
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j)
      P[4 * i + j] = dot3_fma(T[i][0], T[i][1], T[i][2], j == 0 ? 1.f : 0.f, j == 1 ? 1.f : 0.f,
                              j == 2 ? 1.f : 0.f);
    P[4 * i + 3] = dot3_fma(T[i][0], T[i][1], T[i][2], -m[3], -m[7], -m[11]);
  }
}

// inverse of a zero-skew pinhole K = [[fx,0,cx],[0,fy,cy],[0,0,1]] as torch.linalg.inv returns it
MDI_HD void pinhole_inverse(const float* k, float* ki) {
  ki[0] = fdiv(1.0f, k[0]); ki[1] = 0.f; ki[2] = fdiv(-k[2], k[0]);
  ki[3] = 0.f; ki[4] = fdiv(1.0f, k[4]); ki[5] = fdiv(-k[5], k[4]);
  ki[6] = 0.f; ki[7] = 0.f; ki[8] = 1.0f;
}

// aligned = depth * scale + shift, two roundings (lstsqrs.py:52, ransacs.py:187)
MDI_HD float align_depth(float d, float s, float t) { return fadd(fmul(d, s), t); }

// squared residual (h0*d + h1 - g)**2 (ransacs.py:94-97): mul, add, sub, square
MDI_HD float residual_sq(float h0, float h1, float d, float g) {
  float r = fsub(fadd(fmul(h0, d), h1), g);
  return fmul(r, r);
}

// ---- order-preserving float <-> uint key (radix select of torch.quantile's sort) ----------
MDI_HD uint32_t float_to_key(float f) {
  uint32_t u;
#if defined(__CUDA_ARCH__)
  u = __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; u = c.u;
#endif
  if (f != f) return 0xFFFFFFFFu;  // NaN sorts last (torch.sort)
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
MDI_HD float key_to_float(uint32_t k) {
  uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

// ---- torch.quantile(x, q) rank arithmetic (ATen Sorting.cpp quantile_compute) --------------
// rank = q * (n - 1) evaluated in fp32 (q is a float32 scalar tensor), lo = floor, hi = ceil,
// weight = rank - lo (fp32).
struct QuantileRank { int64_t lo, hi; float w; };
MDI_HD QuantileRank quantile_rank(float q, int64_t n) {
  QuantileRank r;
  float rank = fmul(q, (float)(n - 1));
  float fl = floorf(rank);
  r.lo = (int64_t)fl;
  r.hi = (int64_t)ceilf(rank);
  r.w = fsub(rank, fl);
  return r;
}
// at::lerp for scalars: w < 0.5 ? a + w*(b-a) : b - (b-a)*(1-w)   (ATen Lerp.h)
MDI_HD float torch_lerp(float a, float b, float w) {
  float diff = fsub(b, a);
  return (w < 0.5f) ? fadd(a, fmul(w, diff)) : fsub(b, fmul(diff, fsub(1.0f, w)));
}

// iqr bounds + clamping against min/max (adaptive_subsampling.py:82-95).  Python's
// max(a, b) returns b only if b > a, min(a, b) returns b only if b < a (NaN keeps a).
MDI_HD void iqr_range(float q1, float q3, float mn, float mx, float* lo, float* hi) {
  float iqr = fsub(q3, q1);
  float lo_b = fsub(q1, fmul(1.5f, iqr));
  float hi_b = fadd(q3, fmul(1.5f, iqr));
  *lo = (lo_b > mn) ? lo_b : mn;
  *hi = (hi_b < mx) ? hi_b : mx;
}

// Per-pixel adaptive stride (adaptive_subsampling.py:12-17,89-98,105-120 and :63-76).
//   m  = clamp((a - lo) / (hi - lo), 0, 1)       (1*m + 0 is the identity up to -0)
//   x  = 1 - m
//   f  = clamp(10 * ((x - 0) / 1) + 5, fmin, fmax) -> trunc to int ; 0 -> 1
// Returns INT32_MIN for a NaN factor (x86 cvttss2si semantics of .to(int); then only
// row == 0 && col == 0 satisfies row % f == 0 under torch's floor-mod).
MDI_HD int adaptive_factor(float a, float lo, float hi, int fmin, int fmax) {
  float m = fdiv(fsub(a, lo), fsub(hi, lo));
  m = fadd(fmul(1.0f, m), 0.0f);
  // torch.clamp propagates NaN
  if (m == m) m = (m < 0.0f) ? 0.0f : ((m > 1.0f) ? 1.0f : m);
  float x = fsub(1.0f, m);
  float t = fdiv(fsub(x, 0.0f), 1.0f);
  float f = fadd(fmul((float)(fmax - fmin), t), (float)fmin);
  if (f != f) return INT32_MIN;
  f = (f < (float)fmin) ? (float)fmin : ((f > (float)fmax) ? (float)fmax : f);
  int fi = (int)f;  // truncation
  return fi == 0 ? 1 : fi;
}

// torch's floor-mod test `idx % f == 0` for idx >= 0 and any int f != 0
MDI_HD bool divides(int idx, int f) {
  if (f == INT32_MIN) return idx == 0;
  return (idx % f) == 0;  // remainder zero is sign independent
}

// ---- 2x2 normal equations with torch.linalg.pinv semantics (lstsqrs.py:21-26) --------------
// A = [[sdd, sd], [sd, n]], b = [sdg, sg].  pinv drops singular values <= rtol * sigma_max with
// rtol = max(m, n) * eps_fp32 = 2 * 1.1920929e-7 (torch default for a 2x2 float32 matrix).
// The rank decision is taken on the fp32-rounded matrix (what the reference factorises); the
// solve itself uses the unrounded fp64 sums.
MDI_HD void solve_normal_eq(double sdd, double sd, double n, double sdg, double sg, float* h0,
                            float* h1) {
  const double rtol = 2.0 * 1.1920928955078125e-7;
  double a = (double)(float)sdd, b = (double)(float)sd, c = (double)(float)n;
  double half_tr = 0.5 * (a + c);
  double half_df = 0.5 * (a - c);
  double rad = sqrt(half_df * half_df + b * b);
  double l1 = half_tr + rad;                      // sigma_max (A is PSD)
  double det32 = a * c - b * b;
  double l2 = (l1 > 0.0) ? det32 / l1 : 0.0;      // sigma_min, stable form
  if (!(l1 > 0.0)) {                               // zero (or NaN) matrix -> pinv = 0
    bool isnan_in = (sdd != sdd) || (sd != sd) || (sdg != sdg) || (sg != sg);
    float v = isnan_in ? NAN : 0.0f;
    *h0 = v; *h1 = v;
    return;
  }
  if (l2 > rtol * l1) {
    double det = sdd * n - sd * sd;
    *h0 = (float)((n * sdg - sd * sg) / det);
    *h1 = (float)((sdd * sg - sd * sdg) / det);
    return;
  }
  // rank 1: pinv = v v^T / l1 with v the dominant eigenvector of the (unrounded) matrix
  double A = sdd, B = sd, C = n;
  double ht = 0.5 * (A + C), hd = 0.5 * (A - C);
  double r = sqrt(hd * hd + B * B);
  double lam = ht + r;
  // eigenvector of lam: (B, lam - A) or (lam - C, B); pick the better conditioned one
  double vx, vy;
  if (fabs(lam - A) > fabs(lam - C)) { vx = B; vy = lam - A; } else { vx = lam - C; vy = B; }
  double nn = vx * vx + vy * vy;
  if (!(nn > 0.0)) { *h0 = 0.0f; *h1 = 0.0f; return; }
  double proj = (vx * sdg + vy * sg) / (nn * lam);
  *h0 = (float)(vx * proj);
  *h1 = (float)(vy * proj);
}

// ---- RANSAC stopping rule (ransacs.py:79-91) ------------------------------------------------
// inlier_ratio = int64 tensor / int -> float32 ; ratio**m and 1 - x in float32 ; the logs in
// double (math.log of a 0-dim tensor).  ZeroDivisionError (ratio == 0 -> log(1) = 0) and
// ValueError (ratio == 1 -> log(0)) both return 0.
MDI_HD double required_samples(int64_t inliers, int64_t total, int sample_size, double confidence) {
  if (total == 0) return 0.0;  // 0/0 -> nan ratio: log(nan) = nan, `nan <= it` is False
  float ratio = fdiv((float)inliers, (float)total);  // true_divide of int64 by int -> float32
  double p = 1.0;
  for (int i = 0; i < sample_size; ++i) p *= (double)ratio;
  float pw = (float)p;               // float32 pow
  float om = fsub(1.0f, pw);
  if (inliers == 0) return 0.0;      // python int 0 / tensor 0 -> log(1 - 0) = 0 -> ZeroDivisionError
  if (om <= 0.0f) return 0.0;        // math.log(0) -> ValueError
  double den = log((double)om);
  if (den == 0.0) return 0.0;        // ZeroDivisionError
  return log(1.0 - confidence) / den;
}

// ---- patch geometry of the SfM-density mask (num_sfm_points_mask.py:7-35) ------------------
MDI_HD void patch_geometry(int H, int W, int n_small, int* ph, int* pw, int* gh, int* gw) {
  // np.argmin([H, W]) returns 0 on ties
  bool small_is_h = (H <= W);
  int small = small_is_h ? H : W, large = small_is_h ? W : H;
  int ps_small = small / n_small;
  if (ps_small <= 0) { *ph = *pw = 1; *gh = *gw = 0; return; }  // reference divides by zero
  int n_large = (large + ps_small - 1) / ps_small;
  int ps_large = large / n_large;
  if (small_is_h) { *gh = n_small; *gw = n_large; *ph = ps_small; *pw = ps_large; }
  else            { *gh = n_large; *gw = n_small; *ph = ps_large; *pw = ps_small; }
}

}  // namespace mdi
