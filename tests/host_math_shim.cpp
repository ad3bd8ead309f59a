// TEST INFRASTRUCTURE: compiles csrc/mdi_math.cuh as plain C++ (g++ -ffp-contract=off) so the
// scalar numerics of the kernels can be pinned on the CPU.  Never linked into the product.
#include "../3dgs_monocular_depth_init_b200/csrc/mdi_math.cuh"

extern "C" {
void t_solve(double sdd, double sd, double n, double sdg, double sg, float* h) {
  mdi::solve_normal_eq(sdd, sd, n, sdg, sg, h, h + 1);
}
double t_required(long long inl, long long tot, int ss, double conf) {
  return mdi::required_samples(inl, tot, ss, conf);
}
void t_quantile_rank(float q, long long n, long long* lo, long long* hi, float* w) {
  mdi::QuantileRank r = mdi::quantile_rank(q, n);
  *lo = r.lo; *hi = r.hi; *w = r.w;
}
float t_lerp(float a, float b, float w) { return mdi::torch_lerp(a, b, w); }
void t_iqr(float q1, float q3, float mn, float mx, float* lo, float* hi) { mdi::iqr_range(q1, q3, mn, mx, lo, hi); }
void t_factor(const float* a, long long n, float lo, float hi, int fmin, int fmax, int* out) {
  for (long long i = 0; i < n; ++i) out[i] = mdi::adaptive_factor(a[i], lo, hi, fmin, fmax);
}
void t_patch(int H, int W, int n, int* out) { mdi::patch_geometry(H, W, n, out, out + 1, out + 2, out + 3); }
void t_keys(const float* a, long long n, unsigned* k, float* back) {
  for (long long i = 0; i < n; ++i) { k[i] = mdi::float_to_key(a[i]); back[i] = mdi::key_to_float(k[i]); }
}
void t_project(const float* P, const float* X, long long n, float* out) {
  for (long long i = 0; i < n; ++i)
    for (int r = 0; r < 3; ++r)
      out[3 * i + r] = mdi::dot4_fma(P[4 * r], P[4 * r + 1], P[4 * r + 2], P[4 * r + 3], X[3 * i], X[3 * i + 1], X[3 * i + 2]);
}
void t_unproject(const float* Ki, const float* M, const float* z, const int* row, const int* col, long long n, float* out) {
  for (long long i = 0; i < n; ++i) {
    float px = mdi::fmul(mdi::fadd((float)col[i], 0.5f), z[i]), py = mdi::fmul(mdi::fadd((float)row[i], 0.5f), z[i]);
    float cx = mdi::dot3_kinv(Ki[0], Ki[1], Ki[2], px, py, z[i]);
    float cy = mdi::dot3_kinv(Ki[3], Ki[4], Ki[5], px, py, z[i]);
    float cz = mdi::dot3_kinv(Ki[6], Ki[7], Ki[8], px, py, z[i]);
    for (int r = 0; r < 3; ++r)
      out[3 * i + r] = mdi::dot4_fma(M[4 * r], M[4 * r + 1], M[4 * r + 2], M[4 * r + 3], cx, cy, cz);
  }
}
void t_camera(const float* K, const float* c2w, float* P) { mdi::projection_matrix(K, c2w, P); }
void t_pinhole_inv(const float* K, long long n, float* out) { for (long long i = 0; i < n; ++i) mdi::pinhole_inverse(K + 9 * i, out + 9 * i); }
float t_residual(float h0, float h1, float d, float g) { return mdi::residual_sq(h0, h1, d, g); }
}
