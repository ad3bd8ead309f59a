// k8_interp.cu -- dense scale map of the `interp` aligner: piecewise-linear
// interpolation of per-SfM-point scale factors over a Delaunay triangulation, evaluated at every
// pixel and applied to the pre-aligned depth.
//
// Reference: depth_alignment/alignment/interp.py:75-110 (`linear_interpolation`: scipy Delaunay +
// LinearNDInterpolator over the full pixel grid, float64) and :355-361 (scale_map * aligned depth).
// The triangulation itself (Qhull) is built on the host exactly like the reference does; what runs
// here is LinearNDInterpolator's evaluation (scipy/interpolate/interpnd.pyx `_do_evaluate`, third
// party, restated): for the simplex containing x, with scipy's own barycentric transform T
//     c_i = sum_j T[i][j] * (x_j - T[2][j])  (i < 2),  c_2 = 1 - c_0 - c_1,
//     value = ((0 + c_0 v_0) + c_1 v_1) + c_2 v_2          (float64, no FMA)
// and `fill_value` outside the hull.  Point location is a rasterisation instead of scipy's directed
// walk: one CTA per triangle tests the pixels of its bounding box (inside <=> every c_i >= -eps,
// eps = 100 * DBL_EPSILON as in scipy), and a pixel on a shared edge goes to the triangle of lowest
// index (atomicMin), so the result does not depend on scheduling.  Either neighbour gives the same
// value up to float64 rounding -- far inside the 1e-5 bar after the cast to float32.
#include <float.h>

#include <mutex>

#include "common.cuh"

namespace mdi {

constexpr int kInterpThreads = 128;
constexpr int32_t kNoTriangle = 0x7f7f7f7f;   // memset pattern: "no triangle covers this pixel"

__device__ __forceinline__ void barycentric(const double* __restrict__ T, double x, double y,
                                            double* c0, double* c1, double* c2) {
  // interpnd.pyx / qhull.pyx `_barycentric_coordinates`, ndim = 2, row-major T[3][2]
  const double dx = __dsub_rn(x, T[4]), dy = __dsub_rn(y, T[5]);
  const double a = __dadd_rn(__dadd_rn(0.0, __dmul_rn(T[0], dx)), __dmul_rn(T[1], dy));
  const double b = __dadd_rn(__dadd_rn(0.0, __dmul_rn(T[2], dx)), __dmul_rn(T[3], dy));
  *c0 = a; *c1 = b;
  *c2 = __dsub_rn(__dsub_rn(1.0, a), b);
}

// one CTA per triangle: claim the pixels of its bounding box that it contains
__global__ void __launch_bounds__(kInterpThreads)
interp_owner_kernel(const double* __restrict__ transform, const int32_t* __restrict__ simplices,
                    const double* __restrict__ points, int n_tri, int H, int W,
                    int32_t* __restrict__ owner) {
  const int tri = blockIdx.x;
  if (tri >= n_tri) return;
  const double* T = transform + 6 * (int64_t)tri;
  double xmin = DBL_MAX, xmax = -DBL_MAX, ymin = DBL_MAX, ymax = -DBL_MAX;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int64_t p = simplices[3 * (int64_t)tri + k];
    const double x = points[2 * p], y = points[2 * p + 1];
    xmin = fmin(xmin, x); xmax = fmax(xmax, x); ymin = fmin(ymin, y); ymax = fmax(ymax, y);
  }
  if (!(xmin <= xmax) || !(ymin <= ymax)) return;   // NaN vertex
  const int x0 = max(0, (int)floor(xmin)), x1 = min(W - 1, (int)ceil(xmax));
  const int y0 = max(0, (int)floor(ymin)), y1 = min(H - 1, (int)ceil(ymax));
  if (x1 < x0 || y1 < y0) return;
  const int bw = x1 - x0 + 1;
  const int64_t n_px = (int64_t)bw * (y1 - y0 + 1);
  const double eps = 100.0 * DBL_EPSILON;   // interpnd.pyx: eps = 100 * DBL_EPSILON
  for (int64_t i = threadIdx.x; i < n_px; i += kInterpThreads) {
    const int y = y0 + (int)(i / bw), x = x0 + (int)(i % bw);
    double c0, c1, c2;
    barycentric(T, (double)x, (double)y, &c0, &c1, &c2);
    // qhull.pyx `_barycentric_inside`: -eps <= c_i <= 1 + eps (NaN transforms of degenerate simplices fail)
    if (c0 >= -eps && c0 <= 1.0 + eps && c1 >= -eps && c1 <= 1.0 + eps && c2 >= -eps && c2 <= 1.0 + eps)
      atomicMin(&owner[(int64_t)y * W + x], tri);
  }
}

// one thread per pixel: value of its triangle (or fill), times the pre-aligned depth
__global__ void __launch_bounds__(256)
interp_apply_kernel(const double* __restrict__ transform, const int32_t* __restrict__ simplices,
                    const double* __restrict__ values, double fill, int H, int W,
                    const int32_t* __restrict__ owner, const float* __restrict__ base,
                    const float* __restrict__ scale_shift, float* __restrict__ out) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= (int64_t)H * W) return;
  const int32_t tri = owner[p];
  double v = fill;
  if (tri != kNoTriangle) {
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    double c0, c1, c2;
    barycentric(transform + 6 * (int64_t)tri, (double)x, (double)y, &c0, &c1, &c2);
    const int32_t* s = simplices + 3 * (int64_t)tri;
    v = __dadd_rn(0.0, __dmul_rn(c0, values[s[0]]));
    v = __dadd_rn(v, __dmul_rn(c1, values[s[1]]));
    v = __dadd_rn(v, __dmul_rn(c2, values[s[2]]));
  }
  const float scale = (float)v;   // torch.from_numpy(float64).to(float32)
  if (base) {
    const float s = scale_shift ? scale_shift[0] : 1.f, t = scale_shift ? scale_shift[1] : 0.f;
    const float a = scale_shift ? align_depth(base[p], s, t) : base[p];
    out[p] = fmul(scale, a);      // interp.py:361 scale_map * prealigned.aligned_depth
  } else {
    out[p] = scale;
  }
}

}  // namespace mdi

extern "C" int64_t mdi_interp_workspace_bytes(int H, int W) {
  return H > 0 && W > 0 ? (int64_t)H * W * 4 : 0;
}

extern "C" int mdi_interp_scale_map(const double* transform, const int32_t* simplices,
                                    const double* points, const double* values, int n_tri,
                                    double fill_value, int H, int W, const float* base,
                                    const float* scale_shift, void* workspace, float* out,
                                    void* stream) {
  using namespace mdi;
  MDI_REQUIRE(H > 0 && W > 0 && n_tri >= 0, "mdi_interp_scale_map: bad size");
  MDI_REQUIRE(workspace && out, "mdi_interp_scale_map: null pointer");
  MDI_REQUIRE(n_tri == 0 || (transform && simplices && points && values), "mdi_interp_scale_map: null triangulation");
  cudaStream_t st = (cudaStream_t)stream;
  int32_t* owner = static_cast<int32_t*>(workspace);
  MDI_CUDA(cudaMemsetAsync(owner, 0x7f, (size_t)H * W * 4, st), "mdi_interp_scale_map memset");
  if (n_tri > 0) {
    mdi::prof_mark("interp_owner_kernel", st);
    interp_owner_kernel<<<n_tri, kInterpThreads, 0, st>>>(transform, simplices, points, n_tri, H, W, owner);
  }
  mdi::prof_mark("interp_apply_kernel", st);
  const int64_t n = (int64_t)H * W;
  interp_apply_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(transform, simplices, values, fill_value, H, W,
                                                                    owner, base, scale_shift, out);
  MDI_CHECK_LAUNCH("mdi_interp_scale_map");
  return MDI_OK;
}

// ---- allocation-free variant --------------------------------------------------------------------
// No counterpart in the reference: added for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The
// per-pixel owner map lives in a file-scope device array instead of a caller-provided workspace;
// every GPU that loads the library holds its MDI_INTERP_STATIC_MAX_PX * 4 bytes until exit.
// In host code `g_interp_owner_scratch` is only a stand-in for the symbol, never a device address:
// the address is resolved once per device with cudaGetSymbolAddress and passed to the kernels.
static __device__ __align__(256) int32_t g_interp_owner_scratch[MDI_INTERP_STATIC_MAX_PX];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
int32_t* g_scratch_addr[kMaxDevices] = {};
}  // namespace

extern "C" int mdi_interp_scale_map_static(const double* transform, const int32_t* simplices,
                                           const double* points, const double* values, int n_tri,
                                           double fill_value, int H, int W, const float* base,
                                           const float* scale_shift, float* out, void* stream) {
  MDI_REQUIRE(H > 0 && W > 0, "mdi_interp_scale_map_static: bad size");
  MDI_REQUIRE((int64_t)H * W <= (int64_t)MDI_INTERP_STATIC_MAX_PX,
              "mdi_interp_scale_map_static: %d x %d = %lld pixels exceed the static scratch of %d pixels; "
              "use mdi_interp_scale_map with a caller-provided workspace", H, W, (long long)H * W,
              MDI_INTERP_STATIC_MAX_PX);
  int dev = -1;
  MDI_CUDA(cudaGetDevice(&dev), "mdi_interp_scale_map_static");
  MDI_REQUIRE(dev >= 0 && dev < kMaxDevices, "mdi_interp_scale_map_static: device ordinal %d out of range", dev);
  int32_t* scratch;
  {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {   // once per device, with that device current
      void* p = nullptr;
      MDI_CUDA(cudaGetSymbolAddress(&p, g_interp_owner_scratch), "mdi_interp_scale_map_static symbol address");
      g_scratch_addr[dev] = static_cast<int32_t*>(p);
    }
    scratch = g_scratch_addr[dev];
  }
  // the existing entry point with the static array as its workspace: it refills the owner map
  // (memset + kernels, no allocation) on every call, so stale contents of earlier calls never leak
  return mdi_interp_scale_map(transform, simplices, points, values, n_tri, fill_value, H, W, base, scale_shift,
                              scratch, out, stream);
}
