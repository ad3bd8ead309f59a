// k6_extents.cu -- minimal Gaussian extent per point (a13).  Reference:
// native_modules/subsampling/src/impl.cpp:17-35 (visibility + depth) and :70-126 (min over
// cameras of 2 * depth / min(fx, fy); -1 when no camera sees the point).
// One thread per point, cameras streamed through shared memory in chunks.  FP32-ALU bound
// (N * C projections); min is order independent, the per-pair arithmetic keeps the reference
// op order (no contraction): Eigen evaluates P * p.homogeneous() as leftCols(3) * p + col(3) with
// the 3-term dot product reduced as a0 + (a1 + a2) (oracle/native_oracle.c states the rules).
#include <float.h>

#include "common.cuh"

namespace mdi {

constexpr int kExtThreads = 256;
constexpr int kCamChunk = 256;

struct CamRec { float P[12]; float inv_f_dummy; float f; float w, h; };

__global__ void __launch_bounds__(kExtThreads)
min_extents_kernel(const float* __restrict__ pts, int64_t n, const float* __restrict__ K,
                   const float* __restrict__ P, const int32_t* __restrict__ sizes, int n_cams,
                   float* __restrict__ extents) {
  __shared__ CamRec s_cam[kCamChunk];
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float x = 0.f, y = 0.f, z = 0.f;
  if (i < n) { x = pts[3 * i]; y = pts[3 * i + 1]; z = pts[3 * i + 2]; }
  float best = FLT_MAX;
  for (int c0 = 0; c0 < n_cams; c0 += kCamChunk) {
    const int nc = min(kCamChunk, n_cams - c0);
    __syncthreads();
    for (int c = threadIdx.x; c < nc; c += blockDim.x) {
      CamRec r;
      for (int k = 0; k < 12; ++k) r.P[k] = P[(int64_t)(c0 + c) * 12 + k];
      const float fx = K[(int64_t)(c0 + c) * 9 + 0], fy = K[(int64_t)(c0 + c) * 9 + 4];
      r.f = fy < fx ? fy : fx;  // std::min(fx, fy)
      r.inv_f_dummy = 0.f;
      r.w = (float)sizes[2 * (c0 + c)];
      r.h = (float)sizes[2 * (c0 + c) + 1];
      s_cam[c] = r;
    }
    __syncthreads();
    if (i < n) {
      for (int c = 0; c < nc; ++c) {
        const CamRec& r = s_cam[c];
        const float qz = fadd(fadd(fmul(r.P[8], x), fadd(fmul(r.P[9], y), fmul(r.P[10], z))), r.P[11]);
        if (!(qz > 0.f)) continue;  // depth <= 0 (NaN falls through in the reference)
        const float qx = fadd(fadd(fmul(r.P[0], x), fadd(fmul(r.P[1], y), fmul(r.P[2], z))), r.P[3]);
        const float qy = fadd(fadd(fmul(r.P[4], x), fadd(fmul(r.P[5], y), fmul(r.P[6], z))), r.P[7]);
        const float u = fdiv(qx, qz), vv = fdiv(qy, qz);
        if (u < 0.f || u >= r.w || vv < 0.f || vv >= r.h) continue;
        const float e = fmul(2.0f, fdiv(qz, r.f));
        best = e < best ? e : best;  // std::min(best, e)
      }
    }
  }
  if (i < n) extents[i] = (best != FLT_MAX) ? best : -1.0f;
}

}  // namespace mdi

extern "C" int mdi_min_extents(const float* pts, int64_t n, const float* K, const float* P,
                               const int32_t* sizes, int n_cams, float* extents, void* stream) {
  MDI_REQUIRE(n >= 0 && n_cams >= 0, "mdi_min_extents: negative size");
  if (n == 0) return MDI_OK;
  MDI_REQUIRE(pts && extents && (n_cams == 0 || (K && P && sizes)), "mdi_min_extents: null pointer");
  const int64_t blocks = (n + mdi::kExtThreads - 1) / mdi::kExtThreads;
  MDI_REQUIRE(blocks < (1LL << 31), "mdi_min_extents: too many points");
  mdi::prof_mark("min_extents_kernel", (cudaStream_t)stream);
  mdi::min_extents_kernel<<<(unsigned)blocks, mdi::kExtThreads, 0, (cudaStream_t)stream>>>(
      pts, n, K, P, sizes, n_cams, extents);
  MDI_CHECK_LAUNCH("mdi_min_extents");
  return MDI_OK;
}
