// k1_project_fit.cu -- SfM-point projection + filtering (a2), normal-equation accumulation with
// pinv solve (a3) and the SfM-density patch histogram (a8).  One CTA per view; the per-view
// point sets are a few thousand points, so everything is warp-shuffle / block reductions.
#include "common.cuh"

namespace mdi {

constexpr int kK1Threads = 512;

// a1.  P = (K @ R^T... precisely the reference's `K @ R @ hstack([eye(3), -C[:, None]])` with
// R = c2w[:3,:3].T, C = c2w[:3,3] (points_from_depth.py:231-233): the first product (right
// operand a transposed view) as separately rounded products, the second as an FMA chain over k,
// which is how the CPU sgemm evaluates them (bit-identical on every golden view).  Also copies Kinv and the top 3 rows of c2w into the
// descriptors.  One thread per view.
__device__ __forceinline__ void setup_camera(MdiView* views, int v, const float* K, const float* c2w,
                                             const float* Kinv, float* P_out, int keep_P) {
  const float* k = K + 9 * (int64_t)v;
  const float* m = c2w + 16 * (int64_t)v;
  float P[12];
  if (keep_P) {  // the caller filled the descriptors' P already
    for (int q = 0; q < 12; ++q) P[q] = views[v].P[q];
  } else {
    projection_matrix(k, m, P);
  }
  for (int q = 0; q < 12; ++q) {
    views[v].P[q] = P[q];
    views[v].c2w[q] = m[q];
    if (P_out) P_out[12 * (int64_t)v + q] = P[q];
  }
  if (Kinv) {
    for (int q = 0; q < 9; ++q) views[v].Kinv[q] = Kinv[9 * (int64_t)v + q];
  } else {
    // zero-skew pinhole K: torch.linalg.inv (LU + back substitution) is exactly
    // [[1/fx, 0, -cx/fx], [0, 1/fy, -cy/fy], [0, 0, 1]] with true divisions (pinned on the CPU
    // against torch for 5000 random intrinsics, tests/test_host_math.py)
    float ki[9];
    pinhole_inverse(k, ki);
    for (int q = 0; q < 9; ++q) views[v].Kinv[q] = ki[q];
  }
}

__global__ void setup_cameras_kernel(MdiView* __restrict__ views, int n_views,
                                     const float* __restrict__ K, const float* __restrict__ c2w,
                                     const float* __restrict__ Kinv, float* __restrict__ P_out,
                                     int keep_P) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_views) return;
  setup_camera(views, v, K, c2w, Kinv, P_out, keep_P);
}

// a2.  Reference: points_from_depth.py:145-162 (project, round half-even, cast) and :111-142
// (bounds / z >= 0 validity, the "< 1/4 in bounds" rule, mask lookup, ordered filter).
// kSetup: the CTA first fills its own view's camera fields (a1, mdi_setup_project_filter) -- one launch
// less in front of a kernel that is a chain of dependent round trips anyway.
template <bool kSetup>
__global__ void __launch_bounds__(kK1Threads)
project_filter_kernel(MdiView* views, const float* __restrict__ sfm_xyz,
                      const int32_t* __restrict__ sfm_index, int64_t slots,
                      int32_t* __restrict__ uv, float* __restrict__ z_out,
                      int32_t* __restrict__ n_inbounds, int32_t* __restrict__ n_valid,
                      int32_t* __restrict__ status, const float* __restrict__ K,
                      const float* __restrict__ c2w, const float* __restrict__ Kinv,
                      float* __restrict__ P_out, int keep_P) {
  __shared__ int scan_smem[33];
  __shared__ int s_inb;
  const int v = blockIdx.x;
  if (kSetup) {
    if (threadIdx.x == 0) setup_camera(views, v, K, c2w, Kinv, P_out, keep_P);
    __syncthreads();
  }
  const MdiView vw = views[v];
  const int M = vw.sfm_count;
  const int64_t base = vw.sfm_begin;
  if (threadIdx.x == 0) s_inb = 0;
  __syncthreads();

  // Four consecutive points per thread and step: the index, coordinate and mask loads of a step are
  // dependent gathers (index -> xyz -> pixel of the mask), so the four chains of a thread and the
  // 2048 of the CTA are in flight together and a 6000-point view takes three round trips per stage
  // instead of twelve.  The ordered compaction ranks a thread's points behind the block scan of the
  // per-thread counts.
  constexpr int kPer = 4;
  const bool idx_vec = sfm_index && ((reinterpret_cast<uintptr_t>(sfm_index + base) & 15) == 0);
  int out_base = 0, inb_local = 0;
  for (int start = 0; start < M; start += kPer * (int)blockDim.x) {
    const int i0 = start + kPer * (int)threadIdx.x;
    int64_t src[kPer];
    if (idx_vec && i0 + kPer <= M) {
      const int4 q = *reinterpret_cast<const int4*>(sfm_index + base + i0);
      src[0] = q.x; src[1] = q.y; src[2] = q.z; src[3] = q.w;
    } else {
#pragma unroll
      for (int k = 0; k < kPer; ++k)
        src[k] = i0 + k < M ? (sfm_index ? (int64_t)sfm_index[base + i0 + k] : (base + i0 + k)) : 0;
    }
    float X[kPer], Y[kPer], Z[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const bool in = i0 + k < M;
      X[k] = in ? sfm_xyz[3 * src[k] + 0] : 0.f;
      Y[k] = in ? sfm_xyz[3 * src[k] + 1] : 0.f;
      Z[k] = in ? sfm_xyz[3 * src[k] + 2] : 0.f;
    }
    bool ok[kPer];
    int iu[kPer], iv[kPer];
    float zc[kPer];
    uint8_t mb[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      // P @ [X;1] as the sgemm of the reference evaluates it: one FMA chain over k = 0..3
      const float* P = vw.P;
      const float cx = dot4_fma(P[0], P[1], P[2], P[3], X[k], Y[k], Z[k]);
      const float cy = dot4_fma(P[4], P[5], P[6], P[7], X[k], Y[k], Z[k]);
      zc[k] = dot4_fma(P[8], P[9], P[10], P[11], X[k], Y[k], Z[k]);
      const float fu = rintf(fdiv(cx, zc[k]));  // torch.round = half to even
      const float fv = rintf(fdiv(cy, zc[k]));
      // 0 <= u < W, 0 <= v < H, z >= 0 ; NaN / inf / out-of-int64 all fail the comparisons
      const bool inb = (i0 + k < M) && (fu >= 0.f) && (fu < (float)vw.W) && (fv >= 0.f) && (fv < (float)vw.H) &&
                       (zc[k] >= 0.f);
      iu[k] = inb ? (int)fu : 0;
      iv[k] = inb ? (int)fv : 0;
      ok[k] = inb;
      if (inb) ++inb_local;
    }
#pragma unroll
    for (int k = 0; k < kPer; ++k) mb[k] = (ok[k] && vw.mask) ? vw.mask[(int64_t)iv[k] * vw.W + iu[k]] : 1;
    int mine = 0;
#pragma unroll
    for (int k = 0; k < kPer; ++k) { ok[k] = ok[k] && mb[k] != 0; mine += ok[k] ? 1 : 0; }
    int total;
    int pos = block_excl_scan(mine, scan_smem, &total);
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      if (ok[k]) {
        const int64_t o = base + out_base + pos;
        uv[o] = iu[k];
        uv[slots + o] = iv[k];
        z_out[o] = zc[k];
        ++pos;
      }
    }
    out_base += total;
  }
  int inb_w = warp_sum_i32(inb_local);
  if (lane_id() == 0 && inb_w) atomicAdd(&s_inb, inb_w);
  __syncthreads();
  if (threadIdx.x == 0) {
    n_inbounds[v] = s_inb;
    n_valid[v] = out_base;
    // torch.sum(valid) < M / 4  (true division)  <=>  4 * n < M
    status[v] = (4LL * (int64_t)s_inb < (int64_t)M) ? MDI_STATUS_LOW_CONFIDENCE : MDI_STATUS_OK;
  }
}

// a3 + a8.  Reference: lstsqrs.py:39-50 (gather depth[v,u]), :21-26 (normal equations + pinv);
// num_sfm_points_mask.py:45-63 (points per patch, > threshold blocks the patch).
__global__ void __launch_bounds__(kK1Threads)
gather_fit_kernel(const MdiView* __restrict__ views, int64_t slots,
                  const int32_t* __restrict__ uv, const float* __restrict__ z,
                  const int32_t* __restrict__ n_valid, float* __restrict__ d_out,
                  double* __restrict__ sums, float* __restrict__ scale_shift, int write_fit,
                  int patch_threshold, uint8_t* __restrict__ patch_blocked) {
  __shared__ int s_hist[MDI_PATCH_STRIDE];
  __shared__ double s_red[5][kK1Threads / 32];
  const int v = blockIdx.x;
  const MdiView vw = views[v];
  const int n = n_valid[v];
  const int64_t base = vw.sfm_begin;
  const int n_patches = vw.grid_h * vw.grid_w;
  if (patch_blocked)
    for (int p = threadIdx.x; p < MDI_PATCH_STRIDE; p += blockDim.x) s_hist[p] = 0;
  __syncthreads();

  double sdd = 0, sd = 0, sdg = 0, sg = 0;
  // r / patch_h as a multiply-high: exact because r, patch_h < 2^16 (x * d < 2^32 with M = floor(2^32 / d) + 1)
  const bool small = vw.H < 65536 && vw.W < 65536;
  const uint32_t mh = 0xFFFFFFFFu / (uint32_t)max(vw.patch_h, 1) + 1u, mw = 0xFFFFFFFFu / (uint32_t)max(vw.patch_w, 1) + 1u;
  // four depth gathers of a thread in flight together (the loop is a chain of DRAM round trips otherwise);
  // the sums keep the order i = threadIdx.x, + blockDim.x, ... of the plain loop
  constexpr int kPer = 4;
  for (int i0 = threadIdx.x; i0 < n; i0 += kPer * (int)blockDim.x) {
    int iu[kPer], iv[kPer];
    float g[kPer], dv[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = i0 + k * (int)blockDim.x;
      const bool in = i < n;
      iu[k] = in ? uv[base + i] : 0; iv[k] = in ? uv[slots + base + i] : 0;
      g[k] = in ? z[base + i] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < kPer; ++k)
      dv[k] = (i0 + k * (int)blockDim.x < n) ? vw.depth[(int64_t)iv[k] * vw.W + iu[k]] : 0.f;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = i0 + k * (int)blockDim.x;
      if (i >= n) break;
      d_out[base + i] = dv[k];
      const double dd = (double)dv[k], gd = (double)g[k];
      sdd += dd * dd; sd += dd; sdg += dd * gd; sg += gd;
      if (patch_blocked) {
        int pr, pc;
        if (small) {
          pr = vw.patch_h <= 1 ? iv[k] : (int)__umulhi((uint32_t)iv[k], mh);
          pc = vw.patch_w <= 1 ? iu[k] : (int)__umulhi((uint32_t)iu[k], mw);
        } else {
          pr = iv[k] / vw.patch_h; pc = iu[k] / vw.patch_w;
        }
        if (pr < vw.grid_h && pc < vw.grid_w) atomicAdd(&s_hist[pr * vw.grid_w + pc], 1);
      }
    }
  }
  sdd = warp_sum_f64(sdd); sd = warp_sum_f64(sd); sdg = warp_sum_f64(sdg); sg = warp_sum_f64(sg);
  if (lane_id() == 0) {
    const int w = warp_id();
    s_red[0][w] = sdd; s_red[1][w] = sd; s_red[2][w] = sdg; s_red[3][w] = sg;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0, c = 0, e = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      a += s_red[0][w]; b += s_red[1][w]; c += s_red[2][w]; e += s_red[3][w];
    }
    sums[5 * v + 0] = a; sums[5 * v + 1] = b; sums[5 * v + 2] = (double)n;
    sums[5 * v + 3] = c; sums[5 * v + 4] = e;
    if (write_fit) {
      float h0, h1;
      solve_normal_eq(a, b, (double)n, c, e, &h0, &h1);
      scale_shift[2 * v + 0] = h0;
      scale_shift[2 * v + 1] = h1;
    }
  }
  if (patch_blocked) {
    for (int p = threadIdx.x; p < MDI_PATCH_STRIDE; p += blockDim.x)
      patch_blocked[(int64_t)v * MDI_PATCH_STRIDE + p] =
          (p < n_patches && s_hist[p] > patch_threshold) ? 1 : 0;
  }
}

__global__ void apply_scale_shift_kernel(const float* __restrict__ depth, int64_t n,
                                         const float* __restrict__ scale_shift,
                                         float* __restrict__ out) {
  const float s = scale_shift[0], t = scale_shift[1];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = align_depth(depth[i], s, t);
}

// Stitch of the region-wise alignment (pipeline.py:252-283): pixel p of region r gets
// depth * s_r + t_r; pixels of a region without a fit (no SfM point in it) and pixels outside every
// aligned region keep the sentinel -42.0.  region int32 [n]: index into the table, < 0 = none.
__global__ void apply_region_scale_shift_kernel(const float* __restrict__ depth, const int32_t* __restrict__ region,
                                                int64_t n, const float* __restrict__ table,
                                                const uint8_t* __restrict__ valid, int n_regions,
                                                float* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int r = region[i];
    float a = -42.0f;
    if (r >= 0 && r < n_regions && valid[r]) a = align_depth(depth[i], table[2 * r], table[2 * r + 1]);
    out[i] = a;
  }
}

// small host table -> device, read by the SMs from mapped pinned memory (see mdi_upload_table)
__global__ void upload_table_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, int64_t n_words) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += stride) dst[i] = src[i];
}

}  // namespace mdi

extern "C" int mdi_upload_table(void* dst, const void* src_pinned, int64_t bytes, void* stream) {
  MDI_REQUIRE(bytes >= 0 && bytes % 4 == 0, "mdi_upload_table: size %lld is not a multiple of 4", (long long)bytes);
  if (bytes == 0) return MDI_OK;
  MDI_REQUIRE(dst && src_pinned && ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src_pinned)) & 3) == 0,
              "mdi_upload_table: null or unaligned pointer");
  const int64_t n_words = bytes / 4;
  const unsigned grid = (unsigned)((n_words + 255) / 256 < 296 ? (n_words + 255) / 256 : 296);
  mdi::prof_mark("upload_table_kernel", (cudaStream_t)stream);
  mdi::upload_table_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      static_cast<uint32_t*>(dst), static_cast<const uint32_t*>(src_pinned), n_words);
  MDI_CHECK_LAUNCH("mdi_upload_table");
  return MDI_OK;
}

extern "C" int mdi_setup_cameras(MdiView* views, int n_views, const float* K, const float* c2w,
                                 const float* Kinv, float* P_out, int keep_P, void* stream) {
  MDI_REQUIRE(n_views >= 0, "mdi_setup_cameras: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && K && c2w, "mdi_setup_cameras: null pointer");
  mdi::prof_mark("setup_cameras_kernel", (cudaStream_t)stream);
  mdi::setup_cameras_kernel<<<(n_views + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      views, n_views, K, c2w, Kinv, P_out, keep_P);
  MDI_CHECK_LAUNCH("mdi_setup_cameras");
  return MDI_OK;
}

extern "C" int mdi_project_filter(const MdiView* views, int n_views, const float* sfm_xyz,
                                  const int32_t* sfm_index, int64_t slots, int32_t* uv, float* z,
                                  int32_t* n_inbounds, int32_t* n_valid, int32_t* status,
                                  void* stream) {
  MDI_REQUIRE(n_views >= 0 && slots >= 0, "mdi_project_filter: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && sfm_xyz && uv && z && n_inbounds && n_valid && status,
              "mdi_project_filter: null pointer");
  mdi::prof_mark("project_filter_kernel", (cudaStream_t)stream);
  mdi::project_filter_kernel<false><<<n_views, mdi::kK1Threads, 0, (cudaStream_t)stream>>>(
      const_cast<MdiView*>(views), sfm_xyz, sfm_index, slots, uv, z, n_inbounds, n_valid, status, nullptr, nullptr,
      nullptr, nullptr, 0);
  MDI_CHECK_LAUNCH("mdi_project_filter");
  return MDI_OK;
}

extern "C" int mdi_setup_project_filter(MdiView* views, int n_views, const float* K, const float* c2w,
                                        const float* Kinv, float* P_out, int keep_P, const float* sfm_xyz,
                                        const int32_t* sfm_index, int64_t slots, int32_t* uv, float* z,
                                        int32_t* n_inbounds, int32_t* n_valid, int32_t* status,
                                        void* stream) {
  MDI_REQUIRE(n_views >= 0 && slots >= 0, "mdi_setup_project_filter: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && K && c2w && sfm_xyz && uv && z && n_inbounds && n_valid && status,
              "mdi_setup_project_filter: null pointer");
  mdi::prof_mark("project_filter_kernel", (cudaStream_t)stream);
  mdi::project_filter_kernel<true><<<n_views, mdi::kK1Threads, 0, (cudaStream_t)stream>>>(
      views, sfm_xyz, sfm_index, slots, uv, z, n_inbounds, n_valid, status, K, c2w, Kinv, P_out, keep_P);
  MDI_CHECK_LAUNCH("mdi_setup_project_filter");
  return MDI_OK;
}

extern "C" int mdi_gather_fit(const MdiView* views, int n_views, int64_t slots, const int32_t* uv,
                              const float* z, const int32_t* n_valid, float* d, double* sums,
                              float* scale_shift, int write_fit, int patch_threshold,
                              uint8_t* patch_blocked, void* stream) {
  MDI_REQUIRE(n_views >= 0 && slots >= 0, "mdi_gather_fit: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && uv && z && n_valid && d && sums, "mdi_gather_fit: null pointer");
  MDI_REQUIRE(!write_fit || scale_shift, "mdi_gather_fit: write_fit without scale_shift");
  mdi::prof_mark("gather_fit_kernel", (cudaStream_t)stream);
  mdi::gather_fit_kernel<<<n_views, mdi::kK1Threads, 0, (cudaStream_t)stream>>>(
      views, slots, uv, z, n_valid, d, sums, scale_shift, write_fit, patch_threshold,
      patch_blocked);
  MDI_CHECK_LAUNCH("mdi_gather_fit");
  return MDI_OK;
}

extern "C" int mdi_apply_region_scale_shift(const float* depth, const int32_t* region, int64_t n,
                                            const float* table, const uint8_t* valid, int n_regions,
                                            float* out, void* stream) {
  MDI_REQUIRE(n >= 0 && n_regions >= 0, "mdi_apply_region_scale_shift: negative size");
  if (n == 0) return MDI_OK;
  MDI_REQUIRE(depth && region && out && (n_regions == 0 || (table && valid)), "mdi_apply_region_scale_shift: null pointer");
  int blocks = (int)((n + 1023) / 1024);
  if (blocks > mdi::kSMs * 16) blocks = mdi::kSMs * 16;
  mdi::prof_mark("apply_region_scale_shift_kernel", (cudaStream_t)stream);
  mdi::apply_region_scale_shift_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(depth, region, n, table, valid,
                                                                                 n_regions, out);
  MDI_CHECK_LAUNCH("mdi_apply_region_scale_shift");
  return MDI_OK;
}

extern "C" int mdi_apply_scale_shift(const float* depth, int64_t n, const float* scale_shift,
                                     float* out, void* stream) {
  MDI_REQUIRE(n >= 0, "mdi_apply_scale_shift: negative size");
  if (n == 0) return MDI_OK;
  MDI_REQUIRE(depth && scale_shift && out, "mdi_apply_scale_shift: null pointer");
  int blocks = (int)((n + 1023) / 1024);
  if (blocks > mdi::kSMs * 16) blocks = mdi::kSMs * 16;
  mdi::prof_mark("apply_scale_shift_kernel", (cudaStream_t)stream);
  mdi::apply_scale_shift_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(depth, n, scale_shift,
                                                                          out);
  MDI_CHECK_LAUNCH("mdi_apply_scale_shift");
  return MDI_OK;
}
