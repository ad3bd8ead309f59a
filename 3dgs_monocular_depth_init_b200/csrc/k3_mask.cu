// k3_mask.cu -- keep-mask predicate (a5-a10): alignment applied on the fly, sentinel / >= 0 /
// predicted-mask tests, SfM-density patch lookup, optional gradient mask, static or adaptive
// stride test.  One CTA per 4096-pixel tile; every lane owns 4 consecutive pixels per step
// (one 128-bit depth load + one 32-bit mask load), the warp assembles 4 bit words per step
// with redux.or, the tile count feeds the order-preserving compaction (mdi_scan_tiles /
// mdi_emit_points).  HBM-bound: 5 B/px read, 1 B/px (+1/8 B/px bits) written.
#include <mutex>

#include "common.cuh"

namespace mdi {

constexpr int kMaskThreads = 256;
constexpr int kDivTabSize = 1 << 16;

// divtab[x] bit k (1 <= k <= 31) = (x % k == 0)
__global__ void build_divtab_kernel(uint32_t* tab, int n) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n) return;
  uint32_t m = 0;
  for (int k = 1; k < 32; ++k) m |= ((x % k) == 0 ? 1u : 0u) << k;
  tab[x] = m;
}

static std::mutex g_tab_mutex;
static uint32_t* g_divtab[64] = {nullptr};

// lazily built per device; contents are a pure function of the index, so sharing is safe
static int get_divtab(cudaStream_t stream, const uint32_t** out) {
  int dev = 0;
  MDI_CUDA(cudaGetDevice(&dev), "divtab");
  MDI_REQUIRE(dev >= 0 && dev < 64, "divtab: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_tab_mutex);
  if (!g_divtab[dev]) {
    uint32_t* p = nullptr;
    MDI_CUDA(cudaMalloc(&p, sizeof(uint32_t) * kDivTabSize), "divtab alloc");
    build_divtab_kernel<<<kDivTabSize / 256, 256, 0, stream>>>(p, kDivTabSize);
    MDI_CHECK_LAUNCH("divtab build");
    MDI_CUDA(cudaStreamSynchronize(stream), "divtab sync");
    g_divtab[dev] = p;
  }
  *out = g_divtab[dev];
  return MDI_OK;
}

__device__ __forceinline__ bool divides_tab(const uint32_t* __restrict__ tab, int idx, int f) {
  if (f > 0 && f < 32 && idx < kDivTabSize) return (__ldg(&tab[idx]) >> f) & 1u;
  return divides(idx, f);
}

// |d(r,c) - d(r,c-1)| + |d(r,c) - d(r-1,c)| on the aligned depth (points_from_depth.py:205-209)
__device__ __forceinline__ float grad_at(const MdiView& vw, int64_t p, int row, int col, float a,
                                         float s, float t) {
  float g = 0.f;
  if (col >= 1) g = fadd(g, fabsf(fsub(a, align_depth(__ldg(&vw.depth[p - 1]), s, t))));
  if (row >= 1) g = fadd(g, fabsf(fsub(a, align_depth(__ldg(&vw.depth[p - vw.W]), s, t))));
  return g;
}

struct MaskParams {
  int mode, static_factor, fmin, fmax, flags;
  float grad_thresh;
};

__global__ void __launch_bounds__(kMaskThreads)
keep_mask_kernel(const MdiView* __restrict__ views, int n_views,
                 const float* __restrict__ scale_shift, const int32_t* __restrict__ status,
                 const float* __restrict__ lo_hi, const float* __restrict__ grad_range,
                 const uint8_t* __restrict__ patch_blocked, const uint32_t* __restrict__ divtab,
                 MaskParams prm, uint8_t* __restrict__ keep, uint32_t* __restrict__ bits,
                 int32_t* __restrict__ tile_counts) {
  __shared__ int s_count;
  const int64_t tile = blockIdx.x;
  const int v = find_view(views, n_views, tile);
  const MdiView vw = views[v];
  const int64_t HW = (int64_t)vw.H * vw.W;
  const int64_t tile_px0 = (tile - vw.tile_begin) * MDI_TILE_PX;
  const bool skip = status && status[v] != MDI_STATUS_OK;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  float lo = 0.f, hi = 0.f, gmin = 0.f, gden = 1.f;
  if (prm.mode == MDI_SUBSAMPLE_ADAPTIVE && lo_hi) { lo = lo_hi[2 * v]; hi = lo_hi[2 * v + 1]; }
  if ((prm.flags & MDI_KEEP_GRAD) && grad_range) {
    gmin = grad_range[2 * v];
    // (g - min) / ((max - min) + 1e-8)   (points_from_depth.py:210-211)
    gden = fadd(fsub(grad_range[2 * v + 1], gmin), 1e-8f);
  }
  const uint8_t* pb = (prm.flags & MDI_KEEP_PATCH) && patch_blocked
                          ? patch_blocked + (int64_t)v * MDI_PATCH_STRIDE : nullptr;
  uint8_t* keep_v = keep ? keep + vw.px_begin : nullptr;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(vw.depth) & 15) == 0) &&
                      (!vw.mask || (reinterpret_cast<uintptr_t>(vw.mask) & 3) == 0) &&
                      (!keep_v || (reinterpret_cast<uintptr_t>(keep_v) & 3) == 0);
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();

  const int lane = lane_id(), warp = warp_id();
  int my_count = 0;
#pragma unroll 1
  for (int it = 0; it < 4; ++it) {
    const int local = warp * 512 + it * 128 + lane * 4;
    const int64_t p0 = tile_px0 + local;
    uint32_t nib = 0;
    if (!skip && p0 < HW) {
      float dv[4];
      uint32_t mk = 0x01010101u;
      const bool full = (p0 + 4 <= HW);
      if (full && vec_ok) {
        const float4 d4 = __ldg(reinterpret_cast<const float4*>(vw.depth + p0));
        dv[0] = d4.x; dv[1] = d4.y; dv[2] = d4.z; dv[3] = d4.w;
        if (vw.mask) mk = __ldg(reinterpret_cast<const uint32_t*>(vw.mask + p0));
      } else {
        mk = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const bool in = (p0 + k < HW);
          dv[k] = in ? __ldg(&vw.depth[p0 + k]) : 0.f;
          const uint32_t m = in ? (vw.mask ? (uint32_t)__ldg(&vw.mask[p0 + k]) : 1u) : 0u;
          mk |= (m ? 1u : 0u) << (8 * k);
        }
      }
      int row = (int)(p0 / vw.W);
      int col = (int)(p0 - (int64_t)row * vw.W);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (p0 + k < HW) {
          const float a = align_depth(dv[k], s, t);
          bool ok = true;
          if (prm.flags & MDI_KEEP_PRED_MASK) ok = ((mk >> (8 * k)) & 0xffu) != 0;
          if (prm.flags & MDI_KEEP_SENTINEL) ok = ok && (a != -42.0f);
          const bool in_pop = ok;  // population of the subsampler's `mask` argument
          if (prm.flags & MDI_KEEP_NONNEG) ok = ok && (a >= 0.f);
          if (ok && pb) {
            const int pr = row / vw.patch_h, pc = col / vw.patch_w;
            if (pr < vw.grid_h && pc < vw.grid_w && __ldg(&pb[pr * vw.grid_w + pc])) ok = false;
          }
          if (ok && (prm.flags & MDI_KEEP_GRAD)) {
            const float g = grad_at(vw, p0 + k, row, col, a, s, t);
            ok = fdiv(fsub(g, gmin), gden) <= prm.grad_thresh;
          }
          if (ok && (prm.flags & MDI_KEEP_SUBSAMPLE)) {
            int f;
            if (prm.mode == MDI_SUBSAMPLE_STATIC) {
              f = prm.static_factor;
            } else {
              // invalid pixels get multiplier 0.5 but are removed by `& mask` anyway
              f = in_pop ? adaptive_factor(a, lo, hi, prm.fmin, prm.fmax) : 1;
            }
            ok = divides_tab(divtab, row, f) && divides_tab(divtab, col, f);
          }
          nib |= (ok ? 1u : 0u) << k;
        }
        if (++col == vw.W) { col = 0; ++row; }
      }
    }
    // keep bytes
    if (keep_v && p0 < HW) {
      if (vec_ok && p0 + 4 <= HW) {
        const uint32_t bytes = (nib & 1u) | ((nib & 2u) << 7) | ((nib & 4u) << 14) | ((nib & 8u) << 21);
        *reinterpret_cast<uint32_t*>(keep_v + p0) = bytes;
      } else {
        for (int k = 0; k < 4; ++k)
          if (p0 + k < HW) keep_v[p0 + k] = (nib >> k) & 1u;
      }
    }
    // 4 bit words of this 128-pixel step: word j <- nibbles of lanes 8j .. 8j+7
    const uint32_t shifted = nib << (4 * (lane & 7));
    uint32_t mine = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t w = __reduce_or_sync(0xffffffffu, (lane >> 3) == j ? shifted : 0u);
      if (lane == j) mine = w;
    }
    if (lane < 4) {
      bits[tile * MDI_TILE_WORDS + warp * 16 + it * 4 + lane] = mine;
      my_count += __popc(mine);
    }
  }
  my_count = warp_sum_i32(my_count);
  if (lane == 0 && my_count) atomicAdd(&s_count, my_count);
  __syncthreads();
  if (threadIdx.x == 0) tile_counts[tile] = s_count;
}

}  // namespace mdi

extern "C" int mdi_keep_mask(const MdiView* views, int n_views, int64_t total_tiles,
                             const float* scale_shift, const int32_t* status, const float* lo_hi,
                             const float* grad_range, const uint8_t* patch_blocked,
                             const MdiSubsampleConfig* cfg, uint8_t* keep, uint32_t* bits,
                             int32_t* tile_counts, void* stream) {
  MDI_REQUIRE(n_views >= 0 && total_tiles >= 0, "mdi_keep_mask: negative size");
  if (n_views == 0 || total_tiles == 0) return MDI_OK;
  MDI_REQUIRE(views && cfg && bits && tile_counts, "mdi_keep_mask: null pointer");
  MDI_REQUIRE(total_tiles < (1LL << 31), "mdi_keep_mask: too many tiles");
  mdi::MaskParams prm{cfg->mode, cfg->static_factor, cfg->factor_min, cfg->factor_max, cfg->flags,
                      cfg->grad_thresh};
  if (prm.flags & MDI_KEEP_SUBSAMPLE) {
    if (prm.mode == MDI_SUBSAMPLE_STATIC)
      MDI_REQUIRE(prm.static_factor != 0, "mdi_keep_mask: static_factor == 0 (modulo by zero)");
    else {
      MDI_REQUIRE(prm.mode == MDI_SUBSAMPLE_ADAPTIVE, "mdi_keep_mask: unknown mode %d", prm.mode);
      MDI_REQUIRE(lo_hi, "mdi_keep_mask: adaptive mode needs lo_hi");
    }
  }
  MDI_REQUIRE(!(prm.flags & MDI_KEEP_GRAD) || grad_range, "mdi_keep_mask: grad flag needs grad_range");
  const uint32_t* tab = nullptr;
  int rc = mdi::get_divtab((cudaStream_t)stream, &tab);
  if (rc != MDI_OK) return rc;
  mdi::keep_mask_kernel<<<(unsigned)total_tiles, mdi::kMaskThreads, 0, (cudaStream_t)stream>>>(
      views, n_views, scale_shift, status, lo_hi, grad_range, patch_blocked, tab, prm, keep, bits,
      tile_counts);
  MDI_CHECK_LAUNCH("mdi_keep_mask");
  return MDI_OK;
}
