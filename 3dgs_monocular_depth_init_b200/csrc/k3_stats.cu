// k3_stats.cu -- per-view statistics of the aligned depth for adaptive subsampling (a7):
// masked min / max and the exact order statistics behind torch.quantile(x, 0.25 / 0.75)
// (adaptive_subsampling.py:82-95), by a 12/12/8-bit radix select on order-preserving keys.
// Three streaming passes over depth + mask (HBM/L2-bound, shared-memory histograms), each
// followed by a one-CTA-per-view bin search.  Also the gradient-map range for a9.
#include "common.cuh"

namespace mdi {

constexpr int kStatsThreads = 256;
constexpr int kStatsTilesPerCta = 16;  // 64K pixels per CTA amortise histogram zero/flush
constexpr int kBins = 4096;
constexpr int kTargets = 4;            // q1.lo, q1.hi, q3.lo, q3.hi

struct StatsView {
  uint32_t hist[kTargets][kBins];
  uint32_t min_key, max_key, count, n_slots;
  uint32_t slot_prefix[kTargets];  // prefix each histogram slot selects on
  uint32_t tgt_prefix[kTargets];
  uint32_t tgt_slot[kTargets];
  long long tgt_rank[kTargets];
  float w[2];
  uint32_t gmin_key, gmax_key;     // gradient range (mdi_grad_range)
  uint32_t _pad[2];
};

__device__ __forceinline__ bool in_population(uint32_t maskbyte, float a, int flags) {
  bool ok = true;
  if (flags & MDI_KEEP_PRED_MASK) ok = maskbyte != 0;
  if (flags & MDI_KEEP_SENTINEL) ok = ok && (a != -42.0f);
  return ok;
}

// PASS 0: bin = key >> 20 over everything (+ min / max / count)
// PASS 1: bin = (key >> 8) & 0xfff over keys whose top 12 bits match a slot prefix
// PASS 2: bin = key & 0xff over keys whose top 24 bits match a slot prefix
template <int PASS>
__global__ void __launch_bounds__(kStatsThreads)
stats_pass_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                  const int32_t* __restrict__ status, int flags, StatsView* __restrict__ ws) {
  extern __shared__ uint32_t s_hist[];  // n_slots * bins
  __shared__ uint32_t s_min, s_max, s_cnt;
  const int v = blockIdx.y;
  const MdiView vw = views[v];
  const int first_tile = blockIdx.x * kStatsTilesPerCta;
  if (first_tile >= vw.n_tiles) return;
  if (status && (status[v] & MDI_STATUS_LOW_CONFIDENCE)) return;
  StatsView* sv = ws + v;
  constexpr int bins = (PASS == 2) ? 256 : kBins;
  const int n_slots = (PASS == 0) ? 1 : (int)sv->n_slots;
  if (PASS != 0 && n_slots == 0) return;
  uint32_t pre[kTargets];
#pragma unroll
  for (int k = 0; k < kTargets; ++k) pre[k] = (PASS != 0 && k < n_slots) ? sv->slot_prefix[k] : 0;
  for (int i = threadIdx.x; i < n_slots * bins; i += blockDim.x) s_hist[i] = 0;
  if (threadIdx.x == 0) { s_min = 0xFFFFFFFFu; s_max = 0u; s_cnt = 0u; }
  __syncthreads();

  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const int64_t HW = (int64_t)vw.H * vw.W;
  const int64_t px_begin = (int64_t)first_tile * MDI_TILE_PX;
  int64_t px_end = px_begin + (int64_t)kStatsTilesPerCta * MDI_TILE_PX;
  if (px_end > HW) px_end = HW;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(vw.depth) & 15) == 0) &&
                      (!vw.mask || (reinterpret_cast<uintptr_t>(vw.mask) & 3) == 0);
  uint32_t kmin = 0xFFFFFFFFu, kmax = 0u, cnt = 0;
  // run-length combining of equal bins per thread cuts shared atomics on smooth depth
  int run_bin = -1; uint32_t run_len = 0;
  auto flush = [&]() { if (run_len) atomicAdd(&s_hist[run_bin], run_len); run_len = 0; };
  auto feed = [&](float dv, uint32_t mbyte) {
    const float a = align_depth(dv, s, t);
    if (!in_population(mbyte, a, flags)) return;
    const uint32_t key = float_to_key(a);
    if (PASS == 0) {
      kmin = min(kmin, key); kmax = max(kmax, key); ++cnt;
      const int b = (int)(key >> 20);
      if (b == run_bin) ++run_len; else { flush(); run_bin = b; run_len = 1; }
    } else {
      const uint32_t top = (PASS == 1) ? (key >> 20) : (key >> 8);
      const int b = (PASS == 1) ? (int)((key >> 8) & 0xfffu) : (int)(key & 0xffu);
#pragma unroll
      for (int k = 0; k < kTargets; ++k)
        if (k < n_slots && top == pre[k]) atomicAdd(&s_hist[k * bins + b], 1u);
    }
  };
  for (int64_t p0 = px_begin + (int64_t)threadIdx.x * 4; p0 < px_end;
       p0 += (int64_t)kStatsThreads * 4) {
    if (vec_ok && p0 + 4 <= px_end) {
      const float4 d4 = __ldg(reinterpret_cast<const float4*>(vw.depth + p0));
      const uint32_t mk = vw.mask ? __ldg(reinterpret_cast<const uint32_t*>(vw.mask + p0))
                                  : 0x01010101u;
      feed(d4.x, mk & 0xffu); feed(d4.y, (mk >> 8) & 0xffu);
      feed(d4.z, (mk >> 16) & 0xffu); feed(d4.w, mk >> 24);
    } else {
      for (int k = 0; k < 4; ++k)
        if (p0 + k < px_end)
          feed(__ldg(&vw.depth[p0 + k]), vw.mask ? (uint32_t)__ldg(&vw.mask[p0 + k]) : 1u);
    }
  }
  if (PASS == 0) {
    flush();
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane_id() == 0 && cnt) { atomicMin(&s_min, kmin); atomicMax(&s_max, kmax); atomicAdd(&s_cnt, cnt); }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_slots * bins; i += blockDim.x) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&sv->hist[i / bins][i % bins], c);
  }
  if (PASS == 0 && threadIdx.x == 0 && s_cnt) {
    atomicMin(&sv->min_key, s_min); atomicMax(&sv->max_key, s_max); atomicAdd(&sv->count, s_cnt);
  }
}

// One CTA per view: locate each target rank in its slot histogram, extend the prefixes,
// deduplicate them into the next pass's slots and clear the histograms.  After PASS 2 the
// prefixes are the keys of the order statistics: finish the quantiles and the range.
template <int PASS>
__global__ void __launch_bounds__(kStatsThreads)
stats_select_kernel(int32_t* __restrict__ status, StatsView* __restrict__ ws,
                    float* __restrict__ lo_hi, float* __restrict__ quantiles) {
  __shared__ int scan_smem[33];
  __shared__ uint32_t s_bin[kTargets];
  __shared__ long long s_rank[kTargets];
  const int v = blockIdx.x;
  StatsView* sv = ws + v;
  constexpr int bins = (PASS == 2) ? 256 : kBins;
  constexpr int shift = (PASS == 2) ? 8 : 12;
  constexpr int per_thread = bins / kStatsThreads;  // 16 or 1
  const bool low_conf = status && (status[v] & MDI_STATUS_LOW_CONFIDENCE);
  const uint32_t n = sv->count;
  if (low_conf || n == 0) {
    if (threadIdx.x == 0) {
      sv->n_slots = 0;
      if (PASS == 2) {
        if (lo_hi) { lo_hi[2 * v] = 0.f; lo_hi[2 * v + 1] = 0.f; }
        if (quantiles) for (int k = 0; k < 4; ++k) quantiles[4 * v + k] = 0.f;
        if (!low_conf && status) status[v] |= MDI_STATUS_EMPTY_STATS;
      }
    }
    return;
  }
  if (PASS == 0 && threadIdx.x == 0) {
    const QuantileRank r1 = quantile_rank(0.25f, (int64_t)n);
    const QuantileRank r3 = quantile_rank(0.75f, (int64_t)n);
    sv->tgt_rank[0] = r1.lo; sv->tgt_rank[1] = r1.hi; sv->tgt_rank[2] = r3.lo; sv->tgt_rank[3] = r3.hi;
    sv->w[0] = r1.w; sv->w[1] = r3.w;
    for (int k = 0; k < kTargets; ++k) { sv->tgt_prefix[k] = 0; sv->tgt_slot[k] = 0; }
  }
  __syncthreads();
  for (int j = 0; j < kTargets; ++j) {
    const uint32_t* h = sv->hist[sv->tgt_slot[j]];
    const long long rank = sv->tgt_rank[j];
    uint32_t c[per_thread];
    int mine = 0;
#pragma unroll
    for (int k = 0; k < per_thread; ++k) { c[k] = h[threadIdx.x * per_thread + k]; mine += (int)c[k]; }
    int tot;
    long long cum = block_excl_scan(mine, scan_smem, &tot);
    if (rank >= cum && rank < cum + mine) {
#pragma unroll
      for (int k = 0; k < per_thread; ++k) {
        if (rank >= cum && rank < cum + c[k]) { s_bin[j] = threadIdx.x * per_thread + k; s_rank[j] = rank - cum; }
        cum += c[k];
      }
    }
    __syncthreads();
  }
  // all reads of the histograms are done: clear them for the next pass
  for (int i = threadIdx.x; i < kTargets * kBins; i += blockDim.x) (&sv->hist[0][0])[i] = 0;
  if (threadIdx.x == 0) {
    uint32_t pre[kTargets];
    for (int j = 0; j < kTargets; ++j) {
      pre[j] = (sv->tgt_prefix[j] << shift) | s_bin[j];
      sv->tgt_prefix[j] = pre[j];
      sv->tgt_rank[j] = s_rank[j];
    }
    int ns = 0;
    for (int j = 0; j < kTargets; ++j) {
      int found = -1;
      for (int i = 0; i < ns; ++i) if (sv->slot_prefix[i] == pre[j]) found = i;
      if (found < 0) { sv->slot_prefix[ns] = pre[j]; found = ns++; }
      sv->tgt_slot[j] = found;
    }
    sv->n_slots = ns;
    if (PASS == 2) {
      const bool has_nan = sv->max_key == 0xFFFFFFFFu;
      float val[kTargets];
      for (int j = 0; j < kTargets; ++j) val[j] = key_to_float(pre[j]);
      float q1 = torch_lerp(val[0], val[1], sv->w[0]);
      float q3 = torch_lerp(val[2], val[3], sv->w[1]);
      float mn = key_to_float(sv->min_key), mx = key_to_float(sv->max_key);
      if (has_nan) { q1 = q3 = mn = mx = NAN; }
      float lo, hi;
      iqr_range(q1, q3, mn, mx, &lo, &hi);
      if (lo_hi) { lo_hi[2 * v] = lo; lo_hi[2 * v + 1] = hi; }
      if (quantiles) { quantiles[4 * v] = q1; quantiles[4 * v + 1] = q3; quantiles[4 * v + 2] = mn; quantiles[4 * v + 3] = mx; }
    }
  }
}

// a9: range of the gradient map over ALL pixels (points_from_depth.py:205-211)
__global__ void __launch_bounds__(kStatsThreads)
grad_range_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                  const int32_t* __restrict__ status, uint32_t* __restrict__ keys) {
  const int v = blockIdx.y;
  const MdiView vw = views[v];
  const int first_tile = blockIdx.x * kStatsTilesPerCta;
  if (first_tile >= vw.n_tiles) return;
  if (status && status[v] != MDI_STATUS_OK) return;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const int64_t HW = (int64_t)vw.H * vw.W;
  const int64_t px_begin = (int64_t)first_tile * MDI_TILE_PX;
  int64_t px_end = px_begin + (int64_t)kStatsTilesPerCta * MDI_TILE_PX;
  if (px_end > HW) px_end = HW;
  uint32_t kmin = 0xFFFFFFFFu, kmax = 0u;
  for (int64_t p = px_begin + threadIdx.x; p < px_end; p += kStatsThreads) {
    const int row = (int)(p / vw.W), col = (int)(p - (int64_t)row * vw.W);
    const float a = align_depth(__ldg(&vw.depth[p]), s, t);
    float g = 0.f;
    if (col >= 1) g = fadd(g, fabsf(fsub(a, align_depth(__ldg(&vw.depth[p - 1]), s, t))));
    if (row >= 1) g = fadd(g, fabsf(fsub(a, align_depth(__ldg(&vw.depth[p - vw.W]), s, t))));
    const uint32_t key = float_to_key(g);
    kmin = min(kmin, key); kmax = max(kmax, key);
  }
  kmin = __reduce_min_sync(0xffffffffu, kmin);
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  if (lane_id() == 0) { atomicMin(&keys[2 * v], kmin); atomicMax(&keys[2 * v + 1], kmax); }
}

__global__ void grad_range_init_kernel(uint32_t* keys, int n_views) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n_views) { keys[2 * v] = 0xFFFFFFFFu; keys[2 * v + 1] = 0u; }
}
__global__ void grad_range_finish_kernel(float* out, int n_views) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_views) return;
  const uint32_t* keys = reinterpret_cast<const uint32_t*>(out);
  const uint32_t kmn = keys[2 * v], kmx = keys[2 * v + 1];
  float mn = key_to_float(kmn), mx = key_to_float(kmx);
  if (kmx == 0xFFFFFFFFu) mn = mx = NAN;       // torch min / max propagate NaN
  if (kmn == 0xFFFFFFFFu && kmx == 0u) mn = mx = 0.f;  // skipped view
  out[2 * v] = mn; out[2 * v + 1] = mx;
}

__global__ void stats_init_kernel(StatsView* ws, int n_views) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n_views) { ws[v].min_key = 0xFFFFFFFFu; ws[v].max_key = 0u; }
}

static unsigned max_chunks(const MdiView* views_host_unused, int64_t total_tiles) {
  // upper bound usable without reading device descriptors: no view has more tiles than the total
  int64_t c = (total_tiles + kStatsTilesPerCta - 1) / kStatsTilesPerCta;
  return (unsigned)(c < 1 ? 1 : c);
}

}  // namespace mdi

extern "C" int64_t mdi_depth_range_workspace_bytes(int n_views) {
  return (int64_t)sizeof(mdi::StatsView) * (n_views > 0 ? n_views : 0);
}

extern "C" int mdi_depth_range(const MdiView* views, int n_views, int64_t max_tiles_per_view,
                               const float* scale_shift, int32_t* status, int flags,
                               void* workspace, float* lo_hi, float* quantiles, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n_views >= 0 && max_tiles_per_view >= 0, "mdi_depth_range: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && workspace && (lo_hi || quantiles), "mdi_depth_range: null pointer");
  MDI_REQUIRE(n_views <= 65535, "mdi_depth_range: more than 65535 views per call");
  cudaStream_t st = (cudaStream_t)stream;
  StatsView* ws = reinterpret_cast<StatsView*>(workspace);
  MDI_CUDA(cudaMemsetAsync(ws, 0, sizeof(StatsView) * (size_t)n_views, st), "mdi_depth_range");
  stats_init_kernel<<<(n_views + 255) / 256, 256, 0, st>>>(ws, n_views);
  static bool attr_done = false;
  if (!attr_done) {
    MDI_CUDA(cudaFuncSetAttribute(stats_pass_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  kTargets * kBins * 4), "mdi_depth_range attr");
    attr_done = true;
  }
  dim3 grid(max_chunks(views, max_tiles_per_view), (unsigned)n_views);
  stats_pass_kernel<0><<<grid, kStatsThreads, kBins * 4, st>>>(views, scale_shift, status, flags, ws);
  stats_select_kernel<0><<<n_views, kStatsThreads, 0, st>>>(status, ws, lo_hi, quantiles);
  stats_pass_kernel<1><<<grid, kStatsThreads, kTargets * kBins * 4, st>>>(views, scale_shift, status, flags, ws);
  stats_select_kernel<1><<<n_views, kStatsThreads, 0, st>>>(status, ws, lo_hi, quantiles);
  stats_pass_kernel<2><<<grid, kStatsThreads, kTargets * 256 * 4, st>>>(views, scale_shift, status, flags, ws);
  stats_select_kernel<2><<<n_views, kStatsThreads, 0, st>>>(status, ws, lo_hi, quantiles);
  MDI_CHECK_LAUNCH("mdi_depth_range");
  return MDI_OK;
}

extern "C" int mdi_grad_range(const MdiView* views, int n_views, int64_t max_tiles_per_view,
                              const float* scale_shift, const int32_t* status, float* grad_range,
                              void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n_views >= 0 && max_tiles_per_view >= 0, "mdi_grad_range: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && grad_range, "mdi_grad_range: null pointer");
  MDI_REQUIRE(n_views <= 65535, "mdi_grad_range: more than 65535 views per call");
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* keys = reinterpret_cast<uint32_t*>(grad_range);
  grad_range_init_kernel<<<(n_views + 255) / 256, 256, 0, st>>>(keys, n_views);
  dim3 grid(max_chunks(views, max_tiles_per_view), (unsigned)n_views);
  grad_range_kernel<<<grid, kStatsThreads, 0, st>>>(views, scale_shift, status, keys);
  grad_range_finish_kernel<<<(n_views + 255) / 256, 256, 0, st>>>(grad_range, n_views);
  MDI_CHECK_LAUNCH("mdi_grad_range");
  return MDI_OK;
}
