// k2_ransac.cu -- LO-RANSAC / MSAC scale-shift fit (a4).  Reference: ransacs.py:100-189.
//
// Every per-hypothesis quantity of the reference loop (sample fit, loss, inlier set, local
// optimisation refit, its loss and inlier count) depends only on the hypothesis' sample
// indices, never on loop state; the loop state is a scalar scan on top.  So: one CTA per
// view, the (d, g) pairs of the view staged once into shared memory with a 1-D bulk (TMA)
// copy completing on an mbarrier, one warp per hypothesis scoring all points from shared
// memory, and after each round of kWarps hypotheses one thread replays the reference's state
// machine (first-wins `<` comparisons, adaptive stopping rule) in iteration order.
// The local-optimisation pass is only run for hypotheses whose sample loss beats the best
// sample loss at the start of the round (a superset of those the sequential loop refits,
// because that bound only decreases).  Issue-bound, not HBM-bound.
#include <type_traits>

#include "common.cuh"

namespace mdi {

constexpr int kRansacWarps = 16;
constexpr int kRansacThreads = kRansacWarps * 32;
constexpr int kHyp = 4;                               // hypotheses scored together by one warp
constexpr int kRound = kRansacWarps * kHyp;           // hypotheses per round
constexpr int kMaxSample = 8;

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

struct Sums { double dd, d, n, dg, g; };

__device__ __forceinline__ void solve(const Sums& s, float* h0, float* h1) {
  solve_normal_eq(s.dd, s.d, s.n, s.dg, s.g, h0, h1);
}

template <bool USE_SMEM>
__global__ void __launch_bounds__(kRansacThreads)
ransac_kernel(const MdiView* __restrict__ views, const float* __restrict__ d_all,
              const float* __restrict__ g_all, const int32_t* __restrict__ n_valid,
              const int32_t* __restrict__ status, const int32_t* __restrict__ sample_idx,
              MdiRansacConfig cfg, int smem_points, float* __restrict__ scale_shift,
              uint8_t* __restrict__ inliers, int32_t* __restrict__ n_iter,
              int32_t* __restrict__ n_inliers) {
  extern __shared__ __align__(16) float s_pts[];  // d[smem_points] then g[smem_points]
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ double s_loss_sample[kRound], s_loss_lo[kRound];
  __shared__ float s_hlo[kRound][2];
  __shared__ int s_ninl_lo[kRound];
  __shared__ double s_best_sample, s_best_lo;
  __shared__ float s_hbest[2];
  __shared__ int s_have, s_done, s_last_it, s_ninl;
  __shared__ double s_req;  // iteration at which the stopping rule fires if the best model stays

  const int v = blockIdx.x;
  const int n = n_valid[v];
  const int64_t base = views[v].sfm_begin;
  const int lane = lane_id(), warp = warp_id();
  const bool skip = status && status[v] != MDI_STATUS_OK;
  const float thr = cfg.inlier_threshold;
  const int ss = cfg.sample_size;
  const bool msac_loss = cfg.loss_kind == MDI_LOSS_MSAC;
  const float* dg = d_all + base;
  const float* gg = g_all + base;

  if (threadIdx.x == 0) {
    s_best_sample = INFINITY; s_best_lo = INFINITY;
    s_hbest[0] = NAN; s_hbest[1] = NAN;
    s_have = 0; s_done = 0; s_last_it = cfg.max_iters - 1; s_ninl = 0; s_req = INFINITY;
  }
  if (skip) {
    if (threadIdx.x == 0) {
      scale_shift[2 * v] = NAN; scale_shift[2 * v + 1] = NAN;
      n_iter[v] = 0; n_inliers[v] = 0;
    }
    return;
  }

  const float* dp = dg;
  const float* gp = gg;
  if (USE_SMEM) {
    float* sd = s_pts;
    float* sg = s_pts + smem_points;
    const uint32_t bytes = (uint32_t)(((n + 3) & ~3) * 4);
    const bool tma_ok = bytes > 0 && ((reinterpret_cast<uintptr_t>(dg) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(gg) & 15) == 0);
    if (tma_ok) {
      if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                     ::"r"(smem_u32(&s_bar)), "r"(2 * bytes) : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
            ::"r"(smem_u32(sd)), "l"(dg), "r"(bytes), "r"(smem_u32(&s_bar)) : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
            ::"r"(smem_u32(sg)), "l"(gg), "r"(bytes), "r"(smem_u32(&s_bar)) : "memory");
      }
      // all threads wait on phase 0
      uint32_t ok = 0;
      while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(smem_u32(&s_bar)), "r"(0u) : "memory");
      }
    } else {
      for (int i = threadIdx.x; i < n; i += blockDim.x) { sd[i] = dg[i]; sg[i] = gg[i]; }
    }
    dp = sd; gp = sg;
  }
  __syncthreads();

  const int n_samp = ss < n ? ss : n;
  // One round = KH hypotheses per warp.  The first round takes one per warp (16 hypotheses): with a
  // healthy inlier ratio the stopping rule fires after a handful of iterations, and a 64-wide
  // round would score 50+ hypotheses nobody asks for; later rounds take 1, 2 or kHyp per warp,
  // whatever covers the iterations the stopping rule still requires.
  auto round = [&](auto kh_c, const int round0) {
    constexpr int KH = decltype(kh_c)::value;
    constexpr int HR = kRansacWarps * KH;
    const int it0 = round0 + warp * KH;   // this warp's hypotheses: it0 .. it0 + KH - 1
    const double cutoff = s_best_sample;
    // ---- sample fits: lane k < KH fits hypothesis it0 + k, then broadcasts it --------------
    float hs0 = 0.f, hs1 = 0.f;
    if (lane < KH && it0 + lane < cfg.max_iters) {
      const int it = it0 + lane;
      Sums sm = {0, 0, 0, 0, 0};
      int chosen[kMaxSample];
      for (int k = 0; k < n_samp; ++k) {
        int idx;
        if (sample_idx) {
          idx = sample_idx[((int64_t)v * cfg.max_iters + it) * ss + k];
        } else {
          // k-th of n_samp distinct uniform indices (randperm(n)[:ss] in distribution)
          uint64_t ctr = 0;
          bool dup;
          do {
            const uint64_t r = splitmix64(cfg.seed ^ splitmix64(((uint64_t)(uint32_t)(v + cfg.view_base) << 40) ^ ((uint64_t)it << 8) ^ (uint64_t)k ^ (ctr << 56)));
            idx = (int)(r % (uint64_t)n);
            dup = false;
            for (int j = 0; j < k; ++j) dup |= (chosen[j] == idx);
            ++ctr;
          } while (dup);
        }
        chosen[k] = idx;
        if (idx >= 0 && idx < n) {
          const double dd = (double)dp[idx], gd = (double)gp[idx];
          sm.dd += dd * dd; sm.d += dd; sm.n += 1.0; sm.dg += dd * gd; sm.g += gd;
        }
      }
      solve(sm, &hs0, &hs1);
    }
    float h0[KH], h1[KH];
    bool act[KH];
#pragma unroll
    for (int k = 0; k < KH; ++k) {
      h0[k] = __shfl_sync(0xffffffffu, hs0, k);
      h1[k] = __shfl_sync(0xffffffffu, hs1, k);
      act[k] = it0 + k < cfg.max_iters;
    }
    // ---- phase A: sample loss of the KH hypotheses; inlier normal equations of the candidates ---
    // Only a hypothesis whose sample loss beats the best one so far is ever refitted, and after
    // the first round that is a few per round: the 5 double-precision sums per (point, hypothesis)
    // are left to a second pass over the candidates alone.  While there is nothing to beat yet
    // (first round) every hypothesis is a candidate and one fused pass does both.
    const bool fused = !(cutoff < (double)INFINITY);
    Sums in[KH];
    int n_out[KH];
    double msac[KH];
#pragma unroll
    for (int k = 0; k < KH; ++k) { in[k] = {0, 0, 0, 0, 0}; n_out[k] = 0; msac[k] = 0.0; }
    if (act[0]) {
      if (fused) {
        for (int i = lane; i < n; i += 32) {
          const float dv = dp[i], gv = gp[i];
          const double dd = (double)dv, gd = (double)gv;
          const double ddd = dd * dd, ddg = dd * gd;
#pragma unroll
          for (int k = 0; k < KH; ++k) {
            const float r = residual_sq(h0[k], h1[k], dv, gv);
            if (r < thr) { in[k].dd += ddd; in[k].d += dd; in[k].n += 1.0; in[k].dg += ddg; in[k].g += gd; }
            if (r >= thr) ++n_out[k];
            if (msac_loss) msac[k] += (r != r) ? (double)NAN : (double)fminf(r, thr);
          }
        }
      } else {
        for (int i = lane; i < n; i += 32) {
          const float dv = dp[i], gv = gp[i];
#pragma unroll
          for (int k = 0; k < KH; ++k) {
            const float r = residual_sq(h0[k], h1[k], dv, gv);
            if (r >= thr) ++n_out[k];
            if (msac_loss) msac[k] += (r != r) ? (double)NAN : (double)fminf(r, thr);
          }
        }
      }
    }
    bool candk[KH];
    bool any_cand = false;
    double loss_s_mine = INFINITY;
#pragma unroll
    for (int k = 0; k < KH; ++k) {
      double ls;
      if (msac_loss) ls = (double)(float)warp_sum_f64(msac[k]);
      else ls = (double)warp_sum_i32(n_out[k]);
      candk[k] = act[k] && ls < cutoff;   // warp-uniform
      any_cand = any_cand || candk[k];
      if (lane == k) loss_s_mine = act[k] ? ls : INFINITY;
    }
    if (!fused && any_cand) {
      for (int i = lane; i < n; i += 32) {
        const float dv = dp[i], gv = gp[i];
        const double dd = (double)dv, gd = (double)gv;
        const double ddd = dd * dd, ddg = dd * gd;
#pragma unroll
        for (int k = 0; k < KH; ++k) {
          if (!candk[k]) continue;
          const float r = residual_sq(h0[k], h1[k], dv, gv);
          if (r < thr) { in[k].dd += ddd; in[k].d += dd; in[k].n += 1.0; in[k].dg += ddg; in[k].g += gd; }
        }
      }
    }
    float hl0 = 0.f, hl1 = 0.f;   // lane k keeps the local-optimisation fit of hypothesis k
#pragma unroll
    for (int k = 0; k < KH; ++k) {
      if (!candk[k]) continue;
      Sums t;
      t.dd = warp_sum_f64(in[k].dd); t.d = warp_sum_f64(in[k].d); t.n = warp_sum_f64(in[k].n);
      t.dg = warp_sum_f64(in[k].dg); t.g = warp_sum_f64(in[k].g);
      if (lane == k) solve(t, &hl0, &hl1);
    }
    if (lane < KH && it0 + lane < cfg.max_iters) {
      s_loss_sample[warp * KH + lane] = loss_s_mine;
      s_hlo[warp * KH + lane][0] = hl0; s_hlo[warp * KH + lane][1] = hl1;
    }
    // ---- phase B: local-optimisation loss of the candidates --------------------------------------
#pragma unroll
    for (int k = 0; k < KH; ++k) {
      h0[k] = __shfl_sync(0xffffffffu, hl0, k);
      h1[k] = __shfl_sync(0xffffffffu, hl1, k);
      n_out[k] = 0; msac[k] = 0.0;
    }
    int n_in[KH];
#pragma unroll
    for (int k = 0; k < KH; ++k) n_in[k] = 0;
    if (any_cand) {
      for (int i = lane; i < n; i += 32) {
        const float dv = dp[i], gv = gp[i];
#pragma unroll
        for (int k = 0; k < KH; ++k) {
          if (!candk[k]) continue;  // warp-uniform
          const float r = residual_sq(h0[k], h1[k], dv, gv);
          if (r < thr) ++n_in[k];
          if (r >= thr) ++n_out[k];
          if (msac_loss) msac[k] += (r != r) ? (double)NAN : (double)fminf(r, thr);
        }
      }
#pragma unroll
      for (int k = 0; k < KH; ++k) {
        if (!candk[k]) continue;
        const int ni = warp_sum_i32(n_in[k]);
        double loss_lo;
        if (msac_loss) loss_lo = (double)(float)warp_sum_f64(msac[k]);
        else loss_lo = (double)warp_sum_i32(n_out[k]);
        if (lane == 0) { s_loss_lo[warp * KH + k] = loss_lo; s_ninl_lo[warp * KH + k] = ni; }
      }
    }
    __syncthreads();
    // ---- replay of the sequential state machine (ransacs.py:141-162) --------------------------
    if (threadIdx.x == 0) {
      // the stopping bound only changes when the best model does
      double req = s_have ? required_samples((int64_t)s_ninl, (int64_t)n, ss, cfg.confidence) : 0.0;
      for (int w = 0; w < HR; ++w) {
        const int itw = round0 + w;
        if (itw >= cfg.max_iters) break;
        if (s_loss_sample[w] < s_best_sample) {
          if (s_loss_lo[w] < s_best_lo) {
            s_hbest[0] = s_hlo[w][0]; s_hbest[1] = s_hlo[w][1];
            s_best_lo = s_loss_lo[w];
            s_best_sample = s_loss_sample[w];
            s_ninl = s_ninl_lo[w];
            s_have = 1;
            req = required_samples((int64_t)s_ninl, (int64_t)n, ss, cfg.confidence);
          }
        }
        if (s_have && itw >= cfg.min_iters && req <= (double)itw) {
          s_done = 1; s_last_it = itw;
          break;
        }
      }
      s_req = s_have ? fmax(req, (double)cfg.min_iters) : INFINITY;
    }
    __syncthreads();
  };
  int round0 = 0;
  if (cfg.max_iters > 0) { round(std::integral_constant<int, 1>{}, 0); round0 = kRansacWarps; }
  while (!s_done && round0 < cfg.max_iters) {
    // size the round to what the stopping rule still asks for with the current best model
    const double need = s_req - (double)round0 + 1.0;
    if (need <= (double)kRansacWarps) {
      round(std::integral_constant<int, 1>{}, round0); round0 += kRansacWarps;
    } else if (need <= (double)(2 * kRansacWarps)) {
      round(std::integral_constant<int, 2>{}, round0); round0 += 2 * kRansacWarps;
    } else {
      round(std::integral_constant<int, kHyp>{}, round0); round0 += kRound;
    }
  }

  const float h0 = s_hbest[0], h1 = s_hbest[1];
  if (inliers)
    for (int i = threadIdx.x; i < n; i += blockDim.x)
      inliers[base + i] = residual_sq(h0, h1, dp[i], gp[i]) < thr ? 1 : 0;
  if (threadIdx.x == 0) {
    scale_shift[2 * v] = h0; scale_shift[2 * v + 1] = h1;
    n_iter[v] = s_last_it; n_inliers[v] = s_ninl;
  }
}

}  // namespace mdi

extern "C" int mdi_ransac(const MdiView* views, int n_views, const float* d, const float* z,
                          const int32_t* n_valid, const int32_t* status, const int32_t* sample_idx,
                          const MdiRansacConfig* cfg, int max_points, float* scale_shift,
                          uint8_t* inliers, int32_t* n_iter, int32_t* n_inliers, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n_views >= 0, "mdi_ransac: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && d && z && n_valid && cfg && scale_shift && n_iter && n_inliers,
              "mdi_ransac: null pointer");
  MDI_REQUIRE(cfg->sample_size >= 1 && cfg->sample_size <= kMaxSample,
              "mdi_ransac: sample_size %d not in 1..%d", cfg->sample_size, kMaxSample);
  MDI_REQUIRE(cfg->max_iters >= 0 && max_points >= 0, "mdi_ransac: negative max_iters/max_points");
  MDI_REQUIRE(cfg->view_base >= 0, "mdi_ransac: negative view_base");
  MDI_REQUIRE(cfg->loss_kind == MDI_LOSS_RANSAC || cfg->loss_kind == MDI_LOSS_MSAC,
              "mdi_ransac: unknown loss kind %d", cfg->loss_kind);
  cudaStream_t st = (cudaStream_t)stream;
  const int pts = (max_points + 3) & ~3;
  const size_t smem = (size_t)pts * 8;
  if (smem <= 200 * 1024) {
    // static + dynamic may not exceed 48 KB by default.  The opt-in is a per-device (context)
    // attribute: set on every call that needs it (microseconds), never cached process-wide
    if (smem > 32 * 1024)
      MDI_CUDA(cudaFuncSetAttribute(ransac_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    200 * 1024), "mdi_ransac attr");
    mdi::prof_mark("ransac_kernel", st);
    ransac_kernel<true><<<n_views, kRansacThreads, smem, st>>>(
        views, d, z, n_valid, status, sample_idx, *cfg, pts, scale_shift, inliers, n_iter, n_inliers);
  } else {
    mdi::prof_mark("ransac_kernel", st);
    ransac_kernel<false><<<n_views, kRansacThreads, 0, st>>>(
        views, d, z, n_valid, status, sample_idx, *cfg, pts, scale_shift, inliers, n_iter, n_inliers);
  }
  MDI_CHECK_LAUNCH("mdi_ransac");
  return MDI_OK;
}
