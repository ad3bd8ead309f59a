// k2_ransac.cu -- LO-RANSAC / MSAC scale-shift fit (a4).  Reference: ransacs.py:100-189.
//
// The reference loop is sequential: draw 4 points, fit, score ALL points; only when the sample
// loss beats the best sample loss so far, refit on the sample model's inliers (local optimisation),
// score again, maybe accept; then test the stopping rule.  The sample fit and its loss depend only
// on the hypothesis' sample indices, never on loop state, so a round scores 16 / 32 / 64 hypotheses
// at once: one CTA per view (two CTAs per SM: a 200-view batch is resident at once), the (d, g)
// pairs of the view staged once into shared memory with a 1-D bulk (TMA) copy completing on an
// mbarrier, one warp per 1, 2 or 4 hypotheses, four points per lane and step (128-bit
// shared-memory loads).  The reference's state machine is then replayed over the round in
// iteration order -- by EVERY thread, redundantly, on registers (warp ballots over the round's
// sample losses find the next hypothesis the reference would refit, the stopping iteration follows
// from the bound in closed form), so the control flow stays uniform across the CTA and nothing is
// broadcast.  Whenever the replay reaches a hypothesis the reference would refit -- sample loss <
// best sample loss AT THAT ITERATION, exactly the reference's test -- the whole CTA does that refit
// together, along with up to three likely followers (later hypotheses of the round on the falling
// chain of sample losses: the ones the loop refits next if this one is accepted): 512 threads
// accumulate the inlier normal equations in fp64 over ~10 points each, a block reduction, one
// solve per candidate (one warp each, with the stopping bound of its inlier count), a second short
// pass for the refit's loss and inlier count.  A view needs ~5 refits in all, so the scoring
// passes are what remains.  Issue-bound, not HBM-bound.
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

constexpr int kLoBatch = 4;                           // refits the CTA does together

template <bool USE_SMEM>
__global__ void __launch_bounds__(kRansacThreads, 2)
ransac_kernel(const MdiView* __restrict__ views, const float* __restrict__ d_all,
              const float* __restrict__ g_all, const int32_t* __restrict__ n_valid,
              const int32_t* __restrict__ status, const int32_t* __restrict__ sample_idx,
              MdiRansacConfig cfg, int smem_points, float* __restrict__ scale_shift,
              uint8_t* __restrict__ inliers, int32_t* __restrict__ n_iter,
              int32_t* __restrict__ n_inliers) {
  extern __shared__ __align__(16) float s_pts[];  // d[smem_points] then g[smem_points]
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ double s_loss_sample[kRound];        // sample loss of the round's hypotheses
  __shared__ float s_hs[kRound][2];               // their sample fits
  __shared__ double s_loss_lo[kRound], s_req_lo[kRound];   // refit loss / stopping bound, where refitted
  __shared__ float s_hlo[kRound][2];
  __shared__ int s_ninl_lo[kRound];
  __shared__ double s_red[kLoBatch][kRansacWarps][6];   // per-warp partial sums of a refit batch
  __shared__ int s_redi[kLoBatch][kRansacWarps][2];

  const int v = blockIdx.x;
  const int n = n_valid[v];
  const int64_t base = views[v].sfm_begin;
  const int tid = threadIdx.x, lane = lane_id(), warp = warp_id();
  const bool skip = status && status[v] != MDI_STATUS_OK;
  const float thr = cfg.inlier_threshold;
  const int ss = cfg.sample_size;
  const bool msac_loss = cfg.loss_kind == MDI_LOSS_MSAC;
  const float* dg = d_all + base;
  const float* gg = g_all + base;

  if (skip) {
    if (tid == 0) {
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
      if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      __syncthreads();
      if (tid == 0) {
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
      for (int i = tid; i < n; i += blockDim.x) { sd[i] = dg[i]; sg[i] = gg[i]; }
    }
    dp = sd; gp = sg;
  }
  __syncthreads();
  // four points per lane and step where both arrays can be read as 16-byte vectors (always in
  // shared memory; the global-memory fallback of very large views only if the slots happen to be aligned)
  const bool vec = ((reinterpret_cast<uintptr_t>(dp) | reinterpret_cast<uintptr_t>(gp)) & 15) == 0;

  const int n_samp = ss < n ? ss : n;
  // ---- scoring pass of one round: KH hypotheses per warp --------------------------------------
  // The first round takes one per warp (16 hypotheses): with a healthy inlier ratio the stopping
  // rule fires after a handful of iterations, and a 64-wide round would score 50+ hypotheses
  // nobody asks for; later rounds take 1, 2 or kHyp per warp, whatever covers the iterations the
  // stopping rule still requires.
  auto score_round = [&](auto kh_c, const int round0) {
    constexpr int KH = decltype(kh_c)::value;
    const int it0 = round0 + warp * KH;   // this warp's hypotheses: it0 .. it0 + KH - 1
    // sample fits: lane k < KH fits hypothesis it0 + k, then broadcasts it
    float hs0 = 0.f, hs1 = 0.f;
    if (lane < KH && it0 + lane < cfg.max_iters) {
      const int it = it0 + lane;
      Sums sm = {0, 0, 0, 0, 0};
      // k-th of n_samp distinct uniform indices (randperm(n)[:ss] in distribution): draw number
      // `ctr` for slot k of hypothesis `it` of batch view v; a draw that repeats an earlier slot
      // is redrawn with the next ctr
      auto draw = [&](const int k, const uint64_t ctr) {
        const uint64_t r = splitmix64(cfg.seed ^ splitmix64(((uint64_t)(uint32_t)(v + cfg.view_base) << 40) ^ ((uint64_t)it << 8) ^ (uint64_t)k ^ (ctr << 56)));
        return (int)(((r >> 32) * (uint64_t)(uint32_t)n) >> 32);   // uniform in 0 .. n-1 (multiply-high)
      };
      auto add_point = [&](const int idx) {
        if (idx >= 0 && idx < n) {
          const double dd = (double)dp[idx], gd = (double)gp[idx];
          sm.dd += dd * dd; sm.d += dd; sm.n += 1.0; sm.dg += dd * gd; sm.g += gd;
        }
      };
      bool drawn = false;
      if (!sample_idx && n_samp == 4) {
        // the usual case: four independent hash chains, and almost never a repeated index
        const int i0 = draw(0, 0), i1 = draw(1, 0), i2 = draw(2, 0), i3 = draw(3, 0);
        if (i0 != i1 && i0 != i2 && i0 != i3 && i1 != i2 && i1 != i3 && i2 != i3) {
          const float d0 = dp[i0], d1 = dp[i1], d2 = dp[i2], d3 = dp[i3];
          const float g0 = gp[i0], g1 = gp[i1], g2 = gp[i2], g3 = gp[i3];
          const double e0 = d0, e1 = d1, e2 = d2, e3 = d3, f0 = g0, f1 = g1, f2 = g2, f3 = g3;
          // same order of additions as the general loop below
          sm.dd = ((e0 * e0 + e1 * e1) + e2 * e2) + e3 * e3;
          sm.d = ((e0 + e1) + e2) + e3;
          sm.n = 4.0;
          sm.dg = ((e0 * f0 + e1 * f1) + e2 * f2) + e3 * f3;
          sm.g = ((f0 + f1) + f2) + f3;
          drawn = true;
        }
      }
      if (!drawn) {
        int chosen[kMaxSample];
        for (int k = 0; k < n_samp; ++k) {
          int idx;
          if (sample_idx) {
            idx = sample_idx[((int64_t)v * cfg.max_iters + it) * ss + k];
          } else {
            uint64_t ctr = 0;
            bool dup;
            do {
              idx = draw(k, ctr);
              dup = false;
              for (int j = 0; j < k; ++j) dup |= (chosen[j] == idx);
              ++ctr;
            } while (dup);
          }
          chosen[k] = idx;
          add_point(idx);
        }
      }
      solve(sm, &hs0, &hs1);
      s_hs[warp * KH + lane][0] = hs0; s_hs[warp * KH + lane][1] = hs1;
    }
    float h0[KH], h1[KH];
    bool act[KH];
#pragma unroll
    for (int k = 0; k < KH; ++k) {
      h0[k] = __shfl_sync(0xffffffffu, hs0, k);
      h1[k] = __shfl_sync(0xffffffffu, hs1, k);
      act[k] = it0 + k < cfg.max_iters;
    }
    int n_out[KH];
    double msac[KH];
#pragma unroll
    for (int k = 0; k < KH; ++k) { n_out[k] = 0; msac[k] = 0.0; }
    auto score_point = [&](const float dv, const float gv) {
#pragma unroll
      for (int k = 0; k < KH; ++k) {
        const float r = residual_sq(h0[k], h1[k], dv, gv);
        if (r >= thr) ++n_out[k];
        if (msac_loss) msac[k] += (r != r) ? (double)NAN : (double)fminf(r, thr);
      }
    };
    if (act[0]) {
      if (vec) {
        const int n4 = n & ~3;
        for (int i0 = 4 * lane; i0 < n4; i0 += 128) {
          const float4 d4 = *reinterpret_cast<const float4*>(dp + i0);
          const float4 g4 = *reinterpret_cast<const float4*>(gp + i0);
          score_point(d4.x, g4.x); score_point(d4.y, g4.y); score_point(d4.z, g4.z); score_point(d4.w, g4.w);
        }
        if (n4 + lane < n) score_point(dp[n4 + lane], gp[n4 + lane]);   // last n % 4 points
      } else {
        for (int i = lane; i < n; i += 32) score_point(dp[i], gp[i]);
      }
    }
#pragma unroll
    for (int k = 0; k < KH; ++k) {
      double ls;
      if (msac_loss) ls = (double)(float)warp_sum_f64(msac[k]);
      else ls = (double)warp_sum_i32(n_out[k]);
      if (lane == 0) s_loss_sample[warp * KH + k] = act[k] ? ls : (double)INFINITY;
    }
  };

  // ---- local optimisation of hypotheses b[0..nb) of the round, by the whole CTA (ransacs.py:142-146):
  // least squares over the inliers of the sample model, then loss / inlier count of the refit.
  // b / nb are the same in every thread; four barriers, all in uniform control flow ------------
  auto refit_batch = [&](const int (&bt)[kLoBatch], const int nb) {
#pragma unroll 1
    for (int c = 0; c < nb; ++c) {
      const int w = bt[c];
      const float a0 = s_hs[w][0], a1 = s_hs[w][1];
      Sums in = {0, 0, 0, 0, 0};
      for (int i = tid; i < n; i += kRansacThreads) {
        const float dv = dp[i], gv = gp[i];
        if (residual_sq(a0, a1, dv, gv) < thr) {
          const double dd = (double)dv, gd = (double)gv;
          in.dd += dd * dd; in.d += dd; in.n += 1.0; in.dg += dd * gd; in.g += gd;
        }
      }
      in.dd = warp_sum_f64(in.dd); in.d = warp_sum_f64(in.d); in.n = warp_sum_f64(in.n);
      in.dg = warp_sum_f64(in.dg); in.g = warp_sum_f64(in.g);
      if (lane == 0) {
        double* o = s_red[c][warp];
        o[0] = in.dd; o[1] = in.d; o[2] = in.n; o[3] = in.dg; o[4] = in.g;
      }
    }
    __syncthreads();
    if (warp < nb) {   // warp c adds up candidate c's partial sums (fixed order) and solves, all lanes alike
      const bool has = lane < kRansacWarps;
      Sums t;
      t.dd = warp_sum_f64(has ? s_red[warp][lane][0] : 0.0);
      t.d = warp_sum_f64(has ? s_red[warp][lane][1] : 0.0);
      t.n = warp_sum_f64(has ? s_red[warp][lane][2] : 0.0);
      t.dg = warp_sum_f64(has ? s_red[warp][lane][3] : 0.0);
      t.g = warp_sum_f64(has ? s_red[warp][lane][4] : 0.0);
      float b0, b1;
      solve(t, &b0, &b1);
      if (lane == 0) { s_hlo[bt[warp]][0] = b0; s_hlo[bt[warp]][1] = b1; }
    }
    __syncthreads();
#pragma unroll 1
    for (int c = 0; c < nb; ++c) {
      const int w = bt[c];
      const float b0 = s_hlo[w][0], b1 = s_hlo[w][1];
      int n_in = 0, n_out = 0;
      double ms = 0.0;
      for (int i = tid; i < n; i += kRansacThreads) {
        const float r = residual_sq(b0, b1, dp[i], gp[i]);
        if (r < thr) ++n_in;
        if (r >= thr) ++n_out;
        if (msac_loss) ms += (r != r) ? (double)NAN : (double)fminf(r, thr);
      }
      n_in = warp_sum_i32(n_in); n_out = warp_sum_i32(n_out);
      if (msac_loss) ms = warp_sum_f64(ms);
      if (lane == 0) { s_redi[c][warp][0] = n_in; s_redi[c][warp][1] = n_out; s_red[c][warp][5] = ms; }
    }
    __syncthreads();
    if (warp < nb) {
      const bool has = lane < kRansacWarps;
      const int ni = warp_sum_i32(has ? s_redi[warp][lane][0] : 0);
      const int no = warp_sum_i32(has ? s_redi[warp][lane][1] : 0);
      double ms = 0.0;
      if (msac_loss) ms = warp_sum_f64(has ? s_red[warp][lane][5] : 0.0);
      const double rq = required_samples((int64_t)ni, (int64_t)n, ss, cfg.confidence);
      if (lane == 0) {
        const int w = bt[warp];
        s_loss_lo[w] = msac_loss ? (double)(float)ms : (double)no;
        s_ninl_lo[w] = ni;
        s_req_lo[w] = rq;
      }
    }
    __syncthreads();
  };

  // ---- state of the reference's loop (ransacs.py:120-162), the same in every thread ------------
  double best_sample = INFINITY, best_lo = INFINITY;
  double req = 0.0;             // required_samples() of the current best model
  float hb0 = NAN, hb1 = NAN;
  int have = 0, ninl = 0, last_it = cfg.max_iters - 1;
  bool done = false;

  int round0 = 0;
  bool first = true;
  while (!done && round0 < cfg.max_iters) {
    // size the round to what the stopping rule still asks for with the current best model (the
    // first round: one hypothesis per warp)
    const double stop_at = have ? fmax(req, (double)cfg.min_iters) : (double)INFINITY;
    const double need = first ? 0.0 : stop_at - (double)round0 + 1.0;
    first = false;
    int hr;
    if (need <= (double)kRansacWarps) {
      score_round(std::integral_constant<int, 1>{}, round0); hr = kRansacWarps;
    } else if (need <= (double)(2 * kRansacWarps)) {
      score_round(std::integral_constant<int, 2>{}, round0); hr = 2 * kRansacWarps;
    } else {
      score_round(std::integral_constant<int, kHyp>{}, round0); hr = kRound;
    }
    __syncthreads();
    // ---- replay of the sequential loop over the round's hr hypotheses (ransacs.py:141-162) ----
    // lane l holds the sample losses of hypotheses l and l + 32 (+inf beyond the round / max_iters)
    const double ls_a = lane < hr ? s_loss_sample[lane] : (double)INFINITY;
    const double ls_b = lane + 32 < hr ? s_loss_sample[lane + 32] : (double)INFINITY;
    auto loss_of = [&](const int w) {   // sample loss of hypothesis w of the round, w uniform
      const double x = (w & 32) ? ls_b : ls_a;
      return __hiloint2double(__shfl_sync(0xffffffffu, __double2hiint(x), w & 31),
                              __shfl_sync(0xffffffffu, __double2loint(x), w & 31));
    };
    auto below = [&](const double bound, const int from) {   // hypotheses >= from with sample loss < bound
      const uint64_t m = ((uint64_t)__ballot_sync(0xffffffffu, ls_b < bound) << 32) | __ballot_sync(0xffffffffu, ls_a < bound);
      return from >= 64 ? 0ull : (m & (~0ull << from));
    };
    uint64_t lo_mask = 0;   // hypotheses of this round whose refit is in s_*_lo
    int pos = 0;
#pragma unroll 1
    for (;;) {
      // next hypothesis the reference refits, and the first iteration at which its stopping rule
      // fires while the best model stays what it is: `req <= it` is monotone in it
      const uint64_t cand = below(best_sample, pos);
      const int w_c = cand ? __ffsll((long long)cand) - 1 : hr;
      int w_s = hr;
      if (have && req == req) {
        const double t = fmax(fmax(ceil(req), (double)cfg.min_iters) - (double)round0, (double)pos);
        if (t < (double)hr) w_s = (int)t;
      }
      if (w_s < w_c) {   // the loop stops before it refits again (iterations past max_iters do not exist)
        if (round0 + w_s < cfg.max_iters) { done = true; last_it = round0 + w_s; }
        break;
      }
      if (w_c >= hr) break;
      if (!((lo_mask >> w_c) & 1ull)) {
        // the whole CTA refits it, with the later hypotheses of the round that the loop refits too if
        // this one is accepted (falling chain of sample losses)
        int bt[kLoBatch];
        int nb = 1;
        bt[0] = w_c;
#pragma unroll
        for (int c = 1; c < kLoBatch; ++c) bt[c] = 0;
        double m = loss_of(w_c);
        int last = w_c;
#pragma unroll 1
        for (int tries = 0; tries < 8 && nb < kLoBatch; ++tries) {
          const uint64_t f = below(m, last + 1);
          if (!f) break;
          last = __ffsll((long long)f) - 1;
          m = loss_of(last);
          if (!((lo_mask >> last) & 1ull)) {
#pragma unroll
            for (int c = 1; c < kLoBatch; ++c) if (c == nb) bt[c] = last;
            ++nb;
          }
        }
        refit_batch(bt, nb);
#pragma unroll
        for (int c = 0; c < kLoBatch; ++c) if (c < nb) lo_mask |= 1ull << bt[c];
      }
      const int itw = round0 + w_c;
      if (s_loss_lo[w_c] < best_lo) {
        hb0 = s_hlo[w_c][0]; hb1 = s_hlo[w_c][1];
        best_lo = s_loss_lo[w_c];
        best_sample = loss_of(w_c);
        ninl = s_ninl_lo[w_c];
        have = 1;
        req = s_req_lo[w_c];
      }
      if (have && itw >= cfg.min_iters && req <= (double)itw) { done = true; last_it = itw; break; }
      pos = w_c + 1;
    }
    round0 += hr;
    __syncthreads();   // the next round rewrites s_loss_sample / s_hs
  }

  if (inliers)
    for (int i = tid; i < n; i += blockDim.x)
      inliers[base + i] = residual_sq(hb0, hb1, dp[i], gp[i]) < thr ? 1 : 0;
  if (tid == 0) {
    scale_shift[2 * v] = hb0; scale_shift[2 * v + 1] = hb1;
    n_iter[v] = last_it; n_inliers[v] = ninl;
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
