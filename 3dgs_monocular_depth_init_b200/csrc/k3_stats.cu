// k3_stats.cu -- per-view statistics of the aligned depth for adaptive subsampling (a7): masked
// min / max and the exact order statistics behind torch.quantile(x, 0.25 / 0.75)
// (adaptive_subsampling.py:82-95).  Also the gradient-map range for a9.
//
// Fast path ("sample-select", ONE streaming pass over depth + mask, HBM-bound):
//   1. stats_sample_kernel   reads one hashed 128-byte line (32 pixels) out of every 16: HW/16 keys
//   2. stats_bracket_kernel  radix-selects, in that sample, key brackets [L, U] around each
//                            quantile wide enough to hold the true order statistics
//   3. stats_main_kernel     the one full pass: min / max / count, the number of keys below each
//                            bracket, a 2048-bin histogram of the keys inside each bracket, and
//                            those keys (a few %) appended to a per-view candidate list through a
//                            shared-memory stage
//   4. stats_finish_kernel   rank - #below located in the bracket histogram, the candidates of
//                            that one bin gathered into shared memory and ranked exactly, then
//                            torch.quantile's fp32 rank / lerp arithmetic
// The result is exact whenever step 4 verifies; if a bracket missed (or a list overflowed, e.g. a
// constant depth map) the view is flagged MDI_STATUS_STATS_RETRY and the host re-runs it through
// the exact path below.
//
// Exact path (three streaming passes): 12/12/8-bit radix select on order-preserving keys over
// all pixels, shared-memory histograms, one-CTA-per-view bin search after each pass.
#include <stdlib.h>
#include <string.h>

#include <type_traits>

#include "common.cuh"

namespace mdi {

constexpr int kStatsThreads = 256;
constexpr int kStatsTilesPerCta = 16;  // 64K pixels per CTA amortise histogram zero/flush
constexpr int kBins = 4096;
constexpr int kTargets = 4;            // q1.lo, q1.hi, q3.lo, q3.hi
constexpr int kSampleGroup = 128;      // one 8-pixel sector sampled per group of 128 pixels
constexpr int kBrBins = 2048;          // histogram bins per quantile bracket
constexpr int kListCap = 2048;         // keys of one bracket bin ranked in shared memory
constexpr int kSelThreads = 1024;
constexpr int kSelBins = 2048;

struct StatsHdr {
  // sample-select path
  uint32_t n_sample, n_cand;
  uint32_t nmin_key, max_key, count;   // nmin_key = ~(smallest key): a zero-filled header needs no init kernel
  uint32_t below[2];
  uint32_t br_lo[2], br_hi[2], br_shift[2];
  uint32_t run_exact;  // this view goes through the exact path in this call
  // exact path
  uint32_t x_min_key, x_max_key, x_count, n_slots;
  uint32_t slot_prefix[kTargets];  // prefix each histogram slot selects on
  uint32_t tgt_prefix[kTargets];
  uint32_t tgt_slot[kTargets];
  uint32_t smp_nmin, smp_max, fail_code;  // smp_nmin = ~(smallest sample key); fail_code: why the one-pass path asked for a retry
  long long tgt_rank[kTargets];
  float w[2];
  uint32_t gmin_key, gmax_key;
  // depth-domain form of the brackets (stats_bracket_kernel -> stats_main_kernel): with a finite
  // scale > 0 and a finite shift, d -> fl(fl(d * s) + t) is monotone, so the tests on the aligned key
  // become float comparisons on the raw depth
  uint32_t fast;        // 1: d_lo / d_hi / ds_* are valid and the view takes the depth-domain loop
  uint32_t force_retry; // set by the depth-domain loop: NaN / inf in the population or a pixel that may hit the sentinel
  float d_lo[2], d_hi[2];
  float ds_lo, ds_hi;   // depths whose aligned value is exactly -42.0 (empty when ds_lo > ds_hi)
  // MDI_STATS_RAW: what stats_finish_kernel found on the raw depth, for stats_map_kernel
  uint32_t raw_state;   // kRaw*
  float raw_val[kTargets], raw_min, raw_max;
};
constexpr uint32_t kRawNone = 0, kRawOk = 1, kRawEmpty = 2, kRawFail = 3, kRawNan = 4;

struct StatsWs {
  StatsHdr* hdr;       // [V]
  uint32_t* hist;      // [V][kTargets][kBins]   (exact path)
  uint32_t* sample;    // [V][s_cap]
  uint32_t* cand;      // [V][c_cap]
  int64_t s_cap, c_cap;
};

static int64_t sample_cap(int64_t max_tiles) {
  return max_tiles * (MDI_TILE_PX / kSampleGroup) * 8;
}
static int64_t cand_cap(int64_t max_tiles) {
  const int64_t px = max_tiles * MDI_TILE_PX;
  int64_t c = px / 4;
  const int64_t floor_c = px < 16384 ? px : 16384;
  return c > floor_c ? c : floor_c;
}
static StatsWs carve(void* workspace, int n_views, int64_t max_tiles) {
  StatsWs w;
  char* p = reinterpret_cast<char*>(workspace);
  w.hdr = reinterpret_cast<StatsHdr*>(p);
  p += ((sizeof(StatsHdr) * (size_t)n_views + 255) / 256) * 256;
  w.hist = reinterpret_cast<uint32_t*>(p);
  p += sizeof(uint32_t) * kTargets * kBins * (size_t)n_views;
  w.s_cap = sample_cap(max_tiles);
  w.c_cap = cand_cap(max_tiles);
  w.sample = reinterpret_cast<uint32_t*>(p);
  p += sizeof(uint32_t) * (size_t)w.s_cap * n_views;
  w.cand = reinterpret_cast<uint32_t*>(p);
  return w;
}

__device__ __forceinline__ bool in_population(uint32_t maskbyte, float a, int flags) {
  bool ok = true;
  if (flags & MDI_KEEP_PRED_MASK) ok = maskbyte != 0;
  if (flags & MDI_KEEP_SENTINEL) ok = ok && (a != -42.0f);
  return ok;
}

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

// order-preserving key without the NaN special case: NaNs land above key(+inf) / below key(-inf),
// which is all the one-pass path needs (their presence makes the result NaN; stats_finish_kernel)
__device__ __forceinline__ uint32_t fast_key(float f) {
  const uint32_t u = __float_as_uint(f);
  return u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);
}
constexpr uint32_t kKeyPosInf = 0xFF800000u, kKeyNegInf = 0x007FFFFFu;

// ---- generic block-wide radix select over a key list -------------------------------------------
// Among keys[0..n) with flo[j] <= key <= fhi[j], finds the key of (0-based, ascending) rank
// rank[j], for j < kTargets.  11/11/10-bit passes, histogram updates aggregated per warp with
// match.any (depth keys cluster in a handful of top-digit bins).  Returns a bit mask of the
// targets whose rank lies outside their filtered population.  All kSelThreads threads call it.
struct SelShared {
  uint32_t hist[kTargets][kSelBins];
  int scan[33];
  uint32_t pre[kTargets];
  uint32_t bin[kTargets];
  long long rank[kTargets];
  long long newrank[kTargets];
  uint32_t fail;
};

__device__ uint32_t block_select(const uint32_t* __restrict__ keys, uint32_t n,
                                 const uint32_t* flo, const uint32_t* fhi, const long long* rank,
                                 uint32_t* out, SelShared& sh) {
  const int tid = threadIdx.x, lane = tid & 31;
  uint32_t lo[kTargets], hi[kTargets];
#pragma unroll
  for (int j = 0; j < kTargets; ++j) { lo[j] = flo[j]; hi[j] = fhi[j]; }
  if (tid < kTargets) { sh.pre[tid] = 0; sh.rank[tid] = rank[tid]; }
  if (tid == 0) sh.fail = 0;
  __syncthreads();
#pragma unroll 1
  for (int pass = 0; pass < 3; ++pass) {
    const int pshift = pass == 0 ? 32 : (pass == 1 ? 21 : 10);   // bits above the digit
    const int dshift = pass == 0 ? 21 : (pass == 1 ? 10 : 0);
    const uint32_t dmask = pass == 2 ? 0x3ffu : 0x7ffu;
    for (int i = tid; i < kTargets * kSelBins; i += kSelThreads) (&sh.hist[0][0])[i] = 0;
    uint32_t pre[kTargets];
    int slot[kTargets];
#pragma unroll
    for (int j = 0; j < kTargets; ++j) {
      pre[j] = sh.pre[j];
      slot[j] = j;
      for (int i = 0; i < j; ++i)
        if (slot[j] == j && lo[i] == lo[j] && hi[i] == hi[j] && pre[i] == pre[j]) slot[j] = i;
    }
    __syncthreads();
    for (uint32_t base = 0; base < n; base += kSelThreads) {
      const uint32_t i = base + tid;
      const bool in = i < n;
      const uint32_t key = in ? keys[i] : 0u;
      const uint32_t top = pass == 0 ? 0u : (key >> pshift);
      const uint32_t b = (key >> dshift) & dmask;
#pragma unroll
      for (int j = 0; j < kTargets; ++j) {
        if (slot[j] != j) continue;  // uniform
        const bool p = in && key >= lo[j] && key <= hi[j] && top == pre[j];
        const uint32_t act = __ballot_sync(0xffffffffu, p);
        if (p) {
          const uint32_t peers = __match_any_sync(act, b);
          if (lane == __ffs(peers) - 1) atomicAdd(&sh.hist[j][b], (uint32_t)__popc(peers));
        }
      }
    }
    __syncthreads();
    // locate each rank in its histogram: 2 bins per thread
#pragma unroll 1
    for (int j = 0; j < kTargets; ++j) {
      const uint32_t* h = sh.hist[slot[j]];
      const uint32_t c0 = h[2 * tid], c1 = h[2 * tid + 1];
      int tot;
      const long long cum = block_excl_scan((int)(c0 + c1), sh.scan, &tot);
      const long long r = sh.rank[j];
      if (r >= cum && r < cum + c0) { sh.bin[j] = 2 * tid; sh.newrank[j] = r - cum; }
      else if (r >= cum + c0 && r < cum + c0 + c1) { sh.bin[j] = 2 * tid + 1; sh.newrank[j] = r - cum - c0; }
      if (tid == 0 && (r < 0 || r >= tot)) { atomicOr(&sh.fail, 1u << j); sh.bin[j] = 0; sh.newrank[j] = 0; }
      __syncthreads();
    }
    if (tid < kTargets) {
      const int bits = pass == 2 ? 10 : 11;
      sh.pre[tid] = (sh.pre[tid] << bits) | sh.bin[tid];
      sh.rank[tid] = sh.newrank[tid];
    }
    __syncthreads();
  }
  if (tid < kTargets) out[tid] = sh.pre[tid];
  __syncthreads();
  return sh.fail;
}

// ---- sample-select path --------------------------------------------------------------------------
__global__ void stats_init_kernel(StatsHdr* hdr, int n_views, const int32_t* status, int mode) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_views) return;
  hdr[v].x_min_key = 0xFFFFFFFFu;
  // mode 0: sample-select for everybody; 1: exact for everybody; 2: exact for views flagged RETRY
  const bool retry = status && (status[v] & MDI_STATUS_STATS_RETRY);
  hdr[v].run_exact = mode == 1 ? 1u : (mode == 2 ? (retry ? 1u : 0u) : 0u);
}

// One thread per group of 128 pixels: a hashed 8-pixel sector of the group is sampled (32 B of
// depth, 8 B of mask), so the sample is spread over the whole image at sector granularity.
__global__ void __launch_bounds__(256)
stats_sample_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                    const int32_t* __restrict__ status, int flags, StatsWs ws) {
  const int v = blockIdx.y;
  const MdiView* vp = views + v;
  const int64_t HW = (int64_t)vp->H * vp->W;
  const int64_t group = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if ((int64_t)blockIdx.x * blockDim.x * kSampleGroup >= HW) return;
  if (status && (status[v] & MDI_STATUS_LOW_CONFIDENCE)) return;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const float* __restrict__ depth = vp->depth;
  const uint8_t* __restrict__ pmask = vp->mask;
  // Four consecutive groups (512 pixels) sample the four sectors of ONE hashed 128-byte depth line
  // of the 16 they span: DRAM serves 128-byte lines, so a lone 32-byte sector per group costs a
  // whole line of depth and a whole line of mask (4x the traffic for the same sample size)
  const int64_t super = group >> 2;
  const uint32_t hsh = hash32((uint32_t)super * 2654435761u + (uint32_t)v);
  const int64_t p0 = super * (4 * kSampleGroup) + (int64_t)(hsh & 15u) * 32 + (group & 3) * 8;
  uint32_t keys[8];
  uint32_t valid = 0;  // bit k: keys[k] is in the population
  if (p0 < HW) {
    float dv[8];
    uint32_t mlo = 0x01010101u, mhi = 0x01010101u;  // mask bytes of pixels 0-3 / 4-7
    uint32_t inside = 0xFFu;
    if (p0 + 8 <= HW && (reinterpret_cast<uintptr_t>(depth) & 15) == 0 &&
        (!pmask || (reinterpret_cast<uintptr_t>(pmask) & 3) == 0)) {
      const float4 x = __ldg(reinterpret_cast<const float4*>(depth + p0));
      const float4 y = __ldg(reinterpret_cast<const float4*>(depth + p0 + 4));
      if (pmask) {
        mlo = __ldg(reinterpret_cast<const uint32_t*>(pmask + p0));
        mhi = __ldg(reinterpret_cast<const uint32_t*>(pmask + p0 + 4));
      }
      dv[0] = x.x; dv[1] = x.y; dv[2] = x.z; dv[3] = x.w; dv[4] = y.x; dv[5] = y.y; dv[6] = y.z; dv[7] = y.w;
    } else {
      inside = 0; mlo = 0; mhi = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const bool in = p0 + k < HW;
        dv[k] = in ? __ldg(&depth[p0 + k]) : 0.f;
        const uint32_t mb = in ? (pmask ? (__ldg(&pmask[p0 + k]) ? 1u : 0u) : 1u) : 0u;
        if (k < 4) mlo |= mb << (8 * k); else mhi |= mb << (8 * (k - 4));
        inside |= (in ? 1u : 0u) << k;
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float a = align_depth(dv[k], s, t);
      const uint32_t mb = ((k < 4 ? mlo : mhi) >> (8 * (k & 3))) & 0xffu;
      keys[k] = fast_key(a);
      if (((inside >> k) & 1u) && in_population(mb, a, flags)) valid |= 1u << k;
    }
  }
  uint32_t kmin = 0xFFFFFFFFu, kmax = 0u;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if ((valid >> k) & 1u) { kmin = min(kmin, keys[k]); kmax = max(kmax, keys[k]); }
  // warp-aggregated append
  const int c = __popc(valid);
  const int incl = warp_incl_scan(c);
  const int tot = __shfl_sync(0xffffffffu, incl, 31);
  if (tot) {
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    uint32_t base = 0;
    if (lane_id() == 31) {
      base = atomicAdd(&ws.hdr[v].n_sample, (uint32_t)tot);
      atomicMax(&ws.hdr[v].smp_nmin, ~kmin);
      atomicMax(&ws.hdr[v].smp_max, kmax);
    }
    base = __shfl_sync(0xffffffffu, base, 31);
    uint32_t* dst = ws.sample + (int64_t)v * ws.s_cap + base + (incl - c);
    int o = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if ((valid >> k) & 1u) dst[o++] = keys[k];
  }
}

// Brackets from the sample, one CTA per view.  A 1024-key sub-sample is sorted in shared memory
// to get a robust range [R_lo, R_hi] ~ (q 0.10, q 0.90) of the keys; one pass over the sample
// then builds a 2048-bin linear histogram of that range (+ the counts on either side), and the
// bracket of each quantile is the span of bins between its lower and upper sample ranks.
constexpr int kBracketThreads = 1024;
__global__ void __launch_bounds__(kBracketThreads)
stats_bracket_kernel(const float* __restrict__ scale_shift, const int32_t* __restrict__ status, int sentinel, StatsWs ws) {
  __shared__ uint32_t s_LU[kTargets];      // L0, U0, L1, U1
  __shared__ uint32_t s_dkey[kTargets + 2]; // depth-domain bounds as order keys (+ sentinel interval)
  __shared__ int s_fast;
  __shared__ uint32_t s_sort[1024];
  __shared__ uint32_t s_hist[kBrBins];
  __shared__ uint16_t s_guide[kBrBins + 2];
  __shared__ int scan_smem[33];
  __shared__ uint32_t s_bin[kTargets];
  __shared__ long long s_rank[kTargets];
  const int v = blockIdx.x, tid = threadIdx.x;
  StatsHdr* h = ws.hdr + v;
  if (status && (status[v] & MDI_STATUS_LOW_CONFIDENCE)) return;
  const uint32_t ns = h->n_sample;
  if (ns < 4096) {  // too small to bracket: everything is a candidate (small images fit the list)
    if (tid == 0)
      for (int g = 0; g < 2; ++g) { h->br_lo[g] = 0; h->br_hi[g] = 0xFFFFFFFFu; h->br_shift[g] = 21; }
    return;
  }
  const uint32_t* keys = ws.sample + (int64_t)v * ws.s_cap;
  // bitonic sort of an evenly spaced 1024-key sub-sample
  s_sort[tid] = keys[(uint64_t)tid * ns / 1024];
  for (int i = tid; i < kBrBins; i += kBracketThreads) s_hist[i] = 0;
  __syncthreads();
  for (int k = 2; k <= 1024; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int ixj = tid ^ j;
      if (ixj > tid) {
        const uint32_t x = s_sort[tid], y = s_sort[ixj];
        const bool up = (tid & k) == 0;
        if ((x > y) == up) { s_sort[tid] = y; s_sort[ixj] = x; }
      }
      __syncthreads();
    }
  }
  // equal-mass bins: bin(key) = number of sorted sub-sample keys < key, 0 .. 1024.  (A histogram
  // that is linear in the key is useless when the values straddle zero, where keys are log-spaced.)
  // A 2048-entry guide over the robust range [sort[102], sort[921]] narrows each lower-bound
  // search to a step or two in the common case.
  const uint32_t Rlo = s_sort[102], Rhi = s_sort[921];
  uint32_t gs = 0;
  while (((Rhi - Rlo) >> gs) >= (uint32_t)kBrBins) ++gs;
  auto lower_bound = [&](uint32_t key, int lo, int hi) {
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (s_sort[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
  };
  for (int g = tid; g <= kBrBins; g += kBracketThreads) {
    const unsigned long long e = (unsigned long long)Rlo + ((unsigned long long)g << gs);
    s_guide[g] = (uint16_t)(e > (unsigned long long)Rhi ? lower_bound(Rhi, 0, 1024) + 0 : lower_bound((uint32_t)e, 0, 1024));
  }
  __syncthreads();
  for (uint32_t base = 0; base < ns; base += 4 * kBracketThreads) {
    uint32_t kk[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t i = base + u * kBracketThreads + tid;
      kk[u] = i < ns ? keys[i] : 0xFFFFFFFFu;
    }
    // Only the robust range is binned finely: the bracket ranks (q -/+ delta) sit well inside it.
    // Keys below it are counted into bin 102 (which then holds everything <= sort[102]), keys
    // above it need no count at all (bin 1024, "above the last edge", is what a rank falls into
    // when no bin claims it); a 7-step search for either would make the whole warp wait.
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const bool in = base + u * kBracketThreads + tid < ns;
      const uint32_t key = kk[u];
      const unsigned below = __ballot_sync(0xffffffffu, in && key < Rlo);
      if (below && lane_id() == 0) atomicAdd(&s_hist[102], (uint32_t)__popc(below));
      if (in && !(key < Rlo) && !(key > Rhi)) {
        const uint32_t g = (key - Rlo) >> gs;
        atomicAdd(&s_hist[lower_bound(key, s_guide[g], s_guide[g + 1])], 1u);
      }
    }
  }
  // half-width of the rank bracket: the sample is ns / 8 sectors of 8 adjacent (correlated) pixels,
  // so 4 sigma of a quantile estimated from ns / 8 draws, + 0.25 % (one sector per group is a
  // stratified sample: on C2 the true rank sits within 0.05 % of the bracket centre; a miss only
  // costs the exact-path retry)
  const double delta = 0.0025 + 4.0 * sqrt(0.1875 * 8.0 / (double)ns);
  if (tid == 0) {
    for (int g = 0; g < 2; ++g) {
      const double q = g == 0 ? 0.25 : 0.75;
      long long rl = (long long)floor((q - delta) * (double)ns);
      long long rh = (long long)ceil((q + delta) * (double)ns);
      s_rank[2 * g] = rl < 0 ? 0 : rl;
      s_rank[2 * g + 1] = rh > (long long)ns - 1 ? (long long)ns - 1 : rh;
    }
  }
  __syncthreads();
  // bin of each bracket rank (one bin per thread; bin 1024 = whatever lies above the last key)
  const uint32_t c0 = s_hist[tid];
  int tot;
  const long long cum = block_excl_scan((int)c0, scan_smem, &tot);
  if (tid < kTargets) s_bin[tid] = 1024u;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kTargets; ++j) {
    const long long r = s_rank[j];
    if (r >= cum && r < cum + c0) s_bin[j] = tid;
  }
  __syncthreads();
  if (tid == 0) {
    for (int g = 0; g < 2; ++g) {
      const uint32_t bl = s_bin[2 * g], bh = s_bin[2 * g + 1];
      // keys of bin b satisfy sort[b-1] < key <= sort[b]
      // (merged ends: bin 102 reaches down to the smallest key, bin 1024 down to sort[921])
      const uint32_t below_bl = bl >= 1024u ? s_sort[921] : (bl <= 102u ? 0xFFFFFFFFu : s_sort[bl - 1]);
      uint32_t L = bl <= 102u ? 0u : below_bl + (below_bl == 0xFFFFFFFFu ? 0u : 1u);
      uint32_t U = bh >= 1024u ? 0xFFFFFFFFu : s_sort[bh < 102u ? 102u : bh];
      if (L < kKeyNegInf) L = kKeyNegInf;
      if (U > kKeyPosInf) U = kKeyPosInf;
      s_LU[2 * g] = L; s_LU[2 * g + 1] = U >= L ? U : L;
    }
    s_fast = 1;
  }
  __syncthreads();
  // Depth-domain form of the brackets.  F(d) = key(fl(fl(d * s) + t)) is monotone non-decreasing in
  // the order key of d when s is finite and > 0 and t is finite, so {key(aligned) >= L} = {d >= D_lo}
  // with D_lo the smallest depth whose aligned key reaches L (binary search over the order keys of
  // all non-NaN floats, 32 steps), and likewise for U and for the sentinel value -42.0.
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const bool mono = s > 0.f && s < INFINITY && fabsf(t) < INFINITY;
  if (tid < kTargets + 2 && mono) {
    const uint32_t kSent = fast_key(-42.0f);
    const uint32_t bound = tid < kTargets ? s_LU[tid] : kSent;
    auto F = [&](uint32_t kd) { return fast_key(align_depth(key_to_float(kd), s, t)); };
    uint32_t lo = kKeyNegInf, hi = kKeyPosInf;
    // the answer is usually within a few floats of (bound - t) / s: try a 33-float window first
    const float guess = fdiv(fsub(key_to_float(bound), t), s);
    uint32_t wlo = kKeyNegInf, whi = kKeyPosInf;
    if (guess == guess) {
      const uint32_t kg = fast_key(guess);
      if (kg > kKeyNegInf + 16u && kg < kKeyPosInf - 16u) { wlo = kg - 16u; whi = kg + 16u; }
    }
    if ((tid & 1) == 0) {  // smallest kd with F(kd) >= bound
      if (wlo != kKeyNegInf && F(whi) >= bound && F(wlo) < bound) { lo = wlo + 1u; hi = whi; }
      if (F(hi) < bound) s_fast = 0;
      while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if (F(mid) >= bound) hi = mid; else lo = mid + 1; }
      if (tid < kTargets && lo == 0x80000000u) lo = 0x7FFFFFFFu;  // +0 -> -0: a float test cannot tell them apart
    } else {               // largest kd with F(kd) <= bound
      if (wlo != kKeyNegInf && F(wlo) <= bound && F(whi) > bound) { lo = wlo; hi = whi - 1u; }
      if (F(lo) > bound) s_fast = 0;
      while (lo < hi) { const uint32_t mid = lo + ((hi - lo + 1) >> 1); if (F(mid) <= bound) lo = mid; else hi = mid - 1; }
      if (tid < kTargets && lo == 0x7FFFFFFFu) lo = 0x80000000u;  // -0 -> +0
    }
    s_dkey[tid] = lo;
    if (tid < kTargets) {
      // the depth bounds were only ever widened: make the key bracket cover what they let in
      const uint32_t fk = F(lo);
      if ((tid & 1) == 0) { if (fk < s_LU[tid]) s_LU[tid] = fk; }
      else if (fk > s_LU[tid]) s_LU[tid] = fk;
    }
  }
  __syncthreads();
  if (tid == 0) {
    for (int g = 0; g < 2; ++g) {
      uint32_t L = s_LU[2 * g], U = s_LU[2 * g + 1];
      // stats_main_tma_kernel tests the bracket with float comparisons, which agree with the key
      // order for every non-NaN value as long as the bounds are finite-or-inf keys and L != key(+0),
      // U != key(-0) (the one place where the two orders differ).  Widening a bracket is always safe.
      if (L == 0x80000000u) L = 0x7FFFFFFFu;       // key(+0) -> key(-0)
      if (U == 0x7FFFFFFFu) U = 0x80000000u;       // key(-0) -> key(+0)
      uint32_t bs = 0;
      while (((U - L) >> bs) >= (uint32_t)kBrBins) ++bs;
      h->br_lo[g] = L; h->br_hi[g] = U >= L ? U : L; h->br_shift[g] = bs;
      if (mono) { h->d_lo[g] = key_to_float(s_dkey[2 * g]); h->d_hi[g] = key_to_float(s_dkey[2 * g + 1]); }
    }
    if (mono) { h->ds_lo = key_to_float(s_dkey[kTargets]); h->ds_hi = key_to_float(s_dkey[kTargets + 1]); }
    // 1: no pixel of the SAMPLE comes near the sentinel -- the loop only certifies that none does
    //    (thread-level range check, else force_retry); 2: the loop tests every pixel against ds_*
    const uint32_t kSent = fast_key(-42.0f);
    const bool far = !sentinel || kSent < ~h->smp_nmin || kSent > h->smp_max;
    h->fast = (mono && s_fast) ? (far ? 1u : 2u) : 0u;
  }
}

constexpr int kWarpStage = 768;  // keys staged per warp; flushed once fewer than 512 slots remain

template <bool kUseMask, bool kSentinel>
__global__ void __launch_bounds__(kStatsThreads, 5)  // 48 registers: five CTAs per SM (41 KB of shared memory each)
stats_main_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                  const int32_t* __restrict__ status, StatsWs ws, int prefetch_ahead) {
  __shared__ uint32_t s_stage[kStatsThreads / 32][kWarpStage];
  __shared__ uint32_t s_brhist[2][kBrBins];
  __shared__ uint32_t s_min, s_max, s_cnt, s_below[2];
  const int v = blockIdx.y;
  const MdiView* vp = views + v;
  const int first_tile = blockIdx.x * kStatsTilesPerCta;
  if (first_tile >= vp->n_tiles) return;
  if (status && (status[v] & MDI_STATUS_LOW_CONFIDENCE)) return;
  StatsHdr* h = ws.hdr + v;
  const uint32_t L0 = h->br_lo[0], L1 = h->br_lo[1];
  const uint32_t W0 = h->br_hi[0] - L0, W1 = h->br_hi[1] - L1;  // bracket widths (unsigned)
  const uint32_t sh0 = h->br_shift[0], sh1 = h->br_shift[1];
  for (int i = threadIdx.x; i < 2 * kBrBins; i += kStatsThreads) (&s_brhist[0][0])[i] = 0;
  if (threadIdx.x == 0) { s_min = 0xFFFFFFFFu; s_max = 0u; s_cnt = 0u; s_below[0] = 0; s_below[1] = 0; }
  __syncthreads();
  const float* __restrict__ depth = vp->depth;
  const uint8_t* __restrict__ pmask = kUseMask ? vp->mask : nullptr;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const int64_t HW = (int64_t)vp->H * vp->W;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(depth) & 15) == 0) &&
                      (!pmask || (reinterpret_cast<uintptr_t>(pmask) & 3) == 0);
  uint32_t kmin = 0xFFFFFFFFu, kmax = 0u, cnt = 0, below0 = 0, below1 = 0;
  uint32_t* cand = ws.cand + (int64_t)v * ws.c_cap;
  const int lane = lane_id();
  uint32_t* stage = s_stage[warp_id()];
  uint32_t wcount = 0;  // keys in this warp's stage (warp-uniform)
  const uint32_t lt = (1u << lane) - 1u;

  // straight-line per-pixel work, called by all 32 lanes together
  auto feed = [&](float dv, bool pop) {
    const float a = align_depth(dv, s, t);
    if (kSentinel) pop = pop && (a != -42.0f);
    const uint32_t key = fast_key(a);
    kmin = min(kmin, pop ? key : 0xFFFFFFFFu);
    kmax = max(kmax, pop ? key : 0u);
    cnt += pop ? 1u : 0u;
    below0 += (pop && key < L0) ? 1u : 0u;
    below1 += (pop && key < L1) ? 1u : 0u;
    const bool c = pop && ((key - L0) <= W0 || (key - L1) <= W1);
    const uint32_t m = __ballot_sync(0xffffffffu, c);
    if (c) stage[wcount + __popc(m & lt)] = key;
    wcount += __popc(m);
  };
  // staged keys -> bracket histograms (all lanes busy: only candidates get here) + candidate list
  auto flush = [&]() {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(&h->n_cand, wcount);
    base = __shfl_sync(0xffffffffu, base, 0);
    __syncwarp();
    for (uint32_t i = lane; i < wcount; i += 32) {
      const uint32_t key = stage[i];
      const uint32_t o0 = key - L0, o1 = key - L1;
      if (o0 <= W0) atomicAdd(&s_brhist[0][o0 >> sh0], 1u);
      if (o1 <= W1) atomicAdd(&s_brhist[1][o1 >> sh1], 1u);
      if ((int64_t)base + i < ws.c_cap) cand[base + i] = key;
    }
    __syncwarp();
    wcount = 0;
  };

  const int last_tile = min(first_tile + kStatsTilesPerCta, (int)vp->n_tiles);
  if (h->fast) {
    // Depth-domain loop (stats_bracket_kernel): the bracket tests are float comparisons on the raw
    // depth, a pixel outside the population becomes NaN (fails every comparison, skipped by
    // fminf / fmaxf), the stage holds raw depths and flush_raw() aligns them.  NaN / inf in the
    // population and pixels that may equal the sentinel after alignment are only DETECTED here;
    // such a view goes through the exact path (force_retry).
    const float D0lo = h->d_lo[0], D0hi = h->d_hi[0], D1lo = h->d_lo[1], D1hi = h->d_hi[1];
    float fmn = INFINITY, fmx = -INFINITY, nanacc = 0.f;
    constexpr int kColSlotsF = kWarpStage / 32;  // 24 candidates per lane between flushes
    const uint32_t sp0 = (uint32_t)__cvta_generic_to_shared(stage + lane);
    uint32_t sp = sp0;
    const uint32_t* col = stage + lane;
    const float Slo = h->ds_lo, Shi = h->ds_hi;
    auto feedf = [&](float dv, bool pop, auto count_tag, auto sent_tag) {
      constexpr bool kCount = decltype(count_tag)::value;
      if (decltype(sent_tag)::value) {  // exact sentinel test: aligned == -42.0 <=> Slo <= depth <= Shi
        const bool hit = pop & (dv >= Slo) & (dv <= Shi);
        if (!kCount) cnt -= hit ? 1u : 0u;  // (the caller counted the mask bytes)
        pop = pop & !hit;
      }
      const float x = pop ? dv : __uint_as_float(0x7FC00000u);  // returned: the caller folds it into min / max
      if (kCount) cnt += pop ? 1u : 0u;
      if (pop) nanacc = __fmaf_rn(dv, 0.f, nanacc);
      const bool lt0 = x < D0lo, lt1 = x < D1lo;
      below0 += lt0 ? 1u : 0u;
      below1 += lt1 ? 1u : 0u;
      // no short circuit: a NaN fails the <= tests, so !lt alone letting it through is harmless
      const bool c = (!lt0 & (x <= D0hi)) | (!lt1 & (x <= D1hi));
      // candidates go to the lane's own column of the warp's stage: no vote, no prefix
      // (one 32-bit shared-window address: a generic pointer costs four predicated adds per pixel)
      if (c) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(sp), "f"(dv) : "memory"); sp += 128; }
      return x;
    };
    // three-input min / max (FMNMX3, NaN operands skipped like fminf / fmaxf): two pixels per instruction
    auto min3 = [](float a, float b, float c) { float d; asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; };
    auto max3 = [](float a, float b, float c) { float d; asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; };
    const uint32_t cap32 = (uint32_t)(ws.c_cap < (int64_t)0x7fffffff ? ws.c_cap : (int64_t)0x7fffffff);
    auto flush_raw = [&]() {
      const uint32_t n = (sp - sp0) >> 7;  // this lane's candidates
      const uint32_t maxn = __reduce_max_sync(0xffffffffu, n);
      const uint32_t tot = __reduce_add_sync(0xffffffffu, n);
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(&h->n_cand, tot);
      base = __shfl_sync(0xffffffffu, base, 0);
      for (uint32_t j = 0; j < maxn; ++j) {  // row by row: coalesced appends (each lane reads its own column)
        const bool has = j < n;
        const uint32_t m = __ballot_sync(0xffffffffu, has);
        if (has) {
          const uint32_t key = fast_key(align_depth(__uint_as_float(col[j * 32]), s, t));
          const uint32_t o0 = key - L0, o1 = key - L1;
          if (o0 <= W0) atomicAdd(&s_brhist[0][o0 >> sh0], 1u);
          if (o1 <= W1) atomicAdd(&s_brhist[1][o1 >> sh1], 1u);
          const uint32_t at = base + __popc(m & lt);
          if (at < cap32) cand[at] = key;
        }
        base += __popc(m);
      }
      sp = sp0;
    };
    // L2 prefetch two tiles ahead (one 128-byte line per thread: 128 lines of depth + 32 of mask per
    // tile): the loop holds only one tile of loads per warp in registers, the prefetches keep more
    // DRAM requests in flight without registers, and the register loads then hit L2
    auto prefetch_tile = [&](int tile) {
      if (tile >= last_tile) return;
      const int64_t px0 = (int64_t)tile * MDI_TILE_PX;
      const int t = threadIdx.x;
      if (t < 128) {
        const int64_t p = px0 + t * 32;
        if (p < HW) asm volatile("prefetch.global.L2 [%0];" ::"l"(depth + p));
      } else if (t < 160 && pmask) {
        const int64_t p = px0 + (t - 128) * 128;
        if (p < HW) asm volatile("prefetch.global.L2 [%0];" ::"l"(pmask + p));
      }
    };
    auto stream_tiles = [&](auto sent_tag) {
#pragma unroll 1
      for (int tile = first_tile; tile < last_tile; ++tile) {
        const int64_t px0 = (int64_t)tile * MDI_TILE_PX;
        prefetch_tile(tile + prefetch_ahead);
        if (vec_ok && px0 + MDI_TILE_PX <= HW) {
          // 4 x (128-bit depth + 32-bit mask) loads in flight per thread, then the arithmetic
          float4 d4[4];
          uint32_t mk[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int64_t p = px0 + (int64_t)(k * kStatsThreads + threadIdx.x) * 4;
            d4[k] = __ldg(reinterpret_cast<const float4*>(depth + p));
            mk[k] = pmask ? __ldg(reinterpret_cast<const uint32_t*>(pmask + p)) : 0x01010101u;
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // bit 7 of every non-zero mask byte: the population of 4 pixels counted by one popc
            const uint32_t nz = (mk[k] | ((mk[k] & 0x7f7f7f7fu) + 0x7f7f7f7fu)) & 0x80808080u;
            cnt += __popc(nz);
            const float x0 = feedf(d4[k].x, (nz & 0x80u) != 0, std::false_type{}, sent_tag);
            const float x1 = feedf(d4[k].y, (nz & 0x8000u) != 0, std::false_type{}, sent_tag);
            const float x2 = feedf(d4[k].z, (nz & 0x800000u) != 0, std::false_type{}, sent_tag);
            const float x3 = feedf(d4[k].w, (nz & 0x80000000u) != 0, std::false_type{}, sent_tag);
            fmn = min3(min3(fmn, x0, x1), x2, x3);
            fmx = max3(max3(fmx, x0, x1), x2, x3);
          }
        } else {
          for (int k = threadIdx.x; k < MDI_TILE_PX; k += kStatsThreads) {  // uniform trip count
            const int64_t p = px0 + k;
            const bool in = p < HW;
            const float x = feedf(in ? __ldg(&depth[p]) : 0.f, in && (pmask ? __ldg(&pmask[p]) != 0 : true), std::true_type{}, sent_tag);
            fmn = fminf(fmn, x);
            fmx = fmaxf(fmx, x);
          }
        }
        // a tile adds at most 16 candidates per lane
        if (__any_sync(0xffffffffu, sp - sp0 > (uint32_t)((kColSlotsF - 16) * 128))) flush_raw();
      }
    };
    for (int a = 0; a < prefetch_ahead; ++a) prefetch_tile(first_tile + a);
    const bool per_pixel_sentinel = kSentinel && h->fast == 2u;
    if (per_pixel_sentinel) stream_tiles(std::true_type{}); else stream_tiles(std::false_type{});
    if (__any_sync(0xffffffffu, sp != sp0)) flush_raw();
    // thread-level detection of what this loop does not handle
    bool bad = !(fabsf(nanacc) == 0.f) || (cnt && (fmn == -INFINITY || fmx == INFINITY));
    if (kSentinel && !per_pixel_sentinel) bad = bad || (cnt && Slo <= Shi && fmn <= Shi && fmx >= Slo);
    if (bad) h->force_retry = 1u;
    // monotone alignment: the extreme aligned values are those of the extreme depths
    kmin = fast_key(align_depth(fmn, s, t));
    kmax = fast_key(align_depth(fmx, s, t));
  } else {
#pragma unroll 1
  for (int tile = first_tile; tile < last_tile; ++tile) {
    const int64_t px0 = (int64_t)tile * MDI_TILE_PX;
    if (vec_ok && px0 + MDI_TILE_PX <= HW) {
      // 4 x (128-bit depth + 32-bit mask) loads in flight per thread, then the arithmetic
      float4 d4[4];
      uint32_t mk[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t p = px0 + (int64_t)(k * kStatsThreads + threadIdx.x) * 4;
        d4[k] = __ldg(reinterpret_cast<const float4*>(depth + p));
        mk[k] = pmask ? __ldg(reinterpret_cast<const uint32_t*>(pmask + p)) : 0x01010101u;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        feed(d4[k].x, (mk[k] & 0xffu) != 0); feed(d4[k].y, (mk[k] & 0xff00u) != 0);
        feed(d4[k].z, (mk[k] & 0xff0000u) != 0); feed(d4[k].w, (mk[k] & 0xff000000u) != 0);
      }
    } else {
      for (int k = threadIdx.x; k < MDI_TILE_PX; k += kStatsThreads) {  // uniform trip count
        const int64_t p = px0 + k;
        const bool in = p < HW;
        feed(in ? __ldg(&depth[p]) : 0.f, in && (pmask ? __ldg(&pmask[p]) != 0 : true));
      }
    }
    if (wcount > (uint32_t)(kWarpStage - 512)) flush();  // a tile adds at most 512 keys per warp
  }
  if (wcount) flush();
  }
  kmin = __reduce_min_sync(0xffffffffu, kmin);
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  below0 = __reduce_add_sync(0xffffffffu, below0);
  below1 = __reduce_add_sync(0xffffffffu, below1);
  if (lane == 0 && cnt) {
    atomicMin(&s_min, kmin); atomicMax(&s_max, kmax); atomicAdd(&s_cnt, cnt);
    atomicAdd(&s_below[0], below0); atomicAdd(&s_below[1], below1);
  }
  __syncthreads();
  if (threadIdx.x == 0 && s_cnt) {
    atomicMax(&h->nmin_key, ~s_min); atomicMax(&h->max_key, s_max); atomicAdd(&h->count, s_cnt);
    atomicAdd(&h->below[0], s_below[0]); atomicAdd(&h->below[1], s_below[1]);
  }
  uint32_t* ghist = ws.hist + (int64_t)v * kTargets * kBins;  // [2][kBrBins] of it
  for (int i = threadIdx.x; i < 2 * kBrBins; i += kStatsThreads) {
    const uint32_t c = (&s_brhist[0][0])[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

// ---- the same pass with the depth staged by TMA (experimental, opt-in) ---------------------------
// The register-staged kernel above leaves the load schedule to ptxas (which sinks loads next to
// their uses as soon as the arithmetic gets leaner) and pays ~34 instructions per pixel.  Here every
// warp streams its share of the CTA's depth through its own ring of kTmaStages shared-memory
// chunks with 1-D bulk copies (cp.async.bulk completing on an mbarrier; lane 0 issues, the warp
// waits), three chunks always in flight and no CTA-wide barrier in the loop; the predicted
// mask (whose per-view base is only 4-byte aligned in a stacked [V,H,W] tensor, too little for a
// bulk copy) is read one chunk ahead into registers.  The per-pixel work is done in the float
// domain -- a pixel outside the population becomes NaN, which fails every comparison and is
// skipped by fminf / fmaxf; the bracket bounds are kept off +/-0 by stats_bracket_kernel so that
// float and key comparisons agree -- and every lane stages its candidates in its own
// shared-memory column: no vote, no prefix.
constexpr int kTmaChunk = 256;    // pixels per staged chunk: one warp's step (1 KB of depth)
constexpr int kTmaStages = 4;     // chunks in flight per warp
constexpr int kColSlots = 24;     // candidate slots per lane
constexpr int kTmaSmemBytes = (kStatsThreads / 32) * kTmaStages * kTmaChunk * 4 + (kStatsThreads / 32) * kColSlots * 32 * 4 + 2 * kBrBins * 4;

template <bool kUseMask, bool kSentinel>
__global__ void __launch_bounds__(kStatsThreads)
stats_main_tma_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                      const int32_t* __restrict__ status, StatsWs ws) {
  extern __shared__ __align__(128) unsigned char s_raw_tma[];
  constexpr int kWarps = kStatsThreads / 32;
  float* s_depth = reinterpret_cast<float*>(s_raw_tma);                                            // [warps][stages][chunk]
  uint32_t* s_col = reinterpret_cast<uint32_t*>(s_raw_tma + kWarps * kTmaStages * kTmaChunk * 4); // [warps][slots][32]
  uint32_t* s_brhist = s_col + kWarps * kColSlots * 32;                                            // [2][kBrBins]
  __shared__ __align__(8) unsigned long long s_full[kWarps][kTmaStages];
  __shared__ uint32_t s_min, s_max, s_cnt, s_below[2];
  const int v = blockIdx.y;
  const MdiView* vp = views + v;
  const int first_tile = blockIdx.x * kStatsTilesPerCta;
  if (first_tile >= vp->n_tiles) return;
  if (status && (status[v] & MDI_STATUS_LOW_CONFIDENCE)) return;
  StatsHdr* h = ws.hdr + v;
  const uint32_t L0 = h->br_lo[0], L1 = h->br_lo[1];
  const uint32_t W0 = h->br_hi[0] - L0, W1 = h->br_hi[1] - L1;
  const uint32_t sh0 = h->br_shift[0], sh1 = h->br_shift[1];
  const float fL0 = key_to_float(L0), fU0 = key_to_float(h->br_hi[0]);
  const float fL1 = key_to_float(L1), fU1 = key_to_float(h->br_hi[1]);
  const float* __restrict__ depth = vp->depth;
  const uint8_t* __restrict__ pmask = kUseMask ? vp->mask : nullptr;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const int64_t HW = (int64_t)vp->H * vp->W;
  const int last_tile = min(first_tile + kStatsTilesPerCta, (int)vp->n_tiles);
  const int64_t px_lo = (int64_t)first_tile * MDI_TILE_PX;
  const int64_t px_hi = min((int64_t)last_tile * MDI_TILE_PX, HW);
  const bool staged = ((reinterpret_cast<uintptr_t>(depth) & 15) == 0) &&
                      (!pmask || (reinterpret_cast<uintptr_t>(pmask) & 3) == 0);
  const int n_chunks = staged ? (int)((px_hi - px_lo) / kTmaChunk) : 0;  // whole chunks of the CTA
  const int tid = threadIdx.x, lane = lane_id(), warp = warp_id();
  // warp w takes chunks w, w + kWarps, ...
  const int n_mine = n_chunks > warp ? (n_chunks - warp + kWarps - 1) / kWarps : 0;

  for (int i = tid; i < 2 * kBrBins; i += kStatsThreads) s_brhist[i] = 0;
  if (tid == 0) {
    s_min = 0xFFFFFFFFu; s_max = 0u; s_cnt = 0u; s_below[0] = 0; s_below[1] = 0;
    for (int i = 0; i < kWarps * kTmaStages; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_full[0][0] + i)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  float* my_depth = s_depth + warp * kTmaStages * kTmaChunk;
  auto chunk_px = [&](int i) { return px_lo + (int64_t)(i * kWarps + warp) * kTmaChunk; };
  auto issue = [&](int i) {  // lane 0: the warp's i-th chunk -> stage i % kTmaStages
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_full[warp][i % kTmaStages]);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(my_depth + (i % kTmaStages) * kTmaChunk);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kTmaChunk * 4) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(depth + chunk_px(i)), "r"(kTmaChunk * 4), "r"(bar) : "memory");
  };
  if (lane == 0)
    for (int i = 0; i < min(kTmaStages, n_mine); ++i) issue(i);

  uint32_t cnt = 0, below0 = 0, below1 = 0, lcount = 0;
  float vmin = INFINITY, vmax = -INFINITY;
  bool any_nan = false;
  uint32_t* cand = ws.cand + (int64_t)v * ws.c_cap;
  uint32_t* col = s_col + (warp * kColSlots) * 32 + lane;  // slot j of this lane: col[32 * j]

  auto feed = [&](float a, bool pop) {
    if (kSentinel) pop = pop && (a != -42.0f);
    const float g = pop ? a : __int_as_float(0x7fc00000);
    vmin = fminf(vmin, g);
    vmax = fmaxf(vmax, g);
    cnt += pop ? 1u : 0u;
    const bool lt0 = g < fL0, lt1 = g < fL1;
    below0 += lt0 ? 1u : 0u;
    below1 += lt1 ? 1u : 0u;
    if ((!lt0 && g <= fU0) || (!lt1 && g <= fU1)) { col[32 * lcount] = fast_key(a); ++lcount; }
  };
  auto flush = [&]() {  // lane columns -> bracket histograms + the view's candidate list
    const uint32_t total = __reduce_add_sync(0xffffffffu, lcount);
    if (total == 0) return;
    const uint32_t incl = (uint32_t)warp_incl_scan((int)lcount);
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(&h->n_cand, total);
    base = __shfl_sync(0xffffffffu, base, 0) + incl - lcount;
    for (uint32_t j = 0; j < lcount; ++j) {
      const uint32_t key = col[32 * j];
      const uint32_t o0 = key - L0, o1 = key - L1;
      if (o0 <= W0) atomicAdd(&s_brhist[o0 >> sh0], 1u);
      if (o1 <= W1) atomicAdd(&s_brhist[kBrBins + (o1 >> sh1)], 1u);
      if ((int64_t)base + j < ws.c_cap) cand[base + j] = key;
    }
    __syncwarp();
    lcount = 0;
  };

  // predicted mask of the lane's 8 pixels of a chunk: two 32-bit words, read one chunk ahead
  auto load_mask = [&](int i, uint32_t (&mk)[2]) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
      mk[k] = pmask ? __ldg(reinterpret_cast<const uint32_t*>(pmask + chunk_px(i) + (k * 32 + lane) * 4)) : 0x01010101u;
  };
  uint32_t mk_cur[2] = {0, 0}, mk_next[2] = {0, 0};
  if (n_mine > 0) load_mask(0, mk_cur);
#pragma unroll 1
  for (int i = 0; i < n_mine; ++i) {
    if (i + 1 < n_mine) load_mask(i + 1, mk_next);
    {  // wait for the chunk's bytes
      const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_full[warp][i % kTmaStages]);
      const uint32_t parity = (uint32_t)((i / kTmaStages) & 1);
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    }
    const float4* sd = reinterpret_cast<const float4*>(my_depth + (i % kTmaStages) * kTmaChunk);
    float a[8];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float4 d4 = sd[k * 32 + lane];
      a[4 * k] = align_depth(d4.x, s, t); a[4 * k + 1] = align_depth(d4.y, s, t);
      a[4 * k + 2] = align_depth(d4.z, s, t); a[4 * k + 3] = align_depth(d4.w, s, t);
    }
    const float chk = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
    if (chk != chk) {  // some NaN (or inf - inf): look at the population exactly (rare, rolled)
#pragma unroll 1
      for (int j = 0; j < 8; ++j) {
        const int64_t p = chunk_px(i) + ((j >> 2) * 32 + lane) * 4 + (j & 3);
        const float av = align_depth(__ldg(&depth[p]), s, t);
        if ((pmask ? __ldg(&pmask[p]) != 0 : true) && av != av) any_nan = true;
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      feed(a[4 * k], (mk_cur[k] & 0xffu) != 0); feed(a[4 * k + 1], (mk_cur[k] & 0xff00u) != 0);
      feed(a[4 * k + 2], (mk_cur[k] & 0xff0000u) != 0); feed(a[4 * k + 3], (mk_cur[k] & 0xff000000u) != 0);
    }
    mk_cur[0] = mk_next[0]; mk_cur[1] = mk_next[1];
    __syncwarp();  // every lane is done with this stage: refill it
    if (lane == 0 && i + kTmaStages < n_mine) issue(i + kTmaStages);
    if (__any_sync(0xffffffffu, lcount + 8u > (uint32_t)kColSlots)) flush();  // a chunk adds <= 8 per lane
  }
  // whatever is not a whole staged chunk (ragged end of the view, or a misaligned view)
  for (int64_t p0 = px_lo + (int64_t)n_chunks * kTmaChunk; p0 < px_hi; p0 += kStatsThreads) {  // uniform trips
    const int64_t p = p0 + tid;
    const bool in = p < px_hi;
    const float av = align_depth(in ? __ldg(&depth[p]) : 0.f, s, t);
    const bool pop = in && (pmask ? __ldg(&pmask[p]) != 0 : true);
    feed(av, pop);
    if (pop && av != av) any_nan = true;
    if (__any_sync(0xffffffffu, lcount + 8u > (uint32_t)kColSlots)) flush();
  }
  flush();
  uint32_t kmin = cnt ? fast_key(vmin) : 0xFFFFFFFFu;
  uint32_t kmax = any_nan ? 0xFFFFFFFFu : (cnt ? fast_key(vmax) : 0u);
  kmin = __reduce_min_sync(0xffffffffu, kmin);
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  below0 = __reduce_add_sync(0xffffffffu, below0);
  below1 = __reduce_add_sync(0xffffffffu, below1);
  if (lane == 0 && cnt) {
    atomicMin(&s_min, kmin); atomicMax(&s_max, kmax); atomicAdd(&s_cnt, cnt);
    atomicAdd(&s_below[0], below0); atomicAdd(&s_below[1], below1);
  }
  __syncthreads();
  if (tid == 0 && s_cnt) {
    atomicMax(&h->nmin_key, ~s_min); atomicMax(&h->max_key, s_max); atomicAdd(&h->count, s_cnt);
    atomicAdd(&h->below[0], s_below[0]); atomicAdd(&h->below[1], s_below[1]);
  }
  uint32_t* ghist = ws.hist + (int64_t)v * kTargets * kBins;  // [2][kBrBins] of it
  for (int i = tid; i < 2 * kBrBins; i += kStatsThreads) {
    const uint32_t cbin = s_brhist[i];
    if (cbin) atomicAdd(&ghist[i], cbin);
  }
}

// quantiles -> iqr range, written to the outputs (shared by both paths)
__device__ __forceinline__ void write_range(int v, const float val[kTargets], float w1, float w3,
                                            uint32_t min_key, uint32_t max_key,
                                            float* __restrict__ lo_hi, float* __restrict__ quantiles) {
  const bool has_nan = max_key == 0xFFFFFFFFu;
  float q1 = torch_lerp(val[0], val[1], w1);
  float q3 = torch_lerp(val[2], val[3], w3);
  float mn = key_to_float(min_key), mx = key_to_float(max_key);
  if (has_nan) { q1 = q3 = mn = mx = NAN; }
  float lo, hi;
  iqr_range(q1, q3, mn, mx, &lo, &hi);
  if (lo_hi) { lo_hi[2 * v] = lo; lo_hi[2 * v + 1] = hi; }
  if (quantiles) { quantiles[4 * v] = q1; quantiles[4 * v + 1] = q3; quantiles[4 * v + 2] = mn; quantiles[4 * v + 3] = mx; }
}

__global__ void __launch_bounds__(kSelThreads, 2)
stats_finish_kernel(int32_t* __restrict__ status, StatsWs ws, float* __restrict__ lo_hi,
                    float* __restrict__ quantiles, int64_t* __restrict__ retry_flag, int raw) {
  // the in-bin lists and the radix-select fallback share one buffer
  __shared__ __align__(16) unsigned char s_raw[sizeof(SelShared) > sizeof(uint32_t) * kTargets * kListCap
                                                   ? sizeof(SelShared) : sizeof(uint32_t) * kTargets * kListCap];
  uint32_t (*s_list)[kListCap] = reinterpret_cast<uint32_t (*)[kListCap]>(s_raw);
  __shared__ int scan_smem[33];
  __shared__ uint32_t s_len[kTargets], s_bin[kTargets], s_out[kTargets], s_fail;
  __shared__ uint32_t s_flo[kTargets], s_fhi[kTargets];
  __shared__ long long s_rank[kTargets], s_rank0[kTargets];
  const int v = blockIdx.x, tid = threadIdx.x;
  StatsHdr* h = ws.hdr + v;
  const bool low_conf = status && (status[v] & MDI_STATUS_LOW_CONFIDENCE);
  const uint32_t n = h->count;
  if (raw && n == 0) {   // MDI_STATS_RAW: the verdict is stats_map_kernel's
    if (tid == 0) h->raw_state = kRawEmpty;
    return;
  }
  if (low_conf || n == 0) {
    if (tid == 0) {
      if (lo_hi) { lo_hi[2 * v] = 0.f; lo_hi[2 * v + 1] = 0.f; }
      if (quantiles) for (int k = 0; k < 4; ++k) quantiles[4 * v + k] = 0.f;
      if (!low_conf && status) status[v] |= MDI_STATUS_EMPTY_STATS;
    }
    return;
  }
  const QuantileRank r1 = quantile_rank(0.25f, (int64_t)n);
  const QuantileRank r3 = quantile_rank(0.75f, (int64_t)n);
  const uint32_t n_cand = h->n_cand;
  const uint32_t min_key = ~h->nmin_key;
  const bool has_nan = h->max_key > kKeyPosInf || min_key < kKeyNegInf;
  const uint32_t L[2] = {h->br_lo[0], h->br_lo[1]};
  const uint32_t Wd[2] = {h->br_hi[0] - L[0], h->br_hi[1] - L[1]};
  const uint32_t sft[2] = {h->br_shift[0], h->br_shift[1]};
  if (tid < kTargets) { s_len[tid] = 0; s_bin[tid] = 0xFFFFFFFFu; s_out[tid] = 0; }
  // 1: candidate list overflow, 6: the depth-domain loop met NaN / inf or a possible sentinel hit
  if (tid == 0) s_fail = (int64_t)n_cand > ws.c_cap ? 1u : (h->force_retry ? 6u : 0u);
  __syncthreads();
  if (!has_nan && !s_fail) {
    // 1. bin of each target rank in its bracket histogram (2 bins per thread)
    const uint32_t* ghist = ws.hist + (int64_t)v * kTargets * kBins;
#pragma unroll 1
    for (int j = 0; j < kTargets; ++j) {
      const int g = j >> 1;
      const long long r = ((j & 1) ? (j < 2 ? r1.hi : r3.hi) : (j < 2 ? r1.lo : r3.lo)) - (long long)h->below[g];
      const uint32_t c0 = ghist[g * kBrBins + 2 * tid], c1 = ghist[g * kBrBins + 2 * tid + 1];
      int tot;
      const long long cum = block_excl_scan((int)(c0 + c1), scan_smem, &tot);
      if (r >= cum && r < cum + c0) { s_bin[j] = 2 * tid; s_rank[j] = r - cum; }
      else if (r >= cum + c0 && r < cum + c0 + c1) { s_bin[j] = 2 * tid + 1; s_rank[j] = r - cum - c0; }
      if (tid == 0) {
        s_rank0[j] = r; s_flo[j] = L[g]; s_fhi[j] = L[g] + Wd[g];
        if (r < 0 || r >= tot) s_fail = r < 0 ? 2u : 3u;  // 2 / 3: rank below / above the bracket
      }
      __syncthreads();
    }
  }
  if (!has_nan && !s_fail) {
    // 2. gather the candidates of those bins (targets sharing a bin share list `slot`)
    int slot[kTargets];
    uint32_t bin[kTargets];
#pragma unroll
    for (int j = 0; j < kTargets; ++j) {
      bin[j] = s_bin[j];
      slot[j] = j;
      for (int i = 0; i < j; ++i)
        if (slot[j] == j && (i >> 1) == (j >> 1) && bin[i] == bin[j]) slot[j] = i;
    }
    const uint32_t* cand = ws.cand + (int64_t)v * ws.c_cap;
    for (uint32_t base = 0; base < n_cand; base += 4 * kSelThreads) {
      uint32_t kk[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t i = base + u * kSelThreads + tid;
        kk[u] = i < n_cand ? cand[i] : 0u;
      }
      // one bracket test per key and bracket, then the two bins of that bracket; hits are rare
      // (a bin holds a few dozen of the ~10^5 candidates), so the appends stay off the hot path
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool valid = base + u * kSelThreads + tid < n_cand;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const uint32_t o = kk[u] - L[g];
          const uint32_t b = o >> sft[g];
          const bool inb = valid && o <= Wd[g];
          const bool hit0 = inb && b == bin[2 * g];
          const bool hit1 = inb && !hit0 && b == bin[2 * g + 1];  // only when the two bins differ
          if (hit0 || hit1) {
            const int j = hit0 ? 2 * g : 2 * g + 1;
            const uint32_t at = atomicAdd(&s_len[j], 1u);
            if (at < (uint32_t)kListCap) s_list[j][at] = kk[u];
          }
        }
      }
    }
    __syncthreads();
    // 3. exact rank inside the list: position = #smaller + #equal with a lower index
#pragma unroll 1
    for (int j = 0; j < kTargets; ++j) {
      const uint32_t m = s_len[slot[j]];
      if (m > (uint32_t)kListCap || (long long)m <= s_rank[j]) { if (tid == 0) s_fail = 4u; continue; }  // 4: bin too full
      const uint32_t* lst = s_list[slot[j]];
      for (uint32_t a = tid; a < m; a += kSelThreads) {
        const uint32_t key = lst[a];
        uint32_t pos = 0;
        for (uint32_t b = 0; b < m; ++b) {
          const uint32_t x = lst[b];
          pos += (x < key || (x == key && b < a)) ? 1u : 0u;
        }
        if ((long long)pos == s_rank[j]) s_out[j] = key;
      }
    }
    __syncthreads();
    if (s_fail == 4u) {
      // a bracket bin holds more keys than the list (heavy ties, values straddling zero): exact
      // radix select among the candidates instead
      __syncthreads();
      SelShared& sel = *reinterpret_cast<SelShared*>(s_raw);
      const uint32_t f = block_select(cand, n_cand, s_flo, s_fhi, s_rank0, s_out, sel);
      if (tid == 0) s_fail = f ? 5u : 0u;
      __syncthreads();
    }
  }
  if (tid == 0 && raw) {
    h->fail_code = s_fail;
    if (s_fail) {
      h->raw_state = kRawFail;
    } else {
      for (int j = 0; j < kTargets; ++j) h->raw_val[j] = key_to_float(s_out[j]);
      h->raw_min = key_to_float(min_key); h->raw_max = key_to_float(h->max_key);
      h->raw_state = has_nan ? kRawNan : kRawOk;
    }
  } else if (tid == 0) {
    h->fail_code = s_fail;
    if (s_fail) {
      if (lo_hi) { lo_hi[2 * v] = NAN; lo_hi[2 * v + 1] = NAN; }
      if (status) status[v] |= MDI_STATUS_STATS_RETRY;
      if (retry_flag) *retry_flag = 1;
    } else {
      float val[kTargets];
      for (int j = 0; j < kTargets; ++j) val[j] = key_to_float(s_out[j]);
      write_range(v, val, r1.w, r3.w, min_key, has_nan ? 0xFFFFFFFFu : h->max_key, lo_hi, quantiles);
    }
  }
}

// MDI_STATS_RAW, second half.  Order statistics commute with a monotone map: with a finite scale > 0
// and a finite shift, F(d) = fl(fl(d * s) + t) is non-decreasing in d, so the k-th smallest ALIGNED
// value of a view is F(k-th smallest raw depth), whatever the fit turns out to be -- the whole
// statistics pass can run on the raw depth, beside the alignment kernels, and this kernel (one
// thread per view, after both) maps its six numbers through the fit and finishes like
// stats_finish_kernel: torch.quantile's lerp, the IQR clamp, MDI_STATUS_EMPTY_STATS.  What the raw
// pass cannot know is handed to the exact path (MDI_STATUS_STATS_RETRY, *retry_flag): a fit that is
// not increasing, and a view in which some depth of the population's range aligns to exactly -42.0,
// the sentinel the population excludes (sentinel != 0).
__global__ void stats_map_kernel(int n_views, const float* __restrict__ scale_shift, int32_t* __restrict__ status,
                                 int sentinel, StatsWs ws, float* __restrict__ lo_hi,
                                 float* __restrict__ quantiles, int64_t* __restrict__ retry_flag) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_views) return;
  StatsHdr* h = ws.hdr + v;
  const bool low_conf = status && (status[v] & MDI_STATUS_LOW_CONFIDENCE);
  const uint32_t state = h->raw_state;
  auto put = [&](float lo, float hi, float q1, float q3, float mn, float mx) {
    if (lo_hi) { lo_hi[2 * v] = lo; lo_hi[2 * v + 1] = hi; }
    if (quantiles) { quantiles[4 * v] = q1; quantiles[4 * v + 1] = q3; quantiles[4 * v + 2] = mn; quantiles[4 * v + 3] = mx; }
  };
  if (low_conf || state == kRawEmpty) {
    put(0.f, 0.f, 0.f, 0.f, 0.f, 0.f);
    if (!low_conf && status) status[v] |= MDI_STATUS_EMPTY_STATS;
    return;
  }
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  bool retry = state != kRawOk && state != kRawNan;
  if (!(s > 0.f && s < INFINITY && fabsf(t) < INFINITY)) retry = true;
  auto F = [&](uint32_t kd) { return fast_key(align_depth(key_to_float(kd), s, t)); };
  if (!retry && sentinel && state == kRawOk) {
    // [d_a, d_b] = the depths that align to exactly -42.0: smallest kd with F(kd) >= key(-42), largest
    // with F(kd) <= key(-42); both exist for an increasing F.  A 33-float window around (-42 - t) / s first.
    const uint32_t bound = fast_key(-42.0f);
    const float guess = fdiv(fsub(-42.0f, t), s);
    uint32_t wlo = kKeyNegInf, whi = kKeyPosInf;
    if (guess == guess) {
      const uint32_t kg = fast_key(guess);
      if (kg > kKeyNegInf + 16u && kg < kKeyPosInf - 16u) { wlo = kg - 16u; whi = kg + 16u; }
    }
    uint32_t lo = kKeyNegInf, hi = kKeyPosInf;
    if (wlo != kKeyNegInf && F(whi) >= bound && F(wlo) < bound) { lo = wlo + 1u; hi = whi; }
    while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if (F(mid) >= bound) hi = mid; else lo = mid + 1; }
    const uint32_t ka = lo;
    lo = kKeyNegInf; hi = kKeyPosInf;
    if (wlo != kKeyNegInf && F(wlo) <= bound && F(whi) > bound) { lo = wlo; hi = whi - 1u; }
    while (lo < hi) { const uint32_t mid = lo + ((hi - lo + 1) >> 1); if (F(mid) <= bound) lo = mid; else hi = mid - 1; }
    const uint32_t kb = lo;
    // compared as floats (the raw pass does not tell -0 from +0)
    const float da = key_to_float(ka), db = key_to_float(kb);
    if (da <= db && h->raw_min <= db && h->raw_max >= da) retry = true;
  }
  if (retry) {
    put(NAN, NAN, NAN, NAN, NAN, NAN);
    if (status) status[v] |= MDI_STATUS_STATS_RETRY;
    if (retry_flag) *retry_flag = 1;
    return;
  }
  const uint32_t n = h->count;
  const QuantileRank r1 = quantile_rank(0.25f, (int64_t)n);
  const QuantileRank r3 = quantile_rank(0.75f, (int64_t)n);
  float val[kTargets];
  for (int j = 0; j < kTargets; ++j) val[j] = align_depth(h->raw_val[j], s, t);
  // monotone alignment: the extreme aligned values are those of the extreme depths
  const uint32_t kmn = fast_key(align_depth(h->raw_min, s, t)), kmx = fast_key(align_depth(h->raw_max, s, t));
  write_range(v, val, r1.w, r3.w, kmn, state == kRawNan ? 0xFFFFFFFFu : kmx, lo_hi, quantiles);
}

// ---- exact path: radix select over all pixels ------------------------------------------------------
// PASS 0: bin = key >> 20 over everything (+ min / max / count)
// PASS 1: bin = (key >> 8) & 0xfff over keys whose top 12 bits match a slot prefix
// PASS 2: bin = key & 0xff over keys whose top 24 bits match a slot prefix
template <int PASS>
__global__ void __launch_bounds__(kStatsThreads)
stats_pass_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                  const int32_t* __restrict__ status, int flags, StatsWs ws) {
  extern __shared__ uint32_t s_hist[];  // n_slots * bins
  __shared__ uint32_t s_min, s_max, s_cnt;
  const int v = blockIdx.y;
  const MdiView vw = views[v];
  const int first_tile = blockIdx.x * kStatsTilesPerCta;
  if (first_tile >= vw.n_tiles) return;
  if (status && (status[v] & MDI_STATUS_LOW_CONFIDENCE)) return;
  StatsHdr* sv = ws.hdr + v;
  if (!sv->run_exact) return;
  uint32_t* ghist = ws.hist + (int64_t)v * kTargets * kBins;
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
    if (c) atomicAdd(&ghist[(i / bins) * kBins + (i % bins)], c);
  }
  if (PASS == 0 && threadIdx.x == 0 && s_cnt) {
    atomicMin(&sv->x_min_key, s_min); atomicMax(&sv->x_max_key, s_max); atomicAdd(&sv->x_count, s_cnt);
  }
}

// One CTA per view: locate each target rank in its slot histogram, extend the prefixes,
// deduplicate them into the next pass's slots and clear the histograms.  After PASS 2 the
// prefixes are the keys of the order statistics: finish the quantiles and the range.
template <int PASS>
__global__ void __launch_bounds__(kStatsThreads)
stats_select_kernel(int32_t* __restrict__ status, StatsWs ws, float* __restrict__ lo_hi,
                    float* __restrict__ quantiles) {
  __shared__ int scan_smem[33];
  __shared__ uint32_t s_bin[kTargets];
  __shared__ long long s_rank[kTargets];
  const int v = blockIdx.x;
  StatsHdr* sv = ws.hdr + v;
  if (!sv->run_exact) return;
  uint32_t* ghist = ws.hist + (int64_t)v * kTargets * kBins;
  constexpr int bins = (PASS == 2) ? 256 : kBins;
  constexpr int shift = (PASS == 2) ? 8 : 12;
  constexpr int per_thread = bins / kStatsThreads;  // 16 or 1
  const bool low_conf = status && (status[v] & MDI_STATUS_LOW_CONFIDENCE);
  const uint32_t n = sv->x_count;
  if (low_conf || n == 0) {
    if (threadIdx.x == 0) {
      sv->n_slots = 0;
      if (PASS == 2) {
        if (lo_hi) { lo_hi[2 * v] = 0.f; lo_hi[2 * v + 1] = 0.f; }
        if (quantiles) for (int k = 0; k < 4; ++k) quantiles[4 * v + k] = 0.f;
        if (status) status[v] &= ~MDI_STATUS_STATS_RETRY;
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
    const uint32_t* h = ghist + (int64_t)sv->tgt_slot[j] * kBins;
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
  for (int i = threadIdx.x; i < kTargets * kBins; i += blockDim.x) ghist[i] = 0;
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
      float val[kTargets];
      for (int j = 0; j < kTargets; ++j) val[j] = key_to_float(pre[j]);
      write_range(v, val, sv->w[0], sv->w[1], sv->x_min_key, sv->x_max_key, lo_hi, quantiles);
      if (status) status[v] &= ~MDI_STATUS_STATS_RETRY;
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

static unsigned max_chunks(int64_t tiles) {
  int64_t c = (tiles + kStatsTilesPerCta - 1) / kStatsTilesPerCta;
  return (unsigned)(c < 1 ? 1 : c);
}

}  // namespace mdi

extern "C" int64_t mdi_depth_range_workspace_bytes(int n_views, int64_t max_tiles_per_view) {
  using namespace mdi;
  if (n_views <= 0) return 0;
  if (max_tiles_per_view < 1) max_tiles_per_view = 1;
  const int64_t hdr = ((int64_t)(sizeof(StatsHdr) * (size_t)n_views + 255) / 256) * 256;
  return hdr + (int64_t)n_views * 4 * ((int64_t)kTargets * kBins + sample_cap(max_tiles_per_view) +
                                       cand_cap(max_tiles_per_view));
}

extern "C" int mdi_depth_range(const MdiView* views, int n_views, int64_t max_tiles_per_view,
                               const float* scale_shift, int32_t* status, int flags,
                               void* workspace, float* lo_hi, float* quantiles,
                               int64_t* retry_flag, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n_views >= 0 && max_tiles_per_view >= 0, "mdi_depth_range: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(views && workspace && (lo_hi || quantiles || (flags & MDI_STATS_RAW)), "mdi_depth_range: null pointer");
  MDI_REQUIRE(n_views <= 65535, "mdi_depth_range: more than 65535 views per call");
  if (max_tiles_per_view < 1) max_tiles_per_view = 1;
  cudaStream_t st = (cudaStream_t)stream;
  const StatsWs ws = carve(workspace, n_views, max_tiles_per_view);
  const int mode = (flags & MDI_STATS_EXACT) ? 1 : ((flags & MDI_STATS_RETRY_FLAGGED) ? 2 : 0);
  const bool raw = (flags & MDI_STATS_RAW) != 0;
  MDI_REQUIRE(!raw || mode == 0, "mdi_depth_range: MDI_STATS_RAW goes with the one-pass algorithm only");
  if (raw) { scale_shift = nullptr; status = nullptr; }   // raw depth, every view: the fit and the status may not exist yet
  const int pop = flags & (MDI_KEEP_PRED_MASK | (raw ? 0 : MDI_KEEP_SENTINEL));
  MDI_CUDA(cudaMemsetAsync(ws.hdr, 0, sizeof(StatsHdr) * (size_t)n_views, st), "mdi_depth_range");
  if (mode != 0) {   // the one-pass path starts from the zero-filled headers
    mdi::prof_mark("stats_init_kernel", st);
    stats_init_kernel<<<(n_views + 255) / 256, 256, 0, st>>>(ws.hdr, n_views, status, mode);
  }
  dim3 grid(max_chunks(max_tiles_per_view), (unsigned)n_views);
  MDI_CUDA(cudaMemsetAsync(ws.hist, 0, sizeof(uint32_t) * kTargets * kBins * (size_t)n_views, st),
           "mdi_depth_range");
  if (mode == 0) {
    const int64_t groups = max_tiles_per_view * (MDI_TILE_PX / kSampleGroup);
    mdi::prof_mark("stats_sample_kernel", st);
    stats_sample_kernel<<<dim3((unsigned)((groups + 255) / 256), (unsigned)n_views), 256, 0, st>>>(
        views, scale_shift, status, pop, ws);
    mdi::prof_mark("stats_bracket_kernel", st);
    stats_bracket_kernel<<<n_views, kBracketThreads, 0, st>>>(scale_shift, status, (pop & MDI_KEEP_SENTINEL) ? 1 : 0, ws);
    const bool um = pop & MDI_KEEP_PRED_MASK, se = pop & MDI_KEEP_SENTINEL;
    // 0: register-staged kernel (default), 1: TMA-staged kernel (MDI_STATS_MAIN=tma; bit-identical
    // results, measured slower so far: 0.42 vs 0.33 ms on C2)
    // process-wide choices read once from the environment (thread-safe static initialisation); the
    // per-DEVICE function attributes below are set before every launch that needs them
    static const int main_variant = [] {
      const char* e = getenv("MDI_STATS_MAIN");
      return (e && strcmp(e, "tma") == 0) ? 1 : 0;
    }();
    static const int prefetch_ahead = [] {  // tiles of L2 prefetch in the depth-domain loop (MDI_STATS_PREFETCH, 0 = off)
      const char* e = getenv("MDI_STATS_PREFETCH");
      const int p = e ? atoi(e) : 2;
      return (p < 0 || p > 16) ? 2 : p;
    }();
    if (main_variant == 1) {
      MDI_CUDA(cudaFuncSetAttribute(stats_main_tma_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes), "mdi_depth_range attr");
      MDI_CUDA(cudaFuncSetAttribute(stats_main_tma_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes), "mdi_depth_range attr");
      MDI_CUDA(cudaFuncSetAttribute(stats_main_tma_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes), "mdi_depth_range attr");
      MDI_CUDA(cudaFuncSetAttribute(stats_main_tma_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes), "mdi_depth_range attr");
    }
    mdi::prof_mark("stats_main_kernel", st);
    if (main_variant == 1) {
      if (um && se) stats_main_tma_kernel<true, true><<<grid, kStatsThreads, kTmaSmemBytes, st>>>(views, scale_shift, status, ws);
      else if (um) stats_main_tma_kernel<true, false><<<grid, kStatsThreads, kTmaSmemBytes, st>>>(views, scale_shift, status, ws);
      else if (se) stats_main_tma_kernel<false, true><<<grid, kStatsThreads, kTmaSmemBytes, st>>>(views, scale_shift, status, ws);
      else stats_main_tma_kernel<false, false><<<grid, kStatsThreads, kTmaSmemBytes, st>>>(views, scale_shift, status, ws);
    } else if (um && se) stats_main_kernel<true, true><<<grid, kStatsThreads, 0, st>>>(views, scale_shift, status, ws, prefetch_ahead);
    else if (um) stats_main_kernel<true, false><<<grid, kStatsThreads, 0, st>>>(views, scale_shift, status, ws, prefetch_ahead);
    else if (se) stats_main_kernel<false, true><<<grid, kStatsThreads, 0, st>>>(views, scale_shift, status, ws, prefetch_ahead);
    else stats_main_kernel<false, false><<<grid, kStatsThreads, 0, st>>>(views, scale_shift, status, ws, prefetch_ahead);
    mdi::prof_mark("stats_finish_kernel", st);
    stats_finish_kernel<<<n_views, kSelThreads, 0, st>>>(status, ws, lo_hi, quantiles, retry_flag, raw ? 1 : 0);
  } else {
    // per device (context) attribute: set on every call (microseconds), never cached process-wide
    MDI_CUDA(cudaFuncSetAttribute(stats_pass_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  kTargets * kBins * 4), "mdi_depth_range attr");
    mdi::prof_mark("stats_pass_kernel", st);
    stats_pass_kernel<0><<<grid, kStatsThreads, kBins * 4, st>>>(views, scale_shift, status, pop, ws);
    mdi::prof_mark("stats_select_kernel", st);
    stats_select_kernel<0><<<n_views, kStatsThreads, 0, st>>>(status, ws, lo_hi, quantiles);
    mdi::prof_mark("stats_pass_kernel", st);
    stats_pass_kernel<1><<<grid, kStatsThreads, kTargets * kBins * 4, st>>>(views, scale_shift, status, pop, ws);
    mdi::prof_mark("stats_select_kernel", st);
    stats_select_kernel<1><<<n_views, kStatsThreads, 0, st>>>(status, ws, lo_hi, quantiles);
    mdi::prof_mark("stats_pass_kernel", st);
    stats_pass_kernel<2><<<grid, kStatsThreads, kTargets * 256 * 4, st>>>(views, scale_shift, status, pop, ws);
    mdi::prof_mark("stats_select_kernel", st);
    stats_select_kernel<2><<<n_views, kStatsThreads, 0, st>>>(status, ws, lo_hi, quantiles);
  }
  MDI_CHECK_LAUNCH("mdi_depth_range");
  return MDI_OK;
}

extern "C" int mdi_depth_range_map(int n_views, int64_t max_tiles_per_view, const float* scale_shift,
                                   int32_t* status, int flags, void* workspace, float* lo_hi,
                                   float* quantiles, int64_t* retry_flag, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n_views >= 0 && max_tiles_per_view >= 0, "mdi_depth_range_map: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(workspace && (lo_hi || quantiles), "mdi_depth_range_map: null pointer");
  if (max_tiles_per_view < 1) max_tiles_per_view = 1;
  const StatsWs ws = carve(workspace, n_views, max_tiles_per_view);
  cudaStream_t st = (cudaStream_t)stream;
  mdi::prof_mark("stats_map_kernel", st);
  stats_map_kernel<<<(n_views + 63) / 64, 64, 0, st>>>(n_views, scale_shift, status, (flags & MDI_KEEP_SENTINEL) ? 1 : 0,
                                                       ws, lo_hi, quantiles, retry_flag);
  MDI_CHECK_LAUNCH("mdi_depth_range_map");
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
  mdi::prof_mark("grad_range_init_kernel", st);
  grad_range_init_kernel<<<(n_views + 255) / 256, 256, 0, st>>>(keys, n_views);
  dim3 grid(max_chunks(max_tiles_per_view), (unsigned)n_views);
  mdi::prof_mark("grad_range_kernel", st);
  grad_range_kernel<<<grid, kStatsThreads, 0, st>>>(views, scale_shift, status, keys);
  mdi::prof_mark("grad_range_finish_kernel", st);
  grad_range_finish_kernel<<<(n_views + 255) / 256, 256, 0, st>>>(grad_range, n_views);
  MDI_CHECK_LAUNCH("mdi_grad_range");
  return MDI_OK;
}
