// k3_mask.cu -- keep-mask predicate (a5-a10): alignment applied on the fly, sentinel / >= 0 /
// predicted-mask tests, SfM-density patch lookup, optional gradient mask, static or adaptive
// stride test.
//
// A pixel can only be kept if its stride f divides BOTH its row and its column, so the kept
// pixels of stride f lie on the lattice (f i, f j).  Instead of testing every pixel, the hot
// kernel (keep_lattice_kernel) zero-fills the outputs at copy speed and then walks the lattices
// of the strides of the configured range: sum_f (H/f)(W/f) = 16 % of the pixels for the default
// 5..15 range, 1 % for stride 10; rows that hold no lattice point are never read.  At a lattice
// point of stride f the pixel is kept iff its own stride equals f.  The adaptive stride f(a) is a
// non-increasing step function of the aligned depth a (every fp32 operation of
// adaptive_subsampling.py:12-17,89-120 is monotone), so `f(a) == f` is two comparisons against
// per-view breakpoints found by bisection over the float keys with the exact arithmetic
// (factor_table_kernel); irregular views (NaN bounds, hi <= lo) evaluate the exact arithmetic at
// the lattice points instead.  Kept pixels set their keep byte, their bit in the tile bit words
// and bump the tile count that the order-preserving compaction consumes (mdi_scan_tiles /
// mdi_emit_points).  Configurations without a stride test run the per-pixel kernel.
#include "common.cuh"

namespace mdi {

constexpr int kMaskThreads = 256;

// a7 breakpoints.  Row v of `table` (MDI_FACTOR_TABLE_STRIDE floats): [0] = 1 if the fast path is
// valid for the view; [1 + j], j = 0 .. n (n = fmax - fmin + 1): B[0] = NaN, B[j] = A_{fmin+j-1},
// A_k = smallest float with adaptive_factor(.) <= k, so that
//     adaptive_factor(a) == k   <=>   a >= B[k - fmin + 1]  &&  !(a >= B[k - fmin]).
__global__ void factor_table_kernel(int n_views, const float* __restrict__ lo_hi, int fmin,
                                    int fmax, float* __restrict__ table) {
  const int v = blockIdx.x * (blockDim.x >> 5) + warp_id();
  if (v >= n_views) return;
  const int lane = lane_id();
  float* row = table + (int64_t)v * MDI_FACTOR_TABLE_STRIDE;
  const float lo = lo_hi[2 * v], hi = lo_hi[2 * v + 1];
  const float den = fsub(hi, lo);
  const int n = fmax - fmin + 1;
  const bool valid = (lo == lo) && (hi == hi) && (fabsf(lo) <= 3.4028234e38f) &&
                     (fabsf(den) <= 3.4028234e38f) && (den > 0.f) && fmin >= 1 && fmax >= fmin &&
                     n <= 31;
  if (lane == 0) { row[0] = valid ? 1.f : 0.f; row[1] = NAN; }
  if (!valid) return;
  // lane j (0 <= j < n) owns A_k for k = fmin + j
  if (lane < n) {
    const int k = fmin + lane;
    float A;
    if (k == fmax) {
      A = -INFINITY;
    } else {
      uint32_t klo = float_to_key(-INFINITY), khi = float_to_key(INFINITY);
      while (klo < khi) {
        const uint32_t mid = klo + ((khi - klo) >> 1);
        if (adaptive_factor(key_to_float(mid), lo, hi, fmin, fmax) <= k) khi = mid; else klo = mid + 1;
      }
      A = key_to_float(klo);
    }
    row[2 + lane] = A;
  }
}

// |d(r,c) - d(r,c-1)| + |d(r,c) - d(r-1,c)| on the aligned depth (points_from_depth.py:205-209)
__device__ __forceinline__ float grad_at(const float* __restrict__ depth, int W, int64_t p, int row,
                                         int col, float a, float s, float t) {
  float g = 0.f;
  if (col >= 1) g = fadd(g, fabsf(fsub(a, align_depth(__ldg(&depth[p - 1]), s, t))));
  if (row >= 1) g = fadd(g, fabsf(fsub(a, align_depth(__ldg(&depth[p - W]), s, t))));
  return g;
}

constexpr int kFlatStrides = 32;
struct MaskParams {
  int mode, static_factor, fmin, fmax, flags;
  float grad_thresh;
  // flat grid of the lattice walk: CTAs of stride fmin + k are blockIdx.x in [cta_begin[k], cta_begin[k+1])
  // (sized for the largest view); n_flat == 0 selects the (chunk, stride) 2-D grid instead
  int n_flat;
  int cta_begin[kFlatStrides + 1];
};

// every test but the stride test, for one pixel q = (r, c) with aligned depth a
struct RestCtx {
  const float* depth; const uint8_t* pmask; const uint8_t* pb; const float* grad_range;
  int W, patch_h, patch_w, grid_h, grid_w, flags, v;
  float s, t, grad_thresh;
};
__device__ __forceinline__ bool rest_ok(const RestCtx& x, int64_t q, int r, int c, float a) {
  bool ok = true;
  if (x.flags & MDI_KEEP_PRED_MASK) ok = x.pmask ? (__ldg(&x.pmask[q]) != 0) : true;
  if (x.flags & MDI_KEEP_SENTINEL) ok = ok && (a != -42.0f);
  if (x.flags & MDI_KEEP_NONNEG) ok = ok && (a >= 0.f);
  if (ok && x.pb) {
    const int pr = r / x.patch_h, pc = c / x.patch_w;
    if (pr < x.grid_h && pc < x.grid_w && __ldg(&x.pb[pr * x.grid_w + pc])) ok = false;
  }
  if (ok && (x.flags & MDI_KEEP_GRAD)) {
    const float gmin = x.grad_range[2 * x.v];
    // (g - min) / ((max - min) + 1e-8)   (points_from_depth.py:210-211)
    const float gden = fadd(fsub(x.grad_range[2 * x.v + 1], gmin), 1e-8f);
    const float g = grad_at(x.depth, x.W, q, r, c, a, x.s, x.t);
    ok = fdiv(fsub(g, gmin), gden) <= x.grad_thresh;
  }
  return ok;
}

__device__ __forceinline__ RestCtx make_ctx(const MdiView* vp, int v, const MaskParams& prm,
                                            const uint8_t* patch_blocked, const float* grad_range,
                                            float s, float t) {
  RestCtx x;
  x.depth = vp->depth; x.pmask = vp->mask;
  x.pb = (prm.flags & MDI_KEEP_PATCH) && patch_blocked ? patch_blocked + (int64_t)v * MDI_PATCH_STRIDE : nullptr;
  x.grad_range = grad_range;
  x.W = vp->W; x.patch_h = vp->patch_h; x.patch_w = vp->patch_w; x.grid_h = vp->grid_h; x.grid_w = vp->grid_w;
  x.flags = prm.flags; x.v = v; x.s = s; x.t = t; x.grad_thresh = prm.grad_thresh;
  return x;
}

// mark pixel q of the view as kept (outputs were zero-filled before the launch)
__device__ __forceinline__ void mark_kept(const MdiView* vp, int64_t q, uint8_t* __restrict__ keep,
                                          uint32_t* __restrict__ bits, int32_t* __restrict__ tile_counts) {
  if (keep) keep[vp->px_begin + q] = 1;
  const int64_t tile = vp->tile_begin + (q >> 12);
  const int local = (int)(q & (MDI_TILE_PX - 1));
  atomicOr(&bits[tile * MDI_TILE_WORDS + (local >> 5)], 1u << (local & 31));
  atomicAdd(&tile_counts[tile], 1);
}

constexpr int kLatticeWarps = 8;
constexpr int kBatch = 8;                              // lattice points per lane, loads in flight together
constexpr int kChunk = 32 * kBatch;                    // lattice points per warp
constexpr int kCtaPoints = kChunk * kLatticeWarps;     // lattice points per CTA

// The lattice of stride f of a view, (f i, f j) with i < ceil(H/f), j < ceil(W/f), is flattened
// row-major and cut into chunks of 256 points, one warp each; lane l owns points chunk*256 + 32 u
// + l (u < 8): the 8 depth loads are issued together, then the stride test, then the predicted-
// mask loads of the survivors together, then their patch lookups together.
// grid = (ceil(max lattice points / kCtaPoints), n_strides, n_views)
__global__ void __launch_bounds__(kLatticeWarps * 32, 6)
keep_lattice_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                    const int32_t* __restrict__ status, const float* __restrict__ lo_hi,
                    const float* __restrict__ grad_range, const uint8_t* __restrict__ patch_blocked,
                    const float* __restrict__ factor_table, MaskParams prm,
                    uint8_t* __restrict__ keep, uint32_t* __restrict__ bits,
                    int32_t* __restrict__ tile_counts) {
  const int v = blockIdx.z;
  if (status && status[v] != MDI_STATUS_OK) return;
  const bool adaptive = prm.mode == MDI_SUBSAMPLE_ADAPTIVE;
  int si = (int)blockIdx.y;
  uint32_t bx = blockIdx.x;
  if (prm.n_flat) {  // flat grid: no CTA is launched for chunks a coarse stride does not have
    si = 0;
    while (si + 1 < prm.n_flat && (int)blockIdx.x >= prm.cta_begin[si + 1]) ++si;
    bx -= (uint32_t)prm.cta_begin[si];
  }
  const int f = adaptive ? prm.fmin + si : prm.static_factor;
  const MdiView* vp = views + v;
  const int H = vp->H, W = vp->W;
  // all index arithmetic in 32 bits: the host entry point requires H * W < 2^31 for this kernel
  const uint32_t n_rows = (uint32_t)((H + f - 1) / f), n_cols = (uint32_t)((W + f - 1) / f);
  const uint32_t n_pts = n_rows * n_cols;
  const uint32_t first = (bx * kLatticeWarps + warp_id()) * kChunk + lane_id();
  if (first - lane_id() >= n_pts) return;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const float* __restrict__ depth = vp->depth;
  // stride test: f(a) == f  <=>  a >= b_lo && !(a >= b_hi) when the view has a breakpoint table
  bool table = false;
  float b_lo = 0.f, b_hi = 0.f;
  if (adaptive) {
    const float* row = factor_table ? factor_table + (int64_t)v * MDI_FACTOR_TABLE_STRIDE : nullptr;
    table = row && row[0] != 0.f;
    if (table) { b_lo = row[1 + (f - prm.fmin + 1)]; b_hi = row[1 + (f - prm.fmin)]; }
  }
  const RestCtx ctx = make_ctx(vp, v, prm, patch_blocked, grad_range, s, t);

  // lattice coordinates of the lane's points: one division, then steps of 32 columns
  uint32_t q[kBatch];           // pixel index, 0xFFFFFFFF = outside
  uint16_t li[kBatch], lj[kBatch];
  {
    uint32_t i = first / n_cols, j = first - i * n_cols;
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      q[u] = (first + 32u * u < n_pts) ? (i * (uint32_t)f) * (uint32_t)W + j * (uint32_t)f : 0xFFFFFFFFu;
      li[u] = (uint16_t)i; lj[u] = (uint16_t)j;
      j += 32u;
      while (j >= n_cols) { j -= n_cols; ++i; }
    }
  }
  float av[kBatch];
  bool ok[kBatch];
#pragma unroll
  for (int u = 0; u < kBatch; ++u) av[u] = q[u] != 0xFFFFFFFFu ? __ldg(&depth[q[u]]) : 0.f;
  bool any = false;
#pragma unroll
  for (int u = 0; u < kBatch; ++u) {
    const float a = align_depth(av[u], s, t);
    av[u] = a;
    bool hit = q[u] != 0xFFFFFFFFu;
    if (hit && adaptive) {
      if (table && a == a) {
        hit = (a >= b_lo) && !(a >= b_hi);
      } else {
        // exact arithmetic; a NaN stride (INT32_MIN) keeps index 0 only: visit it once, at fmin
        const int fa = adaptive_factor(a, lo_hi[2 * v], lo_hi[2 * v + 1], prm.fmin, prm.fmax);
        hit = (fa == INT32_MIN) ? (q[u] == 0 && f == prm.fmin) : (fa == f);
      }
    }
    if (prm.flags & MDI_KEEP_SENTINEL) hit = hit && (a != -42.0f);
    if (prm.flags & MDI_KEEP_NONNEG) hit = hit && (a >= 0.f);
    ok[u] = hit;
    any = any || hit;
  }
  if (!any) return;
  // predicted mask of the survivors
  if ((prm.flags & MDI_KEEP_PRED_MASK) && ctx.pmask) {
    uint8_t mb[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) mb[u] = ok[u] ? __ldg(&ctx.pmask[q[u]]) : 0;
#pragma unroll
    for (int u = 0; u < kBatch; ++u) ok[u] = ok[u] && mb[u] != 0;
  }
  // SfM-density patch lookup of the survivors.  r / patch_h as a multiply-high: exact because
  // r, patch_h < 2^16 (x * d < 2^32 with M = floor(2^32 / d) + 1)
  if (ctx.pb) {
    const uint32_t mh = 0xFFFFFFFFu / (uint32_t)max(ctx.patch_h, 1) + 1u, mw = 0xFFFFFFFFu / (uint32_t)max(ctx.patch_w, 1) + 1u;
    uint8_t pv[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      pv[u] = 0;
      if (ok[u]) {
        const uint32_t r = (uint32_t)li[u] * (uint32_t)f, c = (uint32_t)lj[u] * (uint32_t)f;
        const int pr = ctx.patch_h <= 1 ? (int)r : (int)__umulhi(r, mh);
        const int pc = ctx.patch_w <= 1 ? (int)c : (int)__umulhi(c, mw);
        if (pr < ctx.grid_h && pc < ctx.grid_w) pv[u] = __ldg(&ctx.pb[pr * ctx.grid_w + pc]);
      }
    }
#pragma unroll
    for (int u = 0; u < kBatch; ++u) ok[u] = ok[u] && pv[u] == 0;
  }
#pragma unroll
  for (int u = 0; u < kBatch; ++u) {
    if (!ok[u]) continue;
    if (prm.flags & MDI_KEEP_GRAD) {
      const int r = (int)li[u] * f, c = (int)lj[u] * f;
      const float gmin = grad_range[2 * v];
      // (g - min) / ((max - min) + 1e-8)   (points_from_depth.py:210-211)
      const float gden = fadd(fsub(grad_range[2 * v + 1], gmin), 1e-8f);
      const float g = grad_at(depth, W, (int64_t)q[u], r, c, av[u], s, t);
      if (!(fdiv(fsub(g, gmin), gden) <= prm.grad_thresh)) continue;
    }
    mark_kept(vp, (int64_t)q[u], keep, bits, tile_counts);
  }
}

// Per-pixel predicate for configurations the lattice walk does not cover (no stride test,
// strides outside 1..max).  One CTA per 4096-pixel tile, one lane per pixel.
__global__ void __launch_bounds__(kMaskThreads)
keep_pixel_kernel(const MdiView* __restrict__ views, int n_views,
                  const float* __restrict__ scale_shift, const int32_t* __restrict__ status,
                  const float* __restrict__ lo_hi, const float* __restrict__ grad_range,
                  const uint8_t* __restrict__ patch_blocked, MaskParams prm,
                  uint8_t* __restrict__ keep, uint32_t* __restrict__ bits,
                  int32_t* __restrict__ tile_counts) {
  const int64_t tile = blockIdx.x;
  const int v = find_view(views, n_views, tile);
  if (status && status[v] != MDI_STATUS_OK) return;
  const MdiView* vp = views + v;
  const int W = vp->W;
  const int64_t HW = (int64_t)vp->H * W;
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  float lo = 0.f, hi = 0.f;
  if (prm.mode == MDI_SUBSAMPLE_ADAPTIVE && lo_hi) { lo = lo_hi[2 * v]; hi = lo_hi[2 * v + 1]; }
  const RestCtx ctx = make_ctx(vp, v, prm, patch_blocked, grad_range, s, t);
  const int64_t q0 = (tile - vp->tile_begin) * MDI_TILE_PX;
#pragma unroll 1
  for (int k = threadIdx.x; k < MDI_TILE_PX; k += kMaskThreads) {
    const int64_t q = q0 + k;
    if (q >= HW) break;
    const int r = (int)(q / W), c = (int)(q - (int64_t)r * W);
    const float a = align_depth(__ldg(&vp->depth[q]), s, t);
    bool ok = rest_ok(ctx, q, r, c, a);
    if (ok && (prm.flags & MDI_KEEP_SUBSAMPLE)) {
      const int f = (prm.mode == MDI_SUBSAMPLE_STATIC)
                        ? prm.static_factor : adaptive_factor(a, lo, hi, prm.fmin, prm.fmax);
      ok = divides(r, f) && divides(c, f);
    }
    if (ok) mark_kept(vp, q, keep, bits, tile_counts);
  }
}

}  // namespace mdi

extern "C" int mdi_factor_table(int n_views, const float* lo_hi, int factor_min, int factor_max,
                                float* table, void* stream) {
  MDI_REQUIRE(n_views >= 0, "mdi_factor_table: negative size");
  if (n_views == 0) return MDI_OK;
  MDI_REQUIRE(lo_hi && table, "mdi_factor_table: null pointer");
  mdi::prof_mark("factor_table_kernel", (cudaStream_t)stream);
  mdi::factor_table_kernel<<<(n_views + 3) / 4, 128, 0, (cudaStream_t)stream>>>(
      n_views, lo_hi, factor_min, factor_max, table);
  MDI_CHECK_LAUNCH("mdi_factor_table");
  return MDI_OK;
}

extern "C" int mdi_keep_mask(const MdiView* views, int n_views, int64_t total_tiles,
                             int64_t total_px, int max_height, int max_width,
                             const float* scale_shift,
                             const int32_t* status, const float* lo_hi, const float* grad_range,
                             const uint8_t* patch_blocked, const float* factor_table,
                             const MdiSubsampleConfig* cfg, uint8_t* keep, uint32_t* bits,
                             int32_t* tile_counts, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n_views >= 0 && total_tiles >= 0 && total_px >= 0, "mdi_keep_mask: negative size");
  if (n_views == 0 || total_tiles == 0) return MDI_OK;
  MDI_REQUIRE(views && cfg && bits && tile_counts, "mdi_keep_mask: null pointer");
  MDI_REQUIRE(total_tiles < (1LL << 31), "mdi_keep_mask: too many tiles");
  MDI_REQUIRE(n_views <= 65535, "mdi_keep_mask: more than 65535 views per call");
  MaskParams prm{};
  prm.mode = cfg->mode; prm.static_factor = cfg->static_factor; prm.fmin = cfg->factor_min;
  prm.fmax = cfg->factor_max; prm.flags = cfg->flags; prm.grad_thresh = cfg->grad_thresh;
  bool lattice = false;
  int n_strides = 1, min_stride = 1;
  if (prm.flags & MDI_KEEP_SUBSAMPLE) {
    if (prm.mode == MDI_SUBSAMPLE_STATIC) {
      MDI_REQUIRE(prm.static_factor != 0, "mdi_keep_mask: static_factor == 0 (modulo by zero)");
      // row % f == 0 depends on |f| only (|INT32_MIN| overflows: per-pixel path)
      lattice = prm.static_factor != INT32_MIN;
      if (lattice && prm.static_factor < 0) prm.static_factor = -prm.static_factor;
      min_stride = lattice ? prm.static_factor : 1;
    } else {
      MDI_REQUIRE(prm.mode == MDI_SUBSAMPLE_ADAPTIVE, "mdi_keep_mask: unknown mode %d", prm.mode);
      MDI_REQUIRE(lo_hi, "mdi_keep_mask: adaptive mode needs lo_hi");
      lattice = prm.fmin >= 1 && prm.fmax >= prm.fmin && prm.fmax - prm.fmin < 4096;
      n_strides = prm.fmax - prm.fmin + 1;
      min_stride = prm.fmin;
    }
  }
  MDI_REQUIRE(!(prm.flags & MDI_KEEP_GRAD) || grad_range, "mdi_keep_mask: grad flag needs grad_range");
  cudaStream_t st = (cudaStream_t)stream;
  if (!(prm.flags & MDI_KEEP_PREZEROED)) {
    mdi::prof_mark("keep_zero_fill", st);
    if (keep) MDI_CUDA(cudaMemsetAsync(keep, 0, (size_t)total_px, st), "mdi_keep_mask memset");
    MDI_CUDA(cudaMemsetAsync(bits, 0, (size_t)total_tiles * MDI_TILE_WORDS * 4, st), "mdi_keep_mask memset");
    MDI_CUDA(cudaMemsetAsync(tile_counts, 0, (size_t)total_tiles * 4, st), "mdi_keep_mask memset");
  }
  // the lattice kernel indexes pixels and coordinates in 32 / 16 bits
  if (lattice && max_height > 0 && max_width > 0 && max_height < 65536 && max_width < 65536) {
    const int64_t max_pts = (int64_t)((max_height + min_stride - 1) / min_stride) *
                            ((max_width + min_stride - 1) / min_stride);
    MDI_REQUIRE((max_pts + kCtaPoints - 1) / kCtaPoints < (1LL << 31), "mdi_keep_mask: lattice too large");
    dim3 grid((unsigned)((max_pts + kCtaPoints - 1) / kCtaPoints), n_strides, n_views);
    if (n_strides > 1 && n_strides <= kFlatStrides) {
      int64_t total = 0;
      for (int k = 0; k < n_strides; ++k) {
        const int fk = min_stride + k;
        prm.cta_begin[k] = (int)total;
        total += ((int64_t)((max_height + fk - 1) / fk) * ((max_width + fk - 1) / fk) + kCtaPoints - 1) / kCtaPoints;
      }
      prm.cta_begin[n_strides] = (int)total;
      prm.n_flat = n_strides;
      grid = dim3((unsigned)total, 1, n_views);
    }
    mdi::prof_mark("keep_lattice_kernel", st);
    keep_lattice_kernel<<<grid, kLatticeWarps * 32, 0, st>>>(
        views, scale_shift, status, lo_hi, grad_range, patch_blocked, factor_table, prm, keep, bits,
        tile_counts);
  } else {
    mdi::prof_mark("keep_pixel_kernel", st);
    keep_pixel_kernel<<<(unsigned)total_tiles, kMaskThreads, 0, st>>>(
        views, n_views, scale_shift, status, lo_hi, grad_range, patch_blocked, prm, keep, bits,
        tile_counts);
  }
  MDI_CHECK_LAUNCH("mdi_keep_mask");
  return MDI_OK;
}
