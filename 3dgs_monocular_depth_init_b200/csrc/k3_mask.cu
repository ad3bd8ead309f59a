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
#include <stdlib.h>
#include <string.h>

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

// ---- the lattice walk with the rows staged by bulk copies (adaptive strides, the default) ---------
// keep_lattice_kernel gathers one 32-byte sector per lattice point, and 32-byte sectors fetched one
// by one get about 0.42 of the copy bandwidth out of the 64-byte DRAM bursts (the same ceiling the
// other gather kernels sit at).  But the lattices of the strides fmin..fmax touch EVERY sector of the
// rows they visit (a stride is at most 15 floats), and those rows are known up front: row r is
// visited by every stride that divides it (63 % of the rows for 5..15).  So a warp takes sixteen
// consecutive rows, and for each of them that some stride divides it streams the whole depth row and
// the whole predicted-mask row into shared memory with two 1-D bulk copies (TMA) on an mbarrier,
// two rows in flight per warp (rows wider than 2048 pixels go through the slots in column segments);
// the lattice points of all strides dividing the row are then read from shared memory.  DRAM sees long sequential reads, the per-point work has no global load left except
// the SfM-density patch row (one per ~40 image rows, fetched a row ahead), and the warps need no
// CTA-wide barrier.  The per-point decisions are those of keep_lattice_kernel, operation for operation.
constexpr int kRowWarps = 4;
constexpr int kRowsPerWarp = 16;
constexpr int kRowPatchBytes = 128;   // SfM-density patch row cached per warp (grid_w beyond: global loads)
constexpr int kRowSeg = 2048;         // columns staged per piece: a wider row goes through its slot in segments
constexpr int kRowQueue = 256;        // survivors of the stride test queued per warp (columns), drained by full warps

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(kRowWarps * 32)
keep_rows_kernel(const MdiView* __restrict__ views, const float* __restrict__ scale_shift,
                 const int32_t* __restrict__ status, const float* __restrict__ lo_hi,
                 const uint8_t* __restrict__ patch_blocked, const float* __restrict__ factor_table,
                 MaskParams prm, int slot_bytes, int depth_bytes, uint8_t* __restrict__ keep,
                 uint32_t* __restrict__ bits, int32_t* __restrict__ tile_counts) {
  extern __shared__ __align__(16) unsigned char s_dyn[];   // [warp][slot]{depth row, mask row}, then [warp] patch row
  __shared__ __align__(8) unsigned long long s_bar[kRowWarps][2];
  const int v = blockIdx.y;
  if (status && status[v] != MDI_STATUS_OK) return;
  const MdiView* vp = views + v;
  const int H = vp->H, W = vp->W;
  const int warp = warp_id(), lane = lane_id();
  const int r0 = ((int)blockIdx.x * kRowWarps + warp) * kRowsPerWarp;
  if (r0 >= H) return;   // whole warps leave: the kernel has no CTA-wide barrier
  const int n_strides = prm.fmax - prm.fmin + 1;
  // strides dividing each of the warp's rows: lane k < 16 ends up owning the stride mask of row r0 + k
  // (lane k < n_strides holds r0 mod f_k: one division per warp, the sixteen rows follow by counting up)
  const int my_f = prm.fmin + (lane < n_strides ? lane : 0);
  int rem = r0 % my_f;
  uint32_t fm = 0;
#pragma unroll
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const uint32_t b = __ballot_sync(0xffffffffu, lane < n_strides && rem == 0 && r0 + i < H);
    if (lane == i) fm = b;
    if (++rem == my_f) rem = 0;
  }
  uint32_t todo = __ballot_sync(0xffffffffu, fm != 0);   // rows of the warp that hold lattice points
  if (!todo) return;

  const float* __restrict__ depth = vp->depth;
  const bool use_pm = (prm.flags & MDI_KEEP_PRED_MASK) && vp->mask;
  const uint8_t* __restrict__ pmask = use_pm ? vp->mask : nullptr;
  unsigned char* my = s_dyn + (size_t)warp * 2 * slot_bytes;
  uint8_t* s_patch = s_dyn + (size_t)kRowWarps * 2 * slot_bytes + warp * kRowPatchBytes;
  uint16_t* s_queue = reinterpret_cast<uint16_t*>(s_dyn + (size_t)kRowWarps * 2 * slot_bytes + kRowWarps * kRowPatchBytes) +
                      warp * kRowQueue;
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t bar0 = smem_addr(&s_bar[warp][0]);
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  // the 16-byte aligned part of row r that lies inside the view's own arrays: columns [c0, c1) of the
  // depth row / [m0, m1) of the mask row are staged, the few others (first / last row of a view whose
  // arrays are not 16-byte aligned) are read from global memory
  const uintptr_t d_base = reinterpret_cast<uintptr_t>(depth), d_end = d_base + (size_t)H * W * 4;
  const uintptr_t d_lo_lim = (d_base + 15) & ~(uintptr_t)15, d_hi_lim = d_end & ~(uintptr_t)15;
  const uintptr_t m_base = reinterpret_cast<uintptr_t>(pmask), m_end = m_base + (size_t)H * W;
  const uintptr_t m_lo_lim = (m_base + 15) & ~(uintptr_t)15, m_hi_lim = m_end & ~(uintptr_t)15;
  auto span = [&](uintptr_t a, size_t bytes, uintptr_t lo_lim, uintptr_t hi_lim, uintptr_t* lo, uintptr_t* hi) {
    uintptr_t l = a & ~(uintptr_t)15, h = (a + bytes + 15) & ~(uintptr_t)15;
    if (l < lo_lim) l = lo_lim;
    if (h > hi_lim) h = hi_lim;
    if (h < l) h = l;
    *lo = l; *hi = h;
  };
  const int n_seg = (W + kRowSeg - 1) / kRowSeg;
  auto issue = [&](int r, int seg, int slot) {   // lane 0: columns of segment `seg` of image row r, depth and mask, one barrier
    const int seg_lo = seg * kRowSeg, seg_len = min(W - seg_lo, kRowSeg);
    uintptr_t dl, dh, ml = 0, mh_ = 0;
    span(d_base + ((size_t)r * W + seg_lo) * 4, (size_t)seg_len * 4, d_lo_lim, d_hi_lim, &dl, &dh);
    if (pmask) span(m_base + (size_t)r * W + seg_lo, (size_t)seg_len, m_lo_lim, m_hi_lim, &ml, &mh_);
    const uint32_t db = (uint32_t)(dh - dl), mb = (uint32_t)(mh_ - ml);
    const uint32_t bar = bar0 + 8 * slot;
    const uint32_t dst = smem_addr(my + (size_t)slot * slot_bytes);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(db + mb) : "memory");
    if (db)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "l"(dl), "r"(db), "r"(bar) : "memory");
    if (mb)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst + (uint32_t)depth_bytes), "l"(ml), "r"(mb), "r"(bar) : "memory");
  };
  // issue cursor (the same in every lane; lane 0 issues): next (row, segment) piece to put in flight
  uint32_t iss_rows = todo;
  int iss_seg = 0;
  auto issue_next = [&](int slot) {
    if (!iss_rows) return;
    if (lane == 0) issue(r0 + __ffs(iss_rows) - 1, iss_seg, slot);
    if (++iss_seg == n_seg) { iss_seg = 0; iss_rows &= iss_rows - 1; }
  };
  issue_next(0);
  issue_next(1);

  // per-view constants: loaded while the first rows are on their way
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const uint8_t* __restrict__ pb =
      ((prm.flags & MDI_KEEP_PATCH) && patch_blocked) ? patch_blocked + (int64_t)v * MDI_PATCH_STRIDE : nullptr;
  const int patch_h = vp->patch_h, patch_w = vp->patch_w, grid_h = vp->grid_h, grid_w = vp->grid_w;
  const uint32_t mh = 0xFFFFFFFFu / (uint32_t)max(patch_h, 1) + 1u, mw = 0xFFFFFFFFu / (uint32_t)max(patch_w, 1) + 1u;
  // breakpoints of the view: lane L holds B[L] = row[1 + L], L <= n_strides
  const float* trow = factor_table ? factor_table + (int64_t)v * MDI_FACTOR_TABLE_STRIDE : nullptr;
  const bool table = trow && trow[0] != 0.f;
  const float tab = (table && lane <= n_strides) ? trow[1 + lane] : 0.f;
  // with MDI_KEEP_NONNEG a lattice point of stride f_k is a candidate iff a >= max(B[k+1], 0) && !(a >= B[k])
  // (a >= 0 also rules out the sentinel and a NaN): lane k holds that lower bound
  const float my_lo = fmaxf(__shfl_down_sync(0xffffffffu, tab, 1), 0.f);
  const bool fast_cfg = table && (prm.flags & MDI_KEEP_NONNEG);
  const float vlo = lo_hi[2 * v], vhi = lo_hi[2 * v + 1];


  int cached_pr = -1;   // patch row held in s_patch
  int i = 0;            // pieces consumed so far: piece i sits in slot i & 1, (i / 2)-th use of it
  while (todo) {
    const int k = __ffs(todo) - 1;
    todo &= todo - 1;
    const int r = r0 + k;
    const uint32_t fmr = __shfl_sync(0xffffffffu, fm, k);
    // SfM-density patch row of this image row (changes every patch_h rows): fetched before the wait
    const int pr = pb ? (patch_h <= 1 ? r : (int)__umulhi((uint32_t)r, mh)) : -1;
    const bool pr_in = pb && pr < grid_h;
    if (pr_in && pr != cached_pr && grid_w <= kRowPatchBytes) {
      __syncwarp();
      for (int c = lane; c < grid_w; c += 32) s_patch[c] = __ldg(&pb[pr * grid_w + c]);
      cached_pr = pr;
      __syncwarp();
    }
    for (int seg = 0; seg < n_seg; ++seg, ++i) {
    const int slot = i & 1;
    const int seg_lo = seg * kRowSeg, seg_hi = min(W, seg_lo + kRowSeg);
    {   // wait for the piece (phase of this slot's (i / 2)-th use)
      const uint32_t bar = bar0 + 8 * slot, parity = (uint32_t)(i >> 1) & 1u;
      uint32_t ok = 0;
      while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
      }
    }
    const int64_t rowpx = (int64_t)r * W;
    uintptr_t dl, dh, ml = 0, mh_ = 0;
    span(d_base + ((size_t)rowpx + seg_lo) * 4, (size_t)(seg_hi - seg_lo) * 4, d_lo_lim, d_hi_lim, &dl, &dh);
    if (pmask) span(m_base + (size_t)rowpx + seg_lo, (size_t)(seg_hi - seg_lo), m_lo_lim, m_hi_lim, &ml, &mh_);
    // column c of the row sits at srow[c] for c0 <= c < c1 (srow points in front of the slot)
    const intptr_t d_off = (intptr_t)dl - (intptr_t)(d_base + (size_t)rowpx * 4);   // bytes from the row start, multiple of 4
    const float* srow = reinterpret_cast<const float*>(my + (size_t)slot * slot_bytes) - d_off / 4;
    const int c0 = max(seg_lo, (int)(d_off / 4));
    const int c1 = (int)min((intptr_t)seg_hi, ((intptr_t)dh - (intptr_t)(d_base + (size_t)rowpx * 4)) / 4);
    const intptr_t m_off = (intptr_t)ml - (intptr_t)(m_base + (size_t)rowpx);
    const uint8_t* smask = my + (size_t)slot * slot_bytes + depth_bytes - m_off;
    const int m0 = max(seg_lo, (int)m_off);
    const int m1 = (int)min((intptr_t)seg_hi, (intptr_t)mh_ - (intptr_t)(m_base + (size_t)rowpx));
    const float* grow = depth + rowpx;

    // survivors of the stride test: predicted mask, SfM-density patch, then the outputs
    auto finish_point = [&](const int c) {
      if (pmask) {
        const uint8_t mbyte = (c >= m0 && c < m1) ? smask[c] : __ldg(pmask + rowpx + c);
        if (mbyte == 0) return;
      }
      if (pr_in) {
        const int pc = patch_w <= 1 ? c : (int)__umulhi((uint32_t)c, mw);
        if (pc < grid_w) {
          const uint8_t blocked = grid_w <= kRowPatchBytes ? s_patch[pc] : __ldg(&pb[pr * grid_w + pc]);
          if (blocked) return;
        }
      }
      mark_kept(vp, rowpx + c, keep, bits, tile_counts);
    };
    // About a quarter of the lattice points pass the stride test, a few per warp step: finishing them
    // where they are found would run the survivor code at a quarter of the lanes nearly every step.
    // They are queued (column only: the row is fixed) by ballot rank and finished 32 at a time.
    int n_q = 0;   // warp-uniform
    auto drain = [&]() {
      __syncwarp();
      for (int qi = lane; qi < n_q; qi += 32) finish_point((int)s_queue[qi]);
      __syncwarp();
      n_q = 0;
    };
    auto enqueue = [&](const bool hit, const int c) {   // all lanes together
      const uint32_t m = __ballot_sync(0xffffffffu, hit);
      if (hit) s_queue[n_q + __popc(m & lt)] = (uint16_t)c;
      n_q += __popc(m);
    };
    const bool staged = c0 == seg_lo && c1 >= seg_hi;   // the piece's depths are all in shared memory (all but a view's first / last row)
    for (uint32_t left = fmr; left; left &= left - 1) {
      const int ks = __ffs(left) - 1;
      const int f = prm.fmin + ks;
      const float b_lo = __shfl_sync(0xffffffffu, tab, ks + 1), b_hi = __shfl_sync(0xffffffffu, tab, ks);
      const float lo_eff = __shfl_sync(0xffffffffu, my_lo, ks);
      const int step = 32 * f;
      const int c_begin = seg_lo ? ((seg_lo + f - 1) / f) * f : 0;   // first lattice column of the piece
      const int c_end = seg_hi;   // (a lattice column is < W, so "in the piece" is all there is to test)
      // uniform trip count: every lane takes part in the ballots, lanes past the lattice with hit = false
      for (int cb = c_begin; cb < c_end; cb += 2 * step) {
        if (n_q > kRowQueue - 64) drain();
        const int ca = cb + lane * f, cc = ca + step;
        const bool in0 = ca < c_end, in1 = cc < c_end;
        bool h0, h1;
        if (fast_cfg && staged) {
          // (a >= 0 rules out the sentinel and a NaN; the lower bound carries it)
          const float a0 = align_depth(srow[in0 ? ca : seg_lo], s, t), a1 = align_depth(srow[in1 ? cc : seg_lo], s, t);
          h0 = in0 && (a0 >= lo_eff) && !(a0 >= b_hi);
          h1 = in1 && (a1 >= lo_eff) && !(a1 >= b_hi);
        } else {
          auto test = [&](const bool in, const int c) {
            if (!in) return false;
            const float dv = (c >= c0 && c < c1) ? srow[c] : __ldg(grow + c);
            const float a = align_depth(dv, s, t);
            bool hit;
            if (table && a == a) {
              hit = (a >= b_lo) && !(a >= b_hi);
            } else {
              // exact arithmetic; a NaN stride (INT32_MIN) keeps index 0 only: visit it once, at fmin
              const int fa = adaptive_factor(a, vlo, vhi, prm.fmin, prm.fmax);
              hit = (fa == INT32_MIN) ? (r == 0 && c == 0 && f == prm.fmin) : (fa == f);
            }
            if (prm.flags & MDI_KEEP_SENTINEL) hit = hit && (a != -42.0f);
            if (prm.flags & MDI_KEEP_NONNEG) hit = hit && (a >= 0.f);
            return hit;
          };
          h0 = test(in0, ca);
          h1 = test(in1, cc);
        }
        enqueue(h0, ca);
        enqueue(h1, cc);
      }
    }
    drain();
    __syncwarp();   // every lane is done with the slot
    issue_next(slot);
    }
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
    // adaptive strides without the gradient test: rows staged by bulk copies (MDI_KEEP_KERNEL=lattice keeps the gathers)
    static const bool rows_ok = [] {
      const char* e = getenv("MDI_KEEP_KERNEL");
      return !(e && strcmp(e, "lattice") == 0);
    }();
    const bool rows = rows_ok && prm.mode == MDI_SUBSAMPLE_ADAPTIVE && n_strides > 1 && n_strides <= 31 &&
                      !(prm.flags & MDI_KEEP_GRAD) && min_stride >= 1;
    if (rows) {
      const int piece = max_width < kRowSeg ? max_width : kRowSeg;       // columns staged at a time
      const int depth_bytes = ((piece * 4 + 32) + 15) & ~15;            // piece + the aligned-superset slack
      const int mask_bytes = ((piece + 32) + 15) & ~15;
      const int slot_bytes = depth_bytes + mask_bytes;
      const size_t smem = (size_t)kRowWarps * 2 * slot_bytes + kRowWarps * kRowPatchBytes + kRowWarps * kRowQueue * 2;
      MDI_CUDA(cudaFuncSetAttribute(keep_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
               "mdi_keep_mask attr");
      dim3 rgrid((unsigned)((max_height + kRowWarps * kRowsPerWarp - 1) / (kRowWarps * kRowsPerWarp)), (unsigned)n_views);
      mdi::prof_mark("keep_rows_kernel", st);
      keep_rows_kernel<<<rgrid, kRowWarps * 32, smem, st>>>(views, scale_shift, status, lo_hi, patch_blocked,
                                                           factor_table, prm, slot_bytes, depth_bytes, keep, bits,
                                                           tile_counts);
    } else {
    mdi::prof_mark("keep_lattice_kernel", st);
    keep_lattice_kernel<<<grid, kLatticeWarps * 32, 0, st>>>(
        views, scale_shift, status, lo_hi, grad_range, patch_blocked, factor_table, prm, keep, bits,
        tile_counts);
    }
  } else {
    mdi::prof_mark("keep_pixel_kernel", st);
    keep_pixel_kernel<<<(unsigned)total_tiles, kMaskThreads, 0, st>>>(
        views, n_views, scale_shift, status, lo_hi, grad_range, patch_blocked, prm, keep, bits,
        tile_counts);
  }
  MDI_CHECK_LAUNCH("mdi_keep_mask");
  return MDI_OK;
}
