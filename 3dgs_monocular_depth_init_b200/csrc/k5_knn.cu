// k5_knn.cu -- exact k nearest neighbours over the final cloud, and the Local Outlier Factor
// built on them (a12; also the kNN of SURVEY 8f rank 1).  Reference:
//   point_cloud_postprocess/postprocess.py:16-22  sklearn LocalOutlierFactor(n_neighbors=k).fit_predict
//   utils/runner_utils.py:143-147                 sklearn NearestNeighbors(K).kneighbors(x) (self included)
// scikit-learn (1.9.0 in the image) is a third-party dependency of the reference; its published
// algorithm is restated: kNN excluding the sample itself, k-distance = distance to the k-th
// neighbour, reach(p,o) = max(d(p,o), kdist(o)), lrd(p) = 1 / (mean_o reach(p,o) + 1e-10),
// LOF(p) = mean_o lrd(o) / lrd(p), outlier iff -LOF < -1.5 (contamination="auto").  Distances
// follow sklearn's KDTree (its 64-bit tree is the one neighbors/_base.py builds, also for fp32
// input): differences, squares, sum and square root in fp64, then the cast to fp32 that
// _lof.py:299-302 applies to fp32 input (and `knn`'s `.to(x)`); every LOF array stays fp32 as
// numpy keeps it (the reference passes pts.cpu().numpy(), fp32).
//
// Layout: points are counting-sorted by the 24-bit Morton code of their cell in a 256^3 grid over
// the cube bounding box, so that every node of the implicit octree (a Morton prefix) is one
// contiguous slice [start[m << s], start[(m+1) << s]).  A query walks that octree depth first,
// nearest child first, pruning nodes farther than the current k-th distance: exact, no ring
// search, bounded work for isolated points.  One thread per query, queries in sorted order so
// that a warp walks neighbouring paths; the k-heap lives in local memory (L1).
#include <float.h>

#include "common.cuh"

namespace mdi {

constexpr int kKnnThreads = 128;
// Grid depth by point count: 256 cells per axis (16.7 M cells) up to kDenseGridPoints points, 512
// (134 M cells) above -- a C5-sized cloud (5e7 points on surfaces) fills the cells of a 256^3 grid with
// hundreds of points each, and every leaf is scanned whole.  MDI_KNN_BITS=8|9 overrides.
constexpr int kMaxGridBits = 9;
constexpr int64_t kDenseGridPoints = 8 * 1000 * 1000;
static int grid_bits_for(int64_t n) {
  static const int forced = [] {
    const char* e = getenv("MDI_KNN_BITS");
    const int b = e ? atoi(e) : 0;
    return (b == 8 || b == 9) ? b : 0;
  }();
  return forced ? forced : (n > kDenseGridPoints ? 9 : 8);
}
static int64_t grid_cells_for(int64_t n) { return (int64_t)1 << (3 * grid_bits_for(n)); }
constexpr int kLeafPoints = 24;                    // nodes this small are scanned, not split
constexpr int kKnnMaxK = 64;

constexpr int kHistBins = 1024;
struct KnnGrid { float lo[3]; float size; float inv_cell; float slack; int bits; int cells; };

struct KnnWs {
  uint32_t* keys;      // [6] min/max keys of the coordinates
  float* range;        // [3][6] per-axis (lo x3, hi x3): min/max, then two quantile refinements
  uint32_t* hist;      // [2][3][kHistBins] coordinate histograms of the two refinements
  KnnGrid* grid;
  int32_t* start;      // [cells + 1] first sorted position of each Morton cell
  int32_t* cursor;     // [cells] scatter cursors
  int32_t* blocksum;   // [cells / 1024 + 1]
  uint32_t* cell_of;   // [n] Morton cell of every input point
  float4* sorted;      // [n] (x, y, z, original index bits) in Morton order
};

static inline int64_t knn_align(int64_t x) { return (x + 255) & ~(int64_t)255; }

static int64_t knn_carve(KnnWs* w, char* base, int64_t n) {
  int64_t off = 0;
  auto take = [&](int64_t bytes) { char* p = base ? base + off : nullptr; off += knn_align(bytes); return p; };
  KnnWs t;
  t.keys = (uint32_t*)take(64);
  t.range = (float*)take(3 * 6 * 4);
  t.hist = (uint32_t*)take(2 * 3 * kHistBins * 4);
  t.grid = (KnnGrid*)take(sizeof(KnnGrid));
  const int64_t cells = grid_cells_for(n);
  t.start = (int32_t*)take((cells + 1) * 4);
  t.cursor = (int32_t*)take(cells * 4);
  t.blocksum = (int32_t*)take((cells / 1024 + 1) * 4);
  t.cell_of = (uint32_t*)take(n * 4);
  t.sorted = (float4*)take(n * 16);
  if (w) *w = t;
  return off;
}

__device__ __forceinline__ uint32_t spread3(uint32_t v) {  // 10 bits -> every third bit
  v &= 0x000003FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
__device__ __forceinline__ uint32_t compact3(uint32_t v) {
  v &= 0x09249249u;
  v = (v | (v >> 2)) & 0x030C30C3u;
  v = (v | (v >> 4)) & 0x0300F00Fu;
  v = (v | (v >> 8)) & 0x030000FFu;
  v = (v | (v >> 16)) & 0x000003FFu;
  return v;
}

__global__ void knn_bbox_init_kernel(uint32_t* keys) {
  if (threadIdx.x < 3) { keys[threadIdx.x] = 0xFFFFFFFFu; keys[3 + threadIdx.x] = 0u; }
}

__global__ void __launch_bounds__(256) knn_bbox_kernel(const float* __restrict__ pts, int64_t n, uint32_t* keys) {
  uint32_t kmin[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, kmax[3] = {0u, 0u, 0u};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) {
      const float c = pts[3 * i + a];
      if (c == c && fabsf(c) <= FLT_MAX) { const uint32_t k = float_to_key(c); kmin[a] = min(kmin[a], k); kmax[a] = max(kmax[a], k); }
    }
  for (int a = 0; a < 3; ++a) {
    kmin[a] = __reduce_min_sync(0xffffffffu, kmin[a]);
    kmax[a] = __reduce_max_sync(0xffffffffu, kmax[a]);
  }
  if (lane_id() == 0)
    for (int a = 0; a < 3; ++a) { atomicMin(&keys[a], kmin[a]); atomicMax(&keys[3 + a], kmax[a]); }
}

// The grid covers a ROBUST cube, not the bounding cube: a handful of far-away points (a view whose
// alignment went wild) would otherwise stretch the cube and leave the real cloud in a few cells.
// Per axis the 0.1 % .. 99.9 % coordinate range is located by two histogram refinements (1024 bins over
// [min, max], then 1024 bins over the bins picked there); points outside are clamped into the border
// cells, and the search treats border cells as open-ended, so it stays exact.
__global__ void knn_range_init_kernel(const uint32_t* __restrict__ keys, float* __restrict__ range) {
  const int a = threadIdx.x;
  if (a >= 3) return;
  float lo = keys[a] == 0xFFFFFFFFu ? 0.f : key_to_float(keys[a]);
  float hi = keys[3 + a] == 0u ? 0.f : key_to_float(keys[3 + a]);
  if (!(hi >= lo)) { lo = 0.f; hi = 0.f; }
  range[a] = lo; range[3 + a] = hi;
}

__global__ void __launch_bounds__(256)
knn_hist_kernel(const float* __restrict__ pts, int64_t n, const float* __restrict__ range, uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_h[3 * kHistBins];
  for (int i = threadIdx.x; i < 3 * kHistBins; i += blockDim.x) s_h[i] = 0u;
  __syncthreads();
  float lo[3], inv[3];
  for (int a = 0; a < 3; ++a) {
    lo[a] = range[a];
    const float w = range[3 + a] - lo[a];
    inv[a] = (w > 0.f && w <= FLT_MAX) ? (float)kHistBins / w : 0.f;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) {
      const float c = pts[3 * i + a];
      if (c == c && fabsf(c) <= FLT_MAX) {
        const int b = min(kHistBins - 1, max(0, (int)((c - lo[a]) * inv[a])));
        atomicAdd(&s_h[a * kHistBins + b], 1u);
      }
    }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * kHistBins; i += blockDim.x)
    if (s_h[i]) atomicAdd(&hist[i], s_h[i]);
}

// one warp per axis: bins holding the 0.1 % and 99.9 % ranks -> the next, narrower range
__global__ void knn_hist_pick_kernel(const uint32_t* __restrict__ hist, const float* __restrict__ range,
                                     float* __restrict__ next) {
  const int a = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (a >= 3) return;
  const uint32_t* h = hist + a * kHistBins;
  uint32_t part = 0;
  for (int b = lane; b < kHistBins; b += 32) part += h[b];
  const uint32_t total = __reduce_add_sync(0xffffffffu, part);
  if (lane != 0) return;
  const float lo = range[a], hi = range[3 + a];
  int b_lo = 0, b_hi = kHistBins - 1;
  if (total > 0) {
    const uint32_t cut = total / 1000u;   // points allowed outside on each side
    uint32_t acc = 0;
    for (b_lo = 0; b_lo < kHistBins - 1; ++b_lo) { acc += h[b_lo]; if (acc > cut) break; }
    acc = 0;
    for (b_hi = kHistBins - 1; b_hi > b_lo; --b_hi) { acc += h[b_hi]; if (acc > cut) break; }
  }
  const float w = (hi - lo) / (float)kHistBins;
  const float nlo = lo + (float)b_lo * w, nhi = lo + (float)(b_hi + 1) * w;
  next[a] = (w > 0.f && nlo >= lo) ? nlo : lo;
  next[3 + a] = (w > 0.f && nhi <= hi) ? nhi : hi;
}

__global__ void knn_grid_kernel(const float* __restrict__ range, KnnGrid* g, int bits) {
  float size = 0.f;
  for (int a = 0; a < 3; ++a) {
    const float lo = range[a], hi = range[3 + a];
    g->lo[a] = lo;
    size = fmaxf(size, hi - lo);
  }
  if (!(size > 0.f) || !(size <= FLT_MAX)) size = 1.f;
  g->size = size;
  g->bits = bits; g->cells = 1 << (3 * bits);
  g->inv_cell = (float)(1 << bits) / size;
  g->slack = size * 8e-6f;  // covers the rounding of the cell assignment below
}

__device__ __forceinline__ uint32_t cell_of_point(const KnnGrid& g, float x, float y, float z) {
  const int top = (1 << g.bits) - 1;
  const int cx = min(top, max(0, (int)((x - g.lo[0]) * g.inv_cell)));
  const int cy = min(top, max(0, (int)((y - g.lo[1]) * g.inv_cell)));
  const int cz = min(top, max(0, (int)((z - g.lo[2]) * g.inv_cell)));
  return spread3(cx) | (spread3(cy) << 1) | (spread3(cz) << 2);
}

__global__ void __launch_bounds__(256)
knn_count_kernel(const float* __restrict__ pts, int64_t n, const KnnGrid* __restrict__ grid,
                 uint32_t* __restrict__ cell_of, int32_t* __restrict__ start) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const KnnGrid g = *grid;
  const uint32_t c = cell_of_point(g, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
  cell_of[i] = c;
  atomicAdd(&start[c], 1);
}

// exclusive scan of the cell counts: per-1024 block scan, scan of the block totals, add back
__global__ void __launch_bounds__(1024) knn_scan_local_kernel(int32_t* start, int32_t* blocksum) {
  __shared__ int s_scan[33];
  const int i = blockIdx.x * 1024 + threadIdx.x;
  int total;
  const int ex = block_excl_scan(start[i], s_scan, &total);
  start[i] = ex;
  if (threadIdx.x == 0) blocksum[blockIdx.x] = total;
}
__global__ void __launch_bounds__(1024) knn_scan_top_kernel(int32_t* blocksum, int n_blocks, int32_t* start, int cells) {
  __shared__ int s_scan[33];
  int carry = 0;
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = (i < n_blocks) ? blocksum[i] : 0;
    int total;
    const int ex = block_excl_scan(v, s_scan, &total);
    if (i < n_blocks) blocksum[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) start[cells] = carry;
}
__global__ void __launch_bounds__(1024)
knn_scan_add_kernel(int32_t* start, const int32_t* __restrict__ blocksum, int32_t* cursor) {
  const int i = blockIdx.x * 1024 + threadIdx.x;
  const int v = start[i] + blocksum[blockIdx.x];
  start[i] = v;
  cursor[i] = v;
}

__global__ void __launch_bounds__(256)
knn_scatter_kernel(const float* __restrict__ pts, int64_t n, const uint32_t* __restrict__ cell_of,
                   int32_t* __restrict__ cursor, float4* __restrict__ sorted) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int q = atomicAdd(&cursor[cell_of[i]], 1);
  sorted[q] = make_float4(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], __int_as_float((int)i));
}

// (d2, original index) max-heap: the index breaks distance ties so that the neighbour set does not
// depend on the order of the points inside a cell
struct KnnHeap {
  double d[kKnnMaxK];
  int id[kKnnMaxK];
  int n, k;
  __device__ __forceinline__ bool less(double da, int ia, double db, int ib) const {
    return da < db || (da == db && ia < ib);
  }
  __device__ __forceinline__ void push(double dv, int iv) {
    if (n < k) {
      int c = n++;
      while (c > 0) {
        const int p = (c - 1) >> 1;
        if (!less(d[p], id[p], dv, iv)) break;
        d[c] = d[p]; id[c] = id[p]; c = p;
      }
      d[c] = dv; id[c] = iv;
    } else if (less(dv, iv, d[0], id[0])) {
      sift_down(dv, iv, n);
    }
  }
  // places (dv, iv) at the root of a heap of `len` entries and restores the heap
  __device__ __forceinline__ void sift_down(double dv, int iv, int len) {
    int c = 0;
    for (;;) {
      int ch = 2 * c + 1;
      if (ch >= len) break;
      if (ch + 1 < len && less(d[ch], id[ch], d[ch + 1], id[ch + 1])) ++ch;
      if (!less(dv, iv, d[ch], id[ch])) break;
      d[c] = d[ch]; id[c] = id[ch]; c = ch;
    }
    d[c] = dv; id[c] = iv;
  }
};

__global__ void __launch_bounds__(kKnnThreads)
knn_query_kernel(const float4* __restrict__ sorted, const int32_t* __restrict__ start, int64_t n,
                 const KnnGrid* __restrict__ grid, int k, int exclude_self, int32_t* __restrict__ out_idx,
                 float* __restrict__ out_dist, float* __restrict__ out_kdist, int64_t q_begin, int64_t q_end,
                 int32_t* __restrict__ out_self) {
  // queries = sorted positions q_begin .. q_end - 1.  out_self == nullptr: results go to the row of the
  // point's ORIGINAL index; otherwise (query split over ranks) to the compact row p - q_begin, and
  // out_self records the original index of that row
  const int64_t p = q_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= q_end) return;
  const KnnGrid g = *grid;
  const float4 q = sorted[p];
  const int self = __float_as_int(q.w);
  const int64_t row = out_self ? p - q_begin : (int64_t)self;
  if (out_self) out_self[row] = self;
  KnnHeap h;
  h.n = 0; h.k = k;
  uint32_t stk[7 * kMaxGridBits + 2];   // DFS: a popped node pushes at most 8 children
  int sp = 0;
  stk[sp++] = 0u;  // level 0, prefix 0
  while (sp > 0) {
    const uint32_t ent = stk[--sp];
    const int l = (int)(ent >> 28);
    const uint32_t m = ent & 0x0FFFFFFFu;
    const int shift = 3 * (g.bits - l);
    const int s = start[m << shift], e = start[(m + 1u) << shift];
    const float cs = g.size / (float)(1 << l);
    const float bx = g.lo[0] + (float)compact3(m) * cs, by = g.lo[1] + (float)compact3(m >> 1) * cs,
                bz = g.lo[2] + (float)compact3(m >> 2) * cs;
    if (h.n == k) {
      // distance from the query to the node's box; a node on the border of the grid also holds the
      // points clamped into it from outside the cube, so its box is open on that side
      const uint32_t last = (1u << l) - 1u;
      const uint32_t ux = compact3(m), uy = compact3(m >> 1), uz = compact3(m >> 2);
      const float tx = fmaxf(fmaxf(ux == 0u ? -INFINITY : bx - q.x, ux == last ? -INFINITY : q.x - (bx + cs)) - g.slack, 0.f);
      const float ty = fmaxf(fmaxf(uy == 0u ? -INFINITY : by - q.y, uy == last ? -INFINITY : q.y - (by + cs)) - g.slack, 0.f);
      const float tz = fmaxf(fmaxf(uz == 0u ? -INFINITY : bz - q.z, uz == last ? -INFINITY : q.z - (bz + cs)) - g.slack, 0.f);
      const double d2 = (double)(tx * tx + ty * ty + tz * tz) * (1.0 - 1e-6);
      if (d2 > h.d[0]) continue;
    }
    if (l == g.bits || e - s <= kLeafPoints) {
      for (int j = s; j < e; ++j) {
        const float4 c = sorted[j];
        const int ci = __float_as_int(c.w);
        if (exclude_self && ci == self) continue;
        const double dx = (double)q.x - (double)c.x, dy = (double)q.y - (double)c.y, dz = (double)q.z - (double)c.z;
        const double d2 = (dx * dx + dy * dy) + dz * dz;
        if (d2 == d2) h.push(d2, ci);
      }
      continue;
    }
    const float half = 0.5f * cs;
    const uint32_t oct = (q.x >= bx + half ? 1u : 0u) | (q.y >= by + half ? 2u : 0u) | (q.z >= bz + half ? 4u : 0u);
    const int cshift = shift - 3;
    const uint32_t base = m << 3;
#pragma unroll 1
    for (int i = 7; i >= 0; --i) {  // pushed far to near, popped near to far
      const uint32_t c = base | (oct ^ (uint32_t)i);
      if (start[c << cshift] != start[(c + 1u) << cshift]) stk[sp++] = ((uint32_t)(l + 1) << 28) | c;
    }
  }
  // heap sort: ascending (distance, index)
  const int found = h.n;
  for (int len = found; len > 1; --len) {
    const double dv = h.d[len - 1]; const int iv = h.id[len - 1];
    h.d[len - 1] = h.d[0]; h.id[len - 1] = h.id[0];
    h.sift_down(dv, iv, len - 1);
  }
  for (int j = 0; j < k; ++j) {
    const bool ok = j < found;
    out_idx[row * k + j] = ok ? h.id[j] : -1;
    if (out_dist) out_dist[row * k + j] = ok ? (float)sqrt(h.d[j]) : INFINITY;
  }
  if (out_kdist) out_kdist[row] = found == k ? (float)sqrt(h.d[k - 1]) : INFINITY;
}

// numpy's float32 mean over a contiguous axis of k values: for 8 <= k <= 128 eight running sums,
// tree-combined, remainder added in order (pairwise_sum in numpy/core/src/umath/loops_utils.h),
// then one division
__device__ __forceinline__ float numpy_mean_f32(const float* a, int k) {
  float res;
  if (k < 8) {
    res = 0.f;
    for (int i = 0; i < k; ++i) res += a[i];
  } else {
    float r[8];
    for (int j = 0; j < 8; ++j) r[j] = a[j];
    int i = 8;
    for (; i < k - (k % 8); i += 8)
      for (int j = 0; j < 8; ++j) r[j] += a[i + j];
    res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < k; ++i) res += a[i];
  }
  return res / (float)k;
}

// sklearn _lof.py _local_reachability_density on float32 input: every array stays float32
__global__ void __launch_bounds__(kKnnThreads)
lof_lrd_kernel(int64_t n, int k, const int32_t* __restrict__ idx, const float* __restrict__ dist,
               const float* __restrict__ kdist, float* __restrict__ lrd) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float reach[kKnnMaxK];
  for (int j = 0; j < k; ++j) {
    const float d = dist[i * k + j];
    const float kd = kdist[idx[i * k + j]];
    reach[j] = d < kd ? kd : d;  // np.maximum(dist, k-distance of the neighbour)
  }
  lrd[i] = 1.0f / (numpy_mean_f32(reach, k) + 1e-10f);
}

// sklearn _lof.py fit: negative_outlier_factor_ = -mean(lrd[neighbours] / lrd[self]); offset_ = -1.5
__global__ void __launch_bounds__(kKnnThreads)
lof_score_kernel(int64_t n, int k, const int32_t* __restrict__ idx, const float* __restrict__ lrd,
                 const float* __restrict__ lrd_self, float* __restrict__ lof, uint8_t* __restrict__ outlier) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float mine = lrd_self[i];
  float ratio[kKnnMaxK];
  for (int j = 0; j < k; ++j) ratio[j] = lrd[idx[i * k + j]] / mine;
  const float v = numpy_mean_f32(ratio, k);
  if (lof) lof[i] = v;
  outlier[i] = (-v < -1.5f) ? 1 : 0;
}

}  // namespace mdi

extern "C" int64_t mdi_knn_workspace_bytes(int64_t n) {
  if (n < 0) return -1;
  return mdi::knn_carve(nullptr, nullptr, n > 0 ? n : 1);
}

namespace mdi {
// rank boundaries of a query split: bounds[r] = first sorted position of the first cell that starts
// at or after r * ceil(n / world).  Cell-aligned, because the order of the points INSIDE a cell comes
// from atomics and differs from rank to rank: whole cells keep the ranks' query sets a partition.
__global__ void knn_split_kernel(const int32_t* __restrict__ start, int64_t n, int world, int64_t* __restrict__ bounds,
                                 int cells) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r > world) return;
  if (r == 0) { bounds[0] = 0; return; }
  if (r == world) { bounds[world] = n; return; }
  const int64_t per = (n + world - 1) / world;
  const int64_t want = (int64_t)r * per;
  int lo = 0, hi = cells;    // smallest c with start[c] >= want (start[cells] = n)
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((int64_t)start[mid] >= want) hi = mid; else lo = mid + 1;
  }
  bounds[r] = start[lo];
}

}  // namespace mdi

static int knn_check(const float* pts, int64_t n, void* workspace, const char* who) {
  MDI_REQUIRE(n >= 0 && n < (1LL << 31) - 2048, "%s: n out of range", who);
  MDI_REQUIRE(n == 0 || (pts && workspace), "%s: null pointer", who);
  return MDI_OK;
}

// grid build: bounding cube, Morton cell of every point, counting sort into `sorted`
extern "C" int mdi_knn_build(const float* pts, int64_t n, void* workspace, void* stream) {
  using namespace mdi;
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = knn_check(pts, n, workspace, "mdi_knn_build")) return rc;
  if (n == 0) return MDI_OK;
  KnnWs w;
  knn_carve(&w, (char*)workspace, n);
  const unsigned grid_n = (unsigned)((n + 255) / 256);
  const int bits = grid_bits_for(n);
  const int cells = 1 << (3 * bits);
  MDI_CUDA(cudaMemsetAsync(w.start, 0, ((size_t)cells + 1) * 4, st), "mdi_knn_build");
  prof_mark("knn_bbox_kernel", st);
  knn_bbox_init_kernel<<<1, 32, 0, st>>>(w.keys);
  knn_bbox_kernel<<<kSMs * 8, 256, 0, st>>>(pts, n, w.keys);
  MDI_CUDA(cudaMemsetAsync(w.hist, 0, sizeof(uint32_t) * 2 * 3 * kHistBins, st), "mdi_knn_build");
  knn_range_init_kernel<<<1, 32, 0, st>>>(w.keys, w.range);
  for (int r = 0; r < 2; ++r) {
    knn_hist_kernel<<<kSMs * 4, 256, 0, st>>>(pts, n, w.range + 6 * r, w.hist + r * 3 * kHistBins);
    knn_hist_pick_kernel<<<1, 96, 0, st>>>(w.hist + r * 3 * kHistBins, w.range + 6 * r, w.range + 6 * (r + 1));
  }
  knn_grid_kernel<<<1, 1, 0, st>>>(w.range + 12, w.grid, bits);
  prof_mark("knn_count_kernel", st);
  knn_count_kernel<<<grid_n, 256, 0, st>>>(pts, n, w.grid, w.cell_of, w.start);
  prof_mark("knn_scan_kernels", st);
  knn_scan_local_kernel<<<cells / 1024, 1024, 0, st>>>(w.start, w.blocksum);
  knn_scan_top_kernel<<<1, 1024, 0, st>>>(w.blocksum, cells / 1024, w.start, cells);
  knn_scan_add_kernel<<<cells / 1024, 1024, 0, st>>>(w.start, w.blocksum, w.cursor);
  prof_mark("knn_scatter_kernel", st);
  knn_scatter_kernel<<<grid_n, 256, 0, st>>>(pts, n, w.cell_of, w.cursor, w.sorted);
  MDI_CHECK_LAUNCH("mdi_knn_build");
  return MDI_OK;
}

extern "C" int mdi_knn_split(void* workspace, int64_t n, int world, int64_t* bounds, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n >= 0 && world >= 1 && world <= 1024 && bounds, "mdi_knn_split: bad argument");
  if (n == 0) {
    MDI_CUDA(cudaMemsetAsync(bounds, 0, sizeof(int64_t) * ((size_t)world + 1), (cudaStream_t)stream), "mdi_knn_split");
    return MDI_OK;
  }
  MDI_REQUIRE(workspace, "mdi_knn_split: null workspace");
  KnnWs w;
  knn_carve(&w, (char*)workspace, n);
  prof_mark("knn_split_kernel", (cudaStream_t)stream);
  knn_split_kernel<<<(world + 1 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w.start, n, world, bounds,
                                                                              (int)grid_cells_for(n));
  MDI_CHECK_LAUNCH("mdi_knn_split");
  return MDI_OK;
}

// queries at sorted positions q_begin .. q_end - 1 against the grid built by mdi_knn_build
extern "C" int mdi_knn_query(void* workspace, int64_t n, int k, int exclude_self, int64_t q_begin, int64_t q_end,
                             int32_t* self_idx, int32_t* idx, float* dist, float* kdist, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(n >= 0 && n < (1LL << 31) - 2048, "mdi_knn_query: n out of range");
  MDI_REQUIRE(k >= 1 && k <= kKnnMaxK, "mdi_knn: k must be in 1..%d", kKnnMaxK);
  MDI_REQUIRE(q_begin >= 0 && q_begin <= q_end && q_end <= n, "mdi_knn_query: query range %lld..%lld outside 0..%lld",
              (long long)q_begin, (long long)q_end, (long long)n);
  if (n == 0 || q_end == q_begin) return MDI_OK;
  // sklearn: "Expected n_neighbors <= n_samples_fit" (the sample itself does not count when excluded)
  MDI_REQUIRE((int64_t)k + (exclude_self ? 1 : 0) <= n, "mdi_knn: Expected n_neighbors <= n_samples_fit, but "
              "n_neighbors = %d, n_samples_fit = %lld", k + (exclude_self ? 1 : 0), (long long)n);
  MDI_REQUIRE(workspace && idx, "mdi_knn_query: null pointer");
  KnnWs w;
  knn_carve(&w, (char*)workspace, n);
  prof_mark("knn_query_kernel", (cudaStream_t)stream);
  knn_query_kernel<<<(unsigned)((q_end - q_begin + kKnnThreads - 1) / kKnnThreads), kKnnThreads, 0, (cudaStream_t)stream>>>(
      w.sorted, w.start, n, w.grid, k, exclude_self, idx, dist, kdist, q_begin, q_end, self_idx);
  MDI_CHECK_LAUNCH("mdi_knn_query");
  return MDI_OK;
}

extern "C" int mdi_knn(const float* pts, int64_t n, int k, int exclude_self, void* workspace, int32_t* idx,
                       float* dist, float* kdist, void* stream) {
  MDI_REQUIRE(k >= 1 && k <= mdi::kKnnMaxK, "mdi_knn: k must be in 1..%d", mdi::kKnnMaxK);
  if (n == 0) return MDI_OK;
  MDI_REQUIRE(n < 0 || (int64_t)k + (exclude_self ? 1 : 0) <= n, "mdi_knn: Expected n_neighbors <= n_samples_fit, but "
              "n_neighbors = %d, n_samples_fit = %lld", k + (exclude_self ? 1 : 0), (long long)n);
  MDI_REQUIRE(idx, "mdi_knn: null pointer");
  if (int rc = mdi_knn_build(pts, n, workspace, stream)) return rc;
  return mdi_knn_query(workspace, n, k, exclude_self, 0, n, nullptr, idx, dist, kdist, stream);
}

extern "C" int mdi_lof(int64_t n, int k, const int32_t* idx, const float* dist, const float* kdist, float* lrd,
                       float* lof, uint8_t* outlier, void* stream) {
  using namespace mdi;
  cudaStream_t st = (cudaStream_t)stream;
  MDI_REQUIRE(n >= 0 && k >= 1 && k <= kKnnMaxK, "mdi_lof: bad size");
  if (n == 0) return MDI_OK;
  MDI_REQUIRE(idx && dist && kdist && lrd && outlier, "mdi_lof: null pointer");
  const unsigned grid = (unsigned)((n + kKnnThreads - 1) / kKnnThreads);
  prof_mark("lof_lrd_kernel", st);
  lof_lrd_kernel<<<grid, kKnnThreads, 0, st>>>(n, k, idx, dist, kdist, lrd);
  prof_mark("lof_score_kernel", st);
  lof_score_kernel<<<grid, kKnnThreads, 0, st>>>(n, k, idx, lrd, lrd, lof, outlier);
  MDI_CHECK_LAUNCH("mdi_lof");
  return MDI_OK;
}

// The two LOF passes separately, for a query split over ranks: rows are this rank's m
// queries (compact rows of mdi_knn_queries), neighbours are addressed by original index into the
// FULL per-point arrays that the caller gathers between the passes.
extern "C" int mdi_lof_lrd(int64_t m, int k, const int32_t* idx, const float* dist, const float* kdist_full,
                           float* lrd_rows, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(m >= 0 && k >= 1 && k <= kKnnMaxK, "mdi_lof_lrd: bad size");
  if (m == 0) return MDI_OK;
  MDI_REQUIRE(idx && dist && kdist_full && lrd_rows, "mdi_lof_lrd: null pointer");
  prof_mark("lof_lrd_kernel", (cudaStream_t)stream);
  lof_lrd_kernel<<<(unsigned)((m + kKnnThreads - 1) / kKnnThreads), kKnnThreads, 0, (cudaStream_t)stream>>>(
      m, k, idx, dist, kdist_full, lrd_rows);
  MDI_CHECK_LAUNCH("mdi_lof_lrd");
  return MDI_OK;
}

extern "C" int mdi_lof_score(int64_t m, int k, const int32_t* idx, const float* lrd_full, const float* lrd_rows,
                             float* lof, uint8_t* outlier, void* stream) {
  using namespace mdi;
  MDI_REQUIRE(m >= 0 && k >= 1 && k <= kKnnMaxK, "mdi_lof_score: bad size");
  if (m == 0) return MDI_OK;
  MDI_REQUIRE(idx && lrd_full && lrd_rows && outlier, "mdi_lof_score: null pointer");
  prof_mark("lof_score_kernel", (cudaStream_t)stream);
  lof_score_kernel<<<(unsigned)((m + kKnnThreads - 1) / kKnnThreads), kKnnThreads, 0, (cudaStream_t)stream>>>(
      m, k, idx, lrd_full, lrd_rows, lof, outlier);
  MDI_CHECK_LAUNCH("mdi_lof_score");
  return MDI_OK;
}
