// k7_merge.cu -- merge subsampling of the final cloud (a14).  Reference:
// native_modules/subsampling/src/impl.cpp:198-274 (spatial_median), :313-426 (the DFS),
// geometry.h:27-86 (boxes), stack.h (LIFO: the right child is visited first).
//
// The reference walks the tree depth first with one index vector per frame.  Everything a frame
// decides depends only on its box (fixed by the path) and on its points in original index order
// (partitions are stable), so the tree is built level by level instead: all frames of one depth
// are disjoint slices of one permuted element array, a level is a handful of streaming passes
// over that array, and the slices of the leaves -- read from the highest position down -- are
// exactly the reference's emission order (left child = lower part of the parent's slice,
// right child popped first).
//
// Per level:  reduce (tight box + sum|extent| per frame)  ->  decide (merge / emit / split)
//             ->  side bits + two-level prefix  ->  children  ->  stable scatter.
// Sequential fp32 sums (average extent, merged position / colour) are order dependent; they are
// taken by one thread per frame, and for the average only when the merge test is not already
// decided by a rigorous bound: |fl_seq(sum x)| <= sum|x| (1+u)^(n-1), so max_dim(tight) above
// mult * that bound / n cannot merge.  Large frames are always decided by the bound; frames
// that reach the sequential loop are the ones about to merge (a few points).
#include <float.h>

#include "common.cuh"

namespace mdi {

constexpr int kMergeThreads = 256;
constexpr int kBitsThreads = 1024;  // 32 ballot words per CTA

struct MergeNode { int begin, count; float bmin[3], bmax[3]; int prev_axis; };
struct MergeAcc { uint32_t tmin[3], tmax[3]; double sum_abs; };
struct MergeDec { int kind, axis; float split; int child[2]; int gbegin, nleft; };
enum { kLeaf = 0, kSplit = 1, kChildNone = -1, kChildSingle = -2 };

struct MergeWs {
  float4* pe[2];       // (x, y, z, extent) in slice order
  uint32_t* idx[2];    // original index of the element at each position
  int32_t* node_of[2]; // frame of the element at each position, -1 when it has been emitted
  uint32_t* bits;      // ballot words: element goes left / position emits a point
  int32_t* wlocal;     // exclusive prefix of popc(bits) inside a 32-word block
  int32_t* blockpre;   // exclusive prefix over blocks, [n_blocks] = total
  uint8_t* out_flag;   // position emits a point
  float* stage_xyz;    // staged output rows, by position
  float* stage_rgb;
  MergeNode* nodes[2];
  MergeAcc* acc;
  MergeDec* dec;
  int32_t* counters;   // [0] frames of the next level, [1] error flag
  int64_t n_words, n_blocks, max_nodes;
};

static inline int64_t align_up(int64_t x) { return (x + 255) & ~(int64_t)255; }

static int64_t carve(MergeWs* w, char* base, int64_t n) {
  int64_t off = 0;
  auto take = [&](int64_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  const int64_t n_blocks = n / kBitsThreads + 1, n_words = n_blocks * 32, max_nodes = n / 2 + 1;
  MergeWs t;
  for (int i = 0; i < 2; ++i) t.pe[i] = (float4*)take(n * 16);
  for (int i = 0; i < 2; ++i) t.idx[i] = (uint32_t*)take(n * 4);
  for (int i = 0; i < 2; ++i) t.node_of[i] = (int32_t*)take(n * 4);
  t.bits = (uint32_t*)take(n_words * 4);
  t.wlocal = (int32_t*)take(n_words * 4);
  t.blockpre = (int32_t*)take((n_blocks + 1) * 4);
  t.out_flag = (uint8_t*)take(n_blocks * kBitsThreads);
  t.stage_xyz = (float*)take(n * 12);
  t.stage_rgb = (float*)take(n * 12);
  for (int i = 0; i < 2; ++i) t.nodes[i] = (MergeNode*)take(max_nodes * sizeof(MergeNode));
  t.acc = (MergeAcc*)take(max_nodes * sizeof(MergeAcc));
  t.dec = (MergeDec*)take(max_nodes * sizeof(MergeDec));
  t.counters = (int32_t*)take(64);
  t.n_words = n_words; t.n_blocks = n_blocks; t.max_nodes = max_nodes;
  if (w) *w = t;
  return off;
}

__device__ __forceinline__ float std_min(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float std_max(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float max3(const float* d) { return std_max(d[0], std_max(d[1], d[2])); }
__device__ __forceinline__ float min3(const float* d) { return std_min(d[0], std_min(d[1], d[2])); }

__device__ __forceinline__ void acc_reset(MergeAcc* a) {
  for (int k = 0; k < 3; ++k) { a->tmin[k] = 0xFFFFFFFFu; a->tmax[k] = 0u; }
  a->sum_abs = 0.0;
}

// number of set bits before position p
__device__ __forceinline__ int rank_before(const uint32_t* __restrict__ bits, const int32_t* __restrict__ wlocal,
                                           const int32_t* __restrict__ blockpre, int64_t p) {
  const int64_t w = p >> 5;
  return blockpre[w >> 5] + wlocal[w] + __popc(bits[w] & ((1u << (p & 31)) - 1u));
}

__global__ void __launch_bounds__(kMergeThreads)
merge_init_kernel(const float* __restrict__ pts, const float* __restrict__ ext, int64_t n, float4* pe,
                  uint32_t* idx, int32_t* node_of, MergeNode* nodes, MergeAcc* acc, int32_t* counters) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    nodes[0].begin = 0; nodes[0].count = (int)n; nodes[0].prev_axis = 0;
    acc_reset(&acc[0]);
    counters[0] = 0; counters[1] = 0;
  }
  if (p >= n) return;
  pe[p] = make_float4(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2], ext[p]);
  idx[p] = (uint32_t)p;
  node_of[p] = 0;
}

// tight box and sum|extent| of every live frame.  CTAs whose elements all sit in one frame
// (every CTA of a large frame) reduce in shared memory first.
__global__ void __launch_bounds__(kMergeThreads)
merge_reduce_kernel(const float4* __restrict__ pe, const int32_t* __restrict__ node_of, int64_t n,
                    MergeAcc* __restrict__ acc) {
  __shared__ uint32_t s_min[3][kMergeThreads / 32], s_max[3][kMergeThreads / 32];
  __shared__ double s_sum[kMergeThreads / 32];
  __shared__ int s_k0;
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int k = (p < n) ? node_of[p] : -1;
  if (threadIdx.x == 0) s_k0 = k;
  __syncthreads();
  const int k0 = s_k0;
  const int uniform = __syncthreads_and(k == k0);
  if (uniform && k0 < 0) return;
  uint32_t kmin[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, kmax[3] = {0u, 0u, 0u};
  double sa = 0.0;
  if (k >= 0) {
    const float4 e = pe[p];
    const float c[3] = {e.x, e.y, e.z};
    for (int a = 0; a < 3; ++a)
      if (c[a] == c[a]) { kmin[a] = kmax[a] = float_to_key(c[a]); }  // std::min/max skip NaN
    sa = (double)fabsf(e.w);
  }
  if (!uniform) {
    if (k >= 0) {
      for (int a = 0; a < 3; ++a) { atomicMin(&acc[k].tmin[a], kmin[a]); atomicMax(&acc[k].tmax[a], kmax[a]); }
      atomicAdd(&acc[k].sum_abs, sa);
    }
    return;
  }
  for (int a = 0; a < 3; ++a) {
    kmin[a] = __reduce_min_sync(0xffffffffu, kmin[a]);
    kmax[a] = __reduce_max_sync(0xffffffffu, kmax[a]);
  }
  sa = warp_sum_f64(sa);
  if (lane_id() == 0) {
    for (int a = 0; a < 3; ++a) { s_min[a][warp_id()] = kmin[a]; s_max[a][warp_id()] = kmax[a]; }
    s_sum[warp_id()] = sa;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kMergeThreads / 32; ++w) {
      for (int a = 0; a < 3; ++a) { kmin[a] = min(kmin[a], s_min[a][w]); kmax[a] = max(kmax[a], s_max[a][w]); }
      sa += s_sum[w];
    }
    for (int a = 0; a < 3; ++a) { atomicMin(&acc[k0].tmin[a], kmin[a]); atomicMax(&acc[k0].tmax[a], kmax[a]); }
    atomicAdd(&acc[k0].sum_abs, sa);
  }
}

// geometry.h:56-63 cube_from_points, from the root's tight box
__global__ void merge_root_kernel(MergeNode* nodes, const MergeAcc* acc) {
  float lo[3], hi[3], d[3];
  for (int a = 0; a < 3; ++a) {
    lo[a] = acc[0].tmin[a] == 0xFFFFFFFFu ? FLT_MAX : key_to_float(acc[0].tmin[a]);
    hi[a] = acc[0].tmax[a] == 0u ? -FLT_MAX : key_to_float(acc[0].tmax[a]);
    d[a] = hi[a] - lo[a];
  }
  const float half = max3(d) / 2.0f;
  for (int a = 0; a < 3; ++a) {
    const float c = (lo[a] + hi[a]) / 2.0f;
    nodes[0].bmin[a] = c - half;
    nodes[0].bmax[a] = c + half;
  }
}

// impl.cpp:351-404 for one frame per thread: merge, emit both points, or choose the split.
__global__ void __launch_bounds__(kMergeThreads)
merge_decide_kernel(const MergeNode* __restrict__ nodes, const MergeAcc* __restrict__ acc, int n_nodes,
                    const float4* __restrict__ pe, const uint32_t* __restrict__ idx,
                    const float* __restrict__ rgb, float max_ratio, float mult, MergeDec* __restrict__ dec,
                    uint8_t* __restrict__ out_flag, float* __restrict__ stage_xyz,
                    float* __restrict__ stage_rgb) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_nodes) return;
  const MergeNode nd = nodes[k];
  const MergeAcc ac = acc[k];
  const int64_t b = nd.begin, cnt = nd.count;
  float db[3], dt[3];
  for (int a = 0; a < 3; ++a) {
    const float lo = ac.tmin[a] == 0xFFFFFFFFu ? FLT_MAX : key_to_float(ac.tmin[a]);
    const float hi = ac.tmax[a] == 0u ? -FLT_MAX : key_to_float(ac.tmax[a]);
    dt[a] = hi - lo;
    db[a] = nd.bmax[a] - nd.bmin[a];
  }
  const float a_box = max3(db) / min3(db);
  const float a_tight = max3(dt) / min3(dt);
  const float aspect = std_min(a_box, a_tight);
  const float thresh = max3(dt);
  bool merge = false;
  if (aspect <= max_ratio) {
    // |mult * fl(fl_seq(sum)/n)| <= |mult| * sum|x| / n * (1+u)^(n+1), u = 2^-24
    const double bound = fabs((double)mult) * (ac.sum_abs / (double)cnt) *
                         exp((double)(cnt + 2) * 5.9604644775390625e-8) * 1.000001;
    if (!((double)thresh > bound)) {
      float avg = 0.f;  // impl.cpp:223-230, original index order
      for (int64_t p = b; p < b + cnt; ++p) avg += pe[p].w;
      avg = avg / (float)cnt;
      merge = thresh <= mult * avg;
    }
  }
  MergeDec d;
  d.kind = kLeaf; d.axis = 0; d.split = 0.f; d.child[0] = d.child[1] = kChildNone; d.gbegin = 0; d.nleft = 0;
  if (merge) {  // impl.cpp:372-390
    float mp[3] = {0.f, 0.f, 0.f}, mc[3] = {0.f, 0.f, 0.f};
    for (int64_t p = b; p < b + cnt; ++p) {
      const float4 e = pe[p];
      const float* c = rgb + 3 * (int64_t)idx[p];
      mp[0] += e.x; mp[1] += e.y; mp[2] += e.z;
      mc[0] += c[0]; mc[1] += c[1]; mc[2] += c[2];
    }
    for (int a = 0; a < 3; ++a) {
      stage_xyz[3 * b + a] = mp[a] / (float)cnt;
      stage_rgb[3 * b + a] = mc[a] / (float)cnt;
    }
    out_flag[b] = 1;
  } else if (cnt <= 2) {  // impl.cpp:391-398; staged swapped because positions are read downwards
    for (int j = 0; j < 2; ++j) {
      const float4 e = pe[b + 1 - j];
      const float* c = rgb + 3 * (int64_t)idx[b + 1 - j];
      stage_xyz[3 * (b + j)] = e.x; stage_xyz[3 * (b + j) + 1] = e.y; stage_xyz[3 * (b + j) + 2] = e.z;
      for (int a = 0; a < 3; ++a) stage_rgb[3 * (b + j) + a] = c[a];
      out_flag[b + j] = 1;
    }
  } else {  // impl.cpp:400-404, spatial_median :261-263
    d.kind = kSplit;
    d.axis = (nd.prev_axis + 1) % 3;
    d.split = (nd.bmin[d.axis] + nd.bmax[d.axis]) / 2.0f;
  }
  dec[k] = d;
}

// One ballot word per warp: MODE 0 = "element goes to the left child", MODE 1 = "position emits".
// Also the exclusive prefix of the word popcounts inside the CTA and the CTA total.
template <int MODE>
__global__ void __launch_bounds__(kBitsThreads)
merge_bits_kernel(const float4* __restrict__ pe, const int32_t* __restrict__ node_of,
                  const MergeDec* __restrict__ dec, const uint8_t* __restrict__ out_flag, int64_t n,
                  uint32_t* __restrict__ bits, int32_t* __restrict__ wlocal, int32_t* __restrict__ blocksum) {
  __shared__ int s_cnt[32];
  const int64_t p = (int64_t)blockIdx.x * kBitsThreads + threadIdx.x;
  bool on = false;
  if (p < n) {
    if (MODE == 0) {
      const int k = node_of[p];
      if (k >= 0) {
        const MergeDec d = dec[k];
        if (d.kind == kSplit) {
          const float4 e = pe[p];
          const float c = d.axis == 0 ? e.x : (d.axis == 1 ? e.y : e.z);
          on = c < d.split;
        }
      }
    } else {
      on = out_flag[p] != 0;
    }
  }
  const uint32_t word = __ballot_sync(0xffffffffu, on);
  if (lane_id() == 0) s_cnt[warp_id()] = __popc(word);
  __syncthreads();
  if (warp_id() == 0) {
    const int c = s_cnt[lane_id()];
    const int incl = warp_incl_scan(c);
    s_cnt[lane_id()] = incl - c;
    if (lane_id() == 31) blocksum[blockIdx.x] = incl;
  }
  __syncthreads();
  if (lane_id() == 0) {
    const int64_t w = (int64_t)blockIdx.x * 32 + warp_id();
    bits[w] = word;
    wlocal[w] = s_cnt[warp_id()];
  }
}

// exclusive scan of the CTA totals in place, [n_blocks] = grand total; one CTA
__global__ void __launch_bounds__(1024) merge_scan_blocks_kernel(int32_t* blockpre, int64_t n_blocks) {
  __shared__ int s_scan[33];
  int carry = 0;
  for (int64_t base = 0; base < n_blocks; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int v = (i < n_blocks) ? blockpre[i] : 0;
    int total;
    const int ex = block_excl_scan(v, s_scan, &total);
    if (i < n_blocks) blockpre[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) blockpre[n_blocks] = carry;
}

// geometry.h:67-74 + impl.cpp:408-411: the two child frames of every split.  A child with one
// point is emitted by the scatter pass (impl.cpp:355-359), an empty one is dropped (:352-354).
__global__ void __launch_bounds__(kMergeThreads)
merge_children_kernel(const MergeNode* __restrict__ nodes, int n_nodes, MergeDec* __restrict__ dec,
                      const uint32_t* __restrict__ bits, const int32_t* __restrict__ wlocal,
                      const int32_t* __restrict__ blockpre, MergeNode* __restrict__ next,
                      MergeAcc* __restrict__ acc, int32_t* __restrict__ counters) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_nodes) return;
  MergeDec d = dec[k];
  if (d.kind != kSplit) return;
  const MergeNode nd = nodes[k];
  const int g0 = rank_before(bits, wlocal, blockpre, nd.begin);
  const int g1 = rank_before(bits, wlocal, blockpre, (int64_t)nd.begin + nd.count);
  const int cnt[2] = {g1 - g0, nd.count - (g1 - g0)};
  d.gbegin = g0;
  d.nleft = cnt[0];
  for (int s = 0; s < 2; ++s) {
    if (cnt[s] == 0) { d.child[s] = kChildNone; continue; }
    if (cnt[s] == 1) { d.child[s] = kChildSingle; continue; }
    const int id = atomicAdd(&counters[0], 1);
    MergeNode c = nd;
    c.begin = nd.begin + (s ? cnt[0] : 0);
    c.count = cnt[s];
    if (s == 0) c.bmax[d.axis] = d.split; else c.bmin[d.axis] = d.split;
    c.prev_axis = d.axis;
    next[id] = c;
    acc_reset(&acc[id]);
    d.child[s] = id;
  }
  dec[k] = d;
}

// spatial_median's stable partition (impl.cpp:264-271) for all frames at once
__global__ void __launch_bounds__(kMergeThreads)
merge_scatter_kernel(const float4* __restrict__ pe, const uint32_t* __restrict__ idx,
                     const int32_t* __restrict__ node_of, int64_t n, const MergeNode* __restrict__ nodes,
                     const MergeDec* __restrict__ dec, const uint32_t* __restrict__ bits,
                     const int32_t* __restrict__ wlocal, const int32_t* __restrict__ blockpre,
                     const float* __restrict__ rgb, float4* __restrict__ pe_next,
                     uint32_t* __restrict__ idx_next, int32_t* __restrict__ node_next,
                     uint8_t* __restrict__ out_flag, float* __restrict__ stage_xyz,
                     float* __restrict__ stage_rgb) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const int k = node_of[p];
  if (k < 0) { node_next[p] = -1; return; }
  const MergeDec d = dec[k];
  if (d.kind != kSplit) { node_next[p] = -1; return; }
  const int b = nodes[k].begin;
  const bool left = (bits[p >> 5] >> (p & 31)) & 1u;
  const int rl = rank_before(bits, wlocal, blockpre, p) - d.gbegin;
  const int64_t q = left ? (int64_t)b + rl : (int64_t)b + d.nleft + ((int)(p - b) - rl);
  const int c = d.child[left ? 0 : 1];
  const float4 e = pe[p];
  const uint32_t ix = idx[p];
  if (c == kChildSingle) {
    const float* col = rgb + 3 * (int64_t)ix;
    stage_xyz[3 * q] = e.x; stage_xyz[3 * q + 1] = e.y; stage_xyz[3 * q + 2] = e.z;
    stage_rgb[3 * q] = col[0]; stage_rgb[3 * q + 1] = col[1]; stage_rgb[3 * q + 2] = col[2];
    out_flag[q] = 1;
    node_next[q] = -1;
    return;
  }
  pe_next[q] = e;
  idx_next[q] = ix;
  node_next[q] = c;
}

// rows leave in descending position order: the DFS pops the right child first
__global__ void __launch_bounds__(kMergeThreads)
merge_emit_kernel(const uint8_t* __restrict__ out_flag, int64_t n, const uint32_t* __restrict__ bits,
                  const int32_t* __restrict__ wlocal, const int32_t* __restrict__ blockpre, int64_t n_blocks,
                  const float* __restrict__ stage_xyz, const float* __restrict__ stage_rgb,
                  float* __restrict__ out_pts, float* __restrict__ out_rgb) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n || !out_flag[p]) return;
  const int64_t row = (int64_t)blockpre[n_blocks] - 1 - rank_before(bits, wlocal, blockpre, p);
  for (int a = 0; a < 3; ++a) {
    out_pts[3 * row + a] = stage_xyz[3 * p + a];
    out_rgb[3 * row + a] = stage_rgb[3 * p + a];
  }
}

}  // namespace mdi

extern "C" int64_t mdi_merge_workspace_bytes(int64_t n) {
  if (n < 0) return -1;
  return mdi::carve(nullptr, nullptr, n > 0 ? n : 1);
}

extern "C" int mdi_merge_subsample(const float* pts, const float* rgb, const float* extents, int64_t n,
                                   float max_bbox_aspect_ratio, float min_extent_mult, int max_levels,
                                   void* workspace, float* out_pts, float* out_rgb, int64_t* n_out,
                                   int32_t* n_levels, void* stream) {
  using namespace mdi;
  cudaStream_t st = (cudaStream_t)stream;
  MDI_REQUIRE(n >= 0 && n < (1LL << 31) - 2048, "mdi_merge_subsample: n out of range");
  MDI_REQUIRE(n_out != nullptr, "mdi_merge_subsample: n_out is null");
  *n_out = 0;
  if (n_levels) *n_levels = 0;
  if (n == 0) return MDI_OK;
  MDI_REQUIRE(pts && rgb && extents && workspace && out_pts && out_rgb, "mdi_merge_subsample: null pointer");
  if (max_levels <= 0) max_levels = 1024;
  if (n == 1) {  // impl.cpp:355-359
    MDI_CUDA(cudaMemcpyAsync(out_pts, pts, 12, cudaMemcpyDeviceToDevice, st), "mdi_merge_subsample");
    MDI_CUDA(cudaMemcpyAsync(out_rgb, rgb, 12, cudaMemcpyDeviceToDevice, st), "mdi_merge_subsample");
    MDI_CUDA(cudaStreamSynchronize(st), "mdi_merge_subsample");
    *n_out = 1;
    return MDI_OK;
  }
  MergeWs w;
  carve(&w, (char*)workspace, n);
  const unsigned grid_n = (unsigned)((n + kMergeThreads - 1) / kMergeThreads);
  const unsigned grid_b = (unsigned)w.n_blocks;
  MDI_CUDA(cudaMemsetAsync(w.out_flag, 0, (size_t)(w.n_blocks * kBitsThreads), st), "mdi_merge_subsample");
  prof_mark("merge_init_kernel", st);
  merge_init_kernel<<<grid_n, kMergeThreads, 0, st>>>(pts, extents, n, w.pe[0], w.idx[0], w.node_of[0],
                                                      w.nodes[0], w.acc, w.counters);
  prof_mark("merge_reduce_kernel", st);
  merge_reduce_kernel<<<grid_n, kMergeThreads, 0, st>>>(w.pe[0], w.node_of[0], n, w.acc);
  prof_mark("merge_root_kernel", st);
  merge_root_kernel<<<1, 1, 0, st>>>(w.nodes[0], w.acc);
  MDI_CHECK_LAUNCH("mdi_merge_subsample(init)");

  int cur = 0, n_nodes = 1, level = 0;
  for (; n_nodes > 0; ++level) {
    if (level >= max_levels) {
      set_error("mdi_merge_subsample: %d frames still open after %d levels (coincident points whose "
                "merge test fails; the reference does not terminate on this input)", n_nodes, level);
      return MDI_EINVAL;
    }
    const int nx = cur ^ 1;
    const unsigned grid_k = (unsigned)((n_nodes + kMergeThreads - 1) / kMergeThreads);
    if (level > 0) {
      prof_mark("merge_reduce_kernel", st);
      merge_reduce_kernel<<<grid_n, kMergeThreads, 0, st>>>(w.pe[cur], w.node_of[cur], n, w.acc);
    }
    prof_mark("merge_decide_kernel", st);
    merge_decide_kernel<<<grid_k, kMergeThreads, 0, st>>>(w.nodes[cur], w.acc, n_nodes, w.pe[cur], w.idx[cur],
                                                          rgb, max_bbox_aspect_ratio, min_extent_mult, w.dec,
                                                          w.out_flag, w.stage_xyz, w.stage_rgb);
    prof_mark("merge_bits_kernel", st);
    merge_bits_kernel<0><<<grid_b, kBitsThreads, 0, st>>>(w.pe[cur], w.node_of[cur], w.dec, nullptr, n, w.bits,
                                                          w.wlocal, w.blockpre);
    prof_mark("merge_scan_blocks_kernel", st);
    merge_scan_blocks_kernel<<<1, 1024, 0, st>>>(w.blockpre, w.n_blocks);
    prof_mark("merge_children_kernel", st);
    merge_children_kernel<<<grid_k, kMergeThreads, 0, st>>>(w.nodes[cur], n_nodes, w.dec, w.bits, w.wlocal,
                                                            w.blockpre, w.nodes[nx], w.acc, w.counters);
    prof_mark("merge_scatter_kernel", st);
    merge_scatter_kernel<<<grid_n, kMergeThreads, 0, st>>>(w.pe[cur], w.idx[cur], w.node_of[cur], n, w.nodes[cur],
                                                           w.dec, w.bits, w.wlocal, w.blockpre, rgb, w.pe[nx],
                                                           w.idx[nx], w.node_of[nx], w.out_flag, w.stage_xyz,
                                                           w.stage_rgb);
    MDI_CHECK_LAUNCH("mdi_merge_subsample(level)");
    int32_t next_nodes = 0;
    MDI_CUDA(cudaMemcpyAsync(&next_nodes, w.counters, 4, cudaMemcpyDeviceToHost, st), "mdi_merge_subsample");
    MDI_CUDA(cudaMemsetAsync(w.counters, 0, 4, st), "mdi_merge_subsample");
    MDI_CUDA(cudaStreamSynchronize(st), "mdi_merge_subsample");
    n_nodes = next_nodes;
    cur = nx;
  }
  prof_mark("merge_bits_kernel", st);
  merge_bits_kernel<1><<<grid_b, kBitsThreads, 0, st>>>(nullptr, nullptr, nullptr, w.out_flag, n, w.bits,
                                                        w.wlocal, w.blockpre);
  prof_mark("merge_scan_blocks_kernel", st);
  merge_scan_blocks_kernel<<<1, 1024, 0, st>>>(w.blockpre, w.n_blocks);
  prof_mark("merge_emit_kernel", st);
  merge_emit_kernel<<<grid_n, kMergeThreads, 0, st>>>(w.out_flag, n, w.bits, w.wlocal, w.blockpre, w.n_blocks,
                                                      w.stage_xyz, w.stage_rgb, out_pts, out_rgb);
  MDI_CHECK_LAUNCH("mdi_merge_subsample(emit)");
  int32_t total = 0;
  MDI_CUDA(cudaMemcpyAsync(&total, w.blockpre + w.n_blocks, 4, cudaMemcpyDeviceToHost, st), "mdi_merge_subsample");
  MDI_CUDA(cudaStreamSynchronize(st), "mdi_merge_subsample");
  *n_out = total;
  if (n_levels) *n_levels = level;
  return MDI_OK;
}
