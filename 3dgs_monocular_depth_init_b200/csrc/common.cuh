// common.cuh -- error plumbing, view lookup and block/warp primitives shared by the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mdi_b200.h"
#include "mdi_math.cuh"

namespace mdi {

void set_error(const char* fmt, ...);
// per-kernel timing hook: closes the open range (if any) and opens `name` (nullptr: just close)
extern bool g_prof;
void prof_mark(const char* name, cudaStream_t st);

#define MDI_REQUIRE(cond, ...)      \
  do {                              \
    if (!(cond)) {                  \
      mdi::set_error(__VA_ARGS__);  \
      return MDI_EINVAL;            \
    }                               \
  } while (0)

#define MDI_CHECK_LAUNCH(name)                                                       \
  do {                                                                               \
    if (mdi::g_prof) mdi::prof_mark(nullptr, 0);                                     \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess) {                                                        \
      mdi::set_error("%s: CUDA error: %s", name, cudaGetErrorString(e__));           \
      return MDI_ECUDA;                                                              \
    }                                                                                \
  } while (0)

#define MDI_CUDA(call, name)                                                         \
  do {                                                                               \
    cudaError_t e__ = (call);                                                        \
    if (e__ != cudaSuccess) {                                                        \
      mdi::set_error("%s: CUDA error: %s", name, cudaGetErrorString(e__));           \
      return MDI_ECUDA;                                                              \
    }                                                                                \
  } while (0)

constexpr int kSMs = 148;  // B200

// view owning global tile `tile`: largest v with views[v].tile_begin <= tile
__device__ __forceinline__ int find_view(const MdiView* __restrict__ views, int n_views,
                                         int64_t tile) {
  int lo = 0, hi = n_views - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (__ldg(&views[mid].tile_begin) <= tile) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// tiles of equal-sized views: v = tile * n_views / total_tiles is right on the first probe;
// anything else falls back to the binary search
__device__ __forceinline__ int find_view_fast(const MdiView* __restrict__ views, int n_views,
                                              int64_t tile, int64_t total_tiles) {
  const int v = (int)((tile * n_views) / total_tiles);
  const int64_t tb = __ldg(&views[v].tile_begin);
  if (tile >= tb && tile < tb + __ldg(&views[v].n_tiles)) return v;
  return find_view(views, n_views, tile);
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

__device__ __forceinline__ double shfl_xor_f64(double v, int m) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_xor_sync(0xffffffffu, lo, m);
  hi = __shfl_xor_sync(0xffffffffu, hi, m);
  return __hiloint2double(hi, lo);
}
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v += shfl_xor_f64(v, m);
  return v;
}
__device__ __forceinline__ int warp_sum_i32(int v) { return __reduce_add_sync(0xffffffffu, v); }
__device__ __forceinline__ float warp_sum_f32(float v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

// inclusive warp scan
__device__ __forceinline__ int warp_incl_scan(int v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane_id() >= d) v += t;
  }
  return v;
}

// Block-wide exclusive scan of one int per thread.  `smem` holds >= 33 ints.  Returns the
// exclusive prefix of the calling thread; *total receives the block total.  Contains barriers.
__device__ __forceinline__ int block_excl_scan(int v, int* smem, int* total) {
  int incl = warp_incl_scan(v);
  int w = warp_id(), l = lane_id();
  int nw = (blockDim.x + 31) >> 5;
  __syncthreads();  // protect smem reuse across calls
  if (l == 31) smem[w] = incl;
  __syncthreads();
  if (w == 0) {
    int s = (l < nw) ? smem[l] : 0;
    int si = warp_incl_scan(s);
    smem[l] = si - s;
    if (l == 31) smem[32] = si;
  }
  __syncthreads();
  *total = smem[32];
  return smem[w] + incl - v;
}

}  // namespace mdi
