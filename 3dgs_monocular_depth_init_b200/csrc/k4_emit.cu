// k4_emit.cu -- order-preserving compaction offsets (tile scan) and the unprojection + RGB
// gather of the kept pixels (a10, a11).
#include "common.cuh"

namespace mdi {

constexpr int kScanThreads = 256;

// Level 1: one CTA per view, exclusive scan of the view's tile counts -> offsets WITHIN the view
// and the view's point count.
__global__ void __launch_bounds__(kScanThreads)
scan_view_tiles_kernel(const MdiView* __restrict__ views, const int32_t* __restrict__ tile_counts,
                       int64_t* __restrict__ tile_offsets, int64_t* __restrict__ view_counts) {
  __shared__ int scan_smem[33];
  __shared__ long long s_base;
  const int v = blockIdx.x;
  const int64_t tb = views[v].tile_begin;
  const int n = views[v].n_tiles;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int start = 0; start < n; start += kScanThreads * 4) {
    const int i0 = start + threadIdx.x * 4;
    int c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) c[k] = (i0 + k < n) ? tile_counts[tb + i0 + k] : 0;
    int tot;
    const int excl = block_excl_scan(c[0] + c[1] + c[2] + c[3], scan_smem, &tot);
    long long o = s_base + excl;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (i0 + k < n) tile_offsets[tb + i0 + k] = o;
      o += c[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_base += tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) view_counts[v] = s_base;
}

// Level 2: one CTA, exclusive scan of the per-view counts in view order -> view offsets, total.
__global__ void __launch_bounds__(1024)
scan_views_kernel(int n_views, const int64_t* __restrict__ view_counts,
                  int64_t* __restrict__ view_offsets, int64_t* __restrict__ total) {
  __shared__ long long s_warp[32];
  __shared__ long long s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const int lane = lane_id(), warp = warp_id();
  for (int start = 0; start < n_views; start += 1024) {
    const int i = start + threadIdx.x;
    const long long c = i < n_views ? view_counts[i] : 0;
    long long incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const long long y = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += y;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const long long w = s_warp[lane];
      long long wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const long long y = __shfl_up_sync(0xffffffffu, wi, d);
        if (lane >= d) wi += y;
      }
      s_warp[lane] = wi - w;
    }
    __syncthreads();
    const long long excl = s_base + s_warp[warp] + incl - c;
    if (i < n_views && view_offsets) view_offsets[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) s_base = excl + c;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total) *total = s_base;
}

constexpr int kEmitThreads = 256;
constexpr int kEmitTiles = 16;                 // consecutive 4096-pixel tiles per CTA
constexpr int kEmitWords = kEmitTiles * MDI_TILE_WORDS / kEmitThreads;  // bit words per thread (8)
constexpr int kEmitList = 1024;                // kept pixels listed and staged per pass
static_assert(kEmitWords == 8 && kEmitTiles * MDI_TILE_PX == 65536, "a listed pixel is a uint16");

// destination clouds of the emit kernel: this GPU's and (SURVEY 8e exchange step fused into the
// kernel) the peer GPUs' copies, mapped into this address space by the caller
struct EmitDst {
  int n;
  int64_t row_offset;  // first row of this GPU's points in the destination clouds
  float* xyz[MDI_MAX_EMIT_DST];
  float* rgb[MDI_MAX_EMIT_DST];
  // NVSwitch multicast addresses of the same clouds (n == 0): one multimem store reaches every GPU
  float* mc_xyz;
  float* mc_rgb;
  // speculative launch (mdi_emit_points_capped): nothing is written when *total_rows > capacity
  const int64_t* total_rows;
  int64_t capacity;
  // exchange launch (mdi_emit_points_exchange): row_offset is read from the device ([0] of the record
  // written by mdi_exchange_counts; [1] = rows of all ranks, checked against capacity)
  const int64_t* exchange;
  // byte-rgb exchange (uint8 images): peers receive the gathered rgb BYTES (3 B per point instead of
  // 12) in rgb8[d] and expand them themselves (mdi_expand_rgb8: u8 / 255.0f, the same division);
  // the float rgb is written to this GPU's own cloud (index self) only
  uint8_t* rgb8[MDI_MAX_EMIT_DST];
  int self;
};

__device__ __forceinline__ void multicast_store(float* addr, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void multicast_store4(float* addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// per-tile parameters staged in shared memory, read by the point loop as 128-bit loads
struct __align__(16) EmitTile {
  float4 k0, k1;        // Kinv[0..7]
  float4 k2;            // Kinv[8], scale, shift, W (int bits)
  float4 m0, m1, m2;    // cam2world rows
  const float* depth;   // view base pointers
  const void* image;
  int64_t px0;          // first pixel of the tile inside its view
  int32_t small, _pad;  // H*W < 2^31: 32-bit row / column arithmetic
};
static_assert(sizeof(EmitTile) == 128, "EmitTile is eight 16-byte rows");

// One contiguous run of `n_words` floats from shared memory (`s`, word k of the run at s[shift + k]
// with shift = word misalignment of the destination, so both sides of every 16-byte copy are
// aligned) to every destination: scalar head (<= 3 words), 128-bit body, scalar tail.
// one contiguous run of n_bytes bytes from shared memory (byte k at s[shift + k], shift = byte
// misalignment of the destination) to every peer's byte cloud: scalar head / tail, 16-byte body
__device__ __forceinline__ void store_run_bytes(const EmitDst& dst, const uint8_t* s, int shift, int64_t b0,
                                                int n_bytes) {
  const int head = min((16 - shift) & 15, n_bytes);
  const int n_vec = (n_bytes - head) >> 4;
  const int tail0 = head + 16 * n_vec;
  const int tid = threadIdx.x;
  if (tid < head) {
    const uint8_t v = s[shift + tid];
    for (int d = 0; d < dst.n; ++d) if (d != dst.self) dst.rgb8[d][b0 + tid] = v;
  }
  for (int i = tid; i < n_vec; i += kEmitThreads) {
    const uint4 v = *reinterpret_cast<const uint4*>(s + shift + head + 16 * i);
    for (int d = 0; d < dst.n; ++d)
      if (d != dst.self) *reinterpret_cast<uint4*>(dst.rgb8[d] + b0 + head + 16 * i) = v;
  }
  if (tid < n_bytes - tail0) {
    const uint8_t v = s[shift + tail0 + tid];
    for (int d = 0; d < dst.n; ++d) if (d != dst.self) dst.rgb8[d][b0 + tail0 + tid] = v;
  }
}

__device__ __forceinline__ void store_run(const EmitDst& dst, bool is_rgb, const float* s, int shift,
                                          int64_t w0, int n_words, int d_begin = 0, int d_end = MDI_MAX_EMIT_DST) {
  d_end = min(d_end, dst.n);
  const int head = min((4 - shift) & 3, n_words);
  const int n_vec = (n_words - head) >> 2;
  const int tail0 = head + 4 * n_vec;
  const int tid = threadIdx.x;
  if (dst.n == 0) {  // the switch replicates the store into every GPU's cloud (this one included)
    float* g = (is_rgb ? dst.mc_rgb : dst.mc_xyz) + w0;
    if (tid < head) multicast_store(g + tid, s[shift + tid]);
    for (int i = tid; i < n_vec; i += kEmitThreads)
      multicast_store4(g + head + 4 * i, *reinterpret_cast<const float4*>(s + shift + head + 4 * i));
    if (tid < n_words - tail0) multicast_store(g + tail0 + tid, s[shift + tail0 + tid]);
    return;
  }
  if (tid < head) {
    const float v = s[shift + tid];
    for (int d = d_begin; d < d_end; ++d) (is_rgb ? dst.rgb[d] : dst.xyz[d])[w0 + tid] = v;
  }
  for (int i = tid; i < n_vec; i += kEmitThreads) {
    const float4 v = *reinterpret_cast<const float4*>(s + shift + head + 4 * i);
    for (int d = d_begin; d < d_end; ++d)
      *reinterpret_cast<float4*>((is_rgb ? dst.rgb[d] : dst.xyz[d]) + w0 + head + 4 * i) = v;
  }
  if (tid < n_words - tail0) {
    const float v = s[shift + tail0 + tid];
    for (int d = d_begin; d < d_end; ++d) (is_rgb ? dst.rgb[d] : dst.xyz[d])[w0 + tail0 + tid] = v;
  }
}

// Reference: points_from_depth.py:292-312.  pc = ((x+.5) z, (y+.5) z, z); world =
// c2w[:3,:3] (Kinv pc) + c2w[:3,3]; dot products as FMA chains (mdi_math.cuh dot3/dot4_fma).
// RGB: image.view(-1,3)[mask] (monocular_depth_init.py:128,168), image/255.0 as a division.
//
// One CTA per 16 consecutive tiles (65536 pixels, ~900 kept pixels at the default strides).  Output
// rows are consecutive in global tile order (views are concatenated in order), so the CTA owns ONE
// contiguous run of rows: (1) every thread takes 8 bit words (two 128-bit loads), a block scan of
// the popcounts ranks every kept pixel; (2) the kept pixels are listed in row-major order in shared
// memory, 1024 per pass; (3) thread j unprojects / gathers point j, j + 256, ... -- one point per
// thread, all loads independent -- and parks the record in shared memory; (4) the run is written
// out with 16-byte stores, to this GPU's cloud and, in an exchange launch, to every peer's copy
// over NVLink (or once, through the NVSwitch multicast address).
template <int IMAGE_KIND>
__global__ void __launch_bounds__(kEmitThreads)
emit_points_kernel(const MdiView* __restrict__ views, int n_views,
                   const float* __restrict__ scale_shift, const uint32_t* __restrict__ bits,
                   const int64_t* __restrict__ tile_offsets, const int64_t* __restrict__ view_offsets,
                   int64_t total_tiles, EmitDst dst) {
  __shared__ EmitTile s_tile[kEmitTiles];
  __shared__ uint16_t s_list[kEmitList];
  __shared__ __align__(16) float s_xyz[3 * kEmitList + 4];
  __shared__ __align__(16) float s_rgb[3 * kEmitList + 4];
  __shared__ __align__(16) uint8_t s_rgb8[IMAGE_KIND == MDI_IMAGE_U8 ? 3 * kEmitList + 16 : 16];
  __shared__ int s_scan[33];
  __shared__ long long s_out0;
  const bool want_xyz = dst.xyz[0] != nullptr || dst.mc_xyz != nullptr;
  const bool want_rgb = dst.rgb[0] != nullptr || dst.mc_rgb != nullptr;
  const bool rgb_bytes = IMAGE_KIND == MDI_IMAGE_U8 && want_rgb && dst.n > 1 && dst.rgb8[0] != nullptr;
  const int tid = threadIdx.x;
  const int64_t tile0 = (int64_t)blockIdx.x * kEmitTiles;
  if (dst.total_rows && __ldg(dst.total_rows) > dst.capacity) return;
  if (dst.exchange && __ldg(dst.exchange + 1) > dst.capacity) return;
  // (1) bit words of the thread: word index tid * 8 + k of the CTA, i.e. tile tid / 16
  uint32_t w[kEmitWords];
  {
    const int64_t tile = tile0 + (tid >> 4);
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (tile < total_tiles) {
      const uint4* p = reinterpret_cast<const uint4*>(bits + tile0 * MDI_TILE_WORDS) + 2 * tid;
      a = __ldg(p); b = __ldg(p + 1);
    }
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
  }
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < kEmitWords; ++k) cnt += __popc(w[k]);
  int n;
  const int excl = block_excl_scan(cnt, s_scan, &n);
  if (n == 0) return;
  // per-tile parameters (threads 0..15) and the CTA's first output row
  if (tid < kEmitTiles && tile0 + tid < total_tiles) {
    const int64_t tile = tile0 + tid;
    const int v = find_view_fast(views, n_views, tile, total_tiles);
    const MdiView& vw = views[v];
    EmitTile& T = s_tile[tid];
    const float* Ki = vw.Kinv;
    const float* M = vw.c2w;
    T.k0 = make_float4(Ki[0], Ki[1], Ki[2], Ki[3]);
    T.k1 = make_float4(Ki[4], Ki[5], Ki[6], Ki[7]);
    T.k2 = make_float4(Ki[8], scale_shift ? scale_shift[2 * v] : 1.f, scale_shift ? scale_shift[2 * v + 1] : 0.f,
                       __int_as_float(vw.W));
    T.m0 = make_float4(M[0], M[1], M[2], M[3]);
    T.m1 = make_float4(M[4], M[5], M[6], M[7]);
    T.m2 = make_float4(M[8], M[9], M[10], M[11]);
    T.depth = vw.depth; T.image = vw.image;
    T.px0 = (tile - vw.tile_begin) * MDI_TILE_PX;
    T.small = (int64_t)vw.H * vw.W <= 0x7FFFFFFFLL;
    if (tid == 0) {
      const int64_t row0 = dst.exchange ? __ldg(dst.exchange) : dst.row_offset;
      s_out0 = row0 + view_offsets[v] + tile_offsets[tile];
    }
  }
  __syncthreads();
  const int64_t out0 = s_out0;
#pragma unroll 1
  for (int base = 0; base < n; base += kEmitList) {
    // (2) list points base .. base + kEmitList - 1 of the CTA
    int idx = excl;
#pragma unroll
    for (int k = 0; k < kEmitWords; ++k) {
      uint32_t word = w[k];
      const int c = __popc(word);
      if (idx + c > base && idx < base + kEmitList) {
        int i = idx;
        while (word) {
          const int b = __ffs(word) - 1;
          word &= word - 1;
          if (i >= base && i < base + kEmitList) s_list[i - base] = (uint16_t)(((tid * kEmitWords + k) << 5) | b);
          ++i;
        }
      }
      idx += c;
    }
    __syncthreads();
    const int n_here = min(kEmitList, n - base);
    // records of this pass start at word w0 of every destination array; both arrays share the
    // word misalignment of w0 when their base pointers are 16-byte aligned (checked by the host)
    const int64_t w0 = 3 * (out0 + base);
    const float* any_x = dst.n ? dst.xyz[0] : dst.mc_xyz;
    const float* any_r = dst.n ? dst.rgb[0] : dst.mc_rgb;
    const int sh_x = want_xyz ? (int)((reinterpret_cast<uintptr_t>(any_x + w0) >> 2) & 3) : 0;
    const int sh_r = want_rgb ? (int)((reinterpret_cast<uintptr_t>(any_r + w0) >> 2) & 3) : 0;
    const int sh_b = rgb_bytes ? (int)(reinterpret_cast<uintptr_t>(dst.rgb8[0] + w0) & 15) : 0;   // byte w0 = 3 * row
    // (3) one point per thread
    for (int j = tid; j < n_here; j += kEmitThreads) {
      const uint32_t e = s_list[j];
      const EmitTile& T = s_tile[e >> 12];
      const float4 k2 = T.k2;
      const int W = __float_as_int(k2.w);
      const int64_t p = T.px0 + (e & 4095u);
      const int row = T.small ? (int)((uint32_t)p / (uint32_t)W) : (int)(p / W);
      const int col = (int)(p - (int64_t)row * W);
      const float z = align_depth(__ldg(T.depth + p), k2.y, k2.z);
      if (want_rgb) {
        float r, g, bl;
        if (IMAGE_KIND == MDI_IMAGE_U8) {
          const uint8_t* im = reinterpret_cast<const uint8_t*>(T.image) + 3 * p;
          const uint8_t a0 = im[0], a1 = im[1], a2 = im[2];
          r = fdiv((float)a0, 255.0f); g = fdiv((float)a1, 255.0f); bl = fdiv((float)a2, 255.0f);
          if (rgb_bytes) { uint8_t* ob = s_rgb8 + sh_b + 3 * j; ob[0] = a0; ob[1] = a1; ob[2] = a2; }
        } else {
          const float* im = reinterpret_cast<const float*>(T.image) + 3 * p;
          r = im[0]; g = im[1]; bl = im[2];
          if (IMAGE_KIND == MDI_IMAGE_F32_255) {
            r = fdiv(r, 255.0f); g = fdiv(g, 255.0f); bl = fdiv(bl, 255.0f);
          }
        }
        float* o = s_rgb + sh_r + 3 * j;
        o[0] = r; o[1] = g; o[2] = bl;
      }
      if (want_xyz) {
        const float4 k0 = T.k0, k1 = T.k1, m0 = T.m0, m1 = T.m1, m2 = T.m2;
        const float px = fmul(fadd((float)col, 0.5f), z);
        const float py = fmul(fadd((float)row, 0.5f), z);
        // inv(K) @ pc (dot-product kernel) and cam2world @ [cam;1] (FMA chain), see mdi_math.cuh
        const float cx = dot3_kinv(k0.x, k0.y, k0.z, px, py, z);
        const float cy = dot3_kinv(k0.w, k1.x, k1.y, px, py, z);
        const float cz = dot3_kinv(k1.z, k1.w, k2.x, px, py, z);
        float* o = s_xyz + sh_x + 3 * j;
        o[0] = dot4_fma(m0.x, m0.y, m0.z, m0.w, cx, cy, cz);
        o[1] = dot4_fma(m1.x, m1.y, m1.z, m1.w, cx, cy, cz);
        o[2] = dot4_fma(m2.x, m2.y, m2.z, m2.w, cx, cy, cz);
      }
    }
    __syncthreads();
    // (4) the pass is one contiguous run of 3 * n_here words in every destination cloud
    if (want_xyz) store_run(dst, false, s_xyz, sh_x, w0, 3 * n_here);
    if (want_rgb && !rgb_bytes) store_run(dst, true, s_rgb, sh_r, w0, 3 * n_here);
    if (rgb_bytes) {   // floats to the own cloud, bytes to the peers
      store_run(dst, true, s_rgb, sh_r, w0, 3 * n_here, dst.self, dst.self + 1);
      store_run_bytes(dst, s_rgb8, sh_b, w0, 3 * n_here);
    }
    __syncthreads();
  }
}

// Receiver side of the byte-rgb exchange: rows of OTHER ranks arrive as bytes; rgb = u8 / 255.0f, the
// division the emit kernel applies to its own rows (bit-identical).  record: the mdi_exchange_counts
// record of the step ([0] first row of this rank, [1] rows of all ranks, [3 + rank] rows of this rank).
__global__ void __launch_bounds__(256)
expand_rgb8_kernel(const uint8_t* __restrict__ rgb8, float* __restrict__ rgb, const int64_t* __restrict__ record,
                   int rank, int64_t capacity) {
  const int64_t total = record[1];
  if (total > capacity) return;
  const int64_t own0 = 3 * record[0], own1 = own0 + 3 * record[3 + rank], n = 3 * total;
  const int64_t n4 = n >> 2;   // groups of four elements: one 32-bit load, one 128-bit store
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = 4 * g;
    if (i >= own0 && i + 4 <= own1) continue;          // wholly inside this rank's own rows
    const uint32_t w = *reinterpret_cast<const uint32_t*>(rgb8 + i);
    float4 v = make_float4(fdiv((float)(w & 255u), 255.0f), fdiv((float)((w >> 8) & 255u), 255.0f),
                           fdiv((float)((w >> 16) & 255u), 255.0f), fdiv((float)(w >> 24), 255.0f));
    if (i >= own1 || i + 4 <= own0) {
      *reinterpret_cast<float4*>(rgb + i) = v;
    } else {                                           // straddles the boundary of the own rows
      const float e[4] = {v.x, v.y, v.z, v.w};
      for (int k = 0; k < 4; ++k) if (i + k < own0 || i + k >= own1) rgb[i + k] = e[k];
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n & 3)) {
    const int64_t i = 4 * n4 + threadIdx.x;
    if (i < own0 || i >= own1) rgb[i] = fdiv((float)rgb8[i], 255.0f);
  }
}

// Count exchange of a multi-GPU step, entirely on the device (replaces an 8-byte NCCL all-gather
// plus a host synchronisation in front of the emit kernel).  Every rank owns a peer-mapped header
// of 2 * world int64 slots; rank r publishes (epoch, retry flag, its row count) into slot
// [epoch & 1][r] of EVERY rank's header with a system-scope release store, then waits until its own
// header holds the current epoch from every rank, and writes out[0] = first row of this rank in the
// scene cloud, out[1] = rows of all ranks, out[2] = 1 if any rank raised its retry flag,
// out[3 + q] = rows of rank q.  Two slot sets: a rank can be at most one exchange ahead of a peer.
struct ExchangeHdrs { int64_t* p[MDI_MAX_EMIT_DST]; };
constexpr int kEpochShift = 41;   // value = epoch << 41 | retry << 40 | rows
__global__ void exchange_counts_kernel(ExchangeHdrs hdrs, int rank, int world, uint32_t epoch,
                                       const int64_t* __restrict__ local /* [rows, retry] */,
                                       int64_t* __restrict__ out) {
  __shared__ long long s_cnt[MDI_MAX_EMIT_DST];
  __shared__ int s_retry;
  const int t = threadIdx.x;
  if (t == 0) s_retry = 0;
  __syncthreads();
  const uint64_t tag = (uint64_t)(epoch & 0x3FFFFFu);
  const int slot0 = (int)(epoch & 1u) * world;
  if (t < world) {
    const uint64_t mine = (tag << kEpochShift) | ((uint64_t)(local[1] != 0) << 40) | (uint64_t)local[0];
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(hdrs.p[t] + slot0 + rank), "l"(mine) : "memory");
    const int64_t* src = hdrs.p[rank] + slot0 + t;
    uint64_t got;
    // a peer that never publishes (crashed rank) must not hang this GPU: give up after ~20 s
    const long long t0 = clock64();
    bool timed_out = false;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(src) : "memory");
      if ((got >> kEpochShift) == tag) break;
      __nanosleep(200);
      if (clock64() - t0 > 40000000000LL) { timed_out = true; break; }
    }
    s_cnt[t] = timed_out ? 0 : (long long)(got & ((1ull << 40) - 1));
    if (timed_out) atomicOr(&s_retry, 2);
    else if ((got >> 40) & 1ull) atomicOr(&s_retry, 1);
  }
  __syncthreads();
  if (t == 0 && (s_retry & 2)) {   // record[1] < 0: the host raises
    out[0] = 0; out[1] = -1; out[2] = 0;
    return;
  }
  if (t == 0) {
    long long before = 0, total = 0;
    for (int q = 0; q < world; ++q) {
      if (q < rank) before += s_cnt[q];
      total += s_cnt[q];
      out[3 + q] = s_cnt[q];
    }
    out[0] = before; out[1] = total; out[2] = s_retry;
  }
}

}  // namespace mdi

extern "C" int mdi_scan_tiles(const MdiView* views, int n_views, int64_t total_tiles,
                              const int32_t* tile_counts, int64_t* tile_offsets,
                              int64_t* view_counts, int64_t* view_offsets, int64_t* total,
                              void* stream) {
  MDI_REQUIRE(n_views >= 0 && total_tiles >= 0, "mdi_scan_tiles: negative size");
  MDI_REQUIRE(views && tile_counts && tile_offsets && view_counts && view_offsets,
              "mdi_scan_tiles: null pointer");
  if (n_views > 0) {
    mdi::prof_mark("scan_view_tiles_kernel", (cudaStream_t)stream);
    mdi::scan_view_tiles_kernel<<<n_views, mdi::kScanThreads, 0, (cudaStream_t)stream>>>(
        views, tile_counts, tile_offsets, view_counts);
  }
  mdi::prof_mark("scan_views_kernel", (cudaStream_t)stream);
  mdi::scan_views_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(n_views, view_counts, view_offsets, total);
  MDI_CHECK_LAUNCH("mdi_scan_tiles");
  return MDI_OK;
}

static int launch_emit(const MdiView* views, int n_views, int64_t total_tiles, const float* scale_shift,
                       const uint32_t* bits, const int64_t* tile_offsets, const int64_t* view_offsets,
                       int image_kind, const mdi::EmitDst& dst, void* stream, const char* who) {
  MDI_REQUIRE(n_views >= 0 && total_tiles >= 0, "%s: negative size", who);
  if (n_views == 0 || total_tiles == 0) return MDI_OK;
  MDI_REQUIRE(views && bits && tile_offsets && view_offsets, "%s: null pointer", who);
  MDI_REQUIRE(total_tiles < (1LL << 31), "%s: too many tiles", who);
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)((total_tiles + mdi::kEmitTiles - 1) / mdi::kEmitTiles);
  // the 16-byte store path assumes one word misalignment per array across destinations
  for (int d = 1; d < dst.n; ++d)
    MDI_REQUIRE(((reinterpret_cast<uintptr_t>(dst.xyz[d]) ^ reinterpret_cast<uintptr_t>(dst.xyz[0])) & 15) == 0 &&
                ((reinterpret_cast<uintptr_t>(dst.rgb[d]) ^ reinterpret_cast<uintptr_t>(dst.rgb[0])) & 15) == 0,
                "%s: destination clouds must share their 16-byte alignment", who);
  mdi::prof_mark("emit_points_kernel", st);
  switch (image_kind) {
    case MDI_IMAGE_F32_UNIT:
      mdi::emit_points_kernel<MDI_IMAGE_F32_UNIT><<<grid, mdi::kEmitThreads, 0, st>>>(
          views, n_views, scale_shift, bits, tile_offsets, view_offsets, total_tiles, dst);
      break;
    case MDI_IMAGE_F32_255:
      mdi::emit_points_kernel<MDI_IMAGE_F32_255><<<grid, mdi::kEmitThreads, 0, st>>>(
          views, n_views, scale_shift, bits, tile_offsets, view_offsets, total_tiles, dst);
      break;
    case MDI_IMAGE_U8:
      mdi::emit_points_kernel<MDI_IMAGE_U8><<<grid, mdi::kEmitThreads, 0, st>>>(
          views, n_views, scale_shift, bits, tile_offsets, view_offsets, total_tiles, dst);
      break;
    default:
      mdi::prof_mark(nullptr, 0);
      mdi::set_error("%s: unknown image kind %d", who, image_kind);
      return MDI_EINVAL;
  }
  MDI_CHECK_LAUNCH(who);
  return MDI_OK;
}

extern "C" int mdi_emit_points(const MdiView* views, int n_views, int64_t total_tiles,
                               const float* scale_shift, const uint32_t* bits,
                               const int64_t* tile_offsets, const int64_t* view_offsets,
                               int image_kind, float* xyz, float* rgb, void* stream) {
  mdi::EmitDst dst{};
  dst.n = 1; dst.row_offset = 0; dst.xyz[0] = xyz; dst.rgb[0] = rgb;
  return launch_emit(views, n_views, total_tiles, scale_shift, bits, tile_offsets, view_offsets, image_kind,
                     dst, stream, "mdi_emit_points");
}

extern "C" int mdi_emit_points_capped(const MdiView* views, int n_views, int64_t total_tiles,
                                      const float* scale_shift, const uint32_t* bits,
                                      const int64_t* tile_offsets, const int64_t* view_offsets,
                                      int image_kind, float* xyz, float* rgb, const int64_t* total_rows,
                                      int64_t capacity_rows, void* stream) {
  MDI_REQUIRE(total_rows && capacity_rows >= 0, "mdi_emit_points_capped: null total_rows / negative capacity");
  mdi::EmitDst dst{};
  dst.n = 1; dst.row_offset = 0; dst.xyz[0] = xyz; dst.rgb[0] = rgb;
  dst.total_rows = total_rows; dst.capacity = capacity_rows;
  return launch_emit(views, n_views, total_tiles, scale_shift, bits, tile_offsets, view_offsets, image_kind,
                     dst, stream, "mdi_emit_points_capped");
}

extern "C" int mdi_emit_points_multi(const MdiView* views, int n_views, int64_t total_tiles,
                                     const float* scale_shift, const uint32_t* bits,
                                     const int64_t* tile_offsets, const int64_t* view_offsets,
                                     int image_kind, int n_dst, float* const* xyz_dst,
                                     float* const* rgb_dst, int64_t row_offset, void* stream) {
  MDI_REQUIRE(n_dst >= 1 && n_dst <= MDI_MAX_EMIT_DST, "mdi_emit_points_multi: n_dst %d not in 1..%d", n_dst,
              MDI_MAX_EMIT_DST);
  MDI_REQUIRE(row_offset >= 0, "mdi_emit_points_multi: negative row_offset");
  MDI_REQUIRE(xyz_dst || rgb_dst, "mdi_emit_points_multi: no destination");
  mdi::EmitDst dst{};
  dst.n = n_dst; dst.row_offset = row_offset;
  for (int d = 0; d < n_dst; ++d) {
    dst.xyz[d] = xyz_dst ? xyz_dst[d] : nullptr;
    dst.rgb[d] = rgb_dst ? rgb_dst[d] : nullptr;
    MDI_REQUIRE((!xyz_dst || dst.xyz[d]) && (!rgb_dst || dst.rgb[d]), "mdi_emit_points_multi: null destination %d", d);
  }
  return launch_emit(views, n_views, total_tiles, scale_shift, bits, tile_offsets, view_offsets, image_kind,
                     dst, stream, "mdi_emit_points_multi");
}

extern "C" int mdi_emit_points_exchange(const MdiView* views, int n_views, int64_t total_tiles,
                                        const float* scale_shift, const uint32_t* bits,
                                        const int64_t* tile_offsets, const int64_t* view_offsets,
                                        int image_kind, int n_dst, float* const* xyz_dst,
                                        float* const* rgb_dst, float* xyz_mc, float* rgb_mc,
                                        const int64_t* exchange_record, int64_t capacity_rows,
                                        uint8_t* const* rgb8_dst, int self_rank, void* stream) {
  MDI_REQUIRE(n_dst >= 0 && n_dst <= MDI_MAX_EMIT_DST, "mdi_emit_points_exchange: n_dst %d not in 0..%d", n_dst,
              MDI_MAX_EMIT_DST);
  MDI_REQUIRE(exchange_record && capacity_rows >= 0, "mdi_emit_points_exchange: null record / negative capacity");
  mdi::EmitDst dst{};
  dst.n = n_dst; dst.exchange = exchange_record; dst.capacity = capacity_rows;
  if (n_dst == 0) {
    MDI_REQUIRE(xyz_mc || rgb_mc, "mdi_emit_points_exchange: no multicast destination");
    dst.mc_xyz = xyz_mc; dst.mc_rgb = rgb_mc;
  } else {
    MDI_REQUIRE(xyz_dst || rgb_dst, "mdi_emit_points_exchange: no destination");
    for (int d = 0; d < n_dst; ++d) {
      dst.xyz[d] = xyz_dst ? xyz_dst[d] : nullptr;
      dst.rgb[d] = rgb_dst ? rgb_dst[d] : nullptr;
      MDI_REQUIRE((!xyz_dst || dst.xyz[d]) && (!rgb_dst || dst.rgb[d]), "mdi_emit_points_exchange: null destination %d", d);
    }
    if (rgb8_dst && rgb_dst && image_kind == MDI_IMAGE_U8 && n_dst > 1) {
      MDI_REQUIRE(self_rank >= 0 && self_rank < n_dst, "mdi_emit_points_exchange: self_rank %d outside 0..%d", self_rank, n_dst - 1);
      for (int d = 0; d < n_dst; ++d) {
        MDI_REQUIRE(rgb8_dst[d] && ((reinterpret_cast<uintptr_t>(rgb8_dst[d]) ^ reinterpret_cast<uintptr_t>(rgb8_dst[0])) & 15) == 0,
                    "mdi_emit_points_exchange: byte clouds must be non-null and share their 16-byte alignment");
        dst.rgb8[d] = rgb8_dst[d];
      }
      dst.self = self_rank;
    }
  }
  return launch_emit(views, n_views, total_tiles, scale_shift, bits, tile_offsets, view_offsets, image_kind,
                     dst, stream, "mdi_emit_points_exchange");
}

extern "C" int mdi_expand_rgb8(const uint8_t* rgb8, float* rgb, const int64_t* exchange_record, int rank,
                               int64_t capacity_rows, void* stream) {
  MDI_REQUIRE(rgb8 && rgb && exchange_record && rank >= 0 && rank < MDI_MAX_EMIT_DST && capacity_rows >= 0,
              "mdi_expand_rgb8: bad argument");
  MDI_REQUIRE(((reinterpret_cast<uintptr_t>(rgb8) & 3) | (reinterpret_cast<uintptr_t>(rgb) & 15)) == 0,
              "mdi_expand_rgb8: clouds must be 4- / 16-byte aligned");
  mdi::prof_mark("expand_rgb8_kernel", (cudaStream_t)stream);
  mdi::expand_rgb8_kernel<<<mdi::kSMs * 8, 256, 0, (cudaStream_t)stream>>>(rgb8, rgb, exchange_record, rank, capacity_rows);
  MDI_CHECK_LAUNCH("mdi_expand_rgb8");
  return MDI_OK;
}

extern "C" int mdi_exchange_counts(int64_t* const* hdr_ptrs, int rank, int world, uint32_t epoch,
                                   const int64_t* local, int64_t* out, void* stream) {
  MDI_REQUIRE(world >= 1 && world <= MDI_MAX_EMIT_DST && rank >= 0 && rank < world,
              "mdi_exchange_counts: rank %d / world %d out of range (<= %d ranks)", rank, world, MDI_MAX_EMIT_DST);
  MDI_REQUIRE(hdr_ptrs && local && out, "mdi_exchange_counts: null pointer");
  mdi::ExchangeHdrs h{};
  for (int q = 0; q < world; ++q) {
    MDI_REQUIRE(hdr_ptrs[q], "mdi_exchange_counts: null header of rank %d", q);
    h.p[q] = hdr_ptrs[q];
  }
  mdi::prof_mark("exchange_counts_kernel", (cudaStream_t)stream);
  mdi::exchange_counts_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(h, rank, world, epoch, local, out);
  MDI_CHECK_LAUNCH("mdi_exchange_counts");
  return MDI_OK;
}

extern "C" int mdi_emit_points_multicast(const MdiView* views, int n_views, int64_t total_tiles,
                                         const float* scale_shift, const uint32_t* bits,
                                         const int64_t* tile_offsets, const int64_t* view_offsets,
                                         int image_kind, float* xyz_mc, float* rgb_mc, int64_t row_offset,
                                         void* stream) {
  MDI_REQUIRE(row_offset >= 0, "mdi_emit_points_multicast: negative row_offset");
  MDI_REQUIRE(xyz_mc || rgb_mc, "mdi_emit_points_multicast: no destination");
  mdi::EmitDst dst{};
  dst.n = 0; dst.row_offset = row_offset; dst.mc_xyz = xyz_mc; dst.mc_rgb = rgb_mc;
  return launch_emit(views, n_views, total_tiles, scale_shift, bits, tile_offsets, view_offsets, image_kind,
                     dst, stream, "mdi_emit_points_multicast");
}
