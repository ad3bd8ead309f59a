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

constexpr int kEmitWarps = 8;     // one warp per 4096-pixel tile, 4 bit words per lane
constexpr int kEmitList = 1024;   // kept pixels of a tile listed in shared memory per pass

// Reference: points_from_depth.py:292-312.  pc = ((x+.5) z, (y+.5) z, z); world =
// c2w[:3,:3] (Kinv pc) + c2w[:3,3]; dot products as FMA chains (mdi_math.cuh dot3/dot4_fma).
// RGB: image.view(-1,3)[mask] (monocular_depth_init.py:128,168), image/255.0 as a division.
// The warp first lists the tile's kept pixels in row-major order in shared memory (scan of the
// bit words), then lane j takes points j, j + 32, ...: one point per lane with all its loads
// independent, consecutive lanes writing consecutive output records.
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
};

__device__ __forceinline__ void multicast_store(float* addr, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}

template <int IMAGE_KIND>
__global__ void __launch_bounds__(kEmitWarps * 32)
emit_points_kernel(const MdiView* __restrict__ views, int n_views,
                   const float* __restrict__ scale_shift, const uint32_t* __restrict__ bits,
                   const int64_t* __restrict__ tile_offsets, const int64_t* __restrict__ view_offsets,
                   int64_t total_tiles, EmitDst dst) {
  __shared__ uint16_t s_px[kEmitWarps][kEmitList];
  __shared__ float s_rec[kEmitWarps][2][96];  // one warp-iteration of xyz / rgb records
  const bool want_xyz = dst.xyz[0] != nullptr || dst.mc_xyz != nullptr;
  const bool want_rgb = dst.rgb[0] != nullptr || dst.mc_rgb != nullptr;
  const bool direct = dst.n == 1;
  const int64_t tile = (int64_t)blockIdx.x * kEmitWarps + warp_id();
  if (tile >= total_tiles) return;
  const int lane = lane_id();
  const uint4 w4 = __ldg(reinterpret_cast<const uint4*>(bits + tile * MDI_TILE_WORDS) + lane);
  const int cnt = __popc(w4.x) + __popc(w4.y) + __popc(w4.z) + __popc(w4.w);
  const int incl = warp_incl_scan(cnt);
  const int tot = __shfl_sync(0xffffffffu, incl, 31);
  if (tot == 0) return;
  if (dst.total_rows && __ldg(dst.total_rows) > dst.capacity) return;
  const int v = find_view_fast(views, n_views, tile, total_tiles);
  const MdiView& vw = views[v];
  const float s = scale_shift ? scale_shift[2 * v] : 1.f;
  const float t = scale_shift ? scale_shift[2 * v + 1] : 0.f;
  const int W = vw.W;
  const bool small = (int64_t)vw.H * W <= 0x7FFFFFFFLL;
  const float* __restrict__ depth = vw.depth;
  const float* Ki = vw.Kinv;
  const float* M = vw.c2w;
  const int64_t out0 = dst.row_offset + view_offsets[v] + tile_offsets[tile];
  const int64_t px0 = (tile - vw.tile_begin) * MDI_TILE_PX;
  uint16_t* list = s_px[warp_id()];
#pragma unroll 1
  for (int base = 0; base < tot; base += kEmitList) {
    // list points base .. base + kEmitList - 1 of the tile
    int idx = incl - cnt;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint32_t word = k == 0 ? w4.x : k == 1 ? w4.y : k == 2 ? w4.z : w4.w;
      while (word) {
        const int b = __ffs(word) - 1;
        word &= word - 1;
        if (idx >= base && idx < base + kEmitList) list[idx - base] = (uint16_t)(lane * 128 + 32 * k + b);
        ++idx;
      }
    }
    __syncwarp();
    const int n_here = min(kEmitList, tot - base);
    float* rec_xyz = s_rec[warp_id()][0];
    float* rec_rgb = s_rec[warp_id()][1];
    for (int j0 = 0; j0 < n_here; j0 += 32) {
      const int j = j0 + lane;
      if (j < n_here) {
        const int64_t p = px0 + list[j];
        const int row = small ? (int)((uint32_t)p / (uint32_t)W) : (int)(p / W);
        const int col = (int)(p - (int64_t)row * W);
        const float z = align_depth(__ldg(&depth[p]), s, t);
        if (want_rgb) {
          float r, g, bl;
          if (IMAGE_KIND == MDI_IMAGE_U8) {
            const uint8_t* im = reinterpret_cast<const uint8_t*>(vw.image) + 3 * p;
            const uint8_t a0 = im[0], a1 = im[1], a2 = im[2];
            r = fdiv((float)a0, 255.0f); g = fdiv((float)a1, 255.0f); bl = fdiv((float)a2, 255.0f);
          } else {
            const float* im = reinterpret_cast<const float*>(vw.image) + 3 * p;
            r = im[0]; g = im[1]; bl = im[2];
            if (IMAGE_KIND == MDI_IMAGE_F32_255) {
              r = fdiv(r, 255.0f); g = fdiv(g, 255.0f); bl = fdiv(bl, 255.0f);
            }
          }
          if (direct) {
            float* o = dst.rgb[0] + 3 * (out0 + base + j);
            o[0] = r; o[1] = g; o[2] = bl;
          } else {
            rec_rgb[3 * lane] = r; rec_rgb[3 * lane + 1] = g; rec_rgb[3 * lane + 2] = bl;
          }
        }
        if (want_xyz) {
          const float px = fmul(fadd((float)col, 0.5f), z);
          const float py = fmul(fadd((float)row, 0.5f), z);
          // inv(K) @ pc (dot-product kernel) and cam2world @ [cam;1] (FMA chain), see mdi_math.cuh
          const float cx = dot3_kinv(Ki[0], Ki[1], Ki[2], px, py, z);
          const float cy = dot3_kinv(Ki[3], Ki[4], Ki[5], px, py, z);
          const float cz = dot3_kinv(Ki[6], Ki[7], Ki[8], px, py, z);
          const float wx = dot4_fma(M[0], M[1], M[2], M[3], cx, cy, cz);
          const float wy = dot4_fma(M[4], M[5], M[6], M[7], cx, cy, cz);
          const float wz = dot4_fma(M[8], M[9], M[10], M[11], cx, cy, cz);
          if (direct) {
            float* o = dst.xyz[0] + 3 * (out0 + base + j);
            o[0] = wx; o[1] = wy; o[2] = wz;
          } else {
            rec_xyz[3 * lane] = wx; rec_xyz[3 * lane + 1] = wy; rec_xyz[3 * lane + 2] = wz;
          }
        }
      }
      if (direct) continue;  // one local cloud: every lane has stored its own record
      __syncwarp();
      // the records of this iteration are one contiguous run of words in every destination cloud
      // (this GPU's and, through NVLink, its peers'): three fully coalesced stores per array
      const int n_words = 3 * min(32, n_here - j0);
      const int64_t w0 = 3 * (out0 + base + j0);
      if (dst.n == 0) {  // the switch replicates the store into every GPU's cloud (this one included)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int w = 32 * k + lane;
          if (w < n_words) {
            if (dst.mc_xyz) multicast_store(dst.mc_xyz + w0 + w, rec_xyz[w]);
            if (dst.mc_rgb) multicast_store(dst.mc_rgb + w0 + w, rec_rgb[w]);
          }
        }
      }
      for (int d = 0; d < dst.n; ++d) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int w = 32 * k + lane;
          if (w < n_words) {
            if (want_xyz) dst.xyz[d][w0 + w] = rec_xyz[w];
            if (want_rgb) dst.rgb[d][w0 + w] = rec_rgb[w];
          }
        }
      }
      __syncwarp();
    }
    __syncwarp();
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
  const unsigned grid = (unsigned)((total_tiles + mdi::kEmitWarps - 1) / mdi::kEmitWarps);
  mdi::prof_mark("emit_points_kernel", st);
  switch (image_kind) {
    case MDI_IMAGE_F32_UNIT:
      mdi::emit_points_kernel<MDI_IMAGE_F32_UNIT><<<grid, mdi::kEmitWarps * 32, 0, st>>>(
          views, n_views, scale_shift, bits, tile_offsets, view_offsets, total_tiles, dst);
      break;
    case MDI_IMAGE_F32_255:
      mdi::emit_points_kernel<MDI_IMAGE_F32_255><<<grid, mdi::kEmitWarps * 32, 0, st>>>(
          views, n_views, scale_shift, bits, tile_offsets, view_offsets, total_tiles, dst);
      break;
    case MDI_IMAGE_U8:
      mdi::emit_points_kernel<MDI_IMAGE_U8><<<grid, mdi::kEmitWarps * 32, 0, st>>>(
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
