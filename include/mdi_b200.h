/*
 * mdi_b200.h -- C ABI of libmdi_b200.so: the B200 (sm_100a) kernels behind the
 * monocular-depth point-cloud initialisation path of deivse/3dgs_monocular_depth_init.
 *
 * Every entry point is batched over views, asynchronous on the caller's stream, re-entrant,
 * and writes only into caller-allocated DEVICE buffers (plain pointers + sizes; no torch
 * types).  Return value: 0 on success, otherwise a negative MDI_E* code; the message is
 * available from mdi_last_error().  Per-view soft failures (too few SfM points in bounds,
 * empty adaptive-subsampling population) are DATA: status[view] != 0, and the Python host
 * layer turns them into the reference's exceptions.
 *
 * Reference interface replaced by each entry point (paths relative to the reference repo):
 *   mdi_setup_cameras       gs_init_compare/depth_prediction/points_from_depth.py:229-243
 *   mdi_project_filter      gs_init_compare/depth_prediction/points_from_depth.py:111-142,145-162
 *   mdi_gather_fit          gs_init_compare/depth_alignment/alignment/lstsqrs.py:9-54
 *                           gs_init_compare/depth_subsampling/num_sfm_points_mask.py:7-64
 *   mdi_ransac              gs_init_compare/depth_alignment/alignment/ransacs.py:60-189
 *   mdi_apply_scale_shift   lstsqrs.py:51-54 / ransacs.py:186-189 (aligned_depth = depth*s + t)
 *   mdi_apply_region_scale_shift  depth_alignment/pipeline.py:252-283 (per-region stitch)
 *   mdi_depth_range         gs_init_compare/depth_subsampling/adaptive_subsampling.py:82-98
 *   mdi_grad_range          points_from_depth.py:192-212 (min / max of the gradient map)
 *   mdi_factor_table        adaptive_subsampling.py:12-17,89-120 (stride as a function of depth)
 *   mdi_keep_mask           static_subsampler.py:11-22, adaptive_subsampling.py:48-79,105-122,
 *                           pipeline.py:285-288, points_from_depth.py:268-290
 *   mdi_scan_tiles          points_from_depth.py:292-297 (boolean-index compaction order),
 *                           monocular_depth_init.py:170-171,183-184 (view-major concat)
 *   mdi_emit_points         points_from_depth.py:292-312, monocular_depth_init.py:168
 *   mdi_min_extents         native_modules/subsampling/src/impl.cpp:17-35,70-126
 *   mdi_merge_subsample     native_modules/subsampling/src/impl.cpp:198-274,313-426
 *   mdi_knn / mdi_lof       point_cloud_postprocess/postprocess.py:16-22 (sklearn LOF),
 *                           utils/runner_utils.py:143-147 (knn)
 *   mdi_interp_scale_map    depth_alignment/alignment/interp.py:75-110,355-361 (LinearNDInterpolator
 *                           over the pixel grid, scale map * pre-aligned depth)
 */
#ifndef MDI_B200_H_
#define MDI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MDI_ABI_VERSION 5

/* error codes */
#define MDI_OK 0
#define MDI_EINVAL (-1)  /* bad argument */
#define MDI_ECUDA (-2)   /* CUDA runtime error (launch / config) */
#define MDI_ENOGPU (-3)  /* no sm_100 device */

/* per-view status bits written by the kernels */
#define MDI_STATUS_OK 0
#define MDI_STATUS_LOW_CONFIDENCE 1 /* < 1/4 of SfM points in bounds (points_from_depth.py:124) */
#define MDI_STATUS_EMPTY_STATS 2    /* adaptive subsampling over zero valid pixels (torch raises) */
#define MDI_STATUS_STATS_RETRY 4    /* mdi_depth_range's one-pass path could not certify the order
                                       statistics of this view: re-run with MDI_STATS_RETRY_FLAGGED */

/* pixels per tile of the mask / compaction kernels (one CTA per tile) */
#define MDI_TILE_PX 4096
#define MDI_TILE_WORDS (MDI_TILE_PX / 32)

/* max patches of the SfM-density grid per view (20 on the small axis, <= 102 on the large) */
#define MDI_PATCH_STRIDE 2048

/* floats per view of the adaptive-stride breakpoint table (mdi_factor_table) */
#define MDI_FACTOR_TABLE_STRIDE 40

/* image element kinds for the RGB gather */
#define MDI_IMAGE_F32_UNIT 0 /* float32 already in 0..1 (InputImage.data)          */
#define MDI_IMAGE_F32_255 1  /* float32 in 0..255, divided by 255.0f on the fly    */
#define MDI_IMAGE_U8 2       /* uint8, converted and divided by 255.0f on the fly  */

/* subsampling modes */
#define MDI_SUBSAMPLE_STATIC 0
#define MDI_SUBSAMPLE_ADAPTIVE 1

/* mask-composition flags of mdi_keep_mask */
#define MDI_KEEP_PRED_MASK 1  /* & predicted.mask                       (pipeline.py:287)          */
#define MDI_KEEP_SENTINEL 2   /* & aligned != -42.0                     (pipeline.py:253,287)      */
#define MDI_KEEP_NONNEG 4     /* & aligned >= 0                         (points_from_depth.py:270) */
#define MDI_KEEP_PATCH 8      /* & SfM-density patch mask               (points_from_depth.py:276) */
#define MDI_KEEP_GRAD 16      /* & depth-gradient mask                  (points_from_depth.py:271) */
#define MDI_KEEP_SUBSAMPLE 32 /* & subsampler mask                      (points_from_depth.py:290) */
#define MDI_KEEP_PREZEROED 256 /* keep / bits / tile_counts were zero-filled by the caller (e.g. on a
                                 side stream, overlapped with the alignment stages)              */

/* mdi_depth_range algorithm selection (OR-ed into its flags) */
#define MDI_STATS_EXACT 64          /* three-pass radix select over all pixels, every view      */
#define MDI_STATS_RETRY_FLAGGED 128 /* the same, only for views flagged MDI_STATUS_STATS_RETRY  */
#define MDI_STATS_RAW 512           /* one-pass statistics of the RAW depth of every view; completed by
                                       mdi_depth_range_map once the fit exists                    */

/* RANSAC loss kinds */
#define MDI_LOSS_RANSAC 0 /* outlier count       (ransacs.py:60) */
#define MDI_LOSS_MSAC 1   /* truncated quadratic (ransacs.py:64) */

/*
 * One view.  The array of descriptors lives in DEVICE memory (the host builds it once per
 * scene and copies it); all pointers inside are device pointers (the image may be a
 * device-mapped pinned host pointer: only kept pixels are read from it).
 */
typedef struct MdiView {
  const float* depth;    /* [H*W] float32 predicted depth, row-major                   */
  const uint8_t* mask;   /* [H*W] bool/uint8 predicted mask, or NULL = all valid       */
  const void* image;     /* [H*W*3] pixels, element kind given per call                */
  int32_t H, W;
  int32_t sfm_begin;     /* first slot of this view in the per-point arrays            */
  int32_t sfm_count;     /* M: number of SfM points of this view                       */
  int64_t px_begin;      /* first byte of this view in the keep-mask buffer            */
  int64_t tile_begin;    /* first global tile of this view                             */
  int32_t n_tiles;       /* ceil(H*W / MDI_TILE_PX)                                    */
  int32_t patch_h, patch_w, grid_h, grid_w; /* SfM-density patch geometry              */
  int32_t _pad;
  float P[12];           /* 3x4 projection K R^T [I|-C]  (points_from_depth.py:231-233) */
  float Kinv[9];         /* inverse intrinsics           (points_from_depth.py:303)     */
  float c2w[12];         /* first three rows of cam2world (points_from_depth.py:304-309) */
} MdiView;

typedef struct MdiRansacConfig {
  float inlier_threshold; /* applied to the SQUARED residual (ransacs.py:137) */
  int32_t max_iters;
  double confidence;
  int32_t sample_size;    /* must be 4 when sample_idx is generated on device; <= 8 otherwise */
  int32_t min_iters;
  int32_t loss_kind;      /* MDI_LOSS_*                                        */
  int32_t view_base;      /* index of views[0] in the caller's batch: the device sampler draws for
                             view view_base + v, so a batch launched as several view ranges (one per
                             stream) draws the samples of one launch over all views; 0 otherwise */
  uint64_t seed;          /* device sampler seed when sample_idx == NULL       */
} MdiRansacConfig;

typedef struct MdiSubsampleConfig {
  int32_t mode;           /* MDI_SUBSAMPLE_*                                   */
  int32_t static_factor;  /* static stride                                     */
  int32_t factor_min;     /* adaptive range (depth_subsampling/config.py:11-12) */
  int32_t factor_max;
  int32_t flags;          /* MDI_KEEP_* bits                                   */
  float grad_thresh;      /* depth_grad_mask_thresh                            */
} MdiSubsampleConfig;

const char* mdi_last_error(void);
int mdi_abi_version(void);
/* returns 0 if device `device` is an sm_100 part, MDI_ENOGPU otherwise */
int mdi_check_device(int device);
/* sizeof(MdiView), so that the host-side struct packing can be asserted */
int mdi_sizeof_view(void);

/*
 * Optional tracing: with profiling enabled every kernel launch of the library is bracketed by CUDA
 * events on its stream; mdi_profile_collect synchronises, sums the device time by kernel name
 * ('\n'-separated into names_buf), clears the record and returns the number of names.
 */
void mdi_profile_enable(int on);
int mdi_profile_collect(char* names_buf, int buf_len, float* ms, int32_t* calls, int cap);

/*
 * Copies a small host table (descriptors, K, cam2world) to the device with a kernel that reads the
 * page-locked, device-mapped source directly.  Unlike an async memcpy it does not queue behind bulk
 * host->device copies already handed to the copy engine, which is what lets a block of views be
 * processed while the depth maps of the next blocks are still in flight.  bytes % 4 == 0.
 */
int mdi_upload_table(void* dst, const void* src_pinned, int64_t bytes, void* stream);

/*
 * a1: fills P, Kinv and c2w of every descriptor on the device.  K float32 [V,9], c2w float32
 * [V,16] (row-major 4x4), Kinv float32 [V,9] (inv(K) computed by the caller) or NULL when every K
 * is a zero-skew pinhole matrix (then the closed form is used); P_out float32 [V,12] or NULL.
 * keep_P != 0: the descriptors already hold P (it is kept and copied to P_out).
 */
int mdi_setup_cameras(MdiView* views, int n_views, const float* K, const float* c2w,
                      const float* Kinv, float* P_out, int keep_P, void* stream);

/*
 * a2: project the SfM points of every view, filter, compact in order.
 *   sfm_xyz   [Np,3] float32 world points; sfm_index int32 (concatenated per view, slot
 *   sfm_begin..sfm_begin+M) indexes sfm_xyz, or NULL when the points are already gathered
 *   (then slot i reads sfm_xyz[i]).
 * Outputs (slot ranges of each view): uv int32 [2,S] (row 0 = x/u, row 1 = y/v, S = total
 *   slots, plane stride `slots`), z float32 [S]; n_inbounds/n_valid/status int32 [V].
 */
int mdi_project_filter(const MdiView* views, int n_views, const float* sfm_xyz,
                       const int32_t* sfm_index, int64_t slots, int32_t* uv, float* z,
                       int32_t* n_inbounds, int32_t* n_valid, int32_t* status, void* stream);

/*
 * mdi_setup_cameras + mdi_project_filter in one launch (a1 + a2: points_from_depth.py:231-233 and
 * :111-162): the CTA of a view fills the view's camera fields before it projects.  Same arguments
 * and results as the two calls.
 */
int mdi_setup_project_filter(MdiView* views, int n_views, const float* K, const float* c2w,
                             const float* Kinv, float* P_out, int keep_P, const float* sfm_xyz,
                             const int32_t* sfm_index, int64_t slots, int32_t* uv, float* z,
                             int32_t* n_inbounds, int32_t* n_valid, int32_t* status, void* stream);

/*
 * a3 + a8: gather d_i = depth[v_i,u_i], accumulate the normal equations in fp64, solve with
 * torch.linalg.pinv semantics, count points per patch.
 *   sums      float64 [V,5] = (sum d*d, sum d, n, sum d*g, sum g)
 *   scale_shift float32 [V,2] (written only if write_fit != 0)
 *   patch_blocked uint8 [V, MDI_PATCH_STRIDE] or NULL
 */
int mdi_gather_fit(const MdiView* views, int n_views, int64_t slots, const int32_t* uv,
                   const float* z, const int32_t* n_valid, float* d, double* sums,
                   float* scale_shift, int write_fit, int patch_threshold,
                   uint8_t* patch_blocked, void* stream);

/*
 * a4: LO-RANSAC / MSAC, one CTA per view, one warp per hypothesis, (d, g) staged in shared
 * memory by a 1-D bulk (TMA) copy.  sample_idx int32 [V, max_iters, sample_size] or NULL
 * (device sampler).  Outputs: scale_shift float32 [V,2]; inliers uint8 [S] (slot layout) or
 * NULL; n_iter / n_inliers int32 [V] (n_iter = last loop index, as printed at ransacs.py:183).
 * max_points >= every n_valid[v] sizes the shared-memory staging; the slot ranges of d / z
 * must be readable up to the next multiple of 4 elements (the host pads sfm_begin to 4).
 */
int mdi_ransac(const MdiView* views, int n_views, const float* d, const float* z,
               const int32_t* n_valid, const int32_t* status, const int32_t* sample_idx,
               const MdiRansacConfig* cfg, int max_points, float* scale_shift,
               uint8_t* inliers, int32_t* n_iter, int32_t* n_inliers, void* stream);

/* aligned = depth * s + t as two separately rounded fp32 operations; scale_shift on device */
int mdi_apply_scale_shift(const float* depth, int64_t n, const float* scale_shift, float* out,
                          void* stream);

/*
 * Region-wise alignment (pipeline.py:196-283, 8f-4): stitches the per-region fits into one aligned
 * map.  region int32 [n] = index of each pixel's region in `table` (float32 [R,2] scale / shift),
 * negative = no region; valid uint8 [R] = region has a fit.  Pixels without a valid region get the
 * reference's sentinel -42.0 (masked out by the caller, pipeline.py:285-288).
 */
int mdi_apply_region_scale_shift(const float* depth, const int32_t* region, int64_t n,
                                 const float* table, const uint8_t* valid, int n_regions, float* out,
                                 void* stream);

/*
 * a7 (statistics): per view, over pixels with mask & (aligned != -42): min, max and the exact
 * order statistics torch.quantile(., 0.25 / 0.75) needs; writes lo_hi float32 [V,2] =
 * (max(min, q1 - 1.5 iqr), min(max, q3 + 1.5 iqr)) and ORs MDI_STATUS_EMPTY_STATS into status.
 *   flags: MDI_KEEP_PRED_MASK | MDI_KEEP_SENTINEL select the population; MDI_STATS_* the algorithm.
 *   Default algorithm: ONE streaming pass over depth + mask (key brackets around each quantile
 *   from a 1/16 sample, exact select among the bracketed keys, verified by rank counting).  A
 *   view whose brackets cannot be certified (constant depth maps, adversarial layouts) gets
 *   MDI_STATUS_STATS_RETRY in status, NaN in lo_hi and *retry_flag = 1 (int64, device, may be
 *   NULL; never cleared here); the caller then repeats the call with MDI_STATS_RETRY_FLAGGED,
 *   which runs the three-pass exact radix select for the flagged views and clears the bit.
 *   workspace: mdi_depth_range_workspace_bytes(n_views, max_tiles_per_view) bytes.
 */
int64_t mdi_depth_range_workspace_bytes(int n_views, int64_t max_tiles_per_view);
int mdi_depth_range(const MdiView* views, int n_views, int64_t max_tiles_per_view,
                    const float* scale_shift, int32_t* status, int flags, void* workspace,
                    float* lo_hi, float* quantiles /* [V,4] q1,q3,min,max or NULL */,
                    int64_t* retry_flag, void* stream);

/*
 * a7 (statistics) split in two so that the pass over the pixels does not wait for the fit
 * (adaptive_subsampling.py:82-95 evaluated as F(order statistics of the raw depth), F = the
 * alignment, monotone for a finite scale > 0 and a finite shift).  mdi_depth_range with
 * MDI_STATS_RAW ignores scale_shift / status / lo_hi / quantiles / retry_flag (they may be NULL) and
 * leaves, per view, the four order statistics, min, max and count of the raw depth over the pixels
 * the predicted mask admits in `workspace`; it may run on another stream, beside mdi_project_filter /
 * mdi_gather_fit / mdi_ransac.  mdi_depth_range_map (same n_views, max_tiles_per_view, workspace;
 * after both) maps them through the fit and writes lo_hi / quantiles / status bits exactly like
 * mdi_depth_range.  Views it cannot finish this way -- a scale that is not finite and positive, a
 * depth in the view's range that aligns to exactly -42.0 (MDI_KEEP_SENTINEL in flags), brackets the
 * raw pass could not certify -- get MDI_STATUS_STATS_RETRY and *retry_flag = 1, for a
 * mdi_depth_range call with MDI_STATS_RETRY_FLAGGED.
 */
int mdi_depth_range_map(int n_views, int64_t max_tiles_per_view, const float* scale_shift,
                        int32_t* status, int flags, void* workspace, float* lo_hi,
                        float* quantiles, int64_t* retry_flag, void* stream);

/* a9 (statistics): min and max of |dx|+|dy| of the aligned depth per view -> grad_range [V,2] */
int mdi_grad_range(const MdiView* views, int n_views, int64_t max_tiles_per_view,
                   const float* scale_shift, const int32_t* status, float* grad_range,
                   void* stream);

/*
 * a7 (breakpoints): the adaptive stride f(a) = trunc(clamp((fmax-fmin) * (1 - clamp((a-lo)/(hi-lo)))
 * + fmin)) is a non-increasing step function of the aligned depth a; this finds, per view, the
 * exact fp32 breakpoints of every stride by bisection with the reference's arithmetic, so that
 * mdi_keep_mask decides `f(a) == k` with two comparisons.  table float32
 * [V, MDI_FACTOR_TABLE_STRIDE]; entry 0 of a row is 0 when the view is irregular (NaN / empty
 * range, strides > 31) and mdi_keep_mask then evaluates the arithmetic per pixel.
 */
int mdi_factor_table(int n_views, const float* lo_hi, int factor_min, int factor_max,
                     float* table, void* stream);

/*
 * a5-a10 (predicate): composes the keep mask of every pixel.
 *   factor_table: output of mdi_factor_table for the same lo_hi / factor range, or NULL.
 *   total_px = sum H*W (size of keep), max_height / max_width = max H / W: the outputs are zero-filled and the
 *   stride lattices (f i, f j) of the configured strides are walked, 256 points per warp.
 *   keep      uint8 [sum H*W] at px_begin (may be NULL), bits uint32 [total_tiles*128] (bit i of
 *   word w of tile t = pixel t*4096 + 32*w + i), tile_counts int32 [total_tiles].
 */
int mdi_keep_mask(const MdiView* views, int n_views, int64_t total_tiles, int64_t total_px,
                  int max_height, int max_width, const float* scale_shift, const int32_t* status, const float* lo_hi,
                  const float* grad_range, const uint8_t* patch_blocked,
                  const float* factor_table, const MdiSubsampleConfig* cfg, uint8_t* keep,
                  uint32_t* bits,
                  int32_t* tile_counts, void* stream);

/*
 * compaction offsets in (view-major, row-major) order: tile_offsets int64 [total_tiles] = exclusive
 * scan of tile_counts WITHIN each view; view_counts / view_offsets int64 [V] = points per view and
 * their exclusive scan over views; total int64 [1].  A kept pixel's output slot is
 * view_offsets[v] + tile_offsets[tile] + its rank inside the tile.
 */
int mdi_scan_tiles(const MdiView* views, int n_views, int64_t total_tiles,
                   const int32_t* tile_counts, int64_t* tile_offsets, int64_t* view_counts,
                   int64_t* view_offsets, int64_t* total, void* stream);

/*
 * a10 + a11: for every kept pixel in row-major order: unproject to world space and gather RGB.
 *   xyz float32 [N,3], rgb float32 [N,3] (either may be NULL), written at the slots above.
 */
int mdi_emit_points(const MdiView* views, int n_views, int64_t total_tiles,
                    const float* scale_shift, const uint32_t* bits, const int64_t* tile_offsets,
                    const int64_t* view_offsets, int image_kind, float* xyz, float* rgb,
                    void* stream);

/*
 * The same, launched BEFORE the host knows the output size (which is data dependent and otherwise
 * costs a stream sync between mdi_scan_tiles and this call): xyz / rgb hold capacity_rows rows, and
 * the kernel writes nothing at all when *total_rows (device, the `total` of mdi_scan_tiles) exceeds
 * that, in which case the caller re-emits with mdi_emit_points once it has read the size.
 */
int mdi_emit_points_capped(const MdiView* views, int n_views, int64_t total_tiles,
                           const float* scale_shift, const uint32_t* bits, const int64_t* tile_offsets,
                           const int64_t* view_offsets, int image_kind, float* xyz, float* rgb,
                           const int64_t* total_rows, int64_t capacity_rows, void* stream);

/*
 * a10/a11 with the exchange step of a multi-GPU run fused in (replaces the torch.cat of
 * monocular_depth_init.py:183-184 across ranks): every kept point of this GPU is written to n_dst
 * clouds -- its own and its NVLink peers', mapped into this address space by the caller (CUDA IPC /
 * symmetric memory) -- at row row_offset + its local slot, so that after all ranks have run it and
 * met at a barrier every destination holds the whole scene in view order.  xyz_dst / rgb_dst are
 * HOST arrays of n_dst device pointers (either may be NULL to skip that array), n_dst <= 16.
 */
#define MDI_MAX_EMIT_DST 16
int mdi_emit_points_multi(const MdiView* views, int n_views, int64_t total_tiles,
                          const float* scale_shift, const uint32_t* bits, const int64_t* tile_offsets,
                          const int64_t* view_offsets, int image_kind, int n_dst,
                          float* const* xyz_dst, float* const* rgb_dst, int64_t row_offset,
                          void* stream);

/*
 * Same exchange through NVSwitch multicast (NVLS): xyz_mc / rgb_mc are multicast addresses bound to
 * the clouds of every GPU of the group (this one included), so one multimem.st per word reaches
 * all of them -- egress per GPU is its own points once instead of once per peer.
 */
int mdi_emit_points_multicast(const MdiView* views, int n_views, int64_t total_tiles,
                              const float* scale_shift, const uint32_t* bits, const int64_t* tile_offsets,
                              const int64_t* view_offsets, int image_kind, float* xyz_mc, float* rgb_mc,
                              int64_t row_offset, void* stream);

/*
 * Count exchange of a multi-GPU step on the device (no NCCL call, no host synchronisation in front
 * of the emit kernel).  hdr_ptrs: HOST array of `world` device pointers, hdr_ptrs[q] = rank q's
 * peer-mapped header of 2 * world int64 slots (zero-filled once, before the first call, on every
 * rank).  local: device int64 [2] = (rows of this rank, retry flag) -- the `total` / retry_flag pair
 * written by mdi_scan_tiles / mdi_depth_range.  epoch: 1, 2, 3, ... -- the same on every rank for the
 * same step (the call is collective: every rank launches it once per step, in the same order).
 * out: device int64 [3 + world] = (first row of this rank in the scene cloud, rows of all ranks,
 * 1 if any rank's retry flag is set, rows of rank 0, 1, ...).
 */
int mdi_exchange_counts(int64_t* const* hdr_ptrs, int rank, int world, uint32_t epoch,
                        const int64_t* local, int64_t* out, void* stream);

/*
 * mdi_emit_points_multi / _multicast with the row offset taken from the record written by
 * mdi_exchange_counts (device): rows go to exchange_record[0] + local slot.  Nothing is written when
 * exchange_record[1] (rows of all ranks) exceeds capacity_rows; the caller, who reads the record
 * after the launch, then grows the clouds and launches again.  n_dst == 0 selects the multicast
 * addresses.  Destination base pointers must be 16-byte aligned (records leave as 128-bit stores).
 */
int mdi_emit_points_exchange(const MdiView* views, int n_views, int64_t total_tiles,
                             const float* scale_shift, const uint32_t* bits, const int64_t* tile_offsets,
                             const int64_t* view_offsets, int image_kind, int n_dst,
                             float* const* xyz_dst, float* const* rgb_dst, float* xyz_mc, float* rgb_mc,
                             const int64_t* exchange_record, int64_t capacity_rows,
                             uint8_t* const* rgb8_dst, int self_rank, void* stream);

/*
 * Byte-rgb exchange (uint8 images, n_dst > 1, rgb8_dst != NULL): the rgb a peer needs is the three
 * gathered BYTES of a point -- the float rgb is u8 / 255.0f, which the receiver computes itself with
 * the same division.  Peers then receive 12 + 3 instead of 12 + 12 bytes per point over NVLink:
 * mdi_emit_points_exchange writes float rgb into rgb_dst[self_rank] only and the bytes into every
 * other rank's rgb8_dst[d] (uint8 [capacity_rows * 3], peer-mapped, 16-byte aligned), and after the
 * step's barrier every rank calls mdi_expand_rgb8 on its own byte cloud to fill the other ranks' rows
 * of its float rgb cloud.  Bit-identical to exchanging the floats.
 */
int mdi_expand_rgb8(const uint8_t* rgb8, float* rgb, const int64_t* exchange_record, int rank,
                    int64_t capacity_rows, void* stream);

/*
 * a13: minimal Gaussian extent per point over all cameras that see it (else -1).
 *   K float32 [C,9], P float32 [C,12], sizes int32 [C,2] = (W,H).
 */
int mdi_min_extents(const float* pts, int64_t n, const float* K, const float* P,
                    const int32_t* sizes, int n_cams, float* extents, void* stream);

/*
 * a14: merge subsampling of the final cloud (impl.cpp:313-426 with spatial_median :261-274).
 *   pts, rgb float32 [n,3]; extents float32 [n] (mdi_min_extents).  out_pts / out_rgb float32 [n,3]
 *   (worst case: nothing merges) receive *n_out rows in the reference's emission order (DFS, right
 *   child first; merged rows are sequential fp32 means in original index order).
 *   The tree is built level by level; the call synchronises `stream` once per level and on return,
 *   so n_out / n_levels are HOST pointers.  max_levels <= 0 selects 1024; when frames are still
 *   open after max_levels the call fails with MDI_EINVAL (>= 3 coincident points whose merge test
 *   fails -- the reference loops forever on that input).  The debug cloud of the reference
 *   (impl.cpp:322-323,378-388) is not produced.
 */
int64_t mdi_merge_workspace_bytes(int64_t n);
int mdi_merge_subsample(const float* pts, const float* rgb, const float* extents, int64_t n,
                        float max_bbox_aspect_ratio, float min_extent_mult, int max_levels,
                        void* workspace, float* out_pts, float* out_rgb, int64_t* n_out,
                        int32_t* n_levels, void* stream);

/*
 * a12 / 8f-1: exact k nearest neighbours of every point of the cloud among the cloud
 * (sklearn NearestNeighbors(k).kneighbors, utils/runner_utils.py:143-147; LocalOutlierFactor's
 * neighbour search, point_cloud_postprocess/postprocess.py:16-22).
 *   pts float32 [n,3]; idx int32 [n,k] original indices in ascending (distance, index) order;
 *   dist float32 [n,k] or NULL; kdist float32 [n] (distance to the k-th neighbour) or NULL.
 *   exclude_self != 0 leaves the query point itself out (LOF); 0 keeps it (kneighbors(x) of the
 *   fitted x: first column is the point itself at distance 0).  1 <= k <= 64.
 *   Distances: differences, squares, sum and square root in fp64 (sklearn's KDTree), cast to fp32.
 */
int64_t mdi_knn_workspace_bytes(int64_t n);
int mdi_knn(const float* pts, int64_t n, int k, int exclude_self, void* workspace, int32_t* idx,
            float* dist, float* kdist, void* stream);

/*
 * a12: Local Outlier Factor from the neighbours of mdi_knn(exclude_self=1) (sklearn _lof.py):
 *   idx / dist / kdist as written by mdi_knn; lrd float32 [n] (scratch / result), lof float32 [n] or
 *   NULL (= -negative_outlier_factor_), outlier uint8 [n] = (-lof < -1.5), the contamination="auto"
 *   rule of fit_predict.  All arithmetic in fp32 with numpy's summation order, as sklearn does for
 *   fp32 input.
 */
int mdi_lof(int64_t n, int k, const int32_t* idx, const float* dist, const float* kdist, float* lrd,
            float* lof, uint8_t* outlier, void* stream);

/*
 * `interp` aligner (8f-3): evaluates scipy's LinearNDInterpolator at every pixel (x, y) of an H x W
 * view and multiplies the pre-aligned depth by it (interp.py:75-110, 355-361).
 *   transform float64 [T,3,2], simplices int32 [T,3], points float64 [P,2] (x, y): the arrays of the
 *   scipy.spatial.Delaunay object the host built (dt.transform, dt.simplices, dt.points), on the device;
 *   values float64 [P] per-vertex scale factors; fill_value for pixels outside the hull.
 *   base float32 [H*W] predicted depth and scale_shift float32 [2] (device) of the init alignment:
 *   out[p] = float32(value(p)) * (base[p] * s + t); scale_shift NULL: out = float32(value) * base;
 *   base NULL: out = float32(value) (the scale map itself).
 *   workspace: mdi_interp_workspace_bytes(H, W) bytes.
 */
int64_t mdi_interp_workspace_bytes(int H, int W);
int mdi_interp_scale_map(const double* transform, const int32_t* simplices, const double* points,
                         const double* values, int n_tri, double fill_value, int H, int W,
                         const float* base, const float* scale_shift, void* workspace, float* out,
                         void* stream);

/*
 * mdi_interp_scale_map without any allocation by caller or library.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * mdi_interp_scale_map minus `workspace`: the per-pixel owner map lives in a static device array of
 * MDI_INTERP_STATIC_MAX_PX int32 (8 MiB, held by every GPU that loads the library until exit).
 * H * W beyond that limit is rejected with MDI_EINVAL.  The call itself uses no cudaMalloc / cudaFree
 * or allocating routine.  ALL calls on a device share the one array: it is refilled at the start of
 * every call (nothing of an earlier call survives), and calls that overlap on the GPU (different
 * streams, or threads) race on it -- serialise them or use mdi_interp_scale_map.
 */
#define MDI_INTERP_STATIC_MAX_PX (2 * 1024 * 1024)
int mdi_interp_scale_map_static(const double* transform, const int32_t* simplices, const double* points,
                                const double* values, int n_tri, double fill_value, int H, int W,
                                const float* base, const float* scale_shift, float* out, void* stream);

/*
 * Query split of a12 / a13 over ranks ("stages that do not shard by view"): every rank
 * holds the whole cloud (after the exchange) and builds the same grid (mdi_knn_build), but answers
 * only the queries at SORTED positions q_begin .. q_end - 1 (mdi_knn_query; Morton order of the grid,
 * so a rank's queries are spatially coherent).  mdi_knn_split writes the rank boundaries bounds
 * [world + 1] (device int64): cell-aligned cuts near r * ceil(n / world) -- whole cells, because the
 * order of the points inside a cell differs between ranks.  With self_idx != NULL results go to
 * compact rows 0 .. q_end - q_begin - 1 and self_idx int32 [rows] receives the original index of
 * each row (NULL: rows are the original indices, as mdi_knn).  The caller gathers the 4-byte
 * k-distances, then the 4-byte lrd, then the flags of all ranks (mdi_lof_lrd / mdi_lof_score take the
 * gathered FULL arrays for the neighbours and the rank's own rows for the queries).  Every query's
 * answer is independent of the split, so the result is bit-identical for any number of ranks.
 * mdi_min_extents needs no variant: pass the rank's slice of pts.  mdi_knn = build + query(0, n).
 */
int mdi_knn_build(const float* pts, int64_t n, void* workspace, void* stream);
int mdi_knn_split(void* workspace, int64_t n, int world, int64_t* bounds, void* stream);
int mdi_knn_query(void* workspace, int64_t n, int k, int exclude_self, int64_t q_begin, int64_t q_end,
                  int32_t* self_idx, int32_t* idx, float* dist, float* kdist, void* stream);
int mdi_lof_lrd(int64_t m, int k, const int32_t* idx, const float* dist, const float* kdist_full,
                float* lrd_rows, void* stream);
int mdi_lof_score(int64_t m, int k, const int32_t* idx, const float* lrd_full, const float* lrd_rows,
                  float* lof, uint8_t* outlier, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MDI_B200_H_ */
