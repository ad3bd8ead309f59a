/* TEST INFRASTRUCTURE -- CPU restatement, in plain C, of the reference's native module
 * (native_modules/subsampling/src/impl.cpp, geometry.h, stack.h).  NOT product code: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load the library this
 * file builds into (oracle/_build/libmdi_native_oracle.so, see oracle/build_oracle.py).
 *
 * PARITY PIN: tests/test_native_ref_pin.py compares this file, bit for bit (extents, output count,
 * emission order, positions, colours), with the reference's OWN impl.cpp / pointcloud_subsampling.cpp
 * compiled unmodified into oracle/_ref/ by oracle/build_ref_native.py.  That build replaces Eigen
 * 3.4 (third-party, absent here: no Conan, no network) by the stand-in headers of
 * oracle/eigen_standin/, so the reference's control flow, partition rule, stack discipline and
 * emission order are the reference's own code; what remains restated (here and in the stand-in
 * alike) are Eigen's published evaluation rules for the few expressions used:
 *   - `P * point.homogeneous()` (impl.cpp:19-21) evaluates as leftCols(3)*p + col(3)
 *     (Eigen/src/Geometry/Homogeneous.h), each 3-term dot product reduced by the unrolled
 *     non-vectorised redux as a0 + (a1 + a2) (Eigen/src/Core/Redux.h);
 *   - maxCoeff/minCoeff of a 3-vector reduce as op(d0, op(d1, d2)) with std::max/std::min
 *     argument-order semantics; cwiseMin/cwiseMax are std::min/std::max per coefficient;
 *   - `v /= s` is a true division per coefficient.
 * Build with -ffp-contract=off so that no product is fused into an FMA.
 */
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline float std_min(float a, float b) { return (b < a) ? b : a; } /* std::min(a,b) */
static inline float std_max(float a, float b) { return (a < b) ? b : a; } /* std::max(a,b) */
static inline float max3(const float d[3]) { return std_max(d[0], std_max(d[1], d[2])); }
static inline float min3(const float d[3]) { return std_min(d[0], std_min(d[1], d[2])); }

/* impl.cpp:17-35 get_point_depth_in_camera: returns 1 and *depth when the camera sees p. */
static int point_depth_in_camera(const float p[3], const float P[12], int width, int height,
                                 float* depth) {
  float q[3];
  for (int r = 0; r < 3; ++r) {
    const float a0 = P[4 * r + 0] * p[0], a1 = P[4 * r + 1] * p[1], a2 = P[4 * r + 2] * p[2];
    q[r] = (a0 + (a1 + a2)) + P[4 * r + 3];
  }
  const float z = q[2];
  if (z <= 0) return 0;
  const float u = q[0] / z, v = q[1] / z;
  if (u < 0 || u >= (float)width || v < 0 || v >= (float)height) return 0;
  *depth = z;
  return 1;
}

/* impl.cpp:70-126 compute_minimal_gaussian_extents.  K [C,3,3], P [C,3,4] row-major (the
 * pybind11 Eigen caster copies numpy row-major data into the column-major matrices, so
 * element (r,c) is what matters), sizes [C,2] = (width, height). */
void oracle_min_extents(const float* pts, int64_t n, const float* K, const float* P,
                        const int32_t* sizes, int n_cams, float* ext) {
  for (int64_t i = 0; i < n; ++i) {
    float best = FLT_MAX;
    for (int c = 0; c < n_cams; ++c) {
      float depth;
      if (!point_depth_in_camera(pts + 3 * i, P + 12 * (int64_t)c, sizes[2 * c], sizes[2 * c + 1], &depth))
        continue;
      const float fx = K[9 * (int64_t)c + 0], fy = K[9 * (int64_t)c + 4];
      const float f = std_min(fx, fy);
      const float interval = depth / f;
      best = std_min(best, 2.0f * interval);
    }
    ext[i] = (best != FLT_MAX) ? best : -1.0f;
  }
}

typedef struct {
  int64_t begin, end; /* slice of the index buffer; order inside = original index order */
  float bmin[3], bmax[3];
  int prev_axis;
} Frame;

/* impl.cpp:313-426 subsample_pointcloud_impl with split_fn = spatial_median (:261-274).
 * The reference gives every frame its own index vectors; a stable in-place partition of one
 * shared index buffer visits the same indices in the same order.  Returns the number of
 * output points, or -1 when `max_frames` frames were popped without finishing (the reference
 * does not terminate on >= 3 coincident points whose merge test fails). */
int64_t oracle_merge_subsample(const float* pts, const float* rgb, const float* ext, int64_t n,
                               float max_ratio, float mult, float* out_pts, float* out_rgb,
                               int64_t max_frames) {
  if (n == 0) return 0;
  uint32_t* idx = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
  uint32_t* tmp = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
  int64_t cap = 1024, top = 0, n_out = 0, popped = 0;
  Frame* stack = (Frame*)malloc(sizeof(Frame) * (size_t)cap);
  for (int64_t i = 0; i < n; ++i) idx[i] = (uint32_t)i;

  /* geometry.h:56-63 cube_from_points */
  Frame root;
  root.begin = 0; root.end = n; root.prev_axis = 0;
  {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}, d[3];
    for (int64_t i = 0; i < n; ++i)
      for (int k = 0; k < 3; ++k) {
        lo[k] = std_min(lo[k], pts[3 * i + k]);
        hi[k] = std_max(hi[k], pts[3 * i + k]);
      }
    for (int k = 0; k < 3; ++k) d[k] = hi[k] - lo[k];
    const float half = max3(d) / 2.0f;
    for (int k = 0; k < 3; ++k) {
      const float c = (lo[k] + hi[k]) / 2.0f;
      root.bmin[k] = c - half;
      root.bmax[k] = c + half;
    }
  }
  stack[top++] = root;

  while (top > 0) {
    if (++popped > max_frames) { n_out = -1; break; }
    const Frame f = stack[--top];
    const int64_t cnt = f.end - f.begin;
    if (cnt == 0) continue;
    if (cnt == 1) {
      const uint32_t ix = idx[f.begin];
      memcpy(out_pts + 3 * n_out, pts + 3 * (int64_t)ix, 12);
      memcpy(out_rgb + 3 * n_out, rgb + 3 * (int64_t)ix, 12);
      ++n_out;
      continue;
    }
    float avg = 0; /* :223-230 */
    for (int64_t p = f.begin; p < f.end; ++p) avg += ext[idx[p]];
    avg = avg / (float)cnt;

    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int64_t p = f.begin; p < f.end; ++p)
      for (int k = 0; k < 3; ++k) {
        const float x = pts[3 * (int64_t)idx[p] + k];
        lo[k] = std_min(lo[k], x);
        hi[k] = std_max(hi[k], x);
      }
    float db[3], dt[3];
    for (int k = 0; k < 3; ++k) { db[k] = f.bmax[k] - f.bmin[k]; dt[k] = hi[k] - lo[k]; }
    const float a_box = max3(db) / min3(db);
    const float a_tight = max3(dt) / min3(dt);
    const float aspect = std_min(a_box, a_tight);
    const float thresh = max3(dt);

    if (aspect <= max_ratio && thresh <= mult * avg) { /* :372-390 merge */
      float mp[3] = {0, 0, 0}, mc[3] = {0, 0, 0};
      for (int64_t p = f.begin; p < f.end; ++p)
        for (int k = 0; k < 3; ++k) {
          mp[k] += pts[3 * (int64_t)idx[p] + k];
          mc[k] += rgb[3 * (int64_t)idx[p] + k];
        }
      for (int k = 0; k < 3; ++k) {
        out_pts[3 * n_out + k] = mp[k] / (float)cnt;
        out_rgb[3 * n_out + k] = mc[k] / (float)cnt;
      }
      ++n_out;
      continue;
    }
    if (cnt <= 2) { /* :391-398 */
      for (int64_t p = f.begin; p < f.end; ++p) {
        memcpy(out_pts + 3 * n_out, pts + 3 * (int64_t)idx[p], 12);
        memcpy(out_rgb + 3 * n_out, rgb + 3 * (int64_t)idx[p], 12);
        ++n_out;
      }
      continue;
    }
    const int axis = (f.prev_axis + 1) % 3; /* :400 */
    const float split = (f.bmin[axis] + f.bmax[axis]) / 2.0f;
    int64_t nl = 0, nr = 0;
    for (int64_t p = f.begin; p < f.end; ++p) {
      if (pts[3 * (int64_t)idx[p] + axis] < split) idx[f.begin + nl++] = idx[p];
      else tmp[nr++] = idx[p];
    }
    memcpy(idx + f.begin + nl, tmp, sizeof(uint32_t) * (size_t)nr);
    if (top + 2 > cap) { cap *= 2; stack = (Frame*)realloc(stack, sizeof(Frame) * (size_t)cap); }
    Frame l = f, r = f; /* geometry.h:67-74 split; :410-411 push left, then right */
    l.end = f.begin + nl; l.bmax[axis] = split; l.prev_axis = axis;
    r.begin = f.begin + nl; r.bmin[axis] = split; r.prev_axis = axis;
    stack[top++] = l;
    stack[top++] = r;
  }
  free(stack); free(tmp); free(idx);
  return n_out;
}
