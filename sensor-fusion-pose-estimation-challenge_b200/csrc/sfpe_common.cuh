// Shared device-side definitions for libsfpe (sm_100a).
//
// Arithmetic contract: every expression that decides a cell index or a winner is
// evaluated in fp64 with explicitly non-fused IEEE operations (__dmul_rn, __dadd_rn,
// __ddiv_rn, __dsqrt_rn) in the operation order of the reference's numpy code
// (modules/lidar/process/extract_rosbag_lidar.py:32-39,60-61 and numpy's histogram2d /
// arange rules for modules/lidar/common/birds_eye_view_generator.py:43-44), so that the
// only possible deviation from numpy is the last-ulp behaviour of atan2 / asin / log.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sfpe {

constexpr unsigned long long KEY_EMPTY = 0xFFFFFFFFFFFFFFFFull;
constexpr int ERR_ROW_OOB = 1, ERR_COL_OOB = 2, ERR_NAN = 4;

// Host-derived grid constants (sfpe_api.cu: derive_grid), passed by value to kernels.
struct DevGrid {
  double res_rad_h, res_rad_v;  // RES_RAD[1], RES_RAD[0]     globals.py:9
  double x_min, y_min;          // X_MIN, Y_MIN               globals.py:10-11
  double height_bg;             // LIDAR_MIN_HEIGHT           globals.py:4
  double bev_start;             // -max_range
  double bev_delta_x, bev_delta_y;  // np.arange fill delta = (start+step)-start
  int H, W;                     // Y_MAX+1, X_MAX+1           globals.py:12-13
  int nex, ney;                 // number of bin EDGES per axis (bins = edges-1)
  int nbx, nby;
  int idx_bits;                 // low bits of the packed winner key that hold the point index
  int use_extrinsic;
  int n_slices;
  int enable_range, enable_bev;
  double M[12];
  double slice_edges[9];
  // fast path constants (fused kernel)
  double bev_inv_delta_x, bev_inv_delta_y;   // 1/delta (candidate bin only; edges decide)
  float inv_res_h_f, inv_res_v_f;            // 1/RES_RAD[1], 1/RES_RAD[0]
  float neg_x_min_f, neg_y_min_f;            // -X_MIN, -Y_MIN
};

struct PointSrc {
  const char* data;
  long long stride;
  int dtype;  // 0 f32, 1 f64
  int ox, oy, oz, oi;
};

struct Pt {
  double x, y, z, it;
};

enum { SRC_PACKED_F32 = 0, SRC_GENERIC = 1 };

template <int KIND>
__device__ __forceinline__ Pt load_point_raw(const PointSrc& s, long long p) {
  Pt r;
  if (KIND == SRC_PACKED_F32) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(s.data) + p);
    r.x = v.x; r.y = v.y; r.z = v.z; r.it = v.w;
  } else {
    const char* b = s.data + p * s.stride;
    if (s.dtype == 0) {
      r.x = *reinterpret_cast<const float*>(b + s.ox);
      r.y = *reinterpret_cast<const float*>(b + s.oy);
      r.z = *reinterpret_cast<const float*>(b + s.oz);
      r.it = *reinterpret_cast<const float*>(b + s.oi);
    } else {
      r.x = *reinterpret_cast<const double*>(b + s.ox);
      r.y = *reinterpret_cast<const double*>(b + s.oy);
      r.z = *reinterpret_cast<const double*>(b + s.oz);
      r.it = *reinterpret_cast<const double*>(b + s.oi);
    }
  }
  return r;
}

// Optional rigid extrinsic (extension; default off).  Defined as: evaluate the affine map
// in fp64 left to right without fusion, then round the result to float32 -- i.e. the
// transformed cloud is again a float32 cloud.
__device__ __forceinline__ void apply_extrinsic(const DevGrid& g, Pt& p) {
  const double x = p.x, y = p.y, z = p.z;
  double o[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    double a = __dmul_rn(g.M[4 * r + 0], x);
    a = __dadd_rn(a, __dmul_rn(g.M[4 * r + 1], y));
    a = __dadd_rn(a, __dmul_rn(g.M[4 * r + 2], z));
    a = __dadd_rn(a, g.M[4 * r + 3]);
    o[r] = (double)(float)a;
  }
  p.x = o[0]; p.y = o[1]; p.z = o[2];
}

template <int KIND>
__device__ __forceinline__ Pt load_point(const DevGrid& g, const PointSrc& s, long long p) {
  Pt r = load_point_raw<KIND>(s, p);
  if (g.use_extrinsic) apply_extrinsic(g, r);
  return r;
}

// fp64 l2 norm exactly as numpy evaluates sqrt(x**2 + y**2 + z**2)   (:33)
__device__ __forceinline__ double rho_of(const Pt& p) {
  const double s = __dadd_rn(__dadd_rn(__dmul_rn(p.x, p.x), __dmul_rn(p.y, p.y)), __dmul_rn(p.z, p.z));
  return __dsqrt_rn(s);
}
// sqrt(x**2 + y**2)   (:32)
__device__ __forceinline__ double dist_of(const Pt& p) {
  return __dsqrt_rn(__dadd_rn(__dmul_rn(p.x, p.x), __dmul_rn(p.y, p.y)));
}

// Range-scan cell of a point, full fp64 (:34-39, :60-61 + numpy negative-index wrap).
// Returns the error mask (0 = ok); row is the PRE-flip row.
__device__ __forceinline__ int range_cell_f64(const DevGrid& g, const Pt& p, double rho, int& row, int& col) {
  const double xi = __dsub_rn(__ddiv_rn(atan2(-p.y, p.x), g.res_rad_h), g.x_min);
  const double yi = __dsub_rn(__ddiv_rn(asin(__ddiv_rn(p.z, rho)), g.res_rad_v), g.y_min);
  if (xi != xi || yi != yi) return ERR_NAN;
  // astype(int): truncation toward zero.  |xi|,|yi| are bounded (angles), no overflow.
  const int c = __double2int_rz(xi);
  const int r = __double2int_rz(yi);
  int err = 0;
  if (r < -g.H || r >= g.H) err |= ERR_ROW_OOB;
  if (c < -g.W || c >= g.W) err |= ERR_COL_OOB;
  row = r < 0 ? r + g.H : r;
  col = c < 0 ? c + g.W : c;
  return err;
}

// Packed winner key: high (64-idx_bits) bits = leading bits of the fp64 pattern of rho
// (sign bit dropped; rho >= 0 so the pattern is order preserving), low idx_bits = index of the
// point inside its frame.  atomicMin on it selects (smallest rho prefix, lowest index); equal
// prefixes with different full rho are resolved exactly by settle_collision().
__device__ __forceinline__ unsigned long long pack_key(double rho, unsigned idx, int idx_bits) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(rho) << 1;
  const unsigned long long m = (1ull << idx_bits) - 1ull;
  return (b & ~m) | (unsigned long long)idx;
}

// exact order: (rho fp64, index)
__device__ __forceinline__ bool better(double ra, unsigned ia, double rb, unsigned ib) {
  return ra < rb || (ra == rb && ia < ib);
}

// np.arange edge i:  start + i*delta                           (numpy DOUBLE_fill)
__device__ __forceinline__ double bev_edge(double start, double delta, int i) {
  return __dadd_rn(start, __dmul_rn((double)i, delta));
}

// histogram2d bin of v against the n_edges arange edges: left-closed bins, last bin
// right-closed, NaN/inf/out of range -> -1.   (searchsorted(edges, v, 'right') - 1)
__device__ __forceinline__ int bev_bin_f64(double v, double start, double delta, int n_edges) {
  if (!(v >= start)) return -1;  // also NaN
  const double last = bev_edge(start, delta, n_edges - 1);
  if (v >= last) return v == last ? n_edges - 2 : -1;
  double t = __ddiv_rn(__dsub_rn(v, start), delta);
  int i = (int)fmin(t, (double)(n_edges - 2));
  if (i < 0) i = 0;
  while (i > 0 && v < bev_edge(start, delta, i)) --i;
  while (i < n_edges - 2 && v >= bev_edge(start, delta, i + 1)) ++i;
  return i;
}

// ---- fast path helpers (fused kernel) -------------------------------------------------
// Guard bands of the fp32 cell computation: a point whose fp32 cell coordinate lies within
// EPS of a (non-zero) integer is re-done in fp64.  Bounds: atan2f/asinf <= 2 ulp (CUDA math
// API), one fp32 division and one fma -> |xi32 - xi64| < 2.6e-4 cells, |yi32 - yi64| < 3e-5
// cells for |z/rho| < 0.95 ("fp32 fast path"); sfpe_selftest_fastpath measures it.
constexpr float EPS_COL = 1.0e-3f;
constexpr float EPS_ROW = 2.5e-4f;
constexpr float MAX_SIN_EL = 0.95f;

// fp32 range-scan cell.  Returns true when (row, col) is PROVEN equal to the fp64 result and
// inside the image; false -> the caller must use range_cell_f64.
__device__ __forceinline__ bool range_cell_fast(const DevGrid& g, float x, float y, float z, float rho32, int& row,
                                                int& col) {
  const float xi = fmaf(atan2f(-y, x), g.inv_res_h_f, g.neg_x_min_f);
  const float q = z / rho32;
  const float yi = fmaf(asinf(q), g.inv_res_v_f, g.neg_y_min_f);
  const float cx = rintf(xi), cy = rintf(yi);
  // trunc-toward-zero semantics: the cell boundaries are the NON-ZERO integers
  const bool near_x = fabsf(xi - cx) <= EPS_COL && cx != 0.0f;
  const bool near_y = fabsf(yi - cy) <= EPS_ROW && cy != 0.0f;
  col = (int)xi;
  row = (int)yi;
  // written so that NaN anywhere gives false
  return !near_x && !near_y && fabsf(q) < MAX_SIN_EL && xi > -0.5f && yi > -0.5f && col < g.W && row < g.H;
}

// BEV bin with one fp64 multiply for the candidate and the exact np.arange edges for the
// decision; falls back to bev_bin_f64 near the ends or when the candidate is off by one.
__device__ __forceinline__ int bev_bin_fast(double v, double start, double delta, double inv_delta, int n_edges) {
  const double t = (v - start) * inv_delta;
  if (!(t >= 0.0) || !(t < (double)(n_edges - 2))) return bev_bin_f64(v, start, delta, n_edges);
  const int i = (int)t;
  if (v >= bev_edge(start, delta, i) && v < bev_edge(start, delta, i + 1)) return i;
  return bev_bin_f64(v, start, delta, n_edges);
}

}  // namespace sfpe
