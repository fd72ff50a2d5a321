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
  // fp32 BEV candidate: t = v * bev_inv_f + bev_c0_f (in bins), guard band bev_eps_f bins
  float bev_inv_x_f, bev_c0_x_f, bev_eps_x_f;
  float bev_inv_y_f, bev_c0_y_f, bev_eps_y_f;
  // derived once on the host so that the point pass takes them as constant-bank operands (the same
  // fp32 operations the kernel used to repeat per point): range tests |t - c| < h of range_cell_decide,
  // BEV half extents nb/2, nb/2 + 1 and the fraction bound 0.5 - eps
  float rc_cx_f, rc_hx_f, rc_cy_f, rc_hy_f;
  float bev_hx_f, bev_hy_f, bev_hx1_f, bev_hy1_f, bev_fe_x_f, bev_fe_y_f;
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
// The fused kernel decides every cell index in fp32 and PROVES the decision with a guard band:
// a point whose fp32 cell coordinate lies within EPS of a cell boundary (or outside the safe
// domain of the approximations) is re-done with the fp64 functions above.  Error budgets (the
// device self-test sfpe_selftest_fastpath measures the real deviations, tests/ assert them):
//   column  xi = atan2(-y,x)/res_h - X_MIN:  polynomial 1.3e-7 rad + quotient 1.8e-7 + quadrant
//           constants/rounding 2.5e-7 = 5.6e-7 rad = 1.6e-4 cells, + fma rounding and 1/res_h
//           1.1e-4 cells  -> < 3e-4 cells,  EPS_COL = 1e-3
//   row     yi = asin(z/rho)/res_v - Y_MIN, |z/rho| <= 0.55:  < 3e-5 cells, EPS_ROW = 2.5e-4
//   range   t = 64*rho: relative 4e-7 -> guard 1e-6*t + 1e-3 buckets
//   BEV     t = (v+R)/delta: < 1.5 ulp32(number of bins) -> guard 3 ulp32 (host: bev_eps_f)
constexpr float EPS_COL = 1.0e-3f;
constexpr float EPS_ROW = 2.5e-4f;
constexpr float MAX_SIN_EL = 0.55f;

// atan2(yy, xx) in (-pi, pi]: octant reduction to t = min/max in [0,1], odd minimax polynomial
// (degree 15, fitted for abs error 1.3e-7 incl. fp32 Horner evaluation).  0/0, inf/inf and NaN
// give NaN, which every guard test below rejects.
__device__ __forceinline__ float atan2_fast(float yy, float xx) {
  const float ax = fabsf(xx), ay = fabsf(yy);
  const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
  float rc;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(mx));   // 1 ulp; 0 -> inf -> NaN below, rejected by the guards
  const float t = mn * rc;
  const float s = t * t;
  float p = -0.004054493270814419f;
  p = fmaf(p, s, 0.021862683817744255f);
  p = fmaf(p, s, -0.055911920964717865f);
  p = fmaf(p, s, 0.09642166644334793f);
  p = fmaf(p, s, -0.13908617198467255f);
  p = fmaf(p, s, 0.19946563243865967f);
  p = fmaf(p, s, -0.33329859375953674f);
  p = fmaf(p, s, 0.9999993443489075f);
  float r = p * t;
  if (ay > ax) r = 1.57079637f - r;
  if (xx < 0.0f) r = 3.14159274f - r;
  return copysignf(r, yy);
}

// asin(q) for |q| <= MAX_SIN_EL: odd minimax polynomial, abs error 8e-8 incl. evaluation.
__device__ __forceinline__ float asin_fast(float q) {
  const float s = q * q;
  float p = 0.05534907057881355f;
  p = fmaf(p, s, 0.014851636253297329f);
  p = fmaf(p, s, 0.047862760722637177f);
  p = fmaf(p, s, 0.07468582689762115f);
  p = fmaf(p, s, 0.16667962074279785f);
  p = fmaf(p, s, 0.9999998211860657f);
  return p * q;
}

// 1/sqrt(s): MUFU.RSQ (2 ulp) + one Newton step -> ~1 ulp, 5 instructions (the correctly rounded
// __frsqrt_rn costs 23).  s = 0 / inf / NaN give NaN or 0, rejected by the guards.
__device__ __forceinline__ float rsqrt_nr(float s) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
  const float e = fmaf(-s * r, r, 1.0f);
  return fmaf(0.5f * r, e, r);
}

// fp32 cell coordinates (before truncation) of the range scan: xi (column), yi (row);
// q = z/rho as computed from rs = rsqrt(x^2+y^2+z^2).
__device__ __forceinline__ void range_coords_fast(const DevGrid& g, float x, float y, float z, float rs, float& xi,
                                                  float& yi, float& q) {
  xi = fmaf(atan2_fast(-y, x), g.inv_res_h_f, g.neg_x_min_f);
  q = z * rs;
  yi = fmaf(asin_fast(q), g.inv_res_v_f, g.neg_y_min_f);
}

// floor / round-to-nearest of |t| < 2^22 on the FMA pipe (magic-number adds) instead of the
// quarter-rate conversion unit (FRND / F2I); fi = the integer value.  Outside that range (and
// for NaN) the results are garbage -- every caller masks such t out by its own range tests.
__device__ __forceinline__ float floor_small(float t, int& fi) {
  const float m = __fadd_rd(t, 12582912.0f);             // 1.5 * 2^23: ulp 1 for sums in [2^23, 2^24)
  fi = __float_as_int(m) - 0x4B400000;
  return m - 12582912.0f;
}
__device__ __forceinline__ float rint_small(float t) { return __fadd_rn(t, 12582912.0f) - 12582912.0f; }

// Returns true when (row, col) = trunc(yi), trunc(xi) is PROVEN equal to the fp64 result and
// inside the image; false -> the caller must use range_cell_f64.  A coordinate is proven when it
// lies at least EPS away from both neighbouring integers (the fraction is within 0.5 - EPS of 0.5);
// under truncation toward zero the integer 0 is not a cell boundary, but the few points within EPS
// of it are simply left to the fp64 path as well.  Accepted coordinates are > -0.5, where truncation
// toward zero = max(floor, 0).  The range test (-0.5, W) is |xi - c| < h with c = W/2 - 0.25,
// h = W/2 + 0.25 (both exact in fp32, and rounding of the difference is monotone, so a coordinate
// outside the interval can never pass).  Written so that NaN anywhere gives false.
__device__ __forceinline__ bool range_cell_decide(const DevGrid& g, float xi, float yi, float q, int& row, int& col) {
  int fx, fy;
  const float frx = xi - floor_small(xi, fx), fry = yi - floor_small(yi, fy);
  col = max(fx, 0);
  row = max(fy, 0);
  // c = g.rc_c?_f, h = g.rc_h?_f (host: derive_grid); no short circuit: five compares, no branch
  const bool in_fx = fabsf(frx - 0.5f) < 0.5f - EPS_COL, in_fy = fabsf(fry - 0.5f) < 0.5f - EPS_ROW;
  const bool in_q = fabsf(q) <= MAX_SIN_EL;
  const bool in_x = fabsf(xi - g.rc_cx_f) < g.rc_hx_f, in_y = fabsf(yi - g.rc_cy_f) < g.rc_hy_f;
  return in_fx & in_fy & in_q & in_x & in_y;
}

// fp32 BEV bin candidate with guard band; -2 = undecided (caller uses bev_bin_f64, which also
// handles the out-of-range / last-edge / NaN cases).
__device__ __forceinline__ int bev_bin_f32(float v, float inv, float c0, float eps, int n_bins) {
  const float t = fmaf(v, inv, c0);
  const float fl = floorf(t);
  const float fr = t - fl;
  const bool ok = fr > eps && fr < 1.0f - eps && t > 0.0f && t < (float)n_bins;   // NaN -> false
  return ok ? (int)fl : -2;
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
