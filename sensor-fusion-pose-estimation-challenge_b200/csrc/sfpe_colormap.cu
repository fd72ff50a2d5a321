// Colormap stage (reference: matplotlib ScalarMappable.to_rgba(..., bytes=True)[:,:,0:3] at
// process/extract_rosbag_lidar.py:75 and common/birds_eye_view_generator.py:53).
// PARITY UNPINNED -- the definition implemented here is written in include/sfpe.h.
#include "sfpe_internal.h"

namespace sfpe {

// order-preserving map double -> u64 (finite values)
__device__ __forceinline__ unsigned long long ord_of(double v) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double ord_inv(unsigned long long o) {
  const unsigned long long b = (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o;
  return __longlong_as_double((long long)b);
}

template <typename T>
__global__ void __launch_bounds__(256) k_minmax(const T* __restrict__ img, long long n, unsigned long long* mm) {
  unsigned long long lo = ~0ull, hi = 0ull;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double v = (double)img[i];
    if (v == v && fabs(v) != INFINITY) {
      const unsigned long long o = ord_of(v);
      lo = o < lo ? o : lo;
      hi = o > hi ? o : hi;
    }
  }
  for (int d = 16; d; d >>= 1) {
    const unsigned long long l2 = __shfl_xor_sync(0xffffffffu, lo, d), h2 = __shfl_xor_sync(0xffffffffu, hi, d);
    lo = l2 < lo ? l2 : lo;
    hi = h2 > hi ? h2 : hi;
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm, lo);
    atomicMax(mm + 1, hi);
  }
}

__global__ void k_minmax_init(unsigned long long* mm) { mm[0] = ~0ull; mm[1] = 0ull; }

// matplotlib "jet" segment data (_cm.py _jet_data), piecewise linear in x
__device__ __forceinline__ double seg(const double* xs, const double* ys, int n, double x) {
  for (int i = 1; i < n; ++i)
    if (x <= xs[i]) return ys[i - 1] + (ys[i] - ys[i - 1]) * ((x - xs[i - 1]) / (xs[i] - xs[i - 1]));
  return ys[n - 1];
}
__device__ __forceinline__ void jet(double x, double& r, double& g, double& b) {
  const double rx[5] = {0., 0.35, 0.66, 0.89, 1.}, ry[5] = {0., 0., 1., 1., 0.5};
  const double gx[6] = {0., 0.125, 0.375, 0.64, 0.91, 1.}, gy[6] = {0., 0., 1., 1., 0., 0.};
  const double bx[5] = {0., 0.11, 0.34, 0.65, 1.}, by[5] = {0.5, 1., 1., 0., 0.};
  r = seg(rx, ry, 5, x); g = seg(gx, gy, 6, x); b = seg(bx, by, 5, x);
}

template <typename T>
__global__ void __launch_bounds__(256) k_colormap(const T* __restrict__ img, long long n, int cmap, int autoscale,
                                                  double vmin, double vmax, const unsigned long long* mm,
                                                  unsigned char* __restrict__ rgb) {
  __shared__ unsigned char lut[256 * 3];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    if (cmap == 0) {
      lut[3 * i] = lut[3 * i + 1] = lut[3 * i + 2] = (unsigned char)i;
    } else {
      double r, g, b;
      jet((double)i / 255.0, r, g, b);
      lut[3 * i] = (unsigned char)(255.0 * r); lut[3 * i + 1] = (unsigned char)(255.0 * g);
      lut[3 * i + 2] = (unsigned char)(255.0 * b);
    }
  }
  if (autoscale) {
    if (mm[0] > mm[1]) { vmin = 0.0; vmax = 0.0; }  // no finite value
    else { vmin = ord_inv(mm[0]); vmax = ord_inv(mm[1]); }
  }
  __syncthreads();
  const double span = vmax - vmin;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double v = (double)img[i];
    unsigned char* o = rgb + 3 * i;
    if (v != v) { o[0] = o[1] = o[2] = 0; continue; }
    const double t = span == 0.0 ? 0.0 : (v - vmin) / span;
    double k = t * 256.0;
    int idx = k >= 255.0 ? 255 : (k <= 0.0 ? 0 : (int)k);
    o[0] = lut[3 * idx]; o[1] = lut[3 * idx + 1]; o[2] = lut[3 * idx + 2];
  }
}

cudaError_t launch_colormap(const void* img, int dtype, long long n, int cmap, int autoscale, double vmin, double vmax,
                            unsigned long long* mm, unsigned char* rgb, cudaStream_t st, long long* launches) {
  if (n <= 0) return cudaSuccess;
  long long nb = (n + 255) / 256;
  const int blocks = (int)(nb > 148 * 8 ? 148 * 8 : nb);
  if (autoscale) {
    k_minmax_init<<<1, 1, 0, st>>>(mm);
    if (dtype == 0) k_minmax<float><<<blocks, 256, 0, st>>>((const float*)img, n, mm);
    else k_minmax<double><<<blocks, 256, 0, st>>>((const double*)img, n, mm);
    *launches += 2;
  }
  if (dtype == 0) k_colormap<float><<<blocks, 256, 0, st>>>((const float*)img, n, cmap, autoscale, vmin, vmax, mm, rgb);
  else k_colormap<double><<<blocks, 256, 0, st>>>((const double*)img, n, cmap, autoscale, vmin, vmax, mm, rgb);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
