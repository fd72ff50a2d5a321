// Compact transport of the BEV density image to host consumers.
//
// The density image (common/birds_eye_view_generator.py:31-37, normalize) takes only 64 distinct
// values: 255*min(ln(N+1)/ln 64, 1) for N = 0..63 and 255 beyond.  A host consumer of a batch
// (process/extract_rosbag.py:162-187 keeps every image on the host) is bound by the PCIe link:
// 5.75 MB of float32 per sweep at 0.2 m.  k_bev_pack_u8 turns the finished image into one byte per
// cell (the class index min(N, 63)) on the device -- 1.44 MB over the link -- and
// bev_unpack_host rebuilds the identical float image in the caller's host buffer with the same
// table the kernels used (sfpe_density_lut).  The class is found by exact comparison against that
// table, so unpack(pack(img)) == img bit for bit; a value that is not in the table raises a flag.
#include <stdint.h>
#include <string.h>

#include <thread>
#include <vector>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include "sfpe_internal.h"

namespace sfpe {

namespace {

template <typename T>
__device__ __forceinline__ unsigned density_class(T v, const T* lut, int* bad) {
  if (v == T(0)) return 0u;
  // first guess from the inverse of the formula, then exact comparison with the table
  int n = __float2int_rn(exp2f((float)v * (6.0f / 255.0f)) - 1.0f);
  n = max(1, min(n, 63));
  if (lut[n] == v) return (unsigned)n;
  if (n > 1 && lut[n - 1] == v) return (unsigned)(n - 1);
  if (n < 63 && lut[n + 1] == v) return (unsigned)(n + 1);
  for (int k = 1; k < 64; ++k)
    if (lut[k] == v) return (unsigned)k;
  *bad = 1;  // not a density value (NaN, or an image this library did not write)
  return (unsigned)n;
}

template <typename T>
__global__ void __launch_bounds__(256) k_bev_pack_u8(const T* __restrict__ in, long long n, const double* __restrict__ lut64,
                                                     unsigned char* __restrict__ out, int* __restrict__ bad) {
  __shared__ T lut[65];
  if (threadIdx.x < 65) lut[threadIdx.x] = (T)lut64[threadIdx.x];
  __syncthreads();
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += stride) {
    T v[4];
    if constexpr (sizeof(T) == 4) {
      const float4 f = __ldcs(reinterpret_cast<const float4*>(in) + q);
      v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    } else {
      const double2 a = __ldcs(reinterpret_cast<const double2*>(in) + 2 * q);
      const double2 b = __ldcs(reinterpret_cast<const double2*>(in) + 2 * q + 1);
      v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    }
    unsigned w = 0;
    if (v[0] != T(0) || v[1] != T(0) || v[2] != T(0) || v[3] != T(0)) {
#pragma unroll
      for (int k = 0; k < 4; ++k) w |= density_class<T>(v[k], lut, bad) << (8 * k);
    }
    reinterpret_cast<unsigned*>(out)[q] = w;
  }
  // tail (n not a multiple of 4)
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    out[i] = (unsigned char)density_class<T>(in[i], lut, bad);
  }
}

template <typename T>
void unpack_slice(const uint8_t* codes, long long n, const T* table, T* out) {
  long long i = 0;
#if defined(__SSE2__)
  if constexpr (sizeof(T) == 4) {
    // scalar head up to a 16-byte boundary of the output, then 16 cells per iteration with
    // non-temporal stores (the image is written once and read later by someone else)
    while (i < n && (reinterpret_cast<uintptr_t>(out + i) & 15)) { out[i] = table[codes[i]]; ++i; }
    const __m128i zero = _mm_setzero_si128();
    const __m128 zf = _mm_setzero_ps();
    for (; i + 16 <= n; i += 16) {
      const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(codes + i));
      float* o = reinterpret_cast<float*>(out + i);
      if (_mm_movemask_epi8(_mm_cmpeq_epi8(c, zero)) == 0xFFFF) {
        _mm_stream_ps(o, zf); _mm_stream_ps(o + 4, zf); _mm_stream_ps(o + 8, zf); _mm_stream_ps(o + 12, zf);
      } else {
        alignas(16) float t[16];
        for (int k = 0; k < 16; ++k) t[k] = (float)table[codes[i + k]];
        _mm_stream_ps(o, _mm_load_ps(t)); _mm_stream_ps(o + 4, _mm_load_ps(t + 4));
        _mm_stream_ps(o + 8, _mm_load_ps(t + 8)); _mm_stream_ps(o + 12, _mm_load_ps(t + 12));
      }
    }
    _mm_sfence();
  }
#endif
  for (; i < n; ++i) out[i] = table[codes[i]];
}

template <typename T>
void unpack_threads(const uint8_t* codes, long long n, const double* lut65, T* out, int n_threads) {
  T table[256];
  for (int k = 0; k < 256; ++k) table[k] = (T)lut65[k < 64 ? k : 64];
  const long long min_per_thread = 1 << 16;
  long long want = (n + min_per_thread - 1) / min_per_thread;
  if (want < 1) want = 1;
  if (n_threads > want) n_threads = (int)want;
  if (n_threads <= 1) { unpack_slice<T>(codes, n, table, out); return; }
  std::vector<std::thread> th;
  th.reserve(n_threads - 1);
  const long long per = ((n + n_threads - 1) / n_threads + 63) & ~63ll;
  for (int t = 1; t < n_threads; ++t) {
    const long long b = per * t, e = (b + per < n) ? b + per : n;
    if (b >= n) break;
    th.emplace_back([=, &table]() { unpack_slice<T>(codes + b, e - b, table, out + b); });
  }
  unpack_slice<T>(codes, per < n ? per : n, table, out);
  for (auto& t : th) t.join();
}

}  // namespace

cudaError_t launch_bev_pack_u8(const void* bev, int dtype, long long n, const double* lut, unsigned char* codes, int* bad,
                               int sm_count, cudaStream_t st, long long* launches) {
  if (n <= 0) return cudaSuccess;
  long long blocks = ((n >> 2) + 255) / 256;
  const long long cap = (long long)(sm_count > 0 ? sm_count : 148) * 8;   // 8 CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (dtype == 0) k_bev_pack_u8<float><<<(unsigned)blocks, 256, 0, st>>>((const float*)bev, n, lut, codes, bad);
  else k_bev_pack_u8<double><<<(unsigned)blocks, 256, 0, st>>>((const double*)bev, n, lut, codes, bad);
  if (launches) ++*launches;
  return cudaGetLastError();
}

void bev_unpack_host(const unsigned char* codes, long long n, const double* lut65, void* out, int dtype, int n_threads) {
  if (n <= 0) return;
  if (n_threads <= 0) {
    n_threads = (int)std::thread::hardware_concurrency();
    if (n_threads <= 0) n_threads = 1;
  }
  if (dtype == 0) unpack_threads<float>(codes, n, lut65, (float*)out, n_threads);
  else unpack_threads<double>(codes, n, lut65, (double*)out, n_threads);
}

}  // namespace sfpe
