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

#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
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

// ---- persistent host thread pool (one per process, grows on demand, never shrinks).  Several callers
// may run jobs at the same time (HostBatchPipeline expands two chunks side by side): a job is a set of
// slices pushed to one queue; the caller works on its own slices too and then waits for the rest.
class HostPool {
 public:
  static HostPool& get() { static HostPool* p = new HostPool;  return *p; }   // never destroyed: its threads outlive main()
  // run fn(k) for k in [0, n) on up to `threads` threads (the caller included)
  void run(int n, int threads, const std::function<void(int)>& fn) {
    if (n <= 0) return;
    if (threads > n) threads = n;
    if (threads <= 1) { for (int k = 0; k < n; ++k) fn(k); return; }
    Job job;
    job.fn = &fn; job.next = 0; job.n = n; job.left = n;
    {
      std::lock_guard<std::mutex> g(mu_);
      while ((int)workers_.size() < threads - 1 && workers_.size() < 256) { workers_.emplace_back([this] { loop(); }); workers_.back().detach(); }
      for (int t = 0; t < threads - 1; ++t) queue_.push_back(&job);   // one ticket per helper
    }
    cv_.notify_all();
    work(job);
    std::unique_lock<std::mutex> g(mu_);
    for (auto it = queue_.begin(); it != queue_.end();) it = (*it == &job) ? queue_.erase(it) : it + 1;   // unclaimed tickets
    done_.wait(g, [&] { return job.left == 0 && job.active == 0; });
  }

 private:
  struct Job { const std::function<void(int)>* fn; int next, n, left; int active = 0; };
  void work(Job& job) {
    for (;;) {
      int k;
      { std::lock_guard<std::mutex> g(mu_); if (job.next >= job.n) return; k = job.next++; }
      (*job.fn)(k);
      { std::lock_guard<std::mutex> g(mu_); if (--job.left == 0) done_.notify_all(); }
    }
  }
  void loop() {
    for (;;) {
      Job* job;
      {
        std::unique_lock<std::mutex> g(mu_);
        cv_.wait(g, [&] { return !queue_.empty(); });
        job = queue_.front(); queue_.pop_front();
        ++job->active;
      }
      work(*job);
      { std::lock_guard<std::mutex> g(mu_); if (--job->active == 0) done_.notify_all(); }
    }
  }
  HostPool() {}
  std::mutex mu_;
  std::condition_variable cv_, done_;
  std::deque<Job*> queue_;
  std::vector<std::thread> workers_;
};

static int slices_for(long long n, int n_threads, long long* per) {
  const long long min_per_thread = 1 << 16;
  long long want = (n + min_per_thread - 1) / min_per_thread;
  if (want < 1) want = 1;
  if (n_threads > want) n_threads = (int)want;
  if (n_threads < 1) n_threads = 1;
  *per = ((n + n_threads - 1) / n_threads + 63) & ~63ll;
  return n_threads;
}

template <typename T>
void unpack_threads(const uint8_t* codes, long long n, const double* lut65, T* out, int n_threads) {
  T table[256];
  for (int k = 0; k < 256; ++k) table[k] = (T)lut65[k < 64 ? k : 64];
  long long per;
  n_threads = slices_for(n, n_threads, &per);
  HostPool::get().run(n_threads, n_threads, [&](int t) {
    const long long b = per * t, e = (b + per < n) ? b + per : n;
    if (b < n) unpack_slice<T>(codes + b, e - b, table, out + b);
  });
}

// codes -> RGB bytes through a 64-colour table (3 bytes per cell).  16 cells (48 bytes) per step, built
// in six 64-bit registers: they start as the background colour and the few non-empty cells of a group
// (> 95 % of a BEV image is empty) are patched in with shifts and masks -- no byte stores that a wide
// load would have to wait for -- then go out as three non-temporal 16-byte stores.
static inline void put24(uint64_t (&w)[6], int k, uint64_t rgb) {     // bytes 3k..3k+2 of the 48-byte group
  const int bit = 24 * k, q = bit >> 6, sh = bit & 63;
  w[q] = (w[q] & ~(0xFFFFFFull << sh)) | (rgb << sh);
  if (sh > 40) {                                                      // straddles two words
    const int r = 64 - sh;
    w[q + 1] = (w[q + 1] & ~(0xFFFFFFull >> r)) | (rgb >> r);
  }
}
void unpack_rgb_slice(const uint8_t* codes, long long n, const uint8_t* table /* [256][4] */, uint8_t* out) {
  long long i = 0;
#if defined(__SSE2__)
  while (i < n && (reinterpret_cast<uintptr_t>(out + 3 * i) & 15)) {
    const uint8_t* c = table + 4 * codes[i];
    out[3 * i] = c[0]; out[3 * i + 1] = c[1]; out[3 * i + 2] = c[2];
    ++i;
  }
  uint32_t rgb[256];
  for (int k = 0; k < 256; ++k) rgb[k] = (uint32_t)table[4 * k] | ((uint32_t)table[4 * k + 1] << 8) | ((uint32_t)table[4 * k + 2] << 16);
  uint64_t bg[6] = {0, 0, 0, 0, 0, 0};
  for (int k = 0; k < 16; ++k) put24(bg, k, rgb[0]);
  const __m128i b0 = _mm_set_epi64x((long long)bg[1], (long long)bg[0]), b1 = _mm_set_epi64x((long long)bg[3], (long long)bg[2]),
                b2 = _mm_set_epi64x((long long)bg[5], (long long)bg[4]);
  const __m128i zero = _mm_setzero_si128();
  for (; i + 16 <= n; i += 16) {
    const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(codes + i));
    __m128i* o = reinterpret_cast<__m128i*>(out + 3 * i);
    unsigned nz = 0xFFFFu & ~(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(c, zero));
    if (nz == 0u) {
      _mm_stream_si128(o, b0); _mm_stream_si128(o + 1, b1); _mm_stream_si128(o + 2, b2);
    } else {
      uint64_t w[6] = {bg[0], bg[1], bg[2], bg[3], bg[4], bg[5]};
      while (nz) {
        const int k = __builtin_ctz(nz);
        nz &= nz - 1u;
        put24(w, k, rgb[codes[i + k]]);
      }
      _mm_stream_si128(o, _mm_set_epi64x((long long)w[1], (long long)w[0]));
      _mm_stream_si128(o + 1, _mm_set_epi64x((long long)w[3], (long long)w[2]));
      _mm_stream_si128(o + 2, _mm_set_epi64x((long long)w[5], (long long)w[4]));
    }
  }
  _mm_sfence();
#endif
  for (; i < n; ++i) {
    const uint8_t* c = table + 4 * codes[i];
    out[3 * i] = c[0]; out[3 * i + 1] = c[1]; out[3 * i + 2] = c[2];
  }
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

void bev_unpack_rgb_host(const unsigned char* codes, long long n, const unsigned char* rgb65, unsigned char* out, int n_threads) {
  if (n <= 0) return;
  if (n_threads <= 0) {
    n_threads = (int)std::thread::hardware_concurrency();
    if (n_threads <= 0) n_threads = 1;
  }
  uint8_t table[256 * 4];
  for (int k = 0; k < 256; ++k)
    for (int c = 0; c < 3; ++c) table[4 * k + c] = rgb65[3 * (k < 64 ? k : 64) + c];
  long long per;
  n_threads = slices_for(n, n_threads, &per);
  HostPool::get().run(n_threads, n_threads, [&](int t) {
    const long long b = per * t, e = (b + per < n) ? b + per : n;
    if (b < n) unpack_rgb_slice(codes + b, e - b, table, out + 3 * b);
  });
}

}  // namespace sfpe
