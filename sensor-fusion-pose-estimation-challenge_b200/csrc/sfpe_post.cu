// Consumers of the range scan that stay on the device:
//   k_mask_update   create_lidar_invariant_mask.py:44-62  mask &= (reference == frame), every frame
//   k_roll_width    train/loader.py:31-57                 np.roll(image, rotate_by, axis=1) per batch item
//   k_back_project  train/predict.py:230-293              nearest valid pixel to a cluster centroid, then
//                                                         (distance, height, theta) -> x, y, z
// All three are HBM-streaming or tiny; none touches the projection kernels.
#include "sfpe_internal.h"
#include "../../include/sfpe.h"

namespace sfpe {
namespace {

// ---- invariant mask -----------------------------------------------------------------------------------
// One thread owns V consecutive pixels and walks the frames f = blockIdx.y, blockIdx.y + gridDim.y, ...
// The mask only ever goes 1 -> 0, so the frame groups need no atomics: whoever sees a difference stores 0.
template <typename T, int V>
__global__ void __launch_bounds__(128) k_mask_update(const T* __restrict__ ref, const T* __restrict__ frames,
                                                     long long n_frames, long long frame_stride, long long n_pix,
                                                     int pix_stride, unsigned char* __restrict__ mask) {
  const long long p0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (p0 >= n_pix) return;
  T r[V];
  bool same[V];
  if constexpr (V == 4) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(ref + p0));
    r[0] = q.x; r[1] = q.y; r[2] = q.z; r[3] = q.w;
  } else {
    r[0] = __ldg(ref + p0 * pix_stride);
  }
#pragma unroll
  for (int v = 0; v < V; ++v) same[v] = true;
  long long f = blockIdx.y;
  const long long step = gridDim.y;
  if constexpr (V == 4) {
    for (; f + 3 * step < n_frames; f += 4 * step) {   // four independent 16-byte loads in flight
      float4 q[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) q[u] = __ldcs(reinterpret_cast<const float4*>(frames + (f + u * step) * frame_stride + p0));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        same[0] &= q[u].x == r[0]; same[1] &= q[u].y == r[1]; same[2] &= q[u].z == r[2]; same[3] &= q[u].w == r[3];
      }
    }
    for (; f < n_frames; f += step) {
      const float4 q = __ldcs(reinterpret_cast<const float4*>(frames + f * frame_stride + p0));
      same[0] &= q.x == r[0]; same[1] &= q.y == r[1]; same[2] &= q.z == r[2]; same[3] &= q.w == r[3];
    }
  } else {
    for (; f < n_frames; f += step) same[0] &= __ldcs(frames + f * frame_stride + p0 * pix_stride) == r[0];
  }
#pragma unroll
  for (int v = 0; v < V; ++v)
    if (!same[v]) mask[p0 + v] = 0;
}

// ---- roll along the width -----------------------------------------------------------------------------
// Rows are row_words 32-bit words; out_row[(j + shift*pix_words) mod row_words] = in_row[j].  A CTA moves
// ROLL_WORDS consecutive words of the flattened [frame][row][word] array; thread t owns words t, t+256, ...:
// every warp load and store is one contiguous 128-byte request, eight loads are in flight per thread, and
// the (row, column, frame) of a word is found by one division per thread and carried forward after that.
constexpr int ROLL_THREADS = 256, ROLL_PER_THREAD = 8, ROLL_WORDS = ROLL_THREADS * ROLL_PER_THREAD;

__device__ __forceinline__ unsigned roll_shift_words(const int* __restrict__ shift, long long frame, unsigned width,
                                                     unsigned pix_words) {
  int s = __ldg(shift + frame) % (int)width;   // np.roll takes any integer shift
  if (s < 0) s += (int)width;
  return (unsigned)s * pix_words;
}

__global__ void __launch_bounds__(ROLL_THREADS) k_roll_width(const unsigned* __restrict__ in, unsigned* __restrict__ out,
                                                             long long total_words, unsigned row_words,
                                                             unsigned rows_per_frame, unsigned pix_words, unsigned width,
                                                             const int* __restrict__ shift) {
  long long j = (long long)blockIdx.x * ROLL_WORDS + threadIdx.x;
  if (j >= total_words) return;
  long long row;
  if (total_words <= 0xffffffffll) row = (unsigned)j / row_words; else row = j / row_words;
  unsigned col = (unsigned)(j - row * row_words);
  long long frame = total_words <= 0xffffffffll ? (long long)((unsigned)row / rows_per_frame) : row / rows_per_frame;
  unsigned rif = (unsigned)(row - frame * rows_per_frame);
  long long rowbase = row * row_words;
  unsigned sw = roll_shift_words(shift, frame, width, pix_words);
  unsigned v[ROLL_PER_THREAD];
#pragma unroll
  for (int u = 0; u < ROLL_PER_THREAD; ++u) {
    if (j + (long long)u * ROLL_THREADS < total_words) {
      const unsigned src = col >= sw ? col - sw : col + row_words - sw;
      v[u] = __ldcs(in + rowbase + src);
      col += ROLL_THREADS;
      while (col >= row_words && j + (long long)(u + 1) * ROLL_THREADS < total_words) {
        col -= row_words;
        rowbase += row_words;
        if (++rif == rows_per_frame) {
          rif = 0;
          sw = roll_shift_words(shift, ++frame, width, pix_words);
        }
      }
    }
  }
#pragma unroll
  for (int u = 0; u < ROLL_PER_THREAD; ++u)
    if (j + (long long)u * ROLL_THREADS < total_words) __stcs(out + j + (long long)u * ROLL_THREADS, v[u]);
}

// ---- back-projection of cluster centroids ---------------------------------------------------------------
template <typename T>
__device__ __forceinline__ bool pixel_valid(const T* dist, const T* hgt, long long at, double min_height) {
  return (double)dist[at] > 0.0 && (double)hgt[at] > min_height;
}

// One CTA per centroid.  status: 0 ok, 1 centroid moved to the nearest valid pixel of its box, 2 no valid pixel
// (centroid and box zeroed, xyz = 0), 3 centroid or box outside the image (the reference would index out of range
// or wrap around; nothing is written for that item but the status).
template <typename T>
__global__ void __launch_bounds__(128) k_back_project(double* __restrict__ centroids, double* __restrict__ bboxes,
                                                      const T* __restrict__ distance, const T* __restrict__ height,
                                                      long long frame_stride, int pix_stride, int H, int W,
                                                      double min_height, double x_min, double res_rad,
                                                      double* __restrict__ xyz, int* __restrict__ status) {
  const long long i = blockIdx.x;
  const T* dist = distance + i * frame_stride;
  const T* hgt = height + i * frame_stride;
  __shared__ unsigned long long s_best[4];
  double cx = centroids[2 * i], cy = centroids[2 * i + 1];
  const double b0 = bboxes[4 * i], b1 = bboxes[4 * i + 1], b2 = bboxes[4 * i + 2], b3 = bboxes[4 * i + 3];
  __syncthreads();   // every thread has read the item before thread 0 rewrites it
  int st = 0;
  if (!(cx >= 0.0 && cx < (double)W && cy >= 0.0 && cy < (double)H)) {
    if (threadIdx.x == 0) status[i] = 3;
    return;
  }
  int ix = (int)cx, iy = (int)cy;
  const bool valid0 = pixel_valid(dist, hgt, ((long long)iy * W + ix) * pix_stride, min_height);
  if (!valid0 && b0 != 0.0 && b2 != 0.0) {
    const long long L = (long long)b0, R = (long long)b2 + 1, Tp = (long long)b1, Bt = (long long)b3 + 1;
    if (!(b0 >= 0.0 && b1 >= 0.0 && L < R && Tp < Bt && R <= W && Bt <= H)) {
      if (threadIdx.x == 0) status[i] = 3;
      return;
    }
    const int bw = (int)(R - L), n = bw * (int)(Bt - Tp);
    unsigned long long best = ~0ull;   // [squared pixel distance | row-major index inside the box]: first minimum wins
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      const int r = k / bw, c = k - r * bw;
      const int py = (int)Tp + r, px = (int)L + c;
      if (pixel_valid(dist, hgt, ((long long)py * W + px) * pix_stride, min_height)) {
        const long long dx = px - ix, dy = py - iy;
        const unsigned long long key = ((unsigned long long)(dx * dx + dy * dy) << 32) | (unsigned)k;
        best = key < best ? key : best;
      }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
      best = other < best ? other : best;
    }
    if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
    __syncthreads();
    best = s_best[0];
#pragma unroll
    for (int w = 1; w < 4; ++w) best = s_best[w] < best ? s_best[w] : best;
    if (best == ~0ull) {
      st = 2;
      cx = 0.0; cy = 0.0;
      if (threadIdx.x == 0) {
        centroids[2 * i] = 0.0; centroids[2 * i + 1] = 0.0;
        bboxes[4 * i] = 0.0; bboxes[4 * i + 1] = 0.0; bboxes[4 * i + 2] = 0.0; bboxes[4 * i + 3] = 0.0;
      }
    } else {
      st = 1;
      const int k = (int)(unsigned)best;
      const int r = k / bw;
      cx = (double)((int)L + (k - r * bw));
      cy = (double)((int)Tp + r);
      if (threadIdx.x == 0) { centroids[2 * i] = cx; centroids[2 * i + 1] = cy; }
    }
  }
  if (threadIdx.x != 0) return;
  double x = 0.0, y = 0.0, z = 0.0;
  if (!(cx == 0.0 && cy == 0.0)) {
    ix = (int)cx; iy = (int)cy;
    const long long at = ((long long)iy * W + ix) * pix_stride;
    const double d = (double)dist[at] + 0.75;   // towards the centre of the car, not its surface (predict.py:283)
    const double theta = (cx + x_min) * res_rad;
    x = d * cos(theta);
    y = -d * sin(theta);
    z = (double)hgt[at];
  }
  xyz[4 * i] = x; xyz[4 * i + 1] = y; xyz[4 * i + 2] = z; xyz[4 * i + 3] = 0.0;
  status[i] = st;
}

}  // namespace

cudaError_t launch_mask_update(const void* ref, const void* frames, int dtype, long long n_frames, long long frame_stride,
                               long long n_pix, int pix_stride, unsigned char* mask, int sm_count, cudaStream_t st,
                               long long* launches) {
  if (n_frames <= 0 || n_pix <= 0) return cudaSuccess;
  const bool vec = dtype == SFPE_F32 && pix_stride == 1 && n_pix % 4 == 0 && frame_stride % 4 == 0 &&
                   ((uintptr_t)ref % 16) == 0 && ((uintptr_t)frames % 16) == 0;
  const long long threads = vec ? n_pix / 4 : n_pix;
  const unsigned gx = (unsigned)((threads + 127) / 128);
  // enough frame groups for ~16 CTAs of 128 threads per SM
  long long gy = ((long long)sm_count * 16 + gx - 1) / gx;
  gy = gy < 1 ? 1 : (gy > n_frames ? n_frames : (gy > 65535 ? 65535 : gy));
  const dim3 grid(gx, (unsigned)gy);
  if (vec)
    k_mask_update<float, 4><<<grid, 128, 0, st>>>((const float*)ref, (const float*)frames, n_frames, frame_stride, n_pix, 1, mask);
  else if (dtype == SFPE_F32)
    k_mask_update<float, 1><<<grid, 128, 0, st>>>((const float*)ref, (const float*)frames, n_frames, frame_stride, n_pix, pix_stride, mask);
  else
    k_mask_update<double, 1><<<grid, 128, 0, st>>>((const double*)ref, (const double*)frames, n_frames, frame_stride, n_pix, pix_stride, mask);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_roll_width(const void* in, void* out, long long n_frames, int height, int width, int pix_bytes,
                              const int* shift, cudaStream_t st, long long* launches) {
  const unsigned pix_words = (unsigned)pix_bytes / 4, row_words = pix_words * (unsigned)width;
  const long long total = n_frames * height * (long long)row_words;
  if (total <= 0) return cudaSuccess;
  const long long blocks = (total + ROLL_WORDS - 1) / ROLL_WORDS;
  k_roll_width<<<(unsigned)blocks, ROLL_THREADS, 0, st>>>((const unsigned*)in, (unsigned*)out, total, row_words,
                                                          (unsigned)height, pix_words, (unsigned)width, shift);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_back_project(double* centroids, double* bboxes, long long n, const void* distance, const void* height,
                                int dtype, long long frame_stride, int pix_stride, int H, int W, double min_height,
                                double x_min, double res_rad, double* xyz, int* status, cudaStream_t st,
                                long long* launches) {
  if (n <= 0) return cudaSuccess;
  if (dtype == SFPE_F32)
    k_back_project<float><<<(unsigned)n, 128, 0, st>>>(centroids, bboxes, (const float*)distance, (const float*)height,
                                                       frame_stride, pix_stride, H, W, min_height, x_min, res_rad, xyz, status);
  else
    k_back_project<double><<<(unsigned)n, 128, 0, st>>>(centroids, bboxes, (const double*)distance, (const double*)height,
                                                        frame_stride, pix_stride, H, W, min_height, x_min, res_rad, xyz, status);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
