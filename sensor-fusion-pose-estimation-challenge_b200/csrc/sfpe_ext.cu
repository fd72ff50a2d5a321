// Extension E (MV3D 3.1 -- no reference code) next to the fused kernel:
// per height slice the max z of the slice's points of a BEV cell (empty: the slice's lower edge) and the
// intensity of the highest point of the cell within [edge 0, edge M) (ties on z: lowest index; empty: 0),
// both reversed on both axes like the density.  Same definitions, bit for bit, as k_project /
// k_resolve_bev_ext of the generic path; the range scan and the density of such a batch come from
// k_fused_sweep.
//
// The planes are dense ((M + 1) x nbx x nby floats per sweep: 29 MB at 0.2 m and M = 4) and almost empty
// (a sweep touches ~4 % of the cells), so nothing dense is ever READ:
//   k_ext_fill    writes every plane with its "empty" value (a pure store stream);
//   k_ext_project one pass over the points: max z straight into the output plane with an integer atomic on
//                 the float's bit pattern (max for non-negative values, unsigned min for negative ones: the
//                 order of the ordered-u32 key of the generic path; the empty value, the slice's lower edge,
//                 is below every z of the slice), the (z, lowest index) key of the cell's highest point with
//                 a 64-bit atomicMax into a scratch plane, and the point's cell into a per-point scratch;
//   k_ext_finish  one pass over the points: the point whose key won its cell writes its intensity and puts
//                 the scratch word back to 0 -- the scratch is never swept.
#include "sfpe_internal.h"

namespace sfpe {
namespace {

__device__ __forceinline__ unsigned ord_f32_ext(float v) {   // as ord_f32 of sfpe_generic.cu
  const unsigned b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// blockIdx.y = plane of the chunk: frame lf = y / (M + 1), k = y % (M + 1); k < M: height slice k, else intensity
__global__ void __launch_bounds__(256) k_ext_fill(const __grid_constant__ DevGrid g, float* __restrict__ height,
                                                  float* __restrict__ intensity, int frame0) {
  const int M = g.n_slices;
  const int lf = blockIdx.y / (M + 1), k = blockIdx.y - lf * (M + 1);
  const size_t cells = (size_t)g.nbx * g.nby;
  const size_t f = (size_t)(frame0 + lf);
  float* p;
  float v;
  if (k < M) {
    if (!height) return;
    p = height + (f * M + k) * cells;
    v = (float)g.slice_edges[k];
  } else {
    if (!intensity) return;
    p = intensity + f * cells;
    v = 0.0f;
  }
  // scalar head up to the first 16-byte boundary, float4 body, scalar tail
  const size_t head = min((size_t)((4u - (unsigned)(((uintptr_t)p >> 2) & 3u)) & 3u), cells);
  const size_t n4 = (cells - head) >> 2;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  if (t < head) p[t] = v;
  float4* const p4 = reinterpret_cast<float4*>(p + head);
  const float4 v4 = make_float4(v, v, v, v);
  for (size_t i = t; i < n4; i += stride) p4[i] = v4;
  const size_t done = head + 4 * n4;
  if (t < cells - done) p[done + t] = v;
}

template <int KIND>
__global__ void __launch_bounds__(256) k_ext_project(const __grid_constant__ DevGrid g, const __grid_constant__ PointSrc src,
                                                     const FrameSet fs, float* __restrict__ height,
                                                     unsigned long long* __restrict__ top, int* __restrict__ pcell,
                                                     long long max_n) {
  const int lf = blockIdx.y;
  const int f = fs.frame0 + lf;
  long long base, n;
  frame_span(fs, f, base, n);
  const size_t cells = (size_t)g.nbx * g.nby;
  const int M = g.n_slices;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const Pt p = load_point<KIND>(g, src, base + i);
    int mine = -1;
    const int bx = bev_bin_fast(p.x, g.bev_start, g.bev_delta_x, g.bev_inv_delta_x, g.nex);
    const int by = bev_bin_fast(p.y, g.bev_start, g.bev_delta_y, g.bev_inv_delta_y, g.ney);
    if (bx >= 0 && by >= 0) {
      int sl = -1;
      for (int k = 0; k < M; ++k)
        if (p.z >= g.slice_edges[k] && p.z < g.slice_edges[k + 1]) sl = k;
      if (sl >= 0) {
        const size_t cell = (size_t)bx * g.nby + by;
        const float z = (float)p.z;
        const unsigned zb = __float_as_uint(z);
        if (height) {
          float* const a = height + ((size_t)f * M + sl) * cells + (cells - 1 - cell);
          if (zb & 0x80000000u) atomicMin(reinterpret_cast<unsigned*>(a), zb);
          else atomicMax(reinterpret_cast<int*>(a), (int)zb);
        }
        atomicMax(top + (size_t)lf * cells + cell, ((unsigned long long)ord_f32_ext(z) << 32) | (0xFFFFFFFFu - (unsigned)i));
        mine = (int)cell;
      }
    }
    pcell[(size_t)lf * max_n + i] = mine;
  }
}

template <int KIND>
__global__ void __launch_bounds__(256) k_ext_finish(const __grid_constant__ DevGrid g, const __grid_constant__ PointSrc src,
                                                    const FrameSet fs, float* __restrict__ intensity,
                                                    unsigned long long* __restrict__ top, const int* __restrict__ pcell,
                                                    long long max_n) {
  const int lf = blockIdx.y;
  const int f = fs.frame0 + lf;
  long long base, n;
  frame_span(fs, f, base, n);
  const size_t cells = (size_t)g.nbx * g.nby;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int cell = pcell[(size_t)lf * max_n + i];
    if (cell < 0) continue;
    unsigned long long* const t = top + (size_t)lf * cells + cell;
    const unsigned long long key = *t;
    if ((unsigned)(key & 0xFFFFFFFFull) != 0xFFFFFFFFu - (unsigned)i) continue;   // another point of the cell is its highest
    const Pt p = load_point<KIND>(g, src, base + i);
    if (intensity) intensity[(size_t)f * cells + (cells - 1 - cell)] = (float)p.it;
    *t = 0ull;   // only the winner writes: the scratch is idle again
  }
}

inline int blocks_for(size_t n, int threads, int cap) {
  size_t b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > (size_t)cap) b = cap;
  return (int)b;
}

}  // namespace

cudaError_t launch_ext_chunk(const DevGrid& g, const PointSrc& src, const FrameSet& fs, long long max_n,
                             const ExtScratch& xs, const DevOut& out, cudaStream_t st, long long* launches) {
  if (!g.n_slices || fs.n_frames <= 0) return cudaSuccess;
  const bool packed = src.dtype == 0 && src.stride == 16 && src.ox == 0 && src.oy == 4 && src.oz == 8 &&
                      src.oi == 12 && ((uintptr_t)src.data & 15) == 0;
  const size_t cells = (size_t)g.nbx * g.nby;
  const dim3 grid_f(blocks_for(cells >> 2, 256, 148 * 2), fs.n_frames * (g.n_slices + 1));
  k_ext_fill<<<grid_f, 256, 0, st>>>(g, out.bev_height, out.bev_intensity, fs.frame0);
  ++*launches;
  if (max_n > 0) {
    const dim3 grid_p(blocks_for((size_t)max_n, 256, 65535), fs.n_frames);
    if (packed) {
      k_ext_project<SRC_PACKED_F32><<<grid_p, 256, 0, st>>>(g, src, fs, out.bev_height, xs.top, xs.pcell, xs.max_n);
      k_ext_finish<SRC_PACKED_F32><<<grid_p, 256, 0, st>>>(g, src, fs, out.bev_intensity, xs.top, xs.pcell, xs.max_n);
    } else {
      k_ext_project<SRC_GENERIC><<<grid_p, 256, 0, st>>>(g, src, fs, out.bev_height, xs.top, xs.pcell, xs.max_n);
      k_ext_finish<SRC_GENERIC><<<grid_p, 256, 0, st>>>(g, src, fs, out.bev_intensity, xs.top, xs.pcell, xs.max_n);
    }
    *launches += 2;
  }
  return cudaGetLastError();
}

}  // namespace sfpe
