// Generic path: global (L2-resident) scratch, works for every grid size, input layout and
// output option.  Two kernels per chunk of frames:
//   k_project : one pass over the points -> 64-bit atomicMin winner keys (range scan) and
//               32-bit atomicAdd counts (BEV) in scratch
//   k_resolve : one pass over the output cells -> images; scratch is reset in the same pass
// The chunk is sized by the host so that the scratch stays L2 resident.
#include "sfpe_internal.h"

namespace sfpe {

// ---------------------------------------------------------------------------------------
// exact settlement of packed-key prefix collisions (see pack_key): CAS loop that champions
// the best point this thread has seen under the exact order (rho fp64, index).
template <int KIND>
__device__ __noinline__ void settle_collision(const DevGrid& g, const PointSrc& src, long long base,
                                              unsigned long long* addr, unsigned long long my_key, double my_rho,
                                              unsigned my_idx, unsigned long long old_key) {
  const unsigned long long mask = (1ull << g.idx_bits) - 1ull;
  unsigned long long champ_key = my_key;
  double champ_rho = my_rho;
  unsigned champ_idx = my_idx;
  {
    const unsigned oi = (unsigned)(old_key & mask);
    const double orho = rho_of(load_point<KIND>(g, src, base + oi));
    if (better(orho, oi, champ_rho, champ_idx)) {
      champ_key = old_key; champ_rho = orho; champ_idx = oi;
    }
  }
  for (;;) {
    const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(addr);
    if (cur == champ_key) return;
    if ((cur & ~mask) < (champ_key & ~mask)) return;  // strictly nearer prefix
    const unsigned ci = (unsigned)(cur & mask);
    const double crho = rho_of(load_point<KIND>(g, src, base + ci));
    if (better(crho, ci, champ_rho, champ_idx)) return;
    if (atomicCAS(addr, cur, champ_key) == cur) return;
  }
}

// order-preserving u32 encoding of a float (0 is below every encoded value except -NaN)
__device__ __forceinline__ unsigned ord_f32(float v) {
  const unsigned b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float unord_f32(unsigned o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}

template <int KIND>
__global__ void __launch_bounds__(256) k_project(const __grid_constant__ DevGrid g, const __grid_constant__ PointSrc src, const FrameSet fs,
                                                 const Scratch s, int* __restrict__ status,
                                                 int* __restrict__ dropped) {
  const int lf = blockIdx.y;
  const int f = fs.frame0 + lf;
  long long base, n;
  frame_span(fs, f, base, n);
  unsigned long long* keys = s.keys + (size_t)lf * g.H * g.W;
  unsigned* counts = s.counts + (size_t)lf * g.nbx * g.nby;
  int err_acc = 0, ndrop = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const Pt p = load_point<KIND>(g, src, base + i);
    if (g.enable_range) {
      const double rho = rho_of(p);
      int row, col;
      const int err = range_cell_f64(g, p, rho, row, col);
      if (err) {
        err_acc |= err;
        ++ndrop;
      } else {
        unsigned long long* addr = keys + row * g.W + col;
        const unsigned long long key = pack_key(rho, (unsigned)i, g.idx_bits);
        const unsigned long long old = atomicMin(addr, key);
        if (old != KEY_EMPTY && ((old ^ key) >> g.idx_bits) == 0)
          settle_collision<KIND>(g, src, base, addr, key, rho, (unsigned)i, old);
      }
    }
    if (g.enable_bev) {
      const int bx = bev_bin_f64(p.x, g.bev_start, g.bev_delta_x, g.nex);
      const int by = bev_bin_f64(p.y, g.bev_start, g.bev_delta_y, g.ney);
      if (bx >= 0 && by >= 0) {
        const size_t cell = (size_t)bx * g.nby + by;
        atomicAdd(counts + cell, 1u);
        if (g.n_slices) {   // extension: per-slice max z, highest point of the cell
          const size_t cells = (size_t)g.nbx * g.nby;
          int sl = -1;
          for (int k = 0; k < g.n_slices; ++k)
            if (p.z >= g.slice_edges[k] && p.z < g.slice_edges[k + 1]) sl = k;
          if (sl >= 0) {
            const unsigned oz = ord_f32((float)p.z);
            atomicMax(s.slice_max + ((size_t)lf * g.n_slices + sl) * cells + cell, oz);
            atomicMax(s.top + (size_t)lf * cells + cell, ((unsigned long long)oz << 32) | (0xFFFFFFFFu - (unsigned)i));
          }
        }
      }
    }
  }
  if (err_acc) {
    if (status) atomicOr(status + f, err_acc);
    if (dropped) atomicAdd(dropped + f, ndrop);
  }
}

template <typename T>
__device__ __forceinline__ void store_range(T* out, int layout, size_t frame, int H, int W, int orow, int col,
                                            double d, double h, double it) {
  const size_t hw = (size_t)H * W;
  const size_t cell = (size_t)orow * W + col;
  if (layout == 0) {  // CHW
    T* o = out + frame * 3 * hw + cell;
    o[0] = (T)d; o[hw] = (T)h; o[2 * hw] = (T)it;
  } else {            // HWC
    T* o = out + (frame * hw + cell) * 3;
    o[0] = (T)d; o[1] = (T)h; o[2] = (T)it;
  }
}

// One thread per output cell of the range scan.  Output row r shows pre-flip row H-1-r (:74).
template <int KIND>
__global__ void __launch_bounds__(256) k_resolve_range(const __grid_constant__ DevGrid g, const __grid_constant__ PointSrc src, const FrameSet fs,
                                                       const Scratch s, const DevOut out) {
  const int lf = blockIdx.y;
  const int f = fs.frame0 + lf;
  long long base, n;
  frame_span(fs, f, base, n);
  const int hw = g.H * g.W;
  unsigned long long* keys = s.keys + (size_t)lf * hw;
  const unsigned long long mask = (1ull << g.idx_bits) - 1ull;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < hw; c += gridDim.x * blockDim.x) {
    const int orow = c / g.W, col = c - orow * g.W;
    const int src_cell = (g.H - 1 - orow) * g.W + col;
    const unsigned long long key = keys[src_cell];
    keys[src_cell] = KEY_EMPTY;
    double d = 0.0, h = g.height_bg, it = 0.0;
    int win = -1;
    if (key != KEY_EMPTY) {
      win = (int)(key & mask);
      const Pt p = load_point<KIND>(g, src, base + win);
      d = dist_of(p); h = p.z; it = p.it;
    }
    if (out.range) {
      if (out.range_dtype == 0) store_range<float>((float*)out.range, out.range_layout, f, g.H, g.W, orow, col, d, h, it);
      else store_range<double>((double*)out.range, out.range_layout, f, g.H, g.W, orow, col, d, h, it);
    }
    if (out.range_winner) out.range_winner[(size_t)f * hw + c] = win;
  }
}

// One thread per output cell of the BEV: out[ox][oy] = lut(count[nbx-1-ox][nby-1-oy])   (:45-46)
__global__ void __launch_bounds__(256) k_resolve_bev(const DevGrid g, const FrameSet fs, const Scratch s,
                                                     const DevOut out) {
  const int lf = blockIdx.y;
  const int f = fs.frame0 + lf;
  const size_t cells = (size_t)g.nbx * g.nby;
  unsigned* counts = s.counts + (size_t)lf * cells;
  __shared__ double lut[65];
  if (threadIdx.x < 65) lut[threadIdx.x] = s.lut[threadIdx.x];
  __syncthreads();
  for (size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < cells; c += (size_t)gridDim.x * blockDim.x) {
    // reversing both axes of a row-major [nbx][nby] array == reversing the flat index
    const size_t sc = cells - 1 - c;
    const unsigned cnt = counts[sc];
    counts[sc] = 0u;
    const double v = lut[cnt < 64u ? cnt : 64u];
    if (out.bev) {
      if (out.bev_dtype == 0) ((float*)out.bev)[(size_t)f * cells + c] = (float)v;
      else ((double*)out.bev)[(size_t)f * cells + c] = v;
    }
    if (out.bev_counts) out.bev_counts[(size_t)f * cells + sc] = cnt;
  }
}

// Extension outputs (MV3D 3.1; no reference code): per height slice the max z of the
// slice's points of the cell (empty: the slice's lower edge), and the intensity of the highest
// point of the cell within [edge 0, edge M) (ties on z: lowest index; empty: 0).  Both reversed
// on both axes like the density.  Scratch is reset in the same pass.
template <int KIND>
__global__ void __launch_bounds__(256) k_resolve_bev_ext(const __grid_constant__ DevGrid g, const __grid_constant__ PointSrc src,
                                                         const FrameSet fs, const Scratch s, const DevOut out) {
  const int lf = blockIdx.y;
  const int f = fs.frame0 + lf;
  long long base, n;
  frame_span(fs, f, base, n);
  const size_t cells = (size_t)g.nbx * g.nby;
  for (size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < cells; c += (size_t)gridDim.x * blockDim.x) {
    const size_t sc = cells - 1 - c;
    for (int k = 0; k < g.n_slices; ++k) {
      unsigned* a = s.slice_max + ((size_t)lf * g.n_slices + k) * cells + sc;
      const unsigned o = *a;
      *a = 0u;
      if (out.bev_height) out.bev_height[((size_t)f * g.n_slices + k) * cells + c] = o ? unord_f32(o) : (float)g.slice_edges[k];
    }
    unsigned long long* t = s.top + (size_t)lf * cells + sc;
    const unsigned long long key = *t;
    *t = 0ull;
    float it = 0.f;
    if (key) {
      const unsigned idx = 0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull);
      it = (float)load_point<KIND>(g, src, base + idx).it;
    }
    if (out.bev_intensity) out.bev_intensity[(size_t)f * cells + c] = it;
  }
}

__global__ void k_fill_u64(unsigned long long* p, size_t n, unsigned long long v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void k_fill_u32(unsigned* p, size_t n, unsigned v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// normalize(): 255*min(ln(N+1)/ln 64, 1), N>0; 0 stays 0   (birds_eye_view_generator.py:31-37)
__global__ void k_density_lut(double* lut) {
  const int n = threadIdx.x;
  if (n > 64) return;
  double v = 0.0;
  if (n > 0) {
    const double q = __ddiv_rn(log((double)n + 1.0), log(64.0));
    v = __dmul_rn(255.0, fmin(q, 1.0));
  }
  lut[n] = v;
}

// PointCloud2 blob -> packed float4
__global__ void __launch_bounds__(256) k_decode_pc2(const char* __restrict__ blob, long long n, int step, int ox,
                                                    int oy, int oz, int oi, int dtype, float4* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const char* b = blob + i * step;
    float4 v;
    if (dtype == 0) {
      v.x = *reinterpret_cast<const float*>(b + ox);
      v.y = *reinterpret_cast<const float*>(b + oy);
      v.z = *reinterpret_cast<const float*>(b + oz);
      v.w = *reinterpret_cast<const float*>(b + oi);
    } else {
      v.x = (float)*reinterpret_cast<const double*>(b + ox);
      v.y = (float)*reinterpret_cast<const double*>(b + oy);
      v.z = (float)*reinterpret_cast<const double*>(b + oz);
      v.w = (float)*reinterpret_cast<const double*>(b + oi);
    }
    out[i] = v;
  }
}

// ---------------------------------------------------------------------------------------
static inline int blocks_for(size_t n, int threads, int cap) {
  size_t b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > (size_t)cap) b = cap;
  return (int)b;
}

cudaError_t launch_fill_scratch(const DevGrid& g, const Scratch& s, int chunk, cudaStream_t st, long long* launches) {
  const size_t nk = (size_t)chunk * g.H * g.W, nc = (size_t)chunk * g.nbx * g.nby;
  if (s.keys) { k_fill_u64<<<blocks_for(nk, 256, 148 * 8), 256, 0, st>>>(s.keys, nk, KEY_EMPTY); ++*launches; }
  if (s.counts) { k_fill_u32<<<blocks_for(nc, 256, 148 * 8), 256, 0, st>>>(s.counts, nc, 0u); ++*launches; }
  return cudaGetLastError();
}

cudaError_t launch_density_lut(double* d_lut, cudaStream_t st, long long* launches) {
  k_density_lut<<<1, 96, 0, st>>>(d_lut);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_generic_chunk(const DevGrid& g, const PointSrc& src, const FrameSet& fs, long long max_n,
                                 const Scratch& s, const DevOut& out, cudaStream_t st, long long* launches) {
  const bool packed = src.dtype == 0 && src.stride == 16 && src.ox == 0 && src.oy == 4 && src.oz == 8 &&
                      src.oi == 12 && ((uintptr_t)src.data & 15) == 0;
  const dim3 grid_p(blocks_for((size_t)max_n, 256, 65535), fs.n_frames);
  if (max_n > 0) {
    if (packed) k_project<SRC_PACKED_F32><<<grid_p, 256, 0, st>>>(g, src, fs, s, out.status, out.dropped);
    else k_project<SRC_GENERIC><<<grid_p, 256, 0, st>>>(g, src, fs, s, out.status, out.dropped);
    ++*launches;
  }
  if (g.enable_range) {
    const dim3 grid_r(blocks_for((size_t)g.H * g.W, 256, 65535), fs.n_frames);
    if (packed) k_resolve_range<SRC_PACKED_F32><<<grid_r, 256, 0, st>>>(g, src, fs, s, out);
    else k_resolve_range<SRC_GENERIC><<<grid_r, 256, 0, st>>>(g, src, fs, s, out);
    ++*launches;
  }
  if (g.enable_bev) {
    const dim3 grid_b(blocks_for((size_t)g.nbx * g.nby, 256 * 4, 65535), fs.n_frames);
    k_resolve_bev<<<grid_b, 256, 0, st>>>(g, fs, s, out);
    ++*launches;
    if (g.n_slices) {
      if (packed) k_resolve_bev_ext<SRC_PACKED_F32><<<grid_b, 256, 0, st>>>(g, src, fs, s, out);
      else k_resolve_bev_ext<SRC_GENERIC><<<grid_b, 256, 0, st>>>(g, src, fs, s, out);
      ++*launches;
    }
  }
  return cudaGetLastError();
}

cudaError_t launch_decode_pc2(const void* blob, long long n, int step, int ox, int oy, int oz, int oi, int dtype,
                              float* out, cudaStream_t st, long long* launches) {
  if (n <= 0) return cudaSuccess;
  k_decode_pc2<<<blocks_for((size_t)n, 256, 148 * 16), 256, 0, st>>>((const char*)blob, n, step, ox, oy, oz, oi, dtype,
                                                                    (float4*)out);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
