// Fused "one sweep per SM" path (the hot path of bench.py).
//
// A persistent CTA (1024 threads, ~226 KB dynamic shared memory, one CTA per SM) takes whole
// sweeps.  Per sweep:
//   S0  range-scan winner keys (H*W x u32) are initialised in shared memory
//   S1  one coalesced pass over the sweep's float4 points (16 B/point from HBM):
//         - fp64 range rho (exact numpy order), fp32 cell with guard band -> 32-bit packed key
//           [14-bit quantised rho | 18-bit point index] -> shared-memory atomicMin; equal
//           prefixes are settled exactly on the fp64 rho (CAS loop); guard-band / out-of-image
//           points go to a deferred queue that is re-done in fp64 (S1b)
//         - exact BEV bin (np.arange edges in fp64) -> flat output cell written to an entry list
//   S1c range scan resolved from shared memory: winner point gathered, 3 channels written once
//   S2  BEV composed in shared memory as saturating u8 counts, one flat tile of the output at a
//       time: filter the entry list, then write the tile ONCE with aligned 16-byte streaming
//       stores (count -> 255*min(ln(N+1)/ln64,1) through a LUT), re-zeroing the tile in place
// Every output byte is written exactly once from registers; nothing is read back from HBM.
#include "sfpe_internal.h"

namespace sfpe {

constexpr int FUSED_THREADS = 1024;
constexpr int IDX_BITS32 = 18;                       // point index bits of the 32-bit key
constexpr unsigned IDX_MASK32 = (1u << IDX_BITS32) - 1u;
constexpr unsigned KEY32_EMPTY = 0xFFFFFFFFu;
constexpr long long FUSED_MAX_N = (1ll << IDX_BITS32) - 1;  // idx 2^18-1 is reserved for EMPTY
constexpr int SMEM_MISC_WORDS = 128;                 // LUT (65 floats) + counters
constexpr int SMEM_LIMIT = 232448;                   // 227 KB opt-in maximum per CTA

// monotone non-decreasing map rho (fp64, >= 0 or +inf) -> 14 bits: 1/64 m steps, saturating
__device__ __forceinline__ unsigned quant_rho(double rho) {
  const double t = fmin(rho * 64.0, 16383.0);
  return (unsigned)__double2int_rz(t);
}
__device__ __forceinline__ unsigned pack_key32(double rho, unsigned idx) {
  return (quant_rho(rho) << IDX_BITS32) | idx;
}

__device__ __forceinline__ double rho_of_f4(const float4 v) {
  const double x = v.x, y = v.y, z = v.z;
  return __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)));
}

// Exact settlement of equal-prefix keys in shared memory; same algorithm as
// settle_collision() in sfpe_generic.cu (champion = best point seen under (rho fp64, index)).
__device__ __noinline__ void settle32(const float4* __restrict__ frame_pts, unsigned* addr, unsigned my_key,
                                      double my_rho, unsigned my_idx, unsigned old_key) {
  unsigned champ_key = my_key, champ_idx = my_idx;
  double champ_rho = my_rho;
  {
    const unsigned oi = old_key & IDX_MASK32;
    const double orho = rho_of_f4(__ldg(frame_pts + oi));
    if (better(orho, oi, champ_rho, champ_idx)) { champ_key = old_key; champ_rho = orho; champ_idx = oi; }
  }
  for (;;) {
    const unsigned cur = *reinterpret_cast<volatile unsigned*>(addr);
    if (cur == champ_key) return;
    if ((cur & ~IDX_MASK32) < (champ_key & ~IDX_MASK32)) return;
    const unsigned ci = cur & IDX_MASK32;
    const double crho = rho_of_f4(__ldg(frame_pts + ci));
    if (better(crho, ci, champ_rho, champ_idx)) return;
    if (atomicCAS(addr, cur, champ_key) == cur) return;
  }
}

__device__ __forceinline__ void key_min32(const float4* frame_pts, unsigned* keys, int cell, double rho, unsigned idx) {
  unsigned* addr = keys + cell;
  const unsigned key = pack_key32(rho, idx);
  const unsigned old = atomicMin(addr, key);
  if (old != KEY32_EMPTY && ((old ^ key) >> IDX_BITS32) == 0) settle32(frame_pts, addr, key, rho, idx, old);
}

// Saturating u8 count inside a packed word (any count >= 63 renders as 255, so 128..255 are
// equivalent).  Fast form: add, and take it back if the byte was already >= 128.  A byte can
// only wrap into its neighbour if some add observes it at >= 192 first (>= 64 adds in flight on
// one cell); that observation is returned so the caller can redo the tile with the exact form.
__device__ __forceinline__ bool tile_count(unsigned* tile, unsigned r) {
  const unsigned sh = (r & 3u) * 8u;
  const unsigned ob = (atomicAdd(tile + (r >> 2), 1u << sh) >> sh) & 0xFFu;
  if (ob >= 128u) atomicSub(tile + (r >> 2), 1u << sh);
  return ob >= 192u;
}
// Exact form: CAS loop that never lets a byte pass 128.
__device__ __forceinline__ void tile_count_exact(unsigned* tile, unsigned r) {
  const unsigned sh = (r & 3u) * 8u;
  unsigned* a = tile + (r >> 2);
  unsigned cur = *reinterpret_cast<volatile unsigned*>(a);
  while (((cur >> sh) & 0xFFu) < 128u) {
    const unsigned prev = atomicCAS(a, cur, cur + (1u << sh));
    if (prev == cur) break;
    cur = prev;
  }
}

struct FusedArgs {
  const float4* pts;
  FrameSet fs;
  unsigned* entries;   // [grid][per_cta] flat BEV output cell of every point (0xFFFFFFFF = none)
  unsigned* queue;     // [grid][per_cta] deferred point indices
  long long per_cta;
  const double* lut64;
  float* range; int range_layout; int* winner;
  float* bev;
  int* status; int* dropped;
  int tile_cells;      // BEV cells per shared-memory tile (multiple of 16)
};

__global__ void __launch_bounds__(FUSED_THREADS, 1)
k_fused_sweep(const __grid_constant__ DevGrid g, const __grid_constant__ FusedArgs a, const int main_words) {
  extern __shared__ __align__(16) unsigned smem[];
  unsigned* const keys = smem;   // S0..S1c : H*W winner keys
  unsigned* const tile = smem;   // S2      : packed u8 BEV counts
  float* const lut = reinterpret_cast<float*>(smem + main_words);
  unsigned* const misc = smem + main_words + 72;  // [0] deferred count, [1] status bits, [2] dropped, [3] tile redo flag
  const int tid = threadIdx.x;
  const int hw = g.H * g.W;
  const long long cells = (long long)g.nbx * g.nby;
  unsigned* const my_entries = a.entries + (size_t)blockIdx.x * a.per_cta;
  unsigned* const my_queue = a.queue + (size_t)blockIdx.x * a.per_cta;
  const float height_bg = (float)g.height_bg;
  if (tid < 65) lut[tid] = (float)a.lut64[tid];

  for (int f = a.fs.frame0 + blockIdx.x; f < a.fs.frame0 + a.fs.n_frames; f += gridDim.x) {
    long long base, n;
    frame_span(a.fs, f, base, n);
    const float4* __restrict__ fp = a.pts + base;

    // ---- S0
    if (g.enable_range)
      for (int i = tid; i < hw; i += FUSED_THREADS) keys[i] = KEY32_EMPTY;
    if (tid < 4) misc[tid] = 0u;
    __syncthreads();

    // ---- S1: one pass over the points
    for (long long i0 = 0; i0 < n; i0 += 4 * FUSED_THREADS) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = i0 + u * FUSED_THREADS + tid;
        if (i < n) v[u] = __ldg(fp + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = i0 + u * FUSED_THREADS + tid;
        if (i >= n) continue;
        if (g.enable_bev) {
          const int bx = bev_bin_fast((double)v[u].x, g.bev_start, g.bev_delta_x, g.bev_inv_delta_x, g.nex);
          const int by = bev_bin_fast((double)v[u].y, g.bev_start, g.bev_delta_y, g.bev_inv_delta_y, g.ney);
          // both axes reversed == flat index reversed   (birds_eye_view_generator.py:46)
          my_entries[i] = (bx >= 0 && by >= 0) ? (unsigned)(cells - 1 - ((long long)bx * g.nby + by)) : 0xFFFFFFFFu;
        }
        if (g.enable_range) {
          const double rho = rho_of_f4(v[u]);
          int row, col;
          if (range_cell_fast(g, v[u].x, v[u].y, v[u].z, (float)rho, row, col)) {
            key_min32(fp, keys, row * g.W + col, rho, (unsigned)i);
          } else {
            my_queue[atomicAdd(&misc[0], 1u)] = (unsigned)i;
          }
        }
      }
    }
    __syncthreads();

    // ---- S1b: deferred points, full fp64 (guard band, image border, wrap, NaN, errors)
    if (g.enable_range) {
      const unsigned qn = misc[0];
      for (unsigned j = tid; j < qn; j += FUSED_THREADS) {
        const unsigned i = my_queue[j];
        const float4 v = __ldg(fp + i);
        Pt p; p.x = v.x; p.y = v.y; p.z = v.z; p.it = v.w;
        const double rho = rho_of(p);
        int row, col;
        const int err = range_cell_f64(g, p, rho, row, col);
        if (err) {
          atomicOr(&misc[1], (unsigned)err);
          atomicAdd(&misc[2], 1u);
        } else {
          key_min32(fp, keys, row * g.W + col, rho, i);
        }
      }
      __syncthreads();
      if (tid == 0) {
        if (a.status) a.status[f] = (int)misc[1];
        if (a.dropped) a.dropped[f] = (int)misc[2];
      }

      // ---- S1c: resolve the range scan; output row r shows pre-flip row H-1-r   (:74)
      for (int c = tid; c < hw; c += FUSED_THREADS) {
        const int orow = c / g.W, col = c - orow * g.W;
        const unsigned key = keys[(g.H - 1 - orow) * g.W + col];
        float d = 0.f, h = height_bg, it = 0.f;
        int win = -1;
        if (key != KEY32_EMPTY) {
          win = (int)(key & IDX_MASK32);
          const float4 v = __ldg(fp + win);
          const double x = v.x, y = v.y;
          d = (float)__dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)));
          h = v.z; it = v.w;
        }
        if (a.range) {
          if (a.range_layout == 0) {
            float* o = a.range + (size_t)f * 3 * hw + c;
            __stcs(o, d); __stcs(o + hw, h); __stcs(o + 2 * hw, it);
          } else {
            float* o = a.range + ((size_t)f * hw + c) * 3;
            __stcs(o, d); __stcs(o + 1, h); __stcs(o + 2, it);
          }
        }
        if (a.winner) __stcs(a.winner + (size_t)f * hw + c, win);
      }
      __syncthreads();
    } else if (tid == 0) {
      if (a.status) a.status[f] = 0;
      if (a.dropped) a.dropped[f] = 0;
    }

    // ---- S2: BEV, one flat tile of the output at a time
    if (g.enable_bev && a.bev) {
      const long long E0 = (long long)f * cells, E1 = E0 + cells;
      const int T = a.tile_cells, tw = T >> 2;
      for (int w = tid; w < tw; w += FUSED_THREADS) tile[w] = 0u;
      __syncthreads();
      for (long long B = E0 & ~3ll; B < E1; B += T) {
        const long long lo = B > E0 ? B : E0, hi = B + T < E1 ? B + T : E1;
        const unsigned elo = (unsigned)(lo - E0), ehi = (unsigned)(hi - E0);
        const int shift = (int)(E0 - B);
        bool danger = false;
#pragma unroll 8
        for (long long i = tid; i < n; i += FUSED_THREADS) {
          const unsigned e = my_entries[i];
          if (e >= elo && e < ehi) danger |= tile_count(tile, (unsigned)((int)e + shift));
        }
        if (danger) misc[3] = 1u;
        __syncthreads();
        if (misc[3]) {  // a byte may have wrapped: redo this tile with the exact (CAS) form
          __syncthreads();
          if (tid == 0) misc[3] = 0u;
          for (int w = tid; w < tw; w += FUSED_THREADS) tile[w] = 0u;
          __syncthreads();
          for (long long i = tid; i < n; i += FUSED_THREADS) {
            const unsigned e = my_entries[i];
            if (e >= elo && e < ehi) tile_count_exact(tile, (unsigned)((int)e + shift));
          }
          __syncthreads();
        }
        float4* const o4 = reinterpret_cast<float4*>(a.bev + B);
        for (int w = tid; w < tw; w += FUSED_THREADS) {
          const long long E = B + 4ll * w;
          if (E >= hi) break;
          const unsigned word = tile[w];
          float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
          if (word) {
            tile[w] = 0u;
            const unsigned c0 = word & 0xFFu, c1 = (word >> 8) & 0xFFu, c2 = (word >> 16) & 0xFFu, c3 = word >> 24;
            val.x = lut[c0 < 64u ? c0 : 64u]; val.y = lut[c1 < 64u ? c1 : 64u];
            val.z = lut[c2 < 64u ? c2 : 64u]; val.w = lut[c3 < 64u ? c3 : 64u];
          }
          if (E >= lo && E + 4 <= hi) {
            __stcs(o4 + w, val);
          } else {  // frame head / tail: only the cells of THIS frame
            const float vv[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (E + k >= lo && E + k < hi) a.bev[E + k] = vv[k];
          }
        }
        __syncthreads();
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
static int main_words_max() { return SMEM_LIMIT / 4 - SMEM_MISC_WORDS; }

bool fused_supported(const DevGrid& g, const PointSrc& src, const DevOut& out, long long max_n) {
  const bool packed = src.dtype == 0 && src.stride == 16 && src.ox == 0 && src.oy == 4 && src.oz == 8 &&
                      src.oi == 12 && ((uintptr_t)src.data & 15) == 0;
  if (!packed || g.use_extrinsic || g.n_slices || max_n > FUSED_MAX_N) return false;
  if (g.enable_range && (out.range_dtype != 0 || g.H * g.W > main_words_max())) return false;
  if (g.enable_bev && (out.bev_dtype != 0 || out.bev_counts || ((uintptr_t)out.bev & 15) != 0)) return false;
  if (g.enable_bev && (long long)g.nbx * g.nby > 0x7FFFFFF0ll) return false;
  return g.enable_range || g.enable_bev;
}

int fused_grid(int device, int n_frames) {
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms <= 0) sms = 148;
  return n_frames < sms ? n_frames : sms;
}

cudaError_t launch_fused(const DevGrid& g, const PointSrc& src, const FrameSet& fs, const Scratch& s,
                         const DevOut& out, const FusedScratch& fx, cudaStream_t st, long long* launches) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_fused_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const long long cells = (long long)g.nbx * g.nby;
  int tile_words = 0;
  if (g.enable_bev) {
    long long need = (cells + 3 + 3) / 4 + 1;        // whole frame incl. an unaligned head
    tile_words = (int)(need < main_words_max() ? need : main_words_max());
    tile_words &= ~3;                                 // tile_cells multiple of 16
    if (tile_words < 4) tile_words = 4;
  }
  int main_words = g.enable_range ? g.H * g.W : 0;
  if (tile_words > main_words) main_words = tile_words;
  main_words = (main_words + 3) & ~3;
  FusedArgs a;
  a.pts = reinterpret_cast<const float4*>(src.data);
  a.fs = fs;
  a.entries = fx.entries; a.queue = fx.queue; a.per_cta = fx.per_cta;
  a.lut64 = s.lut;
  a.range = (float*)out.range; a.range_layout = out.range_layout; a.winner = out.range_winner;
  a.bev = (float*)out.bev;
  a.status = out.status; a.dropped = out.dropped;
  a.tile_cells = tile_words * 4;
  const size_t smem = (size_t)(main_words + SMEM_MISC_WORDS) * 4;
  k_fused_sweep<<<fx.grid, FUSED_THREADS, smem, st>>>(g, a, main_words);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
