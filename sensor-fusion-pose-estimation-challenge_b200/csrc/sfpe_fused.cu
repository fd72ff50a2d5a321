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
//           (L2 evict_last: it is re-read once per BEV tile)
//   S1c range scan resolved from shared memory, column-major over the image so that the winner
//       gathers of a warp hit neighbouring points; 3 channels written once with 16-byte stores
//   S2  BEV composed in shared memory as saturating u8 counts, one flat tile of the output at a
//       time: filter the entry list, then write the 32-byte sectors that hold a non-empty cell
//   ZF  the BEV image is >95 % zeros and the per-SM store path (~30 B/clk) is the longest single
//       resource of a sweep, so the zero fill of the frame is not a phase: every thread pumps a
//       few 16-byte zero stores per iteration of S1 / S1c / S2-compose, and only catches up (if
//       needed) before the non-empty sectors of a tile are written over the zeros.
// HBM traffic: points read once (+ the winner gather), range written once, BEV written once plus
// the non-empty sectors a second time.
#include "sfpe_internal.h"

namespace sfpe {

constexpr int FUSED_THREADS = 1024;
constexpr int IDX_BITS32 = 18;                       // point index bits of the 32-bit key
constexpr unsigned IDX_MASK32 = (1u << IDX_BITS32) - 1u;
constexpr unsigned KEY32_EMPTY = 0xFFFFFFFFu;
constexpr long long FUSED_MAX_N = (1ll << IDX_BITS32) - 1;  // idx 2^18-1 is reserved for EMPTY
constexpr int SMEM_MISC_WORDS = 128;                 // LUT (65 floats) + counters
constexpr int ENTRY_BATCH = 16;                      // entry loads in flight per thread (S2 compose)
constexpr int SMEM_LIMIT = 232448;                   // 227 KB opt-in maximum per CTA

// ---- L2 residency hints: the entry list is written once and read once per BEV tile
__device__ __forceinline__ unsigned long long l2_policy_evict_last() {
  unsigned long long p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void st_u32_hint(unsigned* ptr, unsigned v, unsigned long long pol) {
  asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(ptr), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ unsigned ld_u32_hint(const unsigned* ptr, unsigned long long pol) {
  unsigned v;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(ptr), "l"(pol));
  return v;
}

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

// Zero-fill pump: up to `cnt` 16-byte zero stores of this thread's share of the frame's BEV
// (word `cur`, then cur + FUSED_THREADS, ...).  Streaming (evict-first) stores.
__device__ __forceinline__ void zf_pump(float4* __restrict__ zbase, int& cur, const int end, const int cnt) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int u = 0; u < cnt; ++u) {
    if (cur < end) __stcs(zbase + cur, z);
    cur += FUSED_THREADS;
  }
}
__device__ __forceinline__ void zf_until(float4* __restrict__ zbase, int& cur, const int end, const int upto) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  const int lim = upto < end ? upto : end;
#pragma unroll 4
  for (; cur < lim; cur += FUSED_THREADS) __stcs(zbase + cur, z);
}

__device__ __forceinline__ float4 lut4(const float* lut, unsigned word) {
  float4 v;
  v.x = lut[min(word & 0xFFu, 64u)]; v.y = lut[min((word >> 8) & 0xFFu, 64u)];
  v.z = lut[min((word >> 16) & 0xFFu, 64u)]; v.w = lut[min(word >> 24, 64u)];
  return v;
}

struct FusedArgs {
  const float4* pts;
  FrameSet fs;
  unsigned* entries;   // [grid][per_cta] flat BEV output cell of every point (0xFFFFFFFF = none)
  unsigned* queue;     // [grid][per_cta] deferred point indices
  long long per_cta;
  const double* lut64;
  float* range; int range_layout; int* winner;
  int range_vec4;      // H*W % 4 == 0 and 16-byte aligned range / winner buffers: vector resolve
  float* bev;
  int* status; int* dropped;
  int tile_cells;      // BEV cells per shared-memory tile (multiple of 16)
  long long* prof;     // optional [grid][8] accumulated SM cycles per phase (sfpe_phase_cycles)
};

// phase stamps: thread 0 adds the cycles since the previous stamp to slot `ph`
#define PROF_STAMP(ph)                                          \
  do {                                                          \
    if (a.prof && tid == 0) {                                   \
      const long long now_ = clock64();                         \
      a.prof[(size_t)blockIdx.x * 8 + (ph)] += now_ - t_prev;   \
      t_prev = now_;                                            \
    }                                                           \
  } while (0)

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
  const unsigned cells_m1 = (unsigned)(cells - 1);
  const unsigned long long keep = l2_policy_evict_last();
  const bool do_bev = g.enable_bev && a.bev;
  if (tid < 65) lut[tid] = (float)a.lut64[tid];
  long long t_prev = clock64();

  for (int f = a.fs.frame0 + blockIdx.x; f < a.fs.frame0 + a.fs.n_frames; f += gridDim.x) {
    long long base, n;
    frame_span(a.fs, f, base, n);
    const float4* __restrict__ fp = a.pts + base;

    // ---- ZF set-up: this frame's BEV is cells [E0, E1) of a.bev; full 16-byte words
    // [W0, W1) are zero-filled by the pump, the (<= 3 + 3) head / tail cells right here.
    const long long E0 = (long long)f * cells, E1 = E0 + cells;
    float4* zbase = nullptr;
    int zcur = tid, zend = 0;
    if (do_bev) {
      const long long W0 = (E0 + 3) >> 2, W1 = E1 >> 2;
      zbase = reinterpret_cast<float4*>(a.bev) + W0;
      zend = W1 > W0 ? (int)(W1 - W0) : 0;
      if (tid < 8) {   // head cells [E0, 4*W0) and tail cells [4*W1, E1), clipped to the frame
        const long long c = tid < 4 ? E0 + tid : 4 * W1 + (tid - 4);
        const bool mine = tid < 4 ? (c < 4 * W0 && c < E1) : (c >= E0 && c < E1);
        if (mine) a.bev[c] = 0.f;
      }
    }
    const int zper = (zend + FUSED_THREADS - 1) / FUSED_THREADS;   // zero stores per thread per frame

    // ---- S0
    if (g.enable_range)
      for (int i = tid; i < hw; i += FUSED_THREADS) keys[i] = KEY32_EMPTY;
    if (tid < 4) misc[tid] = 0u;
    __syncthreads();
    PROF_STAMP(0);

    // ---- S1: one pass over the points (+ ~5/8 of the zero fill)
    {
      const long long iters = (n + 4 * FUSED_THREADS - 1) / (4 * FUSED_THREADS);
      const int k1 = iters ? (int)((((long long)zper * 5 >> 3) + iters - 1) / iters) : 0;
      for (long long i0 = 0; i0 < n; i0 += 4 * FUSED_THREADS) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long i = i0 + u * FUSED_THREADS + tid;
          if (i < n) v[u] = __ldg(fp + i);
        }
        zf_pump(zbase, zcur, zend, k1);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long i = i0 + u * FUSED_THREADS + tid;
          if (i >= n) continue;
          if (g.enable_bev) {
            const int bx = bev_bin_fast((double)v[u].x, g.bev_start, g.bev_delta_x, g.bev_inv_delta_x, g.nex);
            const int by = bev_bin_fast((double)v[u].y, g.bev_start, g.bev_delta_y, g.bev_inv_delta_y, g.ney);
            // both axes reversed == flat index reversed   (birds_eye_view_generator.py:46)
            st_u32_hint(my_entries + i,
                        (bx >= 0 && by >= 0) ? cells_m1 - ((unsigned)bx * (unsigned)g.nby + (unsigned)by) : 0xFFFFFFFFu,
                        keep);
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
    }
    __syncthreads();
    PROF_STAMP(1);

    if (g.enable_range) {
      // ---- S1b: deferred points, full fp64 (guard band, image border, wrap, NaN, errors)
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
      PROF_STAMP(2);
      if (tid == 0) {
        if (a.status) a.status[f] = (int)misc[1];
        if (a.dropped) a.dropped[f] = (int)misc[2];
      }

      // ---- S1c: resolve the range scan; output row r shows pre-flip row H-1-r   (:74)
      if (a.range_vec4) {
        // Task (k, r) = k-th aligned group of 4 flat output cells whose first cell lies in output
        // row r; tasks run k-major / r-minor, so a warp (H = 32: one lane per row) gathers the
        // winners of ~4 neighbouring columns -- neighbouring points of a firing-order sweep.
        const int kmax = ((g.W + 3) >> 2) + 1;
        const int ntask = kmax * g.H;
        const int k2 = (int)(((long long)(zper >> 3) * FUSED_THREADS + ntask - 1) / ntask);
        for (int t = tid; t < ntask; t += FUSED_THREADS) {
          const int k = t / g.H, r = t - k * g.H;
          const int gid = ((r * g.W + 3) >> 2) + k;
          zf_pump(zbase, zcur, zend, k2);
          if (gid >= (((r + 1) * g.W + 3) >> 2)) continue;
          const int c = gid << 2;
          int orow = r, col = c - r * g.W;
          unsigned key[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            key[j] = keys[(g.H - 1 - orow) * g.W + col];
            if (++col == g.W) { col = 0; ++orow; }
          }
          float4 v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (key[j] != KEY32_EMPTY) v[j] = __ldg(fp + (key[j] & IDX_MASK32));
          float d[4], h[4], it[4];
          int win[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            d[j] = 0.f; h[j] = height_bg; it[j] = 0.f; win[j] = -1;
            if (key[j] != KEY32_EMPTY) {
              win[j] = (int)(key[j] & IDX_MASK32);
              const double x = v[j].x, y = v[j].y;
              d[j] = (float)__dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)));
              h[j] = v[j].z; it[j] = v[j].w;
            }
          }
          if (a.range) {
            if (a.range_layout == 0) {
              float* o = a.range + (size_t)f * 3 * hw + c;
              __stcs(reinterpret_cast<float4*>(o), make_float4(d[0], d[1], d[2], d[3]));
              __stcs(reinterpret_cast<float4*>(o + hw), make_float4(h[0], h[1], h[2], h[3]));
              __stcs(reinterpret_cast<float4*>(o + 2 * hw), make_float4(it[0], it[1], it[2], it[3]));
            } else {
              float4* o = reinterpret_cast<float4*>(a.range + ((size_t)f * hw + c) * 3);
              __stcs(o, make_float4(d[0], h[0], it[0], d[1]));
              __stcs(o + 1, make_float4(h[1], it[1], d[2], h[2]));
              __stcs(o + 2, make_float4(it[2], d[3], h[3], it[3]));
            }
          }
          if (a.winner)
            __stcs(reinterpret_cast<int4*>(a.winner + (size_t)f * hw + c), make_int4(win[0], win[1], win[2], win[3]));
        }
      } else {
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
      }
      __syncthreads();
      PROF_STAMP(3);
    } else if (tid == 0) {
      if (a.status) a.status[f] = 0;
      if (a.dropped) a.dropped[f] = 0;
    }

    // ---- S2: BEV, one flat tile of the output at a time.  Tiles start on a 32-byte sector.
    if (do_bev) {
      const int T = a.tile_cells, tw = T >> 2;
      for (int w = tid; w < tw; w += FUSED_THREADS) tile[w] = 0u;
      __syncthreads();
      const long long Bfirst = E0 & ~7ll;
      const int n_tiles = (int)((E1 - Bfirst + T - 1) / T);
      const long long batches = (n + (long long)ENTRY_BATCH * FUSED_THREADS - 1) / ((long long)ENTRY_BATCH * FUSED_THREADS);
      // what is left of the zero fill is spread over all compose batches of the frame
      const long long left = (long long)(zend - zcur + FUSED_THREADS - 1) / FUSED_THREADS;
      const int k3 = (batches && left > 0) ? (int)((left + batches * n_tiles - 1) / (batches * n_tiles)) : 0;
      for (long long B = Bfirst; B < E1; B += T) {
        const long long lo = B > E0 ? B : E0, hi = B + T < E1 ? B + T : E1;
        const unsigned elo = (unsigned)(lo - E0), ehi = (unsigned)(hi - E0);
        const int shift = (int)(E0 - B);
        bool danger = false;
        const unsigned span = ehi - elo;
        // each thread re-reads the entries it wrote in S1 (L2 resident), ENTRY_BATCH loads in flight
        for (long long i0 = 0; i0 < n; i0 += (long long)ENTRY_BATCH * FUSED_THREADS) {
          unsigned e[ENTRY_BATCH];
#pragma unroll
          for (int u = 0; u < ENTRY_BATCH; ++u) {
            const long long i = i0 + u * FUSED_THREADS + tid;
            e[u] = i < n ? ld_u32_hint(my_entries + i, keep) : 0xFFFFFFFFu;
          }
          zf_pump(zbase, zcur, zend, k3);
#pragma unroll
          for (int u = 0; u < ENTRY_BATCH; ++u)
            if (e[u] - elo < span) danger |= tile_count(tile, (unsigned)((int)e[u] + shift));
        }
        if (danger) misc[3] = 1u;
        // the zeros of this tile's cells must be on their way before its non-empty sectors
        zf_until(zbase, zcur, zend, (int)(((hi + 3) >> 2) - ((E0 + 3) >> 2)));
        __syncthreads();
        PROF_STAMP(4);
        if (misc[3]) {  // a byte may have wrapped: redo this tile with the exact (CAS) form
          __syncthreads();
          if (tid == 0) misc[3] = 0u;
          for (int w = tid; w < tw; w += FUSED_THREADS) tile[w] = 0u;
          __syncthreads();
          for (long long i = tid; i < n; i += FUSED_THREADS) {
            const unsigned e = __ldcg(my_entries + i);
            if (e >= elo && e < ehi) tile_count_exact(tile, (unsigned)((int)e + shift));
          }
          __syncthreads();
          PROF_STAMP(5);
        }
        // scan the tile two words (8 cells = one 32-byte sector) at a time; only sectors with a
        // non-empty cell are written (over the zero fill), and their tile words re-zeroed
        {
          const int rel_lo = (int)(lo - B), rel_hi = (int)(hi - B);
          uint2* const tile2 = reinterpret_cast<uint2*>(tile);
          float4* const o4 = reinterpret_cast<float4*>(a.bev + B);
          const int np = (rel_hi + 7) >> 3;
#pragma unroll 4
          for (int p = tid; p < np; p += FUSED_THREADS) {
            const uint2 w = tile2[p];
            if ((w.x | w.y) == 0u) continue;
            tile2[p] = make_uint2(0u, 0u);
            const float4 v0 = lut4(lut, w.x), v1 = lut4(lut, w.y);
            const int c0 = p << 3;
            if (c0 >= rel_lo && c0 + 8 <= rel_hi) {
              __stcs(o4 + 2 * p, v0);
              __stcs(o4 + 2 * p + 1, v1);
            } else {  // frame head / tail: only the cells of THIS frame
              const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
              for (int k = 0; k < 8; ++k)
                if (c0 + k >= rel_lo && c0 + k < rel_hi && vv[k] != 0.f) a.bev[B + c0 + k] = vv[k];
            }
          }
        }
        __syncthreads();
        PROF_STAMP(6);
      }
      zf_until(zbase, zcur, zend, zend);   // n == 0 and rounding leftovers
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
    long long need = ((cells + 7 + 3) / 4 + 3) & ~3ll;  // whole frame incl. an unaligned head (<= 7 cells)
    const int cap = main_words_max() & ~3;
    tile_words = (int)(need < cap ? need : cap);         // tile_cells multiple of 16
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
  a.range_vec4 = ((g.H * g.W) & 3) == 0 && ((uintptr_t)out.range & 15) == 0 && ((uintptr_t)out.range_winner & 15) == 0;
  a.prof = fx.prof;
  const size_t smem = (size_t)(main_words + SMEM_MISC_WORDS) * 4;
  k_fused_sweep<<<fx.grid, FUSED_THREADS, smem, st>>>(g, a, main_words);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
