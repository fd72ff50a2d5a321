// Fused "one sweep per SM" path (the hot path of bench.py).
//
// A persistent CTA (512 threads = 15 compute warps + 1 store warp, 128 registers, ~227 KB
// dynamic shared memory, one CTA per SM) claims whole sweeps from a global counter.  Per sweep
// the compute warps run:
//   S0  range-scan winner keys (H*W x u32) are initialised in shared memory
//   S1  point pass, 4 points per thread and iteration in two register sets (the next iteration's
//       16-byte point loads are in flight while this one is rasterised; L2 prefetch 3 ahead).
//       Branch-free: every decision (range cell, range bucket, BEV bins) is made in fp32 and
//       proven by a guard band; a point with an unproven decision is deferred as a whole.
//         - key [14-bit floor(64 rho) | 18-bit point index] -> shared-memory atomicMin
//         - BEV entry (flat output cell) appended to the bucket of its tile (one shared atomic
//           per (warp, bucket), L2 evict_last: it is read back once in S2)
//   S1b deferred points (~0.4 %) in fp64 exactly as numpy evaluates them
//   S1d equal-bucket keys settled under the exact order (rho fp64, index)
//   S1c range scan resolved from shared memory: a warp takes 8 rows x 16 cells, gathers the
//       winners (prefetched into L2 two tasks ahead, loaded one task ahead) and writes 48
//       contiguous bytes per lane with one 32-byte and one 16-byte store
//   S2  BEV, one flat tile (2^17 cells) of the output at a time: u8 counts composed in shared
//       memory from the tile's bucket; the tile is then scanned and PARKED in L2 in compact
//       form: per 1 KB row of the output (32 sectors of 32 bytes) one bitmap word + the list
//       position of its first non-empty sector, and the u8 counts of the non-empty sectors only
// and the store warp writes the BEV image of the sweep the compute warps have just parked --
// every 32-byte sector exactly once, zeros and values alike, as whole 128-byte-aligned 1 KB rows --
// while the compute warps rasterise the next sweep: fast while they are in S1 / S2 (which hardly
// use the SM's path to L2), slowly while they wait on loads (S1b, S1d) or store themselves
// (S1c), unpaced when they wait for it.  The last sweep of a CTA is written by all 16 warps.
// What bounds it (profiles/README.md): the SM's port to L2 moves one request (<= 128-byte line)
// per ~4 cycles whatever its size, so the phases are tuned for few, full requests, and the
// BEV image's 45 k lines per sweep are placed where the other phases issue the fewest.
// HBM traffic: points read once (+ the winner gather), range written once, BEV written once.
#include <stdio.h>

#include "sfpe_internal.h"

// -DSFPE_CHECKED builds: every index the kernel computes for a write into shared or global scratch is
// checked against its buffer and traps on a violation (compute-sanitizer is not available on the GPU
// pool; tools/checked_run.sh runs the fused-path tests on such a build).  No code in normal builds.
#ifdef SFPE_CHECKED
#define CHK(cond) do { if (!(cond)) { printf("SFPE_CHECKED failed: %s (line %d, block %d thread %d)\n", #cond, __LINE__, (int)blockIdx.x, (int)threadIdx.x); __trap(); } } while (0)
#else
#define CHK(cond) do { } while (0)
#endif


namespace sfpe {

constexpr int FUSED_THREADS = 512;
constexpr int IDX_BITS32 = 18;                       // point index bits of the 32-bit key
constexpr unsigned IDX_MASK32 = (1u << IDX_BITS32) - 1u;
constexpr unsigned KEY32_EMPTY = 0xFFFFFFFFu;
constexpr long long FUSED_MAX_N = (1ll << IDX_BITS32) - 1;  // idx 2^18-1 is reserved for EMPTY
constexpr int SMEM_MISC_WORDS = 256;                 // LUT (132 floats) | 16 counters | 16 profile words | bucket counts | rings
constexpr int ENTRY_BATCH = 3;                       // 16-byte entry loads (4 entries each) in flight per thread (S2 compose)
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
__device__ __forceinline__ void st_u32x4_hint(uint4* ptr, const uint4 v, unsigned long long pol) {
  asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
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

// Winner keys whose quantised range is equal are settled exactly, but not where they are found:
// a point that meets an equal-bucket key at its atomicMin records itself in a per-CTA collision
// list -- and the previous holder too if it has just displaced it (lower index), because a
// displaced point is in nobody's list yet.  So every point of a bucket that holds more than one
// point of a cell is either the holder of the key or recorded.  After the point pass every
// recorded candidate tries to install itself over the current holder under the exact order
// (rho fp64, index) with a CAS loop: the best candidate of the winning bucket ends up in the key
// whatever the schedule -- it replaces any worse holder and no worse candidate can replace it.
__device__ __forceinline__ void note_collision(uint2* __restrict__ cq, unsigned* misc, int cell, unsigned key, unsigned old) {
  if (old != KEY32_EMPTY && ((old ^ key) >> IDX_BITS32) == 0) {
    const bool displaced = key < old;
    const unsigned pos = atomicAdd(&misc[7], displaced ? 2u : 1u);
    cq[pos] = make_uint2((unsigned)cell, key & IDX_MASK32);
    if (displaced) cq[pos + 1] = make_uint2((unsigned)cell, old & IDX_MASK32);
  }
}

// the same for the four points of a point-pass step: one reservation per thread.  The caller passes key 0 for a
// point without a cell (old = EMPTY then): EMPTY ^ key has bucket bits set for key 0 and for every real key (its
// bucket is <= 16381 < 0x3FFF), so "equal bucket" is one XOR and one compare per point, and everything else
// happens only for the few points that did meet an equal-bucket key.
__device__ __forceinline__ void note_collisions4(uint2* __restrict__ cq, unsigned* misc, const int (&cell)[4],
                                                 const unsigned (&key)[4], const unsigned (&old)[4], unsigned cap) {
  const bool c0 = (old[0] ^ key[0]) < (1u << IDX_BITS32), c1 = (old[1] ^ key[1]) < (1u << IDX_BITS32);
  const bool c2 = (old[2] ^ key[2]) < (1u << IDX_BITS32), c3 = (old[3] ^ key[3]) < (1u << IDX_BITS32);
  if (c0 | c1 | c2 | c3) {
    const bool c[4] = {c0, c1, c2, c3};
    unsigned nrec = 0u;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (c[u]) nrec += key[u] < old[u] ? 2u : 1u;
    unsigned pos = atomicAdd(&misc[7], nrec);
    CHK(pos + nrec <= cap);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (c[u]) {
        cq[pos++] = make_uint2((unsigned)cell[u], key[u] & IDX_MASK32);
        if (key[u] < old[u]) cq[pos++] = make_uint2((unsigned)cell[u], old[u] & IDX_MASK32);
      }
  }
}

// Rigid extrinsic (extension, default off; XF instantiation of the kernel only): exactly apply_extrinsic of
// sfpe_common.cuh -- the affine map in fp64, left to right, not fused, rounded to float32 -- so that the
// transformed cloud is again a float32 cloud and every fp32 decision and guard band applies to it unchanged.
template <bool XF>
__device__ __forceinline__ float4 xf_point(const DevGrid& g, float4 v) {
  if (XF) {
    Pt p; p.x = v.x; p.y = v.y; p.z = v.z; p.it = v.w;
    apply_extrinsic(g, p);
    v.x = (float)p.x; v.y = (float)p.y; v.z = (float)p.z;
  }
  return v;
}

// candidate (point idx, coordinates v) against the holder of *addr: `cur` is the key read
// before and hv the coordinates of ITS point (both loaded by the caller, so that the loads of
// several candidates are in flight together); reloaded only when the holder changes under us.
template <bool XF>
__device__ __forceinline__ void settle_candidate(const DevGrid& g, const float4* __restrict__ frame_pts, unsigned* addr,
                                                 const double rho, unsigned idx, unsigned cur, double hrho) {
  // rho: the candidate's range, hrho: the range of the holder `cur` (both computed by the caller, all candidates of
  // a round together, so that the fp64 square roots overlap); recomputed only when the holder changes under us
  const unsigned my_key = (quant_rho(rho) << IDX_BITS32) | idx;
  for (;;) {
    if (cur == my_key) return;
    if ((cur >> IDX_BITS32) < (my_key >> IDX_BITS32)) return;      // a nearer bucket holds the cell
    if ((cur >> IDX_BITS32) == (my_key >> IDX_BITS32)) {
      const unsigned ci = cur & IDX_MASK32;
      if (better(hrho, ci, rho, idx)) return;
    }
    const unsigned seen = atomicCAS(addr, cur, my_key);
    if (seen == cur) return;
    cur = seen;
    hrho = rho_of_f4(xf_point<XF>(g, __ldg(frame_pts + (cur & IDX_MASK32))));
  }
}

// Saturating u8 count inside a packed word (any count >= 63 renders as 255, so 128..255 are
// equivalent).  Fast form: add, and take it back if the byte was already >= 128.  A byte can
// only wrap into its neighbour if some add observes it at >= 192 first (>= 64 adds in flight on
// one cell); that observation is returned so the caller can redo the tile with the exact form.
// The same in two steps, so that many counts can be in flight per thread.  issue: r = position
// in the tile (>= T: not in this tile); returns (r | TC_VALID | previous byte << 24) or 0.
// finish: an add on a byte that was already >= 128 is undone, and the >= 192 observation reported.
constexpr unsigned TC_VALID = 1u << 23;
// Every counted cell also marks its 128-byte output line (32 cells) in the tile's line flags: one
// byte per line, plain stores (everybody writes the same 1).
// An entry is a run: tile coordinate of the cell in the low 30 bits, points - 1 (0..3) on top.
constexpr unsigned ENTRY_CELL = 0x3FFFFFFFu;
__device__ __forceinline__ unsigned tile_count_issue(unsigned* tile, unsigned char* lf, unsigned e, unsigned tlo, unsigned T) {
  const unsigned r = (e & ENTRY_CELL) - tlo;
  if (r >= T) return 0u;
  CHK(T <= (1u << 17));
  lf[r >> 5] = 1;
  const unsigned sh = (r & 3u) * 8u;
  const unsigned ob = (atomicAdd(tile + (r >> 2), ((e >> 30) + 1u) << sh) >> sh) & 0xFFu;
  return r | ((e >> 30) << 21) | TC_VALID | (ob << 24);   // r < 2^17 (T <= 2^17)
}
__device__ __forceinline__ bool tile_count_finish(unsigned* tile, unsigned packed) {
  if (!(packed & TC_VALID)) return false;
  const unsigned ob = packed >> 24, r = packed & 0x1FFFFFu;
  if (ob < 128u) return false;
  atomicSub(tile + (r >> 2), (((packed >> 21) & 3u) + 1u) << ((r & 3u) * 8u));
  return ob >= 192u;
}
// Exact form: CAS loop that never adds to a byte that has reached 128.  Returns true for the first arrival.
__device__ __forceinline__ bool tile_count_exact(unsigned* tile, unsigned r, unsigned c) {
  const unsigned sh = (r & 3u) * 8u;
  unsigned* a = tile + (r >> 2);
  unsigned cur = *reinterpret_cast<volatile unsigned*>(a);
  while (((cur >> sh) & 0xFFu) < 128u) {
    const unsigned prev = atomicCAS(a, cur, cur + (c << sh));
    if (prev == cur) return ((cur >> sh) & 0xFFu) == 0u;
    cur = prev;
  }
  return false;
}

// ---- warp roles.  The last warp of the CTA does nothing but write BEV images (the per-SM store
// path, 32 B/clk, is the longest single resource of a sweep and a warp that waits on it cannot
// compute: tools/ubench_store.cu, ubench_zerofill.cu); the other warps ("compute threads", CT of
// them) rasterise and meet on named barrier 1.
constexpr int STORE_WARPS = 1;
static_assert(STORE_WARPS == 1, "one store warp: its progress word misc[4] counts whole frames");
constexpr int CT = FUSED_THREADS - 32 * STORE_WARPS;   // 480 compute threads
constexpr int CW = CT / 32;                            // 15 compute warps
constexpr int MAX_BUCKETS = 64;                        // BEV entry buckets (consecutive tiles share one beyond 64 tiles)
constexpr int NO_BUCKET = 255;
constexpr long long FILL_IN_IV = 128, FILL_OUT_IV = 256;   // store-warp pacing, cycles per 8 KB group
constexpr int ROW_CELLS = 256;                         // one output row of the store warp: 32 lanes x 32 bytes = 1 KB
constexpr int GROUP_ROWS = 8;                          // rows per claimed group (8 KB)

// ablation bits (timing experiments, tools/ablate.sh) exist only in -DSFPE_ABLATE builds
#ifdef SFPE_ABLATE
#define DBG(bit) ((a.debug & (bit)) != 0)
#else
#define DBG(bit) false
#endif

__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CT) : "memory"); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ uint4 ld_u32x4_hint(const uint4* ptr, unsigned long long pol) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(ptr), "l"(pol));
  return v;
}

// Flags between the compute warps and the store warp live in shared memory: written with release,
// read with acquire semantics at CTA scope (the parked bitmaps they guard are global memory, read
// back from L2 with ld.global.cg).  No fence: a fence would also wait for every store the store
// warp has in flight.
__device__ __forceinline__ unsigned ld_acquire_smem(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_smem(unsigned* p, unsigned v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}

// 32-byte streaming store (STG.256, sm_100+): one L2 request per sector instead of two
__device__ __forceinline__ void st_cs_f32x8(float* p, const float4 a, const float4 b) {
  asm volatile("st.global.cs.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w),
               "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
}

// four u8 counts -> density values; the table has an entry for every value a settled count byte
// can hold (0..131: the saturating add of a run of <= 4 points never leaves a byte above 127 + 4)
__device__ __forceinline__ float4 lut4(const float* lut, unsigned word) {
  float4 v;
  v.x = lut[word & 0xFFu]; v.y = lut[(word >> 8) & 0xFFu];
  v.z = lut[(word >> 16) & 0xFFu]; v.w = lut[word >> 24];
  return v;
}

struct FusedArgs {
  const float4* pts;
  FrameSet fs;
  unsigned* entries;   // [grid][n_buckets][per_cta] BEV entries (output cell + head) by tile bucket
  unsigned* queue;     // [grid][per_cta] deferred point indices
  uint2* cqueue;       // [grid][6*per_cta] collision records (range-scan cell, point index)
  long long per_cta;
  const double* lut64;
  float* range; int range_layout; int* winner;
  int range_vec4;      // H*W % 4 == 0 and 16-byte aligned range / winner buffers: vector resolve
  float* bev;
  int* status; int* dropped;
  int spw_log2;        // log2(beams per firing / 32): the point pass walks the sweep in blocks of four firings
  int tile_log2;       // BEV cells per shared-memory tile = 1 << tile_log2 (>= 10)
  int tiles_per_bucket, n_buckets;
  long long* prof;     // optional [grid][SFPE_N_PHASES] accumulated SM cycles per phase (sfpe_phase_cycles)
  unsigned* counter;   // frames claimed so far (dynamic assignment), zeroed before the launch
  unsigned* rowbits;   // [grid][2][rows_cap / 4] parked: one byte per 1 KB output row, bit l = its 128-byte line l is non-empty
  long long rows_cap;
  // sfpe_params.reserved[0], -DSFPE_ABLATE builds only: ablation bits for timing experiments
  // (tools/ablate.sh); the output is INVALID when any of 1..64, 256, 512 is set.  1 no BEV write,
  // 2 no range stores, 4 no winner gather, 8 no resolve prefetch, 16 no BEV entry append, 32 no
  // range atomics, 64 no point prefetch, 128 report the collision record count in `dropped`,
  // 256 settle: loads only, 512 settle: no point loads
  int debug;
};

// ---- BEV image write-out.  Tile coordinate 0 of a frame is the 1 KB-aligned address at or below
// its first cell, so a "row" (256 cells) is one warp-wide 32-byte store instruction covering eight
// whole 128-byte lines.  Every line of the image is written exactly once: the lines that hold a
// non-empty cell by the compute warps when they scan a composed tile (S2), all the others -- zeros --
// by the store warp (and by every warp of the CTA for its last frame), which learns from the parked
// row bitmaps which lines are whose.
constexpr int BLOCK_ROWS = 32;                         // rows per claimed block (32 KB) = 4 groups
constexpr int BLOCK_GROUPS = BLOCK_ROWS / GROUP_ROWS;
struct BevJob {
  float* bp;               // address of tile coordinate 0 (1 KB aligned; may lie before the frame)
  const unsigned* lines;   // [n_rows / 4] one byte per row: bit l = line l of the row holds a non-empty cell (the compute warps write it)
  int n_rows, n_blocks;
  unsigned tbase;          // tiles this CTA had parked before the frame (misc[5] counts on from there)
  int rpt_log2;            // rows per tile = 1 << rpt_log2
};
__device__ __forceinline__ BevJob make_bev_job(const FusedArgs& a, long long cells, int f, unsigned k, unsigned tbase) {
  BevJob j;
  float* const first = a.bev + (long long)f * cells;
  const unsigned head = (unsigned)(((uintptr_t)first & 1023u) >> 2);
  j.bp = first - head;
  j.n_rows = (int)(((unsigned)cells + head + (ROW_CELLS - 1)) / ROW_CELLS);
  j.n_blocks = (j.n_rows + BLOCK_ROWS - 1) / BLOCK_ROWS;
  j.lines = a.rowbits + ((size_t)blockIdx.x * 2 + (k & 1u)) * (a.rows_cap >> 2);
  j.tbase = tbase;
  j.rpt_log2 = a.tile_log2 - 8;
  return j;
}
// u8 counts of 8 cells -> density values
__device__ __forceinline__ void lut8(const float* lut, const uint2 e, float4& v0, float4& v1) {
  v0 = lut4(lut, e.x);
  v1 = lut4(lut, e.y);
}
// A writer claims blocks of 32 rows from `cursor` (one claimer in the steady state: the store warp;
// all warps for a CTA's last frames, and the compute warps when they would otherwise wait for the
// store warp at the start of a BEV phase); the line bitmaps of the block after next are in flight while
// the current block is written, 8 rows (8 KB) at a time.  A block is claimed once its tiles are
// parked (misc[5]).
// paced: at most one group per FILL_IN_IV cycles while misc[11] says that the compute warps are in
// a phase that hardly uses the SM's path to L2 (point pass, BEV compose: registers and shared
// memory, no waiting on loads), per FILL_OUT_IV while they move data themselves or wait on loads,
// unpaced when they wait for the store warp.  (The optimum is flat: 128..192 / 256..384 within
// +-0.5 %, and one warp does not move more than ~8-14 B/clk unpaced either; profiles/README.md.)
template <bool XF>   // one copy per kernel instantiation (a copy shared by two kernels was assembled wrongly: see __graft_entry__.check_sass)
__device__ __noinline__ void bev_write_frame(const BevJob job, unsigned* cursor, unsigned* misc, const int lane,
                                             const bool paced) {
  const BevJob j = job;   // (by value: in registers, no stack frame)
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  const int myline = lane >> 2;
  // claim the next block (-1: none left) and load its line bitmaps (lane i < 8: rows 4i..4i+3 of the
  // block, one byte each); waits until its tiles are parked
  auto claim = [&](int& blk, unsigned& bm) {
    int b = 0;
    if (lane == 0) b = (int)atomicAdd(cursor, 1u);
    b = __shfl_sync(0xFFFFFFFFu, b, 0);
    bm = ~0u;   // rows past the end of the frame: nothing to write
    if (b >= j.n_blocks) { blk = -1; return; }
    blk = b;
    const int r_last = min((b + 1) * BLOCK_ROWS, j.n_rows) - 1;
    const unsigned need = j.tbase + (unsigned)(r_last >> j.rpt_log2) + 1u;
    while ((int)(ld_acquire_smem(&misc[5]) - need) < 0) __nanosleep(400);
    const int r4 = b * BLOCK_ROWS + 4 * lane;          // first of my four rows
    if (lane < 8 && r4 <= r_last) {
      bm = __ldcg(j.lines + (r4 >> 2));
      const int valid = r_last + 1 - r4;               // rows of the word that exist (1..4, or more)
      if (valid < 4) bm |= ~0u << (8 * valid);
    }
  };
  int blk, blk1, blk2;
  unsigned bm, bm1, bm2;
  claim(blk, bm);
  if (blk < 0) return;
  claim(blk1, bm1);
  long long t_next = clock64();
  while (blk >= 0) {
    claim(blk2, bm2);
#pragma unroll
    for (int s = 0; s < BLOCK_GROUPS; ++s) {
      float* const p = j.bp + ((size_t)blk * BLOCK_ROWS + s * GROUP_ROWS) * ROW_CELLS + lane * 8;
#pragma unroll
      for (int u = 0; u < GROUP_ROWS; ++u) {
        const int row = s * GROUP_ROWS + u;            // row of the block
        const unsigned w = __shfl_sync(0xFFFFFFFFu, bm, row >> 2);
        if (((w >> ((row & 3) * 8 + myline)) & 1u) == 0u) st_cs_f32x8(p + u * ROW_CELLS, z, z);
      }
      if (paced) {
        for (;;) {
          const unsigned wdw = *reinterpret_cast<volatile unsigned*>(&misc[11]);
          const long long rem = t_next + (wdw == 2u ? 0ll : wdw == 1u ? FILL_IN_IV : FILL_OUT_IV) - clock64();
          if (rem <= 0) break;
          __nanosleep((unsigned)min(rem >> 1, 2000ll) + 20u);
        }
        t_next = clock64();
      }
    }
    blk = blk1; bm = bm1;
    blk1 = blk2; bm1 = bm2;
  }
}

// phase stamps: thread 0 adds the cycles since the previous stamp to slot `ph` (accumulated in
// shared memory, written out once at the end of the kernel)
#define PROF_STAMP(ph)                                          \
  do {                                                          \
    if (a.prof && tid == 0) {                                   \
      const long long now_ = clock64();                         \
      prof_acc[(ph)] += now_ - t_prev;                          \
      t_prev = now_;                                            \
    }                                                           \
  } while (0)

template <bool XF>   // XF: apply the extrinsic to every point that is loaded
__global__ void __launch_bounds__(FUSED_THREADS, 1)
k_fused_sweep(const __grid_constant__ DevGrid g, const __grid_constant__ FusedArgs a, const int main_words) {
  extern __shared__ __align__(16) unsigned smem[];
  unsigned* const keys = smem;   // S0..S1c : H*W winner keys
  unsigned* const tile = smem;   // S2      : packed u8 BEV counts
  float* const lut = reinterpret_cast<float*>(smem + main_words);
  // misc: [0] deferred count, [1] status bits, [2] dropped, [3] tile redo flag,
  //       [4] frames whose BEV image the store warp has written, [5] tiles parked so far (all frames),
  //       [6] frames claimed so far, [7] collision records, [9] claim broadcast, [10] parked sectors of the
  //       frame in S2, [11] store-warp pacing window, [12..15] claim ring: frame + 1 (~0: no more frames),
  //       [16..39] phase cycles, [40..103] entries per bucket, [104..107] ring: misc[5] when the frame was
  //       claimed, [108..109] block cursor of the frame being written (by frame parity), [110] frames whose
  //       cursor has been reset (S2 begun)
  unsigned* const misc = smem + main_words + 136;
  const int tid = threadIdx.x;
  const int hw = g.H * g.W;
  const long long cells = (long long)g.nbx * g.nby;
  const bool do_bev = g.enable_bev && a.bev;
  const int f_end = a.fs.frame0 + a.fs.n_frames;
  const bool store_active = do_bev && !DBG(1);   // is there a BEV image to write?
  if (tid < 132) lut[tid] = (float)a.lut64[min(tid, 64)];
  if (tid < 16) misc[tid] = 0u;
  if (tid >= 104 && tid < 112) misc[tid] = 0u;
  long long* const prof_acc = reinterpret_cast<long long*>(misc + 16);   // [SFPE_N_PHASES], thread 0 only
  if (tid == 0)
    for (int k = 0; k < SFPE_N_PHASES; ++k) prof_acc[k] = 0;
  __syncthreads();   // the only CTA-wide barrier: the roles part here

  // ======================================================================== store warp
  if (tid >= CT) {
    if (!store_active) return;
    const int sl = tid - CT;
    for (unsigned k = 0;; ++k) {
      // frame k of this CTA: claimed by the compute warps (ring), written as its tiles get parked
      while ((int)(ld_acquire_smem(&misc[6]) - (k + 1u)) < 0) __nanosleep(1000);
      const unsigned c = *reinterpret_cast<volatile unsigned*>(&misc[12 + (k & 3u)]);
      if (c == ~0u) return;
      // the frame's block cursor still holds the count of frame k - 2 until the compute warps reset it at the
      // start of the frame's S2 (misc[110]); a claim before that would read "no blocks left" and skip the frame
      while ((int)(ld_acquire_smem(&misc[110]) - (k + 1u)) < 0) __nanosleep(1000);
      const BevJob j = make_bev_job(a, cells, (int)c - 1, k, *reinterpret_cast<volatile unsigned*>(&misc[104 + (k & 3u)]));
      bev_write_frame<XF>(j, &misc[108 + (k & 1u)], misc, sl, true);
      __syncwarp();
      if (sl == 0) st_release_smem(&misc[4], k + 1u);   // the parked bitmaps of frame k may be overwritten
    }
  }

  // ======================================================================== compute warps
  unsigned* const my_buckets = a.entries + (size_t)blockIdx.x * a.n_buckets * a.per_cta;   // [bucket][per_cta]
  unsigned* const bcnt = misc + 40;   // entries per bucket [MAX_BUCKETS]
  const int lane = tid & 31;
  const unsigned per_cta32 = (unsigned)a.per_cta;
  unsigned* const my_queue = a.queue + (size_t)blockIdx.x * a.per_cta;
  uint2* const my_cq = a.cqueue + (size_t)blockIdx.x * a.per_cta * 6;   // <= 2n first-level records + <= 2 per live record (S1d)
  const float height_bg = (float)g.height_bg;
  const unsigned cells_m1 = (unsigned)(cells - 1);
  const unsigned long long keep = l2_policy_evict_last();
  const int bucket_shift = a.tile_log2 + 31 - __clz(a.tiles_per_bucket);   // tiles_per_bucket is a power of two
  long long t_prev = clock64();
  unsigned frames_done = 0;

  for (;; ++frames_done) {
    if (tid == 0) {
      // frames are claimed dynamically (CTAs start staggered and drift); the store warp learns the
      // claim through the ring misc[12 + (k & 3)] (it is never more than two frames behind)
      const int fc = a.fs.frame0 + (int)atomicAdd(a.counter, 1u);
      const unsigned c = fc < f_end ? (unsigned)fc + 1u : ~0u;
      misc[9] = c;
      if (store_active) {
        *reinterpret_cast<volatile unsigned*>(&misc[12 + (frames_done & 3u)]) = c;
        *reinterpret_cast<volatile unsigned*>(&misc[104 + (frames_done & 3u)]) = misc[5];
        st_release_smem(&misc[6], frames_done + 1u);
      }
    }
    compute_sync();
    const unsigned claimed = misc[9];
    compute_sync();
    if (claimed == ~0u) break;
    const int f = (int)claimed - 1;
    long long base, n;
    frame_span(a.fs, f, base, n);
    const float4* __restrict__ fp = a.pts + base;
    // tile coordinate 0 of the frame's BEV = the 1 KB-aligned address at or below its first cell
    const unsigned head = do_bev ? (unsigned)(((uintptr_t)(a.bev + (long long)f * cells) & 1023u) >> 2) : 0u;

    // ---- S0
    if (g.enable_range) {   // 16-byte stores (smem is 16-byte aligned, the tail word by word)
      const uint4 e4 = make_uint4(KEY32_EMPTY, KEY32_EMPTY, KEY32_EMPTY, KEY32_EMPTY);
      for (int i = tid; i < (hw >> 2); i += CT) reinterpret_cast<uint4*>(keys)[i] = e4;
      if (tid < (hw & 3)) keys[(hw & ~3) + tid] = KEY32_EMPTY;
    }
    if (tid < 4) misc[tid] = 0u;
    if (tid >= 40 && tid < 40 + MAX_BUCKETS) misc[tid] = 0u;
    if (tid == 0) {
      misc[7] = 0u; misc[8] = 0u; misc[9] = 0u;
      misc[11] = 1u;               // the point pass hardly uses the path to L2: store-warp window open
    }
    // ---- S1: one pass over the points
    // (32-bit indices: n < 2^18 on this path.)  A warp step takes four consecutive FIRINGS of 32
    // beams: lane = beam, and a thread's four points are the same beam at four consecutive azimuths
    // (a.spw_log2: beams per firing = 32 << spw_log2, a hint -- any value rasterises any cloud
    // correctly).  Such neighbours mostly fall into the same BEV cell and almost always into the same
    // BEV tile, so a thread merges them into runs (cell, count); every run is appended to its tile's
    // bucket with one shared-memory atomic.  The points of the next step are loaded while the
    // current ones are rasterised, into the other of two register sets (no copies).
    const int n32 = (int)n;
    const int spw_log2 = a.spw_log2, fstride = 32 << spw_log2;          // points per firing
    const int n_steps = ((n32 + 4 * fstride - 1) / (4 * fstride)) << spw_log2;   // warp steps of this frame
    // first point of warp step q: block of 4 firings (q >> spw_log2), 32-beam part (q & (spw - 1))
    auto step_base = [&](int q) { return (q >> spw_log2) * 4 * fstride + ((q & ((1 << spw_log2) - 1)) << 5); };
    auto prefetch_step = [&](int q) {   // the 16 lines (4 firings x 512 bytes) of step q
      const int i = step_base(q) + (lane >> 2) * fstride + (lane & 3) * 8;
      if (lane < 16 && q < n_steps && i < n32) prefetch_l2(fp + i);
    };
    // the first three steps of every warp are pulled into L2 while the keys are initialised
#pragma unroll
    for (int k = 0; k < 3; ++k) prefetch_step((tid >> 5) + k * CW);
    compute_sync();
    PROF_STAMP(0);

    // (x, y, z only: the intensity is not looked at before the resolve, and three registers per point instead of
    // four leave room in the two register sets of the software pipeline)
    auto ld3 = [&](const float4* p) {
      const float2 xy = __ldg(reinterpret_cast<const float2*>(p));
      return make_float3(xy.x, xy.y, __ldg(reinterpret_cast<const float*>(p) + 2));
    };
    auto load4 = [&](float3 (&v)[4], int q) {
      const int i0 = step_base(q) + lane;
      const float4* const p0 = fp + i0;
      if (i0 - lane + 3 * fstride + 32 <= n32) {   // the whole step lies inside the frame (warp-uniform): no bounds checks
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = ld3(p0 + u * fstride);
      } else {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * fstride;
          v[u] = i < n32 ? ld3(fp + i) : make_float3(0.f, 0.f, 0.f);
        }
      }
    };
    auto rasterise4 = [&](const float3 (&v)[4], const int q) {
      if (!DBG(64)) prefetch_step(q + 3 * CW);   // L2 prefetch three steps ahead
      const int i0 = step_base(q) + lane;
      // Per point, branch-free (so that the four points of a thread interleave): every decision
      // -- range-scan cell, range bucket, BEV bins -- is made in fp32 and PROVEN by its guard
      // band; a point with any unproven decision is deferred as a whole to S1b (fp64).
      int cell[4];        // range-scan cell of the point, -1 = none (tail, deferred, range scan off)
      unsigned key[4];
      unsigned be[4];     // BEV entry = flat output cell of the point + head (tile coordinates)
      int bb[4];          // its bucket, NO_BUCKET = not in the BEV
      bool defer[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * fstride;
        const bool valid = i < n32;
        const float4 w = xf_point<XF>(g, make_float4(v[u].x, v[u].y, v[u].z, 0.f));
        const float x = w.x, y = w.y, z = w.z;
        bool undecided = false;
        be[u] = 0u; bb[u] = NO_BUCKET; cell[u] = -1; key[u] = 0u;
        if (do_bev) {   // uniform
          const float tx = fmaf(x, g.bev_inv_x_f, g.bev_c0_x_f), ty = fmaf(y, g.bev_inv_y_f, g.bev_c0_y_f);
          int ibx, iby;
          const float flx = floor_small(tx, ibx), fly = floor_small(ty, iby);
          const float frx = tx - flx, fry = ty - fly;
          // proven bin: the fraction at least eps away from both edges and 0 < t < nb, as |t - nb/2| < nb/2
          // (nb/2 is exact and the rounding of the difference is monotone); proven outside: |t - nb/2| > nb/2 + 1
          const float dx = fabsf(tx - g.bev_hx_f), dy = fabsf(ty - g.bev_hy_f);
          const bool inx = (fabsf(frx - 0.5f) < g.bev_fe_x_f) & (dx < g.bev_hx_f);
          const bool iny = (fabsf(fry - 0.5f) < g.bev_fe_y_f) & (dy < g.bev_hy_f);
          const bool outx = dx > g.bev_hx1_f;
          const bool outy = dy > g.bev_hy1_f;
          // NaN: neither in nor out -> undecided -> fp64 (which drops it, like histogram2d)
          undecided = !(outx || outy) && !(inx && iny);
          const unsigned e = cells_m1 - ((unsigned)ibx * (unsigned)g.nby + (unsigned)iby) + head;
          const int bk = (int)(e >> bucket_shift);
          const bool inside = valid && inx && iny;
          be[u] = e;
          bb[u] = inside ? bk : NO_BUCKET;
        }
        bool okr = true;
        if (g.enable_range) {   // uniform
          const float s = fmaf(z, z, fmaf(y, y, x * x));
          const float rs = rsqrt_nr(s);
          float xi, yi, q;
          range_coords_fast(g, x, y, z, rs, xi, yi, q);
          int row, col;
          okr = range_cell_decide(g, xi, yi, q, row, col);
          // range bucket floor(64 rho) from rho32 = s * rsqrt(s) (relative error < 4e-7)
          const float t = (s * rs) * 64.0f;
          int ifl;
          const float fl = floor_small(t, ifl);
          const float fr = t - fl;
          const float gq = fmaf(t, 1.0e-6f, 1.0e-4f);
          okr = okr && fr > gq && fr < 1.0f - gq && t < 16382.0f;   // NaN -> false
          cell[u] = row * g.W + col;
          key[u] = ((unsigned)ifl << IDX_BITS32) | (unsigned)i;
        }
        defer[u] = valid && (undecided || !okr);
        if (!valid || defer[u] || !g.enable_range) { cell[u] = -1; key[u] = 0u; }   // (key 0: see note_collisions4)
        if (defer[u]) bb[u] = NO_BUCKET;
      }
      if (defer[0] | defer[1] | defer[2] | defer[3]) {   // rare: one reservation per thread
        unsigned pos = atomicAdd(&misc[0], (unsigned)defer[0] + (unsigned)defer[1] + (unsigned)defer[2] + (unsigned)defer[3]);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (defer[u]) { CHK(pos < per_cta32); my_queue[pos++] = (unsigned)(i0 + u * fstride); }
      }
      // the four shared-memory atomics of the range scan go out back to back; their results are
      // looked at after the BEV append (whose warp-wide steps they overlap)
      unsigned old[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        CHK(cell[u] < hw);
        old[u] = (cell[u] >= 0 && !DBG(32)) ? atomicMin(keys + cell[u], key[u]) : KEY32_EMPTY;
      }
      // BEV entries.  Neighbouring points of the thread in the same cell become one run (the count
      // - 1 rides in the two top bits of the entry).  Each run is appended to the bucket of its tile.
      if (do_bev && !DBG(16)) {
        unsigned cnt[4] = {0u, 0u, 0u, 0u};   // run length - 1
#pragma unroll
        for (int u = 3; u >= 1; --u) {
          const bool same = bb[u] != NO_BUCKET && bb[u - 1] != NO_BUCKET && be[u] == be[u - 1];
          if (same) { cnt[u - 1] += cnt[u] + 1u; bb[u] = NO_BUCKET; }
        }
        // (one shared-memory atomic per run and lane: measured 25 k cycles per sweep faster than electing a
        // leader per (warp, bucket) with __match_any_sync -- the point pass is bound by the latency of its warp-
        // wide steps, not by the shared-memory pipe that serialises the lanes of a bucket)
        unsigned slot[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) slot[u] = bb[u] != NO_BUCKET ? atomicAdd(&bcnt[bb[u]], 1u) : 0u;
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (bb[u] != NO_BUCKET) {
            CHK(bb[u] < a.n_buckets && slot[u] < per_cta32 && (be[u] >> a.tile_log2) < (unsigned)a.n_buckets * a.tiles_per_bucket);
            st_u32_hint(my_buckets + ((unsigned)bb[u] * per_cta32 + slot[u]), be[u] | (cnt[u] << 30), keep);
          }
      }
      note_collisions4(my_cq, misc, cell, key, old, 6u * per_cta32);
    };
    {
      float3 va[4], vb[4];
      int q = tid >> 5;
      load4(va, q);
      while (q < n_steps) {
        load4(vb, q + CW);
        rasterise4(va, q);
        q += CW;
        if (q >= n_steps) break;
        load4(va, q + CW);
        rasterise4(vb, q);
        q += CW;
      }
    }
    compute_sync();
    PROF_STAMP(1);
    if (tid == 0) misc[11] = 0u;   // the fp64 pass and the settlement wait on loads: fill window closed

    // ---- S1b: deferred points, full fp64 (guard bands, image border, wrap, NaN, errors)
    {
      const unsigned qn = misc[0];
      for (unsigned j = tid; j < qn; j += CT) {
        const unsigned i = my_queue[j];
        const float4 v = xf_point<XF>(g, __ldg(fp + i));
        if (do_bev) {
          const int bx = bev_bin_f64((double)v.x, g.bev_start, g.bev_delta_x, g.nex);
          const int by = bev_bin_f64((double)v.y, g.bev_start, g.bev_delta_y, g.ney);
          if (bx >= 0 && by >= 0) {
            const unsigned e = cells_m1 - ((unsigned)bx * (unsigned)g.nby + (unsigned)by) + head;
            const int bk = (int)(e >> bucket_shift);
            const unsigned slot = atomicAdd(&bcnt[bk], 1u);
            CHK(bk < a.n_buckets && slot < per_cta32);
            st_u32_hint(my_buckets + ((unsigned)bk * per_cta32 + slot), e, keep);
          }
        }
        if (g.enable_range) {
          Pt p; p.x = v.x; p.y = v.y; p.z = v.z; p.it = v.w;
          const double rho = rho_of(p);
          int row, col;
          const int err = range_cell_f64(g, p, rho, row, col);
          if (err) {
            atomicOr(&misc[1], (unsigned)err);
            atomicAdd(&misc[2], 1u);
          } else {
            const int cell = row * g.W + col;
            CHK(cell >= 0 && cell < hw);
            const unsigned key = (quant_rho(rho) << IDX_BITS32) | i;
            note_collision(my_cq, misc, cell, key, atomicMin(keys + cell, key));
          }
        }
      }
      compute_sync();
      PROF_STAMP(2);
    }
    if (g.enable_range) {
      // ---- S1d: exact settlement of the recorded equal-bucket candidates, 4 per thread in flight
      // (record -> key from shared memory -> candidate and holder points together -> CAS)
      {
        const unsigned cn = misc[7];
        for (unsigned j0 = 0; j0 < cn; j0 += 4 * CT) {
          uint2 en[4];
          float4 pv[4], hv[4];
          unsigned cur[4];
          bool act[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const unsigned j = j0 + u * CT + tid;
            act[u] = j < cn;
            if (act[u]) en[u] = my_cq[j];
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (!act[u]) continue;
            cur[u] = keys[en[u].x];
            if ((cur[u] & IDX_MASK32) == en[u].y) { act[u] = false; continue; }   // the holder itself has nothing to do
            if (DBG(512)) { pv[u] = make_float4(1.f, 2.f, 3.f, (float)en[u].y); hv[u] = make_float4(1.f, 2.f, 3.f, (float)cur[u]); continue; }
            pv[u] = xf_point<XF>(g, __ldg(fp + en[u].y));
            hv[u] = xf_point<XF>(g, __ldg(fp + (cur[u] & IDX_MASK32)));
          }
          if (DBG(256)) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (act[u] && pv[u].x + hv[u].y == 1.2345f) keys[0] = 0u;
            continue;
          }
          double pr[4], hr[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {   // eight independent square roots
            pr[u] = act[u] ? rho_of_f4(pv[u]) : 0.0;
            hr[u] = act[u] ? rho_of_f4(hv[u]) : 0.0;
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (act[u]) settle_candidate<XF>(g, fp, keys + en[u].x, pr[u], en[u].y, cur[u], hr[u]);
        }
      }
      compute_sync();
      PROF_STAMP(7);
      if (tid == 0) {
        if (a.status) a.status[f] = (int)misc[1];
        if (a.dropped) a.dropped[f] = DBG(128) ? (int)misc[7] : (int)misc[2];
      }

      // ---- S1c: resolve the range scan; output row r shows pre-flip row H-1-r   (:74)
      if (a.range_vec4) {
        // Task (k, r) = k-th aligned group of 4 flat output cells whose first cell lies in output
        // row r; tasks run k-major / r-minor, so a warp (H = 32: one lane per row) gathers the
        // winners of ~4 neighbouring columns -- neighbouring points of a firing-order sweep.
        // One software-pipelined loop: the winners of task i+PF are prefetched into L2, those of
        // task i+1 are gathered into registers, and task i is computed and stored meanwhile.
        const int kmax = ((g.W + 3) >> 2) + 1;
        // H % 8 == 0: a warp takes 8 rows x 4 groups (16 cells) -- in a firing-order sweep its
        // gathers touch 4 lines per instruction and its stores cover 192 contiguous bytes per
        // row; otherwise one lane per row (k-major).  L2 requests, not bytes, bound this phase.
        const bool blk8 = (g.H & 7) == 0;
        const int nrb = g.H >> 3, nkb = (kmax + 3) >> 2;
        const int ntask = blk8 ? nrb * nkb * 32 : kmax * g.H;
        const int hs = (g.H & (g.H - 1)) == 0 ? __ffs(g.H) - 1 : -1;
        constexpr int PF = 2;
        // first cell of task t (a multiple of 4), or -1 when t is past the end / an empty slot
        auto task_cell = [&](int t, int& r) -> int {
          r = 0;
          if (t >= ntask) return -1;
          int k;
          if (blk8) {
            const int w = t >> 5, l = t & 31;
            const int kb = hs >= 0 ? w >> (hs - 3) : w / nrb;
            const int rb = w - kb * nrb;
            r = rb * 8 + (l & 7); k = kb * 4 + (l >> 3);
          } else {
            k = hs >= 0 ? t >> hs : t / g.H;
            r = t - k * g.H;
          }
          const int gid = ((r * g.W + 3) >> 2) + k;
          return gid >= (((r + 1) * g.W + 3) >> 2) ? -1 : gid << 2;
        };
        auto load_keys = [&](int c, int r, unsigned (&key)[4]) {
          int orow = r, col = c - r * g.W;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            key[j] = keys[(g.H - 1 - orow) * g.W + col];
            if (++col == g.W) { col = 0; ++orow; }
          }
        };
        if (!DBG(8)) {
#pragma unroll 1
          for (int s = 1; s < PF; ++s) {   // prologue of the prefetch stream
            int r;
            const int c = task_cell(tid + s * CT, r);
            if (c < 0) continue;
            unsigned key[4];
            load_keys(c, r, key);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (key[j] != KEY32_EMPTY) prefetch_l2(fp + (key[j] & IDX_MASK32));
          }
        }
        unsigned nkey[4];
        float4 nv[4];
        int nr;
        int nc = task_cell(tid, nr);
        if (nc >= 0) {
          load_keys(nc, nr, nkey);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (nkey[j] != KEY32_EMPTY) nv[j] = __ldg(fp + (nkey[j] & IDX_MASK32));
        }
#pragma unroll 1
        for (int t = tid; t < ntask; t += CT) {
          const int c = nc;
          unsigned key[4];
          float4 v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) { key[j] = nkey[j]; v[j] = nv[j]; }
          if (!DBG(8)) {
            int pr;
            const int pc = task_cell(t + PF * CT, pr);
            if (pc >= 0) {
              unsigned pk[4];
              load_keys(pc, pr, pk);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (pk[j] != KEY32_EMPTY) prefetch_l2(fp + (pk[j] & IDX_MASK32));
            }
          }
          nc = task_cell(t + CT, nr);
          if (nc >= 0) {
            load_keys(nc, nr, nkey);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (nkey[j] != KEY32_EMPTY) nv[j] = DBG(4) ? make_float4(1.f, 2.f, 3.f, (float)nkey[j]) : __ldg(fp + (nkey[j] & IDX_MASK32));
          }
          if (c < 0) continue;
          float d[4], h[4], it[4];
          int win[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            d[j] = 0.f; h[j] = height_bg; it[j] = 0.f; win[j] = -1;
            if (key[j] != KEY32_EMPTY) {
              win[j] = (int)(key[j] & IDX_MASK32);
              const float4 w = xf_point<XF>(g, v[j]);
              const double x = w.x, y = w.y;
              d[j] = (float)__dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)));
              h[j] = w.z; it[j] = w.w;
            }
          }
          if (DBG(2)) { if (d[0] + h[1] + it[2] + d[3] == 1.2345f) a.range[0] = 1.f; continue; }
          if (a.range) {
            if (a.range_layout == 0) {
              float* o = a.range + (size_t)f * 3 * hw + c;
              __stcs(reinterpret_cast<float4*>(o), make_float4(d[0], d[1], d[2], d[3]));
              __stcs(reinterpret_cast<float4*>(o + hw), make_float4(h[0], h[1], h[2], h[3]));
              __stcs(reinterpret_cast<float4*>(o + 2 * hw), make_float4(it[0], it[1], it[2], it[3]));
            } else {
              // 48 contiguous bytes per lane: one 32-byte (aligned) + one 16-byte store
              float* o = a.range + ((size_t)f * hw + c) * 3;
              const float4 q0 = make_float4(d[0], h[0], it[0], d[1]), q1 = make_float4(h[1], it[1], d[2], h[2]),
                           q2 = make_float4(it[2], d[3], h[3], it[3]);
              if (((uintptr_t)o & 31) == 0) {
                st_cs_f32x8(o, q0, q1);
                __stcs(reinterpret_cast<float4*>(o + 8), q2);
              } else {
                __stcs(reinterpret_cast<float4*>(o), q0);
                st_cs_f32x8(o + 4, q1, q2);
              }
            }
          }
          if (a.winner)
            __stcs(reinterpret_cast<int4*>(a.winner + (size_t)f * hw + c), make_int4(win[0], win[1], win[2], win[3]));
        }
      } else {
        for (int c = tid; c < hw; c += CT) {
          const int orow = c / g.W, col = c - orow * g.W;
          const unsigned key = keys[(g.H - 1 - orow) * g.W + col];
          float d = 0.f, h = height_bg, it = 0.f;
          int win = -1;
          if (key != KEY32_EMPTY) {
            win = (int)(key & IDX_MASK32);
            const float4 v = xf_point<XF>(g, __ldg(fp + win));
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
      compute_sync();
      PROF_STAMP(3);
    } else if (tid == 0) {
      if (a.status) a.status[f] = 0;
      if (a.dropped) a.dropped[f] = 0;
    }

    // ---- S2: BEV, one flat tile of the output at a time.  Tiles are T = 2^tile_log2 cells of
    // tile coordinates (= frame cell + head; coordinate 0 is 1 KB aligned in the output).  Shared
    // memory now holds the u8 counts of the tile, nothing else.  A composed tile is scanned one
    // output row (32 sectors of 8 cells = 8 lines of 128 bytes) per warp step: the lines with a
    // non-empty cell are converted and written here, whole, and the row's bitmap is parked for the
    // store warp, which writes all the other lines (zeros).
    if (do_bev) {
      const int T = 1 << a.tile_log2, tw = T >> 2;
      const unsigned par = frames_done & 1u;
      unsigned* const bm_k = a.rowbits + ((size_t)blockIdx.x * 2 + par) * (a.rows_cap >> 2);   // one byte per row
      unsigned char* const lf = reinterpret_cast<unsigned char*>(tile + tw);   // line flags of the tile: byte (cell >> 5)
      const int lbw = T >> 7;                  // words of the line flags (8 bytes per row)
      // The parking area and the block cursor of this frame parity are free once frame k-2 is written.  If the
      // store warp is not there yet, the compute warps do not wait for it: they claim blocks of frame k-2 from the
      // same cursor and write them (as all warps do for a CTA's last frames) -- sixteen warps instead of one.
      if (store_active && frames_done >= 2u) {
        if ((int)(ld_acquire_smem(&misc[4]) - (frames_done - 1u)) < 0) {   // (one shared-memory read per warp: uniform)
          if (tid == 0) misc[11] = 2u;   // unpaced
          const unsigned k2 = frames_done - 2u;
          const int fk = (int)*reinterpret_cast<volatile unsigned*>(&misc[12 + (k2 & 3u)]) - 1;
          const BevJob j2 = make_bev_job(a, cells, fk, k2, *reinterpret_cast<volatile unsigned*>(&misc[104 + (k2 & 3u)]));
          bev_write_frame<XF>(j2, &misc[108 + (k2 & 1u)], misc, lane, false);
        }
        compute_sync();   // nobody claims from the cursor any more (the store warp finds it exhausted)
      }
      if (tid == 0) {
        misc[11] = 1u;   // compose works in shared memory: store-warp window open
        if (store_active && frames_done >= 2u && (int)(ld_acquire_smem(&misc[4]) - (frames_done - 1u)) < 0) {
          misc[11] = 2u;   // the store warp is finishing its last blocks of frame k-2
          while ((int)(ld_acquire_smem(&misc[4]) - (frames_done - 1u)) < 0) __nanosleep(200);
          misc[11] = 1u;
        }
        misc[108 + par] = 0u;
        st_release_smem(&misc[110], frames_done + 1u);   // the store warp may claim blocks of this frame
      }
      PROF_STAMP(8);
      for (int w = tid; w < ((tw + lbw) >> 2); w += CT) reinterpret_cast<uint4*>(tile)[w] = make_uint4(0u, 0u, 0u, 0u);   // tw, lbw: multiples of 4
      compute_sync();
      const unsigned span_all = (unsigned)cells + head;
      const int n_tiles = (int)((span_all + (unsigned)T - 1u) >> a.tile_log2);
      // entries of a tile's bucket, ENTRY_BATCH 16-byte loads per thread; slots past the end of
      // the bucket (stale) read as "not in this tile"
      auto load_batch = [&](int t, unsigned q0, uint4 (&e)[ENTRY_BATCH]) {
        const int bkt = t >> (bucket_shift - a.tile_log2);
        const unsigned cnt = bcnt[bkt];
        const uint4* const bucket4 = reinterpret_cast<const uint4*>(my_buckets + (size_t)bkt * a.per_cta);
        const unsigned n4 = (cnt + 3u) >> 2;
#pragma unroll
        for (int u = 0; u < ENTRY_BATCH; ++u) {
          const unsigned q = q0 + u * CT + tid;
          e[u] = q < n4 ? ld_u32x4_hint(bucket4 + q, keep) : make_uint4(~0u, ~0u, ~0u, ~0u);
          if (4u * q + 4u > cnt) {   // last word of the bucket
            if (4u * q + 1u > cnt) e[u].x = ~0u;
            if (4u * q + 2u > cnt) e[u].y = ~0u;
            if (4u * q + 3u > cnt) e[u].z = ~0u;
            e[u].w = ~0u;
          }
        }
      };
      uint4 pe[ENTRY_BATCH];   // first batch of the next tile, loaded while the current tile is scanned
      load_batch(0, 0u, pe);
      for (int t = 0; t < n_tiles; ++t) {
        const unsigned tlo = (unsigned)t << a.tile_log2;                 // tile coordinates of this tile
        const int bkt = t >> (bucket_shift - a.tile_log2);
        const unsigned cnt = bcnt[bkt];
        const unsigned n4 = (cnt + 3u) >> 2;
        bool danger = false;
        if (n4 == 0u && t + 1 < n_tiles) load_batch(t + 1, 0u, pe);
        for (unsigned q0 = 0; q0 < n4; q0 += ENTRY_BATCH * CT) {
          uint4 e[ENTRY_BATCH];
#pragma unroll
          for (int u = 0; u < ENTRY_BATCH; ++u) e[u] = pe[u];
          // the next batch (of this tile, else the first one of the next tile) is loaded now
          if (q0 + ENTRY_BATCH * CT < n4) load_batch(t, q0 + ENTRY_BATCH * CT, pe);
          else if (t + 1 < n_tiles) load_batch(t + 1, 0u, pe);
          // stage 1: all counts go out back to back; every slot keeps (position | valid | old byte << 24)
#pragma unroll
          for (int u = 0; u < ENTRY_BATCH; ++u) {
            e[u].x = tile_count_issue(tile, lf, e[u].x, tlo, (unsigned)T);
            e[u].y = tile_count_issue(tile, lf, e[u].y, tlo, (unsigned)T);
            e[u].z = tile_count_issue(tile, lf, e[u].z, tlo, (unsigned)T);
            e[u].w = tile_count_issue(tile, lf, e[u].w, tlo, (unsigned)T);
          }
          // stage 2: saturation fix-ups (rare)
#pragma unroll
          for (int u = 0; u < ENTRY_BATCH; ++u) {
            danger |= tile_count_finish(tile, e[u].x);
            danger |= tile_count_finish(tile, e[u].y);
            danger |= tile_count_finish(tile, e[u].z);
            danger |= tile_count_finish(tile, e[u].w);
          }
        }
        if (danger) misc[3] = 1u;
        compute_sync();
        PROF_STAMP(4);
        if (misc[3]) {  // a byte may have wrapped: redo this tile with the exact (CAS) form
          compute_sync();
          if (tid == 0) misc[3] = 0u;
          for (int w = tid; w < tw + lbw; w += CT) tile[w] = 0u;
          compute_sync();
          const unsigned* const bucket = my_buckets + (size_t)bkt * a.per_cta;
          for (unsigned i = tid; i < cnt; i += CT) {
            const unsigned e = __ldcg(bucket + i);
            const unsigned r = (e & ENTRY_CELL) - tlo;
            if (r < (unsigned)T) { tile_count_exact(tile, r, (e >> 30) + 1u); lf[r >> 5] = 1; }
          }

          compute_sync();
          PROF_STAMP(5);
        }
        // write-out.  (1) Warp w takes rows [w*chunk, (w+1)*chunk) of the tile, a lane two rows (16 line
        // flags -> 16 bits): the bits are parked for the store warp (which writes every line NOT
        // marked) and expanded into the tile's list of lines (one shared-memory atomic per warp).
        // (2) All warps share the list, eight lines per warp step: the four lanes of a line take one
        // 32-byte sector each (LUT of the counts, zeros included) and write it -- one full-line
        // request per non-empty line -- and re-zero the tile.  The first and the last row of the
        // frame share their lines with the neighbouring frames: all their lines are marked here and
        // written cell by cell.
        {
          const int rpt = T >> 8;                   // rows per tile
          const int chunk = ((rpt + CW - 1) / CW + 3) & ~3;   // <= 64 (T <= 2^17)
          const int rlo = (tid >> 5) * chunk, rhi = min(rlo + chunk, rpt);
          const int n_rows = (int)((span_all + (ROW_CELLS - 1)) / ROW_CELLS);
          uint2* const tile2 = reinterpret_cast<uint2*>(tile);
          unsigned short* const wl = reinterpret_cast<unsigned short*>(tile + tw + lbw);   // [T / 32] lines of the tile
          {
            const int r = rlo + 2 * lane;           // my two rows: r, r + 1 (rlo, rhi are multiples of 4 or the tile's end)
            unsigned b16 = 0u;
            if (r < rhi) {
              const uint4 v = reinterpret_cast<uint4*>(lf)[r >> 1];
              if ((v.x | v.y | v.z | v.w) != 0u) reinterpret_cast<uint4*>(lf)[r >> 1] = make_uint4(0u, 0u, 0u, 0u);
              b16 = ((v.x * 0x01020408u) >> 24) | (((v.y * 0x01020408u) >> 24) << 4) |
                    (((v.z * 0x01020408u) >> 24) << 8) | (((v.w * 0x01020408u) >> 24) << 12);
              const int row = t * rpt + r;
              if (row == 0 || row == n_rows - 1) b16 |= 0x00FFu;
              if (row + 1 == n_rows - 1) b16 |= 0xFF00u;
            }
            // parked: four rows per word
            const unsigned hi = __shfl_down_sync(0xFFFFFFFFu, b16, 1);
            CHK(r >= rhi || (((size_t)t * rpt + r) >> 2) < (size_t)(a.rows_cap >> 2));
            if ((lane & 1) == 0 && r < rhi) bm_k[((size_t)t * rpt + r) >> 2] = b16 | (hi << 16);
            const unsigned cnt = (unsigned)__popc(b16);
            unsigned inc = cnt;                     // inclusive prefix of the line counts over the lanes
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const unsigned o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
              if (lane >= d) inc += o;
            }
            const unsigned wtot = __shfl_sync(0xFFFFFFFFu, inc, 31);
            unsigned off = 0u;
            if (lane == 0 && wtot) off = atomicAdd(&misc[10], wtot);
            off = __shfl_sync(0xFFFFFFFFu, off, 0) + inc - cnt;
            const unsigned base = (unsigned)r << 3; // line index (within the tile) of bit 0
            while (b16) {
              const unsigned b = (unsigned)__ffs(b16) - 1u;
              b16 &= b16 - 1u;
              CHK(off < (unsigned)(T >> 5));
              wl[off++] = (unsigned short)(base + b);
            }
          }
          compute_sync();
          PROF_STAMP(9);
          if (tid == 0) misc[11] = 0u;   // the write-out stores to global memory: store-warp window closed
          const unsigned wn = misc[10];
          float* const bp = a.bev + (long long)f * cells - head;   // tile coordinate 0
          for (unsigned j = (unsigned)(tid >> 2); j < wn; j += CT / 4) {
            const unsigned li = wl[j];
            const int r = (int)(li >> 3), sct = (int)((li & 7u) << 2) + (lane & 3);
            CHK(r < rpt && (size_t)(t * rpt + r) < (size_t)a.rows_cap);
            const uint2 w = tile2[r * 32 + sct];
            tile2[r * 32 + sct] = make_uint2(0u, 0u);
            float4 v0, v1;
            lut8(lut, w, v0, v1);
            const int row = t * rpt + r;
            float* const o = bp + (size_t)row * ROW_CELLS + sct * 8;
            if (row > 0 && row < n_rows - 1) {
              st_cs_f32x8(o, v0, v1);
            } else {
              const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
              const unsigned c0 = (unsigned)row * ROW_CELLS + (unsigned)sct * 8u;
#pragma unroll
              for (int q = 0; q < 8; ++q)
                if (c0 + q >= head && c0 + q < span_all) o[q] = vv[q];
            }
          }
        }
        compute_sync();
        if (tid == 0) {   // the tile's bitmaps are parked: the store warp may write its rows
          st_release_smem(&misc[5], misc[5] + 1u);
          misc[11] = 1u;
          misc[10] = 0u;   // the next tile's list
        }
        PROF_STAMP(6);
      }
    }
  }
  // The compute warps are done; the store warp is up to two frames behind.  All warps now write what
  // is left of the CTA's last two frames (the store warp joins in as it gets there: every block is
  // claimed from the frame's cursor exactly once).
  if (store_active && frames_done > 0u) {
    if (tid == 0) misc[11] = 2u;
    for (unsigned k = frames_done >= 2u ? frames_done - 2u : 0u; k < frames_done; ++k) {
      const int fk = (int)*reinterpret_cast<volatile unsigned*>(&misc[12 + (k & 3u)]) - 1;
      const BevJob j = make_bev_job(a, cells, fk, k, *reinterpret_cast<volatile unsigned*>(&misc[104 + (k & 3u)]));
      bev_write_frame<XF>(j, &misc[108 + (k & 1u)], misc, lane, false);
    }
    PROF_STAMP(10);
  }
  if (a.prof && tid == 0)
    for (int k = 0; k < SFPE_N_PHASES; ++k) a.prof[(size_t)blockIdx.x * SFPE_N_PHASES + k] += prof_acc[k];
}

// ---------------------------------------------------------------------------------------
static int main_words_max() { return SMEM_LIMIT / 4 - SMEM_MISC_WORDS; }


// S2 tile geometry: T = 2^p cells = T bytes (u8 counts)
// bytes of shared memory; at most MAX_BUCKETS entry buckets, consecutive tiles share a bucket.
FusedGeom fused_geometry(const DevGrid& g) {
  FusedGeom o;
  o.tile_log2 = 10; o.tiles_per_bucket = 1; o.n_buckets = 1; o.s2_words = 0; o.rows_cap = 1;
  if (!g.enable_bev) return o;
  const long long cells = (long long)g.nbx * g.nby;
  // tile counts + line flags (a byte per 32 cells) + per-warp lists of non-empty lines (u16 each, 8 per row of the warp's share)
  auto words = [](int p) {
    const long long T = 1ll << p;
    return (int)(T / 4 + (T >> 7) + (T >> 6));
  };
  int p = 10;   // a tile is at least four output rows (1024 cells = one word of the line bitmap)
  while (p < 22 && (1ll << p) < cells + ROW_CELLS && words(p + 1) <= main_words_max()) ++p;
  o.tile_log2 = p;
  o.s2_words = (words(p) + 3) & ~3;
  const long long n_tiles = (cells + (ROW_CELLS - 1) + (1ll << p) - 1) >> p;   // head <= 255
  o.rows_cap = n_tiles << (p - 8);
  o.tiles_per_bucket = 1;   // a power of two: the bucket of an entry is a shift
  while ((n_tiles + o.tiles_per_bucket - 1) / o.tiles_per_bucket > MAX_BUCKETS) o.tiles_per_bucket *= 2;
  o.n_buckets = (int)((n_tiles + o.tiles_per_bucket - 1) / o.tiles_per_bucket);
  return o;
}

bool fused_supported(const DevGrid& g, const PointSrc& src, const DevOut& out, long long max_n) {
  const bool packed = src.dtype == 0 && src.stride == 16 && src.ox == 0 && src.oy == 4 && src.oz == 8 &&
                      src.oi == 12 && ((uintptr_t)src.data & 15) == 0;
  if (!packed || max_n > FUSED_MAX_N) return false;   // (height slices: sfpe_ext.cu, after this kernel; extrinsic: the XF instantiation)
  if (g.enable_range && (out.range_dtype != 0 || g.H * g.W > main_words_max())) return false;
  if (g.enable_bev && (out.bev_dtype != 0 || out.bev_counts)) return false;
  if (g.enable_bev && (long long)g.nbx * g.nby > 0x3FF00000ll) return false;   // entries: 30-bit tile coordinate + 2-bit run length
  return g.enable_range || g.enable_bev;
}

int fused_grid(int device, int n_frames) {
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms <= 0) sms = 148;
  return n_frames < sms ? n_frames : sms;
}

cudaError_t prepare_fused() {
  // per device (per primary context): called from sfpe_create after cudaSetDevice
  const cudaError_t e = cudaFuncSetAttribute(k_fused_sweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k_fused_sweep<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
}

cudaError_t launch_fused(const DevGrid& g, const PointSrc& src, const FrameSet& fs, const Scratch& s,
                         const DevOut& out, const FusedScratch& fx, cudaStream_t st, long long* launches) {
  int s2_words = 0;
  const FusedGeom geo = fused_geometry(g);
  if (g.enable_bev) s2_words = geo.s2_words;
  int main_words = g.enable_range ? g.H * g.W : 0;
  if (s2_words > main_words) main_words = s2_words;
  main_words = (main_words + 3) & ~3;
  FusedArgs a;
  a.pts = reinterpret_cast<const float4*>(src.data);
  a.fs = fs;
  a.entries = fx.entries; a.queue = fx.queue; a.cqueue = reinterpret_cast<uint2*>(fx.cqueue); a.per_cta = fx.per_cta;
  a.rowbits = fx.rowbits; a.rows_cap = geo.rows_cap;
  a.lut64 = s.lut;
  a.counter = fx.counter;
  { cudaError_t e = cudaMemsetAsync(fx.counter, 0, sizeof(unsigned), st); if (e != cudaSuccess) return e; }
  a.range = (float*)out.range; a.range_layout = out.range_layout; a.winner = out.range_winner;
  a.bev = (float*)out.bev;
  a.status = out.status; a.dropped = out.dropped;
  a.spw_log2 = fx.spw_log2;
  a.tile_log2 = geo.tile_log2; a.tiles_per_bucket = geo.tiles_per_bucket; a.n_buckets = geo.n_buckets;
  a.range_vec4 = ((g.H * g.W) & 3) == 0 && ((uintptr_t)out.range & 15) == 0 && ((uintptr_t)out.range_winner & 15) == 0;
  a.prof = fx.prof;
  a.debug = fx.debug;
  const size_t smem = (size_t)(main_words + SMEM_MISC_WORDS) * 4;
  if (g.use_extrinsic) k_fused_sweep<true><<<fx.grid, FUSED_THREADS, smem, st>>>(g, a, main_words);
  else k_fused_sweep<false><<<fx.grid, FUSED_THREADS, smem, st>>>(g, a, main_words);
  ++*launches;
  return cudaGetLastError();
}


// ---------------------------------------------------------------------------------------
// Self-test of the fp32 fast path: largest deviation of each fp32 cell coordinate from its fp64
// definition over a point set, as a FRACTION of the guard band that protects it (< 1 means the
// guard band holds; the tests want < 0.5).  out[0] column, [1] row, [2] range bucket, [3] BEV.
__global__ void __launch_bounds__(256) k_selftest_fastpath(const __grid_constant__ DevGrid g, const float4* __restrict__ pts,
                                                           long long n, unsigned long long* out) {
  double m[4] = {0.0, 0.0, 0.0, 0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = __ldg(pts + i);
    const float x = v.x, y = v.y, z = v.z;
    const float s = fmaf(z, z, fmaf(y, y, x * x));
    const float rs = rsqrt_nr(s);
    float xi, yi, q;
    range_coords_fast(g, x, y, z, rs, xi, yi, q);
    Pt p; p.x = x; p.y = y; p.z = z; p.it = v.w;
    const double rho = rho_of(p);
    const double xi64 = __dsub_rn(__ddiv_rn(atan2(-p.y, p.x), g.res_rad_h), g.x_min);
    const double yi64 = __dsub_rn(__ddiv_rn(asin(__ddiv_rn(p.z, rho)), g.res_rad_v), g.y_min);
    if (xi64 == xi64 && yi64 == yi64 && fabsf(q) <= MAX_SIN_EL) {
      m[0] = fmax(m[0], fabs((double)xi - xi64) / (double)EPS_COL);
      m[1] = fmax(m[1], fabs((double)yi - yi64) / (double)EPS_ROW);
    }
    const float t = (s * rs) * 64.0f;
    if (t < 16382.0f) m[2] = fmax(m[2], fabs((double)t - rho * 64.0) / (double)fmaf(t, 1.0e-6f, 1.0e-4f));
    if (g.enable_bev) {
      const float tx = fmaf(x, g.bev_inv_x_f, g.bev_c0_x_f), ty = fmaf(y, g.bev_inv_y_f, g.bev_c0_y_f);
      if (tx > 0.f && tx < (float)g.nbx) m[3] = fmax(m[3], fabs((double)tx - ((double)x - g.bev_start) / g.bev_delta_x) / (double)g.bev_eps_x_f);
      if (ty > 0.f && ty < (float)g.nby) m[3] = fmax(m[3], fabs((double)ty - ((double)y - g.bev_start) / g.bev_delta_y) / (double)g.bev_eps_y_f);
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) atomicMax(out + k, (unsigned long long)__double_as_longlong(m[k]));
}

cudaError_t launch_selftest_fastpath(const DevGrid& g, const float* pts, long long n, unsigned long long* d_out4,
                                     cudaStream_t st, long long* launches) {
  cudaError_t e = cudaMemsetAsync(d_out4, 0, 4 * sizeof(unsigned long long), st);
  if (e != cudaSuccess || n <= 0) return e;
  k_selftest_fastpath<<<148 * 4, 256, 0, st>>>(g, reinterpret_cast<const float4*>(pts), n, d_out4);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
