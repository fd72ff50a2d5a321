// sfpe_range_scan_static: the generic range-scan kernels (k_project + k_resolve_range of
// sfpe_generic.cu) run with their winner-key scratch in a file-scope device array instead of
// memory from cudaMalloc.  For callers that may not allocate: code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The array is part of the library image: every GPU that loads libsfpe.so holds
// SFPE_STATIC_MAX_CELLS * 8 bytes (2 MiB) until the process exits.  All calls on a device share
// it; it keeps the last call's contents, so each call clears the part it uses first.  Calls that
// overlap on the GPU (two streams, two threads) race on it.
#include <mutex>
#include <stdio.h>

#include "../../include/sfpe.h"
#include "sfpe_internal.h"

namespace sfpe {
int derive_grid_for(const sfpe_params* p, DevGrid* g);  // sfpe_api.cu
}
using namespace sfpe;

static_assert(SFPE_STATIC_MAX_CELLS == (1 << 18), "documented size of the static scratch");

// In host code this name is only a handle for cudaGetSymbolAddress; device code uses it directly.
static __device__ __align__(256) unsigned long long g_static_keys[SFPE_STATIC_MAX_CELLS];

// device code can name the array: idle value of the winner keys for the cells this call uses
__global__ void __launch_bounds__(256) k_static_keys_clear(int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) g_static_keys[i] = KEY_EMPTY;
}

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned long long* g_keys_addr[MAX_DEVICES];  // device address of g_static_keys, per device

// Once per device, with that device current.
int static_keys_for(int device, unsigned long long** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_keys_addr[device]) {
    unsigned long long* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_static_keys);
    if (e != cudaSuccess || !ptr) {
      fprintf(stderr, "libsfpe: cudaGetSymbolAddress(g_static_keys) on device %d failed: %s\n", device,
              cudaGetErrorString(e));
      return SFPE_ERR_CUDA;
    }
    g_keys_addr[device] = ptr;
  }
  *out = g_keys_addr[device];
  return SFPE_OK;
}

int ceil_log2_ll(long long v) {
  int b = 0;
  while ((1ll << b) < v) ++b;
  return b;
}
}  // namespace

#define CKS(call)                                                                   \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess) {                                                        \
      fprintf(stderr, "libsfpe: %s failed: %s\n", #call, cudaGetErrorString(e_));  \
      return SFPE_ERR_CUDA;                                                         \
    }                                                                               \
  } while (0)

extern "C" int sfpe_range_scan_static(int device, const sfpe_params* p, const sfpe_points* pts,
                                      const int64_t* d_offsets, int64_t uniform_n, int64_t max_n, int32_t n_frames,
                                      const sfpe_outputs* o, void* stream) {
  if (!p || !pts || !o || n_frames < 0 || device < 0 || device >= MAX_DEVICES) return SFPE_ERR_INVALID_ARG;
  if (!d_offsets) max_n = uniform_n;
  if (max_n < 0 || max_n >= (1ll << 31) || (!d_offsets && uniform_n < 0)) return SFPE_ERR_INVALID_ARG;
  if (pts->dtype != SFPE_F32 && pts->dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  if (max_n > 0 && !pts->data) return SFPE_ERR_INVALID_ARG;
  const int esz = pts->dtype == SFPE_F32 ? 4 : 8;
  if (pts->stride_bytes < esz || pts->stride_bytes % esz || pts->off_x % esz || pts->off_y % esz ||
      pts->off_z % esz || pts->off_i % esz || ((uintptr_t)pts->data % esz))
    return SFPE_ERR_INVALID_ARG;
  if (o->range_dtype < 0 || o->range_dtype > 1 || o->range_layout < 0 || o->range_layout > 1) return SFPE_ERR_INVALID_ARG;
  // range scan only: the BEV members of the output block have no scratch here
  if (o->bev || o->bev_counts || o->bev_height || o->bev_intensity) return SFPE_ERR_INVALID_ARG;

  sfpe_params q = *p;
  q.enable_range = 1;
  q.enable_bev = 0;
  q.n_slices = 0;
  DevGrid g;
  const int rc = derive_grid_for(&q, &g);
  if (rc) return rc;
  const long long hw = (long long)g.H * g.W;
  if (hw > SFPE_STATIC_MAX_CELLS) return SFPE_ERR_TOO_LARGE;
  if (n_frames == 0) return SFPE_OK;

  const int need = ceil_log2_ll(max_n > 1 ? max_n : 1);
  g.idx_bits = need > 20 ? need : 20;
  if (p->key_prefix_bits > 0) {
    const int ib = 64 - p->key_prefix_bits;
    if (ib < need || ib > 56) return SFPE_ERR_INVALID_ARG;
    g.idx_bits = ib;
  }

  CKS(cudaSetDevice(device));
  Scratch s;
  s.keys = nullptr; s.counts = nullptr; s.top = nullptr; s.slice_max = nullptr; s.lut = nullptr;
  const int krc = static_keys_for(device, &s.keys);
  if (krc) return krc;
  cudaStream_t st = (cudaStream_t)stream;

  PointSrc src;
  src.data = (const char*)pts->data; src.stride = pts->stride_bytes; src.dtype = pts->dtype;
  src.ox = pts->off_x; src.oy = pts->off_y; src.oz = pts->off_z; src.oi = pts->off_i;
  DevOut out;
  out.range = o->range; out.range_dtype = o->range_dtype; out.range_layout = o->range_layout;
  out.range_winner = o->range_winner;
  out.bev = nullptr; out.bev_dtype = 0; out.bev_counts = nullptr; out.bev_height = nullptr; out.bev_intensity = nullptr;
  out.status = o->status; out.dropped = o->dropped;
  if (out.status) CKS(cudaMemsetAsync(out.status, 0, sizeof(int) * n_frames, st));
  if (out.dropped) CKS(cudaMemsetAsync(out.dropped, 0, sizeof(int) * n_frames, st));

  // frames per pass: as many as the array holds
  long long chunk = SFPE_STATIC_MAX_CELLS / hw;
  if (chunk > n_frames) chunk = n_frames;
  if (chunk > 65535) chunk = 65535;  // gridDim.y of the generic kernels
  const int used = (int)(chunk * hw);
  int blocks = (used + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_static_keys_clear<<<blocks, 256, 0, st>>>(used);
  CKS(cudaGetLastError());

  long long launches = 1;
  FrameSet fs;
  fs.offsets = (const long long*)d_offsets; fs.uniform_n = uniform_n;
  for (int f0 = 0; f0 < n_frames; f0 += (int)chunk) {
    fs.frame0 = f0;
    fs.n_frames = n_frames - f0 < chunk ? n_frames - f0 : (int)chunk;
    // k_resolve_range leaves every key it reads idle again, so the next pass starts clean
    CKS(launch_generic_chunk(g, src, fs, max_n, s, out, st, &launches));
  }
  return SFPE_OK;
}
