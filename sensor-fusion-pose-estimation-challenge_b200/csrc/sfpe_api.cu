// C ABI of libsfpe.so (declared in include/sfpe.h).  Host-side logic only: parameter
// derivation, scratch management, chunking, path selection.  No CPU compute fallback.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/sfpe.h"
#include "sfpe_internal.h"

using namespace sfpe;

struct sfpe_ctx {
  int device;
  sfpe_params params;
  DevGrid grid;
  cudaStream_t stream;
  Scratch scratch;
  int chunk;            // frames the scratch is sized for
  double* d_lut;
  long long launches;
  char err[256];
  const char* last_path;
  // staging for sfpe_project_sweep_host
  void* st_points; size_t st_points_cap;
  void* st_range; void* st_bev; int* st_winner; int* st_status;  // status[0], dropped[1]
  unsigned long long* d_mm;  // [0..1] colormap autoscale {min,max}, [2..5] fast-path self-test maxima
  void* st_img; size_t st_img_cap; void* st_rgb; size_t st_rgb_cap;
  void* fused_buf; size_t fused_cap;  // entries + deferred queue of the fused path
  long long* d_prof; int prof_grid;   // phase profile of the fused kernel (sfpe_phase_profile)
  void* ext_buf; size_t ext_cap; size_t ext_top_bytes;   // extension E next to the fused kernel: top keys + per-point cells of a chunk
};

#define CK(ctx, call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      snprintf((ctx)->err, sizeof((ctx)->err), "%s: %s", #call, cudaGetErrorString(e_)); \
      return SFPE_ERR_CUDA;                                                             \
    }                                                                                   \
  } while (0)

extern "C" int sfpe_abi_version(void) { return SFPE_ABI_VERSION; }

extern "C" const char* sfpe_strerror(int s) {
  switch (s) {
    case SFPE_OK: return "ok";
    case SFPE_ERR_INVALID_ARG: return "invalid argument";
    case SFPE_ERR_CUDA: return "CUDA error (see sfpe_last_error)";
    case SFPE_ERR_UNSUPPORTED: return "unsupported configuration";
    case SFPE_ERR_NOMEM: return "out of memory";
    case SFPE_ERR_TOO_LARGE: return "grid larger than the static scratch (H*W > SFPE_STATIC_MAX_CELLS)";
    default: return "unknown status";
  }
}

extern "C" void sfpe_default_params(sfpe_params* p) {
  memset(p, 0, sizeof(*p));
  p->res_v_deg = 1.33;   // process/globals.py:6
  p->res_h_deg = 0.2;
  p->vfov_lo_deg = -30.67;  // :7
  p->vfov_hi_deg = 10.67;
  p->height_bg = -2.0;      // :4
  p->bev_range = 120.0;     // birds_eye_view_generator.py:41
  p->bev_res_x = 0.2;       // extract_rosbag.py:30 class default
  p->bev_res_y = 0.2;
  p->enable_range = 1;
  p->enable_bev = 1;
  for (int r = 0; r < 3; ++r) p->extrinsic[4 * r + r] = 1.0;
}

// np.arange(-R, R, step): length ceil((stop-start)/step), fill delta (start+step)-start
static int arange_len(double R, double step, double* delta) {
  volatile double start = -R, stop = R;
  volatile double first = start + step;
  *delta = first - start;
  const double len = ceil((stop - start) / step);
  if (!(len >= 2.0) || len > 65536.0) return -1;
  return (int)len;
}

static int derive_grid(const sfpe_params* p, DevGrid* g) {
  memset(g, 0, sizeof(*g));
  if (!(p->res_v_deg > 0) || !(p->res_h_deg > 0) || !(p->vfov_lo_deg <= 0) ||
      !(p->vfov_hi_deg > p->vfov_lo_deg))
    return SFPE_ERR_INVALID_ARG;
  const double deg = M_PI / 180.0;                       // globals.py:9
  g->res_rad_v = p->res_v_deg * deg;
  g->res_rad_h = p->res_h_deg * deg;
  g->x_min = -360.0 / p->res_h_deg / 2;                  // :10
  g->y_min = p->vfov_lo_deg / p->res_v_deg;              // :11
  g->W = (int)(360.0 / p->res_h_deg) + 1;                // :12
  g->H = (int)(fabs(p->vfov_lo_deg - p->vfov_hi_deg) / p->res_v_deg) + 1;  // :13
  g->height_bg = p->height_bg;
  g->enable_range = p->enable_range != 0;
  g->enable_bev = p->enable_bev != 0;
  if (g->W < 2 || g->H < 1 || (long long)g->W * g->H > (1ll << 26)) return SFPE_ERR_INVALID_ARG;
  if (g->enable_bev) {
    if (!(p->bev_range > 0) || !(p->bev_res_x > 0) || !(p->bev_res_y > 0)) return SFPE_ERR_INVALID_ARG;
    g->bev_start = -p->bev_range;
    g->nex = arange_len(p->bev_range, p->bev_res_x, &g->bev_delta_x);
    g->ney = arange_len(p->bev_range, p->bev_res_y, &g->bev_delta_y);
    if (g->nex < 2 || g->ney < 2) return SFPE_ERR_INVALID_ARG;
    g->nbx = g->nex - 1;
    g->nby = g->ney - 1;
  }
  g->use_extrinsic = p->use_extrinsic != 0;
  memcpy(g->M, p->extrinsic, sizeof(g->M));
  if (p->n_slices < 0 || p->n_slices > SFPE_MAX_SLICES) return SFPE_ERR_INVALID_ARG;
  g->n_slices = g->enable_bev ? p->n_slices : 0;
  for (int i = 0; i <= p->n_slices; ++i) {
    g->slice_edges[i] = p->slice_edges[i];
    if (i && !(p->slice_edges[i] > p->slice_edges[i - 1])) return SFPE_ERR_INVALID_ARG;
  }
  g->idx_bits = 20;
  g->bev_inv_delta_x = g->enable_bev ? 1.0 / g->bev_delta_x : 0.0;
  g->bev_inv_delta_y = g->enable_bev ? 1.0 / g->bev_delta_y : 0.0;
  g->inv_res_h_f = (float)(1.0 / g->res_rad_h);
  g->inv_res_v_f = (float)(1.0 / g->res_rad_v);
  g->neg_x_min_f = (float)(-g->x_min);
  g->neg_y_min_f = (float)(-g->y_min);
  {
    const float hW = 0.5f * (float)g->W, hH = 0.5f * (float)g->H;   // exact
    g->rc_cx_f = hW - 0.25f; g->rc_hx_f = hW + 0.25f;
    g->rc_cy_f = hH - 0.25f; g->rc_hy_f = hH + 0.25f;
  }
  if (g->enable_bev) {
    // t = (v - start)/delta = v*inv + c0; guard = 3 ulp32 at the largest bin coordinate
    g->bev_inv_x_f = (float)g->bev_inv_delta_x;
    g->bev_inv_y_f = (float)g->bev_inv_delta_y;
    g->bev_c0_x_f = (float)(-g->bev_start * g->bev_inv_delta_x);
    g->bev_c0_y_f = (float)(-g->bev_start * g->bev_inv_delta_y);
    g->bev_eps_x_f = 3.0f * (nextafterf((float)g->nex, INFINITY) - (float)g->nex);
    g->bev_eps_y_f = 3.0f * (nextafterf((float)g->ney, INFINITY) - (float)g->ney);
    g->bev_hx_f = 0.5f * (float)g->nbx; g->bev_hy_f = 0.5f * (float)g->nby;
    g->bev_hx1_f = g->bev_hx_f + 1.0f;  g->bev_hy1_f = g->bev_hy_f + 1.0f;
    g->bev_fe_x_f = 0.5f - g->bev_eps_x_f; g->bev_fe_y_f = 0.5f - g->bev_eps_y_f;
  }
  return SFPE_OK;
}

namespace sfpe {
int derive_grid_for(const sfpe_params* p, DevGrid* g) { return derive_grid(p, g); }  // for sfpe_static.cu
}

extern "C" int sfpe_grid_dims(const sfpe_params* p, int32_t* H, int32_t* W, int32_t* nbx, int32_t* nby) {
  if (!p) return SFPE_ERR_INVALID_ARG;
  DevGrid g;
  const int rc = derive_grid(p, &g);
  if (rc) return rc;
  if (H) *H = g.H;
  if (W) *W = g.W;
  if (nbx) *nbx = g.nbx;
  if (nby) *nby = g.nby;
  return SFPE_OK;
}

static size_t per_frame_scratch(const DevGrid& g) {
  size_t b = 0;
  if (g.enable_range) b += (size_t)g.H * g.W * 8;
  if (g.enable_bev) b += (size_t)g.nbx * g.nby * 4;
  if (g.n_slices) b += (size_t)g.nbx * g.nby * (8 + 4 * (size_t)g.n_slices);
  return b;
}

extern "C" int sfpe_create(int device, const sfpe_params* p, sfpe_ctx** out) {
  if (!p || !out) return SFPE_ERR_INVALID_ARG;
  *out = nullptr;
  DevGrid g;
  const int rc = derive_grid(p, &g);
  if (rc) return rc;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return SFPE_ERR_CUDA;
  sfpe_ctx* c = (sfpe_ctx*)calloc(1, sizeof(sfpe_ctx));
  if (!c) return SFPE_ERR_NOMEM;
  c->device = device;
  c->params = *p;
  c->grid = g;
  c->last_path = "none";
  int status = SFPE_OK;
#define CKC(call)                                                                  \
  if (status == SFPE_OK) {                                                         \
    cudaError_t e_ = (call);                                                       \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "libsfpe: %s failed: %s\n", #call, cudaGetErrorString(e_)); \
      status = e_ == cudaErrorMemoryAllocation ? SFPE_ERR_NOMEM : SFPE_ERR_CUDA;   \
    }                                                                              \
  }
  CKC(cudaSetDevice(device));
  CKC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CKC(prepare_fused());
  // chunk: keep the scratch of one chunk L2 resident (126 MB L2; budget 40 MB)
  const size_t pf = per_frame_scratch(g);
  size_t chunk = pf ? (40u << 20) / pf : 1;
  if (chunk < 1) chunk = 1;
  if (chunk > 32) chunk = 32;
  c->chunk = (int)chunk;
  if (g.enable_range) CKC(cudaMalloc(&c->scratch.keys, chunk * g.H * g.W * 8));
  if (g.enable_bev) CKC(cudaMalloc(&c->scratch.counts, chunk * (size_t)g.nbx * g.nby * 4));
  if (g.n_slices) {
    CKC(cudaMalloc(&c->scratch.top, chunk * (size_t)g.nbx * g.nby * 8));
    CKC(cudaMalloc(&c->scratch.slice_max, chunk * (size_t)g.nbx * g.nby * 4 * g.n_slices));
    CKC(cudaMemsetAsync(c->scratch.top, 0, chunk * (size_t)g.nbx * g.nby * 8, c->stream));
    CKC(cudaMemsetAsync(c->scratch.slice_max, 0, chunk * (size_t)g.nbx * g.nby * 4 * g.n_slices, c->stream));
  }
  CKC(cudaMalloc(&c->d_lut, 65 * sizeof(double)));
  c->scratch.lut = c->d_lut;
  CKC(cudaMalloc(&c->st_status, 2 * sizeof(int)));
  CKC(cudaMalloc(&c->d_mm, 6 * sizeof(unsigned long long)));
  if (status == SFPE_OK) {
    CKC(launch_fill_scratch(g, c->scratch, c->chunk, c->stream, &c->launches));
    CKC(launch_density_lut(c->d_lut, c->stream, &c->launches));
    CKC(cudaStreamSynchronize(c->stream));
  }
#undef CKC
  if (status != SFPE_OK) {
    sfpe_destroy(c);
    return status;
  }
  *out = c;
  return SFPE_OK;
}

extern "C" void sfpe_destroy(sfpe_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  cudaFree(c->scratch.keys); cudaFree(c->scratch.counts); cudaFree(c->scratch.top);
  cudaFree(c->scratch.slice_max); cudaFree(c->d_lut);
  cudaFree(c->st_points); cudaFree(c->st_range); cudaFree(c->st_bev); cudaFree(c->st_winner);
  cudaFree(c->st_status); cudaFree(c->d_mm); cudaFree(c->st_img); cudaFree(c->st_rgb); cudaFree(c->fused_buf); cudaFree(c->d_prof); cudaFree(c->ext_buf);
  if (c->stream) cudaStreamDestroy(c->stream);
  free(c);
}

extern "C" const char* sfpe_last_error(const sfpe_ctx* c) { return c ? c->err : "null ctx"; }
extern "C" int64_t sfpe_launch_count(const sfpe_ctx* c) { return c ? c->launches : 0; }
extern "C" const char* sfpe_last_path(const sfpe_ctx* c) { return c ? c->last_path : "none"; }

static int grow(sfpe_ctx* c, void** p, size_t* cap, size_t need) {
  if (*cap >= need) return SFPE_OK;
  if (*p) CK(c, cudaFree(*p));
  *p = nullptr; *cap = 0;
  CK(c, cudaMalloc(p, need));
  *cap = need;
  return SFPE_OK;
}

static int ceil_log2(long long v) {
  int b = 0;
  while ((1ll << b) < v) ++b;
  return b;
}

extern "C" int sfpe_project_batch(sfpe_ctx* c, const sfpe_points* pts, const int64_t* d_offsets, int64_t uniform_n,
                                  int64_t max_n, int32_t n_frames, const sfpe_outputs* o, void* stream) {
  if (!c || !pts || !o || n_frames < 0) return SFPE_ERR_INVALID_ARG;
  if (!d_offsets) max_n = uniform_n;
  if (max_n < 0 || max_n >= (1ll << 31) || (!d_offsets && uniform_n < 0)) return SFPE_ERR_INVALID_ARG;
  if (pts->dtype != SFPE_F32 && pts->dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  if (max_n > 0 && !pts->data) return SFPE_ERR_INVALID_ARG;
  const int esz = pts->dtype == SFPE_F32 ? 4 : 8;
  if (pts->stride_bytes < esz || pts->stride_bytes % esz || pts->off_x % esz || pts->off_y % esz ||
      pts->off_z % esz || pts->off_i % esz || ((uintptr_t)pts->data % esz))
    return SFPE_ERR_INVALID_ARG;
  if (o->range_dtype < 0 || o->range_dtype > 1 || o->bev_dtype < 0 || o->bev_dtype > 1 || o->range_layout < 0 ||
      o->range_layout > 1)
    return SFPE_ERR_INVALID_ARG;
  if (n_frames == 0) return SFPE_OK;
  CK(c, cudaSetDevice(c->device));
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the CUDA default stream

  DevGrid g = c->grid;
  const int need = ceil_log2(max_n > 1 ? max_n : 1);
  g.idx_bits = need > 20 ? need : 20;
  if (c->params.key_prefix_bits > 0) {
    int ib = 64 - c->params.key_prefix_bits;
    if (ib < need || ib > 56) return SFPE_ERR_INVALID_ARG;
    g.idx_bits = ib;
  }
  PointSrc src;
  src.data = (const char*)pts->data; src.stride = pts->stride_bytes; src.dtype = pts->dtype;
  src.ox = pts->off_x; src.oy = pts->off_y; src.oz = pts->off_z; src.oi = pts->off_i;
  DevOut out;
  out.range = g.enable_range ? o->range : nullptr;
  out.range_dtype = o->range_dtype; out.range_layout = o->range_layout;
  out.range_winner = g.enable_range ? o->range_winner : nullptr;
  out.bev = g.enable_bev ? o->bev : nullptr;
  out.bev_dtype = o->bev_dtype;
  out.bev_counts = g.enable_bev ? o->bev_counts : nullptr;
  out.bev_height = g.n_slices ? o->bev_height : nullptr;
  out.bev_intensity = g.n_slices ? o->bev_intensity : nullptr;
  out.status = o->status; out.dropped = o->dropped;
  if (out.status) CK(c, cudaMemsetAsync(out.status, 0, sizeof(int) * n_frames, st));
  if (out.dropped) CK(c, cudaMemsetAsync(out.dropped, 0, sizeof(int) * n_frames, st));

  FrameSet fs;
  fs.offsets = (const long long*)d_offsets; fs.uniform_n = uniform_n;
  const int want = c->params.path;
  if (want != 1 && fused_supported(g, src, out, max_n)) {
    FusedScratch fx;
    fx.grid = fused_grid(c->device, n_frames);
    fx.per_cta = (max_n + 255) & ~255ll;
    if (fx.per_cta < 256) fx.per_cta = 256;
    const FusedGeom geo = fused_geometry(g);
    // entry buckets + deferred queue + collision records + parked BEV row bitmaps (two frames) + claim counter
    const size_t words = (size_t)fx.grid * fx.per_cta * (geo.n_buckets + 1 + 12) + (size_t)fx.grid * geo.rows_cap * 2;
    const int grc = grow(c, &c->fused_buf, &c->fused_cap, words * 4 + 256);
    if (grc) return grc;
    fx.entries = (unsigned*)c->fused_buf;
    fx.queue = fx.entries + (size_t)fx.grid * fx.per_cta * geo.n_buckets;
    fx.cqueue = fx.queue + (size_t)fx.grid * fx.per_cta;
    fx.rowbits = fx.cqueue + (size_t)fx.grid * fx.per_cta * 12;
    fx.counter = fx.rowbits + (size_t)fx.grid * geo.rows_cap * 2;
    fs.frame0 = 0; fs.n_frames = n_frames;
    fx.debug = c->params.ablation;
    fx.spw_log2 = 0;
    for (int fsd = c->params.firing_stride; fsd >= 64 && fx.spw_log2 < 3; fsd >>= 1) ++fx.spw_log2;
    fx.prof = (c->d_prof && fx.grid <= c->prof_grid) ? c->d_prof : nullptr;
    CK(c, launch_fused(g, src, fs, c->scratch, out, fx, st, &c->launches));
    c->last_path = "fused";
    if (g.n_slices && (out.bev_height || out.bev_intensity)) {
      // extension E: dense fill of the planes + two passes over the points (sfpe_ext.cu), in chunks of
      // <= 64 frames whose top-key scratch stays within 1 GB (three launches per chunk: at 23 frames per chunk the
      // gaps between 135 launches were 10 % of a 1024-sweep batch)
      const size_t cells = (size_t)g.nbx * g.nby;
      size_t xchunk = ((size_t)1 << 30) / (cells * 8);
      if (xchunk < 1) xchunk = 1;
      if (xchunk > 64) xchunk = 64;
      if (xchunk > (size_t)n_frames) xchunk = n_frames;
      const size_t top_bytes = (xchunk * cells * 8 + 255) & ~(size_t)255;
      const size_t need_bytes = top_bytes + xchunk * (size_t)(max_n > 0 ? max_n : 1) * 4;
      if (c->ext_cap < need_bytes) {
        const int xrc = grow(c, &c->ext_buf, &c->ext_cap, need_bytes);
        if (xrc) return xrc;
        CK(c, cudaMemsetAsync(c->ext_buf, 0, top_bytes, st));
      } else if (c->ext_top_bytes != top_bytes) {
        CK(c, cudaMemsetAsync(c->ext_buf, 0, c->ext_cap, st));   // other layout than the last call: all idle again
      }
      c->ext_top_bytes = top_bytes;
      ExtScratch xs;
      xs.top = (unsigned long long*)c->ext_buf;
      xs.pcell = (int*)((char*)c->ext_buf + top_bytes);
      xs.max_n = max_n > 0 ? max_n : 1;
      for (int f0 = 0; f0 < n_frames; f0 += (int)xchunk) {
        fs.frame0 = f0;
        fs.n_frames = n_frames - f0 < (int)xchunk ? n_frames - f0 : (int)xchunk;
        CK(c, launch_ext_chunk(g, src, fs, max_n, xs, out, st, &c->launches));
      }
      c->last_path = "fused+ext";
    }
    return SFPE_OK;
  }
  if (want == 2) return SFPE_ERR_UNSUPPORTED;
  for (int f0 = 0; f0 < n_frames; f0 += c->chunk) {
    fs.frame0 = f0;
    fs.n_frames = n_frames - f0 < c->chunk ? n_frames - f0 : c->chunk;
    CK(c, launch_generic_chunk(g, src, fs, max_n, c->scratch, out, st, &c->launches));
  }
  c->last_path = "generic";
  return SFPE_OK;
}


extern "C" int sfpe_project_sweep_host(sfpe_ctx* c, const void* h_points, const sfpe_points* layout, int64_t n,
                                       void* h_range, int32_t range_dtype, int32_t range_layout, int32_t* h_winner,
                                       void* h_bev, int32_t bev_dtype, int32_t* frame_status,
                                       int32_t* frame_dropped) {
  if (!c || !layout || n < 0 || (n > 0 && !h_points)) return SFPE_ERR_INVALID_ARG;
  if (range_dtype < 0 || range_dtype > 1 || bev_dtype < 0 || bev_dtype > 1) return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  const DevGrid& g = c->grid;
  const size_t hw = (size_t)g.H * g.W, cells = (size_t)g.nbx * g.nby;
  const size_t in_bytes = (size_t)n * layout->stride_bytes;
  int rc = grow(c, &c->st_points, &c->st_points_cap, in_bytes ? in_bytes : 16);
  if (rc) return rc;
  if (g.enable_range && !c->st_range) {
    CK(c, cudaMalloc(&c->st_range, 3 * hw * 8));
    CK(c, cudaMalloc((void**)&c->st_winner, hw * 4));
  }
  if (g.enable_bev && !c->st_bev) CK(c, cudaMalloc(&c->st_bev, cells * 8));
  if (in_bytes) CK(c, cudaMemcpyAsync(c->st_points, h_points, in_bytes, cudaMemcpyHostToDevice, c->stream));
  sfpe_points pts = *layout;
  pts.data = c->st_points;
  sfpe_outputs o;
  memset(&o, 0, sizeof(o));
  o.range = h_range ? c->st_range : nullptr;
  o.range_dtype = range_dtype; o.range_layout = range_layout;
  o.range_winner = h_winner ? c->st_winner : nullptr;
  o.bev = h_bev ? c->st_bev : nullptr;
  o.bev_dtype = bev_dtype;
  o.status = c->st_status; o.dropped = c->st_status + 1;
  rc = sfpe_project_batch(c, &pts, nullptr, n, n, 1, &o, c->stream);
  if (rc) return rc;
  const size_t rsz = range_dtype == SFPE_F32 ? 4 : 8, bsz = bev_dtype == SFPE_F32 ? 4 : 8;
  if (h_range && g.enable_range) CK(c, cudaMemcpyAsync(h_range, c->st_range, 3 * hw * rsz, cudaMemcpyDeviceToHost, c->stream));
  if (h_winner && g.enable_range) CK(c, cudaMemcpyAsync(h_winner, c->st_winner, hw * 4, cudaMemcpyDeviceToHost, c->stream));
  if (h_bev && g.enable_bev) CK(c, cudaMemcpyAsync(h_bev, c->st_bev, cells * bsz, cudaMemcpyDeviceToHost, c->stream));
  int hs[2] = {0, 0};
  CK(c, cudaMemcpyAsync(hs, c->st_status, sizeof(hs), cudaMemcpyDeviceToHost, c->stream));
  CK(c, cudaStreamSynchronize(c->stream));
  if (frame_status) *frame_status = hs[0];
  if (frame_dropped) *frame_dropped = hs[1];
  return SFPE_OK;
}

extern "C" int sfpe_decode_pointcloud2(sfpe_ctx* c, const void* d_blob, int64_t n, int32_t point_step, int32_t off_x,
                                       int32_t off_y, int32_t off_z, int32_t off_i, int32_t dtype, float* d_xyzi,
                                       void* stream) {
  if (!c || n < 0 || point_step <= 0 || (n > 0 && (!d_blob || !d_xyzi))) return SFPE_ERR_INVALID_ARG;
  if (dtype != SFPE_F32 && dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  const int esz = dtype == SFPE_F32 ? 4 : 8;
  if (point_step % esz || off_x % esz || off_y % esz || off_z % esz || off_i % esz || ((uintptr_t)d_blob % esz) ||
      ((uintptr_t)d_xyzi % 16))
    return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the CUDA default stream
  CK(c, launch_decode_pc2(d_blob, n, point_step, off_x, off_y, off_z, off_i, dtype, d_xyzi, st, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_colormap(sfpe_ctx* c, const void* d_img, int32_t dtype, int64_t n, int32_t cmap, int32_t autoscale,
                             double vmin, double vmax, uint8_t* d_rgb, void* stream) {
  if (!c || n < 0 || (n > 0 && (!d_img || !d_rgb))) return SFPE_ERR_INVALID_ARG;
  if ((dtype != SFPE_F32 && dtype != SFPE_F64) || (cmap != SFPE_CMAP_GRAY && cmap != SFPE_CMAP_JET))
    return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the CUDA default stream
  CK(c, launch_colormap(d_img, dtype, n, cmap, autoscale != 0, vmin, vmax, c->d_mm, d_rgb, st, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_colormap_host(sfpe_ctx* c, const void* h_img, int32_t dtype, int64_t n, int32_t cmap,
                                  int32_t autoscale, double vmin, double vmax, uint8_t* h_rgb) {
  if (!c || n < 0 || (n > 0 && (!h_img || !h_rgb))) return SFPE_ERR_INVALID_ARG;
  if (dtype != SFPE_F32 && dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  if (n == 0) return SFPE_OK;
  CK(c, cudaSetDevice(c->device));
  const size_t esz = dtype == SFPE_F32 ? 4 : 8;
  int rc = grow(c, &c->st_img, &c->st_img_cap, (size_t)n * esz);
  if (rc) return rc;
  rc = grow(c, &c->st_rgb, &c->st_rgb_cap, (size_t)n * 3);
  if (rc) return rc;
  CK(c, cudaMemcpyAsync(c->st_img, h_img, (size_t)n * esz, cudaMemcpyHostToDevice, c->stream));
  rc = sfpe_colormap(c, c->st_img, dtype, n, cmap, autoscale, vmin, vmax, (uint8_t*)c->st_rgb, c->stream);
  if (rc) return rc;
  CK(c, cudaMemcpyAsync(h_rgb, c->st_rgb, (size_t)n * 3, cudaMemcpyDeviceToHost, c->stream));
  CK(c, cudaStreamSynchronize(c->stream));
  return SFPE_OK;
}

extern "C" int sfpe_phase_profile(sfpe_ctx* c, int32_t enable) {
  if (!c) return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  if (!enable) {
    if (c->d_prof) CK(c, cudaFree(c->d_prof));
    c->d_prof = nullptr;
    return SFPE_OK;
  }
  if (!c->d_prof) {
    int sms = 0;
    CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    c->prof_grid = sms;
    CK(c, cudaMalloc((void**)&c->d_prof, (size_t)sms * SFPE_N_PHASES * sizeof(long long)));
  }
  CK(c, cudaMemset(c->d_prof, 0, (size_t)c->prof_grid * SFPE_N_PHASES * sizeof(long long)));
  return SFPE_OK;
}

extern "C" int sfpe_phase_cycles(sfpe_ctx* c, int64_t* out8) {
  if (!c || !out8) return SFPE_ERR_INVALID_ARG;
  for (int i = 0; i < SFPE_N_PHASES; ++i) out8[i] = 0;
  if (!c->d_prof) return SFPE_OK;
  CK(c, cudaSetDevice(c->device));
  CK(c, cudaDeviceSynchronize());
  long long* h = (long long*)malloc((size_t)c->prof_grid * SFPE_N_PHASES * sizeof(long long));
  if (!h) return SFPE_ERR_NOMEM;
  cudaError_t e = cudaMemcpy(h, c->d_prof, (size_t)c->prof_grid * SFPE_N_PHASES * sizeof(long long), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess)
    for (int b = 0; b < c->prof_grid; ++b)
      for (int i = 0; i < SFPE_N_PHASES; ++i) out8[i] += h[b * SFPE_N_PHASES + i];
  free(h);
  CK(c, e);
  return SFPE_OK;
}

extern "C" int sfpe_selftest_fastpath(sfpe_ctx* c, const float* d_xyzi, int64_t n, double* out4) {
  if (!c || !out4 || n < 0 || (n > 0 && !d_xyzi) || ((uintptr_t)d_xyzi & 15)) return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  unsigned long long* d = c->d_mm + 2;   // scratch owned by the ctx (no allocation per call)
  CK(c, launch_selftest_fastpath(c->grid, d_xyzi, n, d, c->stream, &c->launches));
  CK(c, cudaMemcpyAsync(out4, d, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CK(c, cudaStreamSynchronize(c->stream));
  return SFPE_OK;
}

// ---- consumers of the range scan ----------------------------------------------

extern "C" int sfpe_invariant_mask_update(sfpe_ctx* c, const void* d_ref, const void* d_frames, int32_t dtype,
                                          int64_t n_frames, int64_t frame_stride, int64_t n_pix, int32_t pix_stride,
                                          uint8_t* d_mask, void* stream) {
  if (!c || n_frames < 0 || n_pix < 0 || pix_stride < 1 || frame_stride < 0) return SFPE_ERR_INVALID_ARG;
  if (dtype != SFPE_F32 && dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  if (n_frames > 0 && n_pix > 0 && (!d_ref || !d_frames || !d_mask)) return SFPE_ERR_INVALID_ARG;
  const size_t esz = dtype == SFPE_F32 ? 4 : 8;
  if (((uintptr_t)d_ref % esz) || ((uintptr_t)d_frames % esz)) return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  int sms = 0;
  CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
  CK(c, launch_mask_update(d_ref, d_frames, dtype, n_frames, frame_stride, n_pix, pix_stride, d_mask, sms,
                           (cudaStream_t)stream, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_roll_width(sfpe_ctx* c, const void* d_in, void* d_out, int64_t n_frames, int32_t height,
                               int32_t width, int32_t pix_bytes, const int32_t* d_shift, void* stream) {
  if (!c || n_frames < 0 || height < 0 || width < 0 || pix_bytes <= 0 || pix_bytes % 4) return SFPE_ERR_INVALID_ARG;
  const long long total = (long long)n_frames * height * width;
  if (total > 0 && (!d_in || !d_out || !d_shift || d_in == d_out)) return SFPE_ERR_INVALID_ARG;
  if (((uintptr_t)d_in % 4) || ((uintptr_t)d_out % 16)) return SFPE_ERR_INVALID_ARG;
  if ((long long)width * (pix_bytes / 4) > 0x7fffffffll) return SFPE_ERR_UNSUPPORTED;
  CK(c, cudaSetDevice(c->device));
  CK(c, launch_roll_width(d_in, d_out, n_frames, height, width, pix_bytes, d_shift, (cudaStream_t)stream, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_density_lut(sfpe_ctx* c, double* h_lut65) {
  if (!c || !h_lut65) return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  CK(c, cudaStreamSynchronize(c->stream));   // the table is filled on the context's stream at creation
  CK(c, cudaMemcpy(h_lut65, c->d_lut, 65 * sizeof(double), cudaMemcpyDeviceToHost));
  return SFPE_OK;
}

extern "C" int sfpe_bev_pack_u8(sfpe_ctx* c, const void* d_bev, int32_t dtype, int64_t n_cells, uint8_t* d_codes,
                                int32_t* d_flag, void* stream) {
  if (!c || n_cells < 0 || (dtype != SFPE_F32 && dtype != SFPE_F64)) return SFPE_ERR_INVALID_ARG;
  if (n_cells > 0 && (!d_bev || !d_codes || !d_flag)) return SFPE_ERR_INVALID_ARG;
  if (((uintptr_t)d_bev % 16) || ((uintptr_t)d_codes % 4)) return SFPE_ERR_INVALID_ARG;
  CK(c, cudaSetDevice(c->device));
  int sms = 0;
  CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
  CK(c, launch_bev_pack_u8(d_bev, dtype, n_cells, c->d_lut, d_codes, d_flag, sms, (cudaStream_t)stream, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_bev_unpack_host(const uint8_t* h_codes, int64_t n_cells, const double* h_lut65, void* h_out,
                                    int32_t dtype, int32_t n_threads) {
  if (n_cells < 0 || (dtype != SFPE_F32 && dtype != SFPE_F64)) return SFPE_ERR_INVALID_ARG;
  if (n_cells > 0 && (!h_codes || !h_lut65 || !h_out)) return SFPE_ERR_INVALID_ARG;
  if ((uintptr_t)h_out % (dtype == SFPE_F32 ? 4 : 8)) return SFPE_ERR_INVALID_ARG;
  bev_unpack_host(h_codes, n_cells, h_lut65, h_out, dtype, n_threads);
  return SFPE_OK;
}

extern "C" int sfpe_bev_unpack_rgb_host(const uint8_t* h_codes, int64_t n_cells, const uint8_t* h_rgb65, uint8_t* h_out,
                                        int32_t n_threads) {
  if (n_cells < 0) return SFPE_ERR_INVALID_ARG;
  if (n_cells > 0 && (!h_codes || !h_rgb65 || !h_out)) return SFPE_ERR_INVALID_ARG;
  bev_unpack_rgb_host(h_codes, n_cells, h_rgb65, h_out, n_threads);
  return SFPE_OK;
}

extern "C" int sfpe_back_project(sfpe_ctx* c, double* d_centroids, double* d_bboxes, int64_t n, const void* d_distance,
                                 const void* d_height, int32_t dtype, int64_t frame_stride, int32_t pix_stride,
                                 int32_t height, int32_t width, double min_height, double x_min, double res_rad,
                                 double* d_xyz, int32_t* d_status, void* stream) {
  if (!c || n < 0 || height <= 0 || width <= 0 || pix_stride < 1 || frame_stride < 0) return SFPE_ERR_INVALID_ARG;
  if (dtype != SFPE_F32 && dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  if (n > 0 && (!d_centroids || !d_bboxes || !d_distance || !d_height || !d_xyz || !d_status)) return SFPE_ERR_INVALID_ARG;
  if ((long long)height * width > 0x7fffffffll || n > 0x7fffffffll) return SFPE_ERR_UNSUPPORTED;
  CK(c, cudaSetDevice(c->device));
  CK(c, launch_back_project(d_centroids, d_bboxes, n, d_distance, d_height, dtype, frame_stride, pix_stride, height, width,
                            min_height, x_min, res_rad, d_xyz, d_status, (cudaStream_t)stream, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_find_bbox_3d(sfpe_ctx* c, const void* d_distance, const void* d_height, int32_t dtype,
                                 int64_t frame_stride, int32_t pix_stride, const void* d_pred, int32_t pred_dtype, int64_t n,
                                 int32_t height, int32_t width, const int32_t* d_bbox2d, const double* d_centroid3d,
                                 double x_min, double y_min, double res_rad_h, double res_rad_v, double* d_boxes,
                                 int32_t* d_counts, int64_t cap, double* d_out_pred, double* d_out_box, int32_t* d_n_boxes,
                                 const uint32_t* d_posbits, void* stream) {
  if (!c || n < 0 || height <= 0 || width <= 0 || pix_stride < 1 || frame_stride < 0 || cap < 0) return SFPE_ERR_INVALID_ARG;
  if ((dtype != SFPE_F32 && dtype != SFPE_F64) || (pred_dtype != SFPE_F32 && pred_dtype != SFPE_F64)) return SFPE_ERR_INVALID_ARG;
  if (n > 0 && (!d_distance || !d_height || !d_pred || !d_bbox2d || !d_centroid3d || !d_out_pred || !d_out_box || !d_n_boxes ||
                (cap > 0 && (!d_boxes || !d_counts))))
    return SFPE_ERR_INVALID_ARG;
  if ((long long)height * width > 0x7fffffffll / 26 || n > 0x7fffffffll || (long long)width + height > 200 * 1024)
    return SFPE_ERR_UNSUPPORTED;
  CK(c, cudaSetDevice(c->device));
  CK(c, launch_find_bbox_3d(d_distance, d_height, dtype, frame_stride, pix_stride, d_pred, pred_dtype, n, height, width,
                            d_bbox2d, d_centroid3d, x_min, y_min, res_rad_h, res_rad_v, d_boxes, d_counts, cap, d_out_pred,
                            d_out_box, d_n_boxes, d_posbits, (cudaStream_t)stream, &c->launches));
  return SFPE_OK;
}

extern "C" int sfpe_find_obstacle(sfpe_ctx* c, const void* d_pred, int32_t pred_dtype, int32_t n_channels, int64_t n,
                                  int32_t height, int32_t width, int32_t* d_scratch, int32_t* d_out, uint32_t* d_posbits,
                                  void* stream) {
  if (!c || n < 0 || height <= 0 || width <= 0 || n_channels < 2) return SFPE_ERR_INVALID_ARG;
  if (pred_dtype != SFPE_F32 && pred_dtype != SFPE_F64) return SFPE_ERR_INVALID_ARG;
  if (n > 0 && (!d_pred || !d_scratch || !d_out)) return SFPE_ERR_INVALID_ARG;
  if (9ll * height * width / 8 > 220 * 1024 || n > 0x7fffffffll) return SFPE_ERR_UNSUPPORTED;   // flags live in shared memory
  CK(c, cudaSetDevice(c->device));
  CK(c, launch_find_obstacle(d_pred, pred_dtype, n_channels, n, height, width, d_scratch, d_out, d_posbits, (cudaStream_t)stream,
                             &c->launches));
  return SFPE_OK;
}
