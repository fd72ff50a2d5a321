// Internal host/device interface between the C-ABI translation unit and the kernel files.
#pragma once
#include "../../include/sfpe.h"
#include "sfpe_common.cuh"

namespace sfpe {

struct FrameSet {
  const long long* offsets;  // device int64[n_frames+1] or nullptr
  long long uniform_n;       // used when offsets == nullptr
  int frame0;                // first frame of this chunk (global frame index)
  int n_frames;              // frames in this chunk
};

__device__ __forceinline__ void frame_span(const FrameSet& fs, int f, long long& base, long long& n) {
  if (fs.offsets) {
    base = fs.offsets[f];
    n = fs.offsets[f + 1] - base;
  } else {
    base = (long long)f * fs.uniform_n;
    n = fs.uniform_n;
  }
}

struct DevOut {
  void* range;
  int range_dtype, range_layout;
  int* range_winner;
  void* bev;
  int bev_dtype;
  unsigned* bev_counts;
  float* bev_height;
  float* bev_intensity;
  int* status;
  int* dropped;
};

// Scratch owned by the ctx, sized for `chunk` frames.
struct Scratch {
  unsigned long long* keys;  // [chunk][H*W]   packed winner keys, KEY_EMPTY when idle
  unsigned* counts;          // [chunk][nbx*nby] BEV counts, 0 when idle
  unsigned long long* top;   // [chunk][nbx*nby] extension: highest-point key, 0 when idle
  unsigned* slice_max;       // [chunk][n_slices][nbx*nby] extension: ordered max z, 0 when idle
  const double* lut;         // [65] density LUT (device computed)
};

// generic (global-scratch) path -- sfpe_generic.cu
cudaError_t launch_fill_scratch(const DevGrid& g, const Scratch& s, int chunk, cudaStream_t st, long long* launches);
cudaError_t launch_density_lut(double* d_lut, cudaStream_t st, long long* launches);
cudaError_t launch_generic_chunk(const DevGrid& g, const PointSrc& src, const FrameSet& fs, long long max_n,
                                 const Scratch& s, const DevOut& out, cudaStream_t st, long long* launches);
cudaError_t launch_decode_pc2(const void* blob, long long n, int step, int ox, int oy, int oz, int oi, int dtype,
                              float* out, cudaStream_t st, long long* launches);

// mm: device double[2] scratch {min,max} (only touched when autoscale)
cudaError_t launch_colormap(const void* img, int dtype, long long n, int cmap, int autoscale, double vmin, double vmax,
                            unsigned long long* mm, unsigned char* rgb, cudaStream_t st, long long* launches);

// consumers of the range scan (invariant mask, roll augmentation, back-projection) -- sfpe_post.cu
cudaError_t launch_mask_update(const void* ref, const void* frames, int dtype, long long n_frames, long long frame_stride,
                               long long n_pix, int pix_stride, unsigned char* mask, int sm_count, cudaStream_t st,
                               long long* launches);
cudaError_t launch_roll_width(const void* in, void* out, long long n_frames, int height, int width, int pix_bytes,
                              const int* shift, cudaStream_t st, long long* launches);
cudaError_t launch_back_project(double* centroids, double* bboxes, long long n, const void* distance, const void* height,
                                int dtype, long long frame_stride, int pix_stride, int H, int W, double min_height,
                                double x_min, double res_rad, double* xyz, int* status, cudaStream_t st,
                                long long* launches);

cudaError_t launch_find_bbox_3d(const void* distance, const void* height, int dtype, long long frame_stride, int pix_stride,
                                const void* y_pred, int pred_dtype, long long n, int H, int W, const int* bbox2d,
                                const double* centroid3d, double x_min, double y_min, double res_h, double res_v,
                                double* boxes, int* counts, long long cap, double* out_pred, double* out_box, int* n_boxes,
                                const unsigned* posbits, cudaStream_t st, long long* launches);

cudaError_t launch_find_obstacle(const void* y_pred, int pred_dtype, int n_ch, long long n, int H, int W, int* scratch,
                                 int* out, unsigned* posbits_out, cudaStream_t st, long long* launches);

// fused one-sweep-per-SM path -- sfpe_fused.cu
struct FusedScratch {
  unsigned* entries;   // [grid][per_cta]
  unsigned* queue;     // [grid][per_cta]
  unsigned* cqueue;    // [grid][12*per_cta] words: collision records
  unsigned* counter;   // frame claim counter (zeroed by launch_fused)
  unsigned* rowbits;   // [grid][2][rows_cap]: parked BEV row bitmaps (non-empty sectors)
  long long per_cta;   // >= max points of a frame
  int grid;            // CTAs (<= SMs)
  int spw_log2;        // log2(firing stride / 32), 0..3
  long long* prof;     // optional [grid][SFPE_N_PHASES] phase cycle accumulators
  int debug;           // sfpe_params.reserved[0]: ablation bits (-DSFPE_ABLATE builds only)
};
struct FusedGeom { int tile_log2, tiles_per_bucket, n_buckets, s2_words; long long rows_cap; };
FusedGeom fused_geometry(const DevGrid& g);
bool fused_supported(const DevGrid& g, const PointSrc& src, const DevOut& out, long long max_n);
int fused_grid(int device, int n_frames);
cudaError_t prepare_fused();   // per device: opt in to the large dynamic shared memory size
cudaError_t launch_fused(const DevGrid& g, const PointSrc& src, const FrameSet& fs, const Scratch& s,
                         const DevOut& out, const FusedScratch& fx, cudaStream_t st, long long* launches);

// extension E next to the fused kernel (height-slice planes + intensity of the highest point) -- sfpe_ext.cu
struct ExtScratch {
  unsigned long long* top;   // [chunk][nbx*nby] key of the cell's highest point, 0 when idle (the winners reset it)
  int* pcell;                // [chunk][max_n]   BEV cell of a point that takes part, -1 otherwise
  long long max_n;
};
cudaError_t launch_ext_chunk(const DevGrid& g, const PointSrc& src, const FrameSet& fs, long long max_n,
                             const ExtScratch& xs, const DevOut& out, cudaStream_t st, long long* launches);

// sfpe_transport.cu: one-byte-per-cell transport of the BEV density image
cudaError_t launch_bev_pack_u8(const void* bev, int dtype, long long n, const double* lut, unsigned char* codes, int* bad,
                               int sm_count, cudaStream_t st, long long* launches);
void bev_unpack_host(const unsigned char* codes, long long n, const double* lut65, void* out, int dtype, int n_threads);
void bev_unpack_rgb_host(const unsigned char* codes, long long n, const unsigned char* rgb65, unsigned char* out, int n_threads);

cudaError_t launch_selftest_fastpath(const DevGrid& g, const float* pts, long long n, unsigned long long* d_out4,
                                     cudaStream_t st, long long* launches);

}  // namespace sfpe
