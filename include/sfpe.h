/* sfpe.h -- C ABI of libsfpe.so: B200 (sm_100a) lidar rasterisation.
 *
 * Drop-in boundary for ONE hot path of J-Rojas/Sensor-Fusion-Pose-Estimation-Challenge:
 * the projection of a Velodyne sweep into (a) the 360-degree cylindrical range scan and
 * (b) the bird's-eye-view density map.  The reference has no FFI layer; it calls plain
 * numpy functions.  Each entry point below names the reference function it replaces
 * (paths relative to modules/lidar/ of the reference):
 *
 *   sfpe_project_batch        process/extract_rosbag_lidar.py:18-86  (lidar_2d_front_view x3 =
 *                             generate_lidar_2d_front_view) and
 *                             common/birds_eye_view_generator.py:31-54 (normalize +
 *                             generate_birds_eye_view), for a batch of independent sweeps
 *   sfpe_project_sweep_host   the same, one sweep from/to HOST buffers -- the call made by
 *                             train/predict.py:335-341 (predict_point_cloud) behind
 *                             team_sf_rosnode/scripts/lidar_predict.py:60-70
 *   sfpe_decode_pointcloud2   sensor_msgs.point_cloud2.read_points + np.array(list(..)) at
 *                             team_sf_rosnode/scripts/lidar_predict.py:64-65,
 *                             process/extract_rosbag.py:165-166, train/predict.py:373-374
 *   sfpe_invariant_mask_update  process/create_lidar_invariant_mask.py:44-62
 *   sfpe_roll_width           train/loader.py:31-57 (data_random_rotate's np.roll)
 *   sfpe_back_project         train/predict.py:230-293 (back_project_2D_2_3D)
 *   sfpe_find_bbox_3d         train/predict.py:94-199 (find_bbox_3d, is_far)
 *   sfpe_find_obstacle        train/predict.py:33-81 (find_obstacle; scipy.ndimage.label)
 *   sfpe_params defaults      process/globals.py:3-13 (RES, VFOV, LIDAR_MIN_HEIGHT, ...)
 *
 * Conventions: every function returns 0 (SFPE_OK) or a negative sfpe_status; nothing throws
 * across the ABI.  All buffers are owned by the caller; the library owns only scratch
 * inside the ctx.  "device" pointers are CUDA device pointers of the ctx's device.
 * `stream` is a cudaStream_t passed as void* (NULL = the CUDA default stream, as in CUDA itself).  One ctx per
 * device per thread; a ctx is not re-entrant.  There is NO CPU fallback: without a CUDA
 * device sfpe_create fails with SFPE_ERR_CUDA.
 */
#ifndef SFPE_H_
#define SFPE_H_

#include <stdint.h>

#if defined(__GNUC__)
#define SFPE_API __attribute__((visibility("default")))
#else
#define SFPE_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define SFPE_ABI_VERSION 1
#define SFPE_MAX_SLICES 8

typedef enum {
  SFPE_OK = 0,
  SFPE_ERR_INVALID_ARG = -1,
  SFPE_ERR_CUDA = -2,
  SFPE_ERR_UNSUPPORTED = -3,
  SFPE_ERR_NOMEM = -4,
  SFPE_ERR_TOO_LARGE = -5 /* sfpe_range_scan_static: H*W beyond SFPE_STATIC_MAX_CELLS */
} sfpe_status;

/* element types of point fields and float outputs */
enum { SFPE_F32 = 0, SFPE_F64 = 1 };

/* range-scan tensor layouts.  Channel order is (distance, height, intensity), the order
 * train/predict.py:339-341 and train/loader.py:197-207 feed the network. */
enum {
  SFPE_RANGE_CHW = 0, /* [frame][3][H][W]  -- three planar images, like the reference dict */
  SFPE_RANGE_HWC = 1  /* [frame][H][W][3]  -- the (32,1801,3) network input */
};

/* per-frame status bits: a non-zero value means the reference would have raised
 * IndexError for that sweep (extract_rosbag_lidar.py:71 fancy assignment). */
enum {
  SFPE_FRAME_ROW_OOB = 1, /* elevation cell outside [-H, H)  */
  SFPE_FRAME_COL_OOB = 2, /* azimuth cell outside [-W, W)    */
  SFPE_FRAME_NAN = 4      /* NaN cell coordinate (NaN input, or the origin) */
};

/* Grid / extent parameters.  sfpe_default_params() fills the reference's values. */
typedef struct {
  double res_v_deg;    /* RES[0]  = 1.33   process/globals.py:6 */
  double res_h_deg;    /* RES[1]  = 0.2                         */
  double vfov_lo_deg;  /* VFOV[0] = -30.67 process/globals.py:7 */
  double vfov_hi_deg;  /* VFOV[1] = 10.67                       */
  double height_bg;    /* LIDAR_MIN_HEIGHT = -2: background of the height image */
  double bev_range;    /* max_range = 120  birds_eye_view_generator.py:41 */
  double bev_res_x;    /* res[1]: step of the x (first histogram axis) bin edges */
  double bev_res_y;    /* res[0]: step of the y bin edges */
  int32_t enable_range; /* compute the range scan */
  int32_t enable_bev;   /* compute the BEV density */
  /* extensions, default off (parity unpinned: no reference code).  Both run with the fused
   * kernel when the batch qualifies for it (packed float32 points, float32 outputs): the extrinsic inside it, the
   * slices in kernels of their own after it (sfpe_last_path "fused+ext"); otherwise on the generic path. */
  int32_t use_extrinsic; /* apply p' = float32(M[3x4] * (x,y,z,1)) before projecting */
  int32_t n_slices;      /* >0: MV3D height-slice maps + intensity-of-highest-point */
  double extrinsic[12];
  double slice_edges[SFPE_MAX_SLICES + 1];
  /* tuning / test hooks */
  int32_t key_prefix_bits; /* 0 = auto.  Bits of the fp64 range kept in the packed winner key;
                              small values force the exact-compare slow path (tests). */
  int32_t path;            /* 0 = auto, 1 = generic global-scratch kernels, 2 = fused cluster kernel */
  int32_t ablation;        /* must be 0 (kernel ablation bits for timing experiments, -DSFPE_ABLATE builds only; results invalid) */
  int32_t firing_stride;   /* 0 = 32.  Points per firing (beams of the sensor: 32, 64, 128, 256) when the sweep is
                              stored firing by firing, as the Velodyne driver delivers it.  A HINT for the fused
                              kernel's point pass (neighbouring azimuths of a beam are merged per thread): any
                              value rasterises any cloud to the same result; other values are rounded down. */
  int32_t reserved[4];     /* must be 0 */
} sfpe_params;

/* Where the points of a batch live (device memory).  Point p of the batch is at
 * data + p*stride_bytes; its fields at the given byte offsets.  The packed hot-path
 * layout is float32 x,y,z,intensity: stride 16, offsets 0,4,8,12, 16-byte aligned. */
typedef struct {
  const void* data;
  int64_t stride_bytes;
  int32_t dtype; /* SFPE_F32 or SFPE_F64 */
  int32_t off_x, off_y, off_z, off_i;
  int32_t reserved;
} sfpe_points;

/* Output buffers (device).  Any pointer may be NULL = not wanted. */
typedef struct {
  void* range;            /* n_frames * 3*H*W elements of range_dtype, layout range_layout */
  int32_t range_dtype;    /* SFPE_F32 / SFPE_F64 */
  int32_t range_layout;   /* SFPE_RANGE_CHW / SFPE_RANGE_HWC */
  int32_t* range_winner;  /* [frame][H][W] index (within the frame) of the winning point, -1 = none */
  void* bev;              /* [frame][nbx][nby] density 255*min(ln(N+1)/ln64,1), both axes reversed */
  int32_t bev_dtype;      /* SFPE_F32 / SFPE_F64 */
  int32_t reserved0;
  uint32_t* bev_counts;   /* [frame][nbx][nby] raw counts, NOT reversed ([x_bin][y_bin]) */
  float* bev_height;      /* [frame][n_slices][nbx][nby], reversed like bev (extension) */
  float* bev_intensity;   /* [frame][nbx][nby] (extension) */
  int32_t* status;        /* [frame] OR of SFPE_FRAME_* bits */
  int32_t* dropped;       /* [frame] number of points that hit a status bit (not rasterised) */
  void* reserved[4];
} sfpe_outputs;

typedef struct sfpe_ctx sfpe_ctx;

SFPE_API int sfpe_abi_version(void);
SFPE_API const char* sfpe_strerror(int status);
SFPE_API void sfpe_default_params(sfpe_params* p);

/* Derived grid sizes for a parameter set (host arithmetic only; mirrors process/globals.py:9-13
 * and np.arange's length rule): H = Y_MAX+1, W = X_MAX+1, nbx/nby = BEV bins per axis. */
SFPE_API int sfpe_grid_dims(const sfpe_params* p, int32_t* H, int32_t* W, int32_t* nbx, int32_t* nby);

SFPE_API int sfpe_create(int device, const sfpe_params* p, sfpe_ctx** out);
SFPE_API void sfpe_destroy(sfpe_ctx* ctx);
SFPE_API const char* sfpe_last_error(const sfpe_ctx* ctx); /* text of the last CUDA failure */

/* Rasterise n_frames independent sweeps.  Frame f owns points
 * [frame_offsets[f], frame_offsets[f+1]) of `pts` (d_frame_offsets: DEVICE int64[n_frames+1]);
 * with d_frame_offsets == NULL every frame has exactly uniform_n points.  max_n bounds the
 * points of any frame (grid sizing; must be < 2^31).  Asynchronous on `stream`. */
SFPE_API int sfpe_project_batch(sfpe_ctx* ctx, const sfpe_points* pts, const int64_t* d_frame_offsets,
                       int64_t uniform_n, int64_t max_n, int32_t n_frames,
                       const sfpe_outputs* out, void* stream);

/* The range scan of sfpe_project_batch WITHOUT any allocation, for callers that may not call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in
 * the reference.  Needs no ctx: `device` and the
 * parameter block are passed per call (enable_bev / n_slices are ignored: range scan only; the BEV
 * members of `out` must be NULL).  Same point layouts, frame offsets, outputs (range, range_winner,
 * status, dropped) and results as sfpe_project_batch; it runs the generic kernels with their
 * winner-key scratch in a static device array of SFPE_STATIC_MAX_CELLS 64-bit keys (2 MiB, held by
 * every GPU that loads the library until the process exits).  A grid with H*W beyond that is refused
 * with SFPE_ERR_TOO_LARGE (the default 32 x 1801 grid uses 22 % of it; frames are processed
 * floor(SFPE_STATIC_MAX_CELLS / (H*W)) at a time).  Calls no cudaMalloc / cudaFree / allocating
 * routine; the first call per device looks the array's address up (cudaGetSymbolAddress), so make
 * one call before capturing a graph.  Leaves `device` current.  Asynchronous on `stream`.  All calls
 * on a device share the array: calls that overlap on the GPU (two streams, two threads) race. */
#define SFPE_STATIC_MAX_CELLS (1 << 18)
SFPE_API int sfpe_range_scan_static(int device, const sfpe_params* p, const sfpe_points* pts,
                           const int64_t* d_frame_offsets, int64_t uniform_n, int64_t max_n,
                           int32_t n_frames, const sfpe_outputs* out, void* stream);

/* One sweep, HOST in / HOST out (pinned or pageable): H2D, kernels, D2H, synchronised.
 * host_points has n points in the layout `layout` describes (layout->data ignored).
 * h_range: 3*H*W elements (range_dtype/range_layout), h_bev: nbx*nby elements, h_winner:
 * H*W int32 -- any may be NULL.  *frame_status receives the SFPE_FRAME_* mask. */
SFPE_API int sfpe_project_sweep_host(sfpe_ctx* ctx, const void* host_points, const sfpe_points* layout,
                            int64_t n, void* h_range, int32_t range_dtype, int32_t range_layout,
                            int32_t* h_winner, void* h_bev, int32_t bev_dtype,
                            int32_t* frame_status, int32_t* frame_dropped);

/* PointCloud2 `data` blob (device) -> packed float32 x,y,z,intensity (device, 16 B/point).
 * n = width*height points of point_step bytes; offsets/dtypes of the 4 fields as in
 * msg.fields (datatype 7 = FLOAT32 -> SFPE_F32, 8 = FLOAT64 -> SFPE_F64). */
SFPE_API int sfpe_decode_pointcloud2(sfpe_ctx* ctx, const void* d_blob, int64_t n, int32_t point_step,
                            int32_t off_x, int32_t off_y, int32_t off_z, int32_t off_i,
                            int32_t dtype, float* d_xyzi, void* stream);

/* Colormap stage (extract_rosbag_lidar.py:41,49-53,75; birds_eye_view_generator.py:48-53:
 * matplotlib ScalarMappable.to_rgba(img, bytes=True)[:,:,0:3]).  PARITY UNPINNED: the byte
 * arithmetic lives in matplotlib, which the reference neither vendors nor pins.  Definition
 * used here: t=(v-vmin)/(vmax-vmin) (autoscale: vmin/vmax = finite min/max of the image; a
 * flat image maps to t=0), index=min(trunc(t*256),255) clamped to [0,255], NaN -> (0,0,0);
 * SFPE_CMAP_GRAY: r=g=b=index; SFPE_CMAP_JET: matplotlib's piecewise-linear "jet" segments
 * sampled at index/255, byte = trunc(255*c).  d_img: n elements of `dtype`; d_rgb: 3*n bytes. */
enum { SFPE_CMAP_GRAY = 0, SFPE_CMAP_JET = 1 };
SFPE_API int sfpe_colormap(sfpe_ctx* ctx, const void* d_img, int32_t dtype, int64_t n, int32_t cmap,
                  int32_t autoscale, double vmin, double vmax, uint8_t* d_rgb, void* stream);
/* same, HOST in / HOST out, synchronised */
SFPE_API int sfpe_colormap_host(sfpe_ctx* ctx, const void* h_img, int32_t dtype, int64_t n, int32_t cmap,
                       int32_t autoscale, double vmin, double vmax, uint8_t* h_rgb);

/* ---- consumers of the range scan, kept on the device ---------------- */

/* Invariant-pixel mask (process/create_lidar_invariant_mask.py:44-62, update_invariant_pixels_new:
 * `np.logical_and(mask, reference == float_pixels, mask)` once per frame):
 *   d_mask[p] = 0 where d_frames[f*frame_stride + p*pix_stride] != d_ref[p*pix_stride] for any f
 * (NaN != NaN as in numpy); other mask bytes are left as they are, so the caller starts from 1s.
 * Strides are in ELEMENTS of `dtype`; pix_stride 1 with n_pix = 32*1801*3 walks an HWC frame whole,
 * pix_stride 3 walks one channel of it.  d_mask: n_pix bytes. */
SFPE_API int sfpe_invariant_mask_update(sfpe_ctx* ctx, const void* d_ref, const void* d_frames, int32_t dtype,
                               int64_t n_frames, int64_t frame_stride, int64_t n_pix, int32_t pix_stride,
                               uint8_t* d_mask, void* stream);

/* Roll augmentation (train/loader.py:31-57, data_random_rotate: `np.roll(image, rotate_by, axis=1)`
 * on (height, width, channels) images and on the reshaped labels):
 *   d_out[f][h][(w + d_shift[f]) mod width] = d_in[f][h][w], pixels of pix_bytes bytes (multiple of 4).
 * Out of place (the reference copies the rolled array back over the input). d_shift: device int32[n_frames]. */
SFPE_API int sfpe_roll_width(sfpe_ctx* ctx, const void* d_in, void* d_out, int64_t n_frames, int32_t height,
                    int32_t width, int32_t pix_bytes, const int32_t* d_shift, void* stream);

/* ---- compact transport of the BEV density image to a host consumer -----------------------------
 * The reference keeps every rendered image on the host (process/extract_rosbag.py:162-187).  The
 * density image of common/birds_eye_view_generator.py:31-37 takes 64 distinct values
 * (255*min(ln(N+1)/ln 64, 1), N = 0..63, and 255 beyond), so a batch can cross PCIe as one byte per
 * cell and be rebuilt on the host bit for bit:
 *   sfpe_density_lut      copies the table the kernels use (65 doubles, [64] = the saturated value);
 *   sfpe_bev_pack_u8      d_codes[i] = class of d_bev[i] (exact comparison against that table, cast to
 *                         the image's dtype); *d_flag (device int32, zeroed by the caller) is set to 1
 *                         if a value is not in the table.  d_bev 16-byte, d_codes 4-byte aligned;
 *   sfpe_bev_unpack_host  h_out[i] = (dtype) h_lut65[min(h_codes[i], 64)] on n_threads host threads
 *                         (<= 0: all cores), non-temporal stores.  Needs no context and no GPU;
 *   sfpe_bev_unpack_rgb_host  the picture the reference's BEV entry returns (uint8 [cell][3],
 *                         birds_eye_view_generator.py:48-54): h_out[3i..3i+2] = h_rgb65[3*min(code,64)..], the
 *                         colours of the 65 density values (sfpe_colormap of the table: parity unpinned).
 * Host threads come from one persistent pool inside the library (grown on demand, shared by concurrent calls). */
SFPE_API int sfpe_density_lut(sfpe_ctx* ctx, double* h_lut65);
SFPE_API int sfpe_bev_pack_u8(sfpe_ctx* ctx, const void* d_bev, int32_t dtype, int64_t n_cells, uint8_t* d_codes,
                     int32_t* d_flag, void* stream);
SFPE_API int sfpe_bev_unpack_host(const uint8_t* h_codes, int64_t n_cells, const double* h_lut65, void* h_out,
                         int32_t dtype, int32_t n_threads);
SFPE_API int sfpe_bev_unpack_rgb_host(const uint8_t* h_codes, int64_t n_cells, const uint8_t* h_rgb65, uint8_t* h_out,
                             int32_t n_threads);

/* Back-projection of cluster centroids (train/predict.py:230-293, back_project_2D_2_3D).  Per item i:
 * if the pixel under the centroid is not valid (distance > 0 and height > min_height) and the box
 * has non-zero left/right, the centroid moves to the valid pixel of the box nearest to it (first in
 * row-major order on ties), or centroid and box are zeroed when the box has none; then, unless the
 * centroid is (0,0), xyz = ((d+0.75) cos t, -(d+0.75) sin t, height, 0) with t = (cx + x_min)*res_rad.
 * d_centroids [n][2] (x=column, y=row) and d_bboxes [n][4] (left, top, right, bottom) are float64 and
 * updated in place like the reference's arrays; d_xyz [n][4] float64; d_status [n]: 0 kept, 1 moved,
 * 2 zeroed, 3 centroid/box outside the image (reference would raise or wrap; item left untouched).
 * d_distance / d_height: item i's image at + i*frame_stride, pixel (r,c) at (r*width+c)*pix_stride,
 * ELEMENTS of `dtype`.  Pixel choice is exact; x,y go through the device's fp64 cos/sin (<= 2 ulp). */
SFPE_API int sfpe_back_project(sfpe_ctx* ctx, double* d_centroids, double* d_bboxes, int64_t n,
                      const void* d_distance, const void* d_height, int32_t dtype, int64_t frame_stride,
                      int32_t pix_stride, int32_t height, int32_t width, double min_height, double x_min,
                      double res_rad, double* d_xyz, int32_t* d_status, void* stream);

/* Per-pixel box decode (train/predict.py:94-199, find_bbox_3d, with is_far :83-92).  Per item i:
 * d_pred is the network output [height*width][26] (background p, car p, 8 corners x 3 in the pixel's
 * rotated frame; `pred_dtype`); every pixel of the window [left-100, right+100) x [top-2, bottom+2)
 * (d_bbox2d[i] = left, top, right, bottom) whose column holds a positive (p >= 0.5) pixel anywhere and
 * whose row holds one anywhere is decoded to 8 corners (rot_z(theta) rot_y(phi) c' + p, fp64); sets
 * with a corner further than (9,3,3) m from d_centroid3d[i] are dropped; the sets with the most
 * neighbours (0 < Frobenius distance < 5 m) are averaged in window order -> d_out_box[i] (8x3);
 * d_out_pred[i] = x, y, z (mean corner), yaw, length, width, height.  d_n_boxes[i] = sets kept
 * (0: d_out_pred = 0 and no box, the reference's `None`; -1: more than `cap` sets, nothing written).
 * Scratch: d_boxes [n][cap][24] float64, d_counts [n][cap] int32.  Images addressed like
 * sfpe_back_project.  d_posbits: NULL, or the positive-pixel bits sfpe_find_obstacle wrote for the
 * same d_pred (saves the second pass over the tensor).  Decisions and sums follow the reference's order; the trigonometry and the
 * 3x3 products are the device's fp64 (test tolerance 1e-9 absolute on metres / radians). */
SFPE_API int sfpe_find_bbox_3d(sfpe_ctx* ctx, const void* d_distance, const void* d_height, int32_t dtype,
                      int64_t frame_stride, int32_t pix_stride, const void* d_pred, int32_t pred_dtype,
                      int64_t n, int32_t height, int32_t width, const int32_t* d_bbox2d,
                      const double* d_centroid3d, double x_min, double y_min, double res_rad_h,
                      double res_rad_v, double* d_boxes, int32_t* d_counts, int64_t cap,
                      double* d_out_pred, double* d_out_box, int32_t* d_n_boxes,
                      const uint32_t* d_posbits, void* stream);

/* Heat map + clustering of the network's positive pixels (train/predict.py:33-81, find_obstacle).
 * d_pred [n][height*width][n_channels] (`pred_dtype`), channel 1 = p(car).  Each pixel with p >= 0.5
 * and row, column >= 2 adds 1 to heatmap[y-2:y+2, x-2:x+2]; cells with heat > 2 are clustered with
 * 4-connectivity (scipy.ndimage.label's default); the cluster with the largest (max-min)*(max-min)
 * box wins, the first in raster order on ties.  d_out [n][8] int32: found (0 = the reference's
 * (None, None, None): area <= 100), centroid x, y, box left, top, right, bottom (shrunk by 2 as in
 * :76), area.  d_scratch: [n][5][height*width] int32.  d_posbits: NULL, or [n][ceil(height*width/32)]
 * words that receive one bit per pixel (p >= 0.5), for sfpe_find_bbox_3d.  Integer work: exact. */
SFPE_API int sfpe_find_obstacle(sfpe_ctx* ctx, const void* d_pred, int32_t pred_dtype, int32_t n_channels,
                       int64_t n, int32_t height, int32_t width, int32_t* d_scratch, int32_t* d_out,
                       uint32_t* d_posbits, void* stream);

/* Kernels launched by this ctx since creation (bench.py's gpu_launches). */
SFPE_API int64_t sfpe_launch_count(const sfpe_ctx* ctx);
/* Name of the code path the last sfpe_project_batch took ("generic", "fused", "fused+ext": the fused kernel
 * followed by the height-slice / intensity kernels of the extension). */
SFPE_API const char* sfpe_last_path(const sfpe_ctx* ctx);

/* Profiling hook (no reference counterpart; the reference's analogue is the cProfile toggle at
 * team_sf_rosnode/scripts/lidar_predict.py:21-23).  enable != 0: the fused kernel accumulates SM
 * cycles per phase (and the counters are reset); 0: off.  sfpe_phase_cycles synchronises the
 * device and returns the sums over all CTAs: out[0..SFPE_N_PHASES-1] = key init, point pass,
 * deferred fp64 pass, range-scan resolve, BEV tile compose, BEV exact redo, BEV line write-out,
 * exact settlement, wait for the store warp, BEV line list, write-out of the last frames by all
 * warps, unused. */
#define SFPE_N_PHASES 12
SFPE_API int sfpe_phase_profile(sfpe_ctx* ctx, int32_t enable);
/* Self-test of the fused kernel's fp32 fast path on n packed float32 x,y,z,intensity points
 * (device): out4[k] = largest deviation of the fp32 cell coordinate from its fp64 definition
 * divided by the guard band that protects it (k = 0 azimuth column, 1 elevation row, 2 range
 * bucket, 3 BEV bin).  A value < 1 means every fp32 decision taken was provably exact. */
SFPE_API int sfpe_selftest_fastpath(sfpe_ctx* ctx, const float* d_xyzi, int64_t n, double* out4);
SFPE_API int sfpe_phase_cycles(sfpe_ctx* ctx, int64_t* out /* [SFPE_N_PHASES] */);

#ifdef __cplusplus
}
#endif
#endif /* SFPE_H_ */
