// Consumers of the range scan that stay on the device:
//   k_find_obstacle train/predict.py:33-81                heat map of positive pixels, 4-connected clusters, largest box
//   k_find_bbox_3d  train/predict.py:94-199               per-pixel decode of the 3-D box corners + neighbour vote
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


// ---- scan of the network output for positive pixels (p(car) >= MIN_PROB, predict.py:28) ---------------------
// y_pred is [pixel][n_ch]; only channel 1 matters, but a 4-byte read at a 4*n_ch-byte stride still drags every
// line of the tensor out of DRAM (ncu: dram bytes read = the whole tensor).  Eight strided loads in flight per
// thread.  Streaming the whole frame with coalesced 16-byte loads and picking channel 1 out of the registers was
// tried and is slower (0.91 vs 0.59 ms for 296 frames): same DRAM traffic, but 26x the elements to test.
template <int NT, typename TP, typename F>
__device__ __forceinline__ void scan_positive_pixels(const TP* __restrict__ yp, int N, int n_ch, int tid, F&& hit) {
  for (int p0 = tid; p0 < N; p0 += 8 * NT) {
    TP v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int p = p0 + u * NT;
      v[u] = p < N ? __ldcs(yp + (long long)p * n_ch + 1) : (TP)0;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if ((double)v[u] >= 0.5) hit(p0 + u * NT);
  }
}

// ---- per-pixel box decode (train/predict.py:94-199, find_bbox_3d) -----------------------------------------
// One CTA per item.  (A) which rows / columns of the image hold a positive pixel anywhere (the reference's
// `img_x in pos[1] and img_y in pos[0]`); (B) walk the window column by column, decode the 8 corners of every
// participating pixel in fp64, drop the sets with a corner far from the centroid, and compact the rest IN ORDER
// into `boxes` (block-wide exclusive scan per 256 window pixels); (C) neighbour counts, all pairs;
// (D) mean of the sets with the highest count, accumulated in index order like numpy's axis-0 reduction;
// (E) one thread derives position, yaw and size.
constexpr int BOX_THREADS = 256;

template <typename T, typename TP>
__global__ void __launch_bounds__(BOX_THREADS) k_find_bbox_3d(
    const T* __restrict__ distance, const T* __restrict__ height, long long frame_stride, int pix_stride,
    const TP* __restrict__ y_pred, int H, int W, const int* __restrict__ bbox2d, const double* __restrict__ centroid3d,
    double x_min, double y_min, double res_h, double res_v, double* __restrict__ boxes_all, int* __restrict__ counts_all,
    long long cap, double* __restrict__ out_pred, double* __restrict__ out_box, int* __restrict__ n_boxes,
    const unsigned* __restrict__ posbits_in) {
  extern __shared__ unsigned char flags[];   // [W] column flags, then [H] row flags
  __shared__ int s_warp[BOX_THREADS / 32];
  __shared__ int s_base, s_max;
  __shared__ double s_cand[24];
  const long long item = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const T* dist = distance + item * frame_stride;
  const T* hgt = height + item * frame_stride;
  const TP* yp = y_pred + item * (long long)H * W * 26;
  double* boxes = boxes_all + item * cap * 24;
  int* counts = counts_all + item * cap;
  unsigned char* colflag = flags;
  unsigned char* rowflag = flags + W;
  for (int k = tid; k < W + H; k += BOX_THREADS) flags[k] = 0;
  if (tid == 0) { s_base = 0; s_max = 0; }
  __syncthreads();
  auto mark = [&](int p) {   // predict.py:96
    const int r = p / W;
    colflag[p - r * W] = 1;
    rowflag[r] = 1;
  };
  if (posbits_in) {   // positive bits left by k_find_obstacle for the same tensor
    const int nw = (H * W + 31) / 32;
    for (int k = tid; k < nw; k += BOX_THREADS)
      for (unsigned b = posbits_in[item * nw + k]; b; b &= b - 1) mark(32 * k + __ffs(b) - 1);
  } else {
    scan_positive_pixels<BOX_THREADS>(yp, H * W, 26, tid, mark);
  }
  __syncthreads();
  const int x0 = max(bbox2d[4 * item] - 100, 0), x1 = min(bbox2d[4 * item + 2] + 100, W);
  const int y0 = max(bbox2d[4 * item + 1] - 2, 0), y1 = min(bbox2d[4 * item + 3] + 2, H);
  const int wh = max(y1 - y0, 0), n_win = max(x1 - x0, 0) * wh;
  const double c3x = centroid3d[3 * item], c3y = centroid3d[3 * item + 1], c3z = centroid3d[3 * item + 2];
  bool overflow = false;
  for (int w0 = 0; w0 < n_win; w0 += BOX_THREADS) {
    const int k = w0 + tid;
    bool keep = false;
    double c[24];
    if (k < n_win) {
      const int wx = k / wh, x = x0 + wx, y = y0 + (k - wx * wh);
      if (colflag[x] && rowflag[y]) {
        const long long at = ((long long)y * W + x) * pix_stride;
        const double d = (double)dist[at], z = (double)hgt[at];
        const double theta = (x + x_min) * res_h, phi = (y + y_min) * res_v;
        double st, ct, sp, cp;
        sincos(theta, &st, &ct);
        sincos(phi, &sp, &cp);
        const double px = d * ct, py = -d * st;
        // rot_z . rot_y (predict.py:119-128)
        const double r00 = ct * cp, r01 = -st, r02 = ct * sp, r10 = st * cp, r11 = ct, r12 = st * sp, r20 = -sp, r22 = cp;
        const TP* q = yp + ((long long)y * W + x) * 26 + 2;
        keep = true;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const double a = (double)q[3 * j], b = (double)q[3 * j + 1], e = (double)q[3 * j + 2];
          const double cx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(r00, a), __dmul_rn(r01, b)), __dmul_rn(r02, e)), px);
          const double cy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(r10, a), __dmul_rn(r11, b)), __dmul_rn(r12, e)), py);
          const double cz = __dadd_rn(__dadd_rn(__dmul_rn(r20, a), __dmul_rn(r22, e)), z);
          c[3 * j] = cx; c[3 * j + 1] = cy; c[3 * j + 2] = cz;
          // is_far (predict.py:83-92): delta = 9, 3, 3
          if (cx > c3x + 9.0 || cx < c3x - 9.0 || cy > c3y + 3.0 || cy < c3y - 3.0 || cz > c3z + 3.0 || cz < c3z - 3.0) keep = false;
        }
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[wid] = __popc(bal);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < wid; ++w) before += s_warp[w];
    const int slot = before + __popc(bal & ((1u << lane) - 1));
    if (keep) {
      if (slot < cap) {
#pragma unroll
        for (int j = 0; j < 24; ++j) boxes[(long long)slot * 24 + j] = c[j];
      } else {
        overflow = true;
      }
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < BOX_THREADS / 32; ++w) tot += s_warp[w];
      s_base += tot;
    }
    __syncthreads();
  }
  const int K = s_base;
  if (__syncthreads_or(overflow) || K > cap) {
    if (tid == 0) n_boxes[item] = -1;
    return;
  }
  if (K == 0) {   // predict.py:137-138: zeros(7), None
    if (tid < 7) out_pred[7 * item + tid] = 0.0;
    if (tid < 24) out_box[24 * item + tid] = 0.0;
    if (tid == 0) n_boxes[item] = 0;
    return;
  }
  // (C) neighbour counts: 0 < ||c - c2|| < MAX_BBX_DIST (predict.py:146-151)
  for (int i = tid; i < K; i += BOX_THREADS) {
    double mine[24];
#pragma unroll
    for (int j = 0; j < 24; ++j) mine[j] = boxes[(long long)i * 24 + j];
    int cnt = 0;
    for (int o = 0; o < K; ++o) {
      double ssq = 0.0;
#pragma unroll
      for (int j = 0; j < 24; ++j) {
        const double df = mine[j] - boxes[(long long)o * 24 + j];
        ssq = __dadd_rn(ssq, __dmul_rn(df, df));
      }
      const double dd = sqrt(ssq);
      cnt += (dd > 0.0 && dd < 5.0) ? 1 : 0;
    }
    counts[i] = cnt;
    atomicMax(&s_max, cnt);
  }
  __syncthreads();
  // (D) mean over the sets with the highest count, summed in index order (predict.py:159-164)
  if (tid < 24) {
    const int mx = s_max;
    double acc = 0.0;
    int m = 0;
    for (int i = 0; i < K; ++i)
      if (counts[i] == mx) { acc = m ? acc + boxes[(long long)i * 24 + tid] : boxes[(long long)i * 24 + tid]; ++m; }
    acc /= (double)m;
    s_cand[tid] = acc;
    out_box[24 * item + tid] = acc;
  }
  __syncthreads();
  if (tid != 0) return;
  // (E) predict.py:166-197
  double* pr = out_pred + 7 * item;
  for (int a = 0; a < 3; ++a) {
    double acc = s_cand[a];
    for (int j = 1; j < 8; ++j) acc += s_cand[3 * j + a];
    pr[a] = acc / 8.0;
  }
  double yaw_s = 0.0, l_s = 0.0, w_s = 0.0, h_s = 0.0;
  const double half_pi = 3.141592653589793 / 2.0;
  for (int i = 0; i < 4; ++i) {
    double dx = s_cand[3 * i] - s_cand[3 * (i + 4)], dy = s_cand[3 * i + 1] - s_cand[3 * (i + 4) + 1];
    const double yaw = atan2(dy, dx);
    const double cyaw = cos(yaw);
    const double bl = yaw != half_pi ? dx / cyaw : dy;
    dx = s_cand[3 * i] - s_cand[3 * (i + 2)];
    dy = s_cand[3 * i + 1] - s_cand[3 * (i + 2) + 1];
    const double bw = yaw != half_pi ? dy / cyaw : dx;
    const double bh = s_cand[3 * i + 2] - s_cand[3 * (i + 1) + 2];
    yaw_s = i ? yaw_s + yaw : yaw;
    l_s = i ? l_s + fabs(bl) : fabs(bl);
    w_s = i ? w_s + fabs(bw) : fabs(bw);
    h_s = i ? h_s + fabs(bh) : fabs(bh);
  }
  pr[3] = yaw_s / 4.0; pr[4] = l_s / 4.0; pr[5] = w_s / 4.0; pr[6] = h_s / 4.0;
  n_boxes[item] = K;
}


// ---- heat map + clustering (train/predict.py:33-81, find_obstacle) ------------------------------------------
// One CTA per frame.  Positive pixels (p >= 0.5) each add 1 to the 4x4 block heatmap[y-2:y+2, x-2:x+2]; a python
// slice that starts below 0 is empty, so positives with y < 2 or x < 2 add nothing.  Cells with heat > 2 are
// clustered with 4-connectivity (scipy.ndimage.label's default): union-find over `labels` in global scratch, root =
// smallest pixel index = first pixel in raster order, which is also scipy's label order.  The cluster with the
// largest (max-min) x (max-min) box wins, the earliest label on ties (`area > max_area`).
constexpr int OBS_THREADS = 1024;

__device__ __forceinline__ int uf_find(const int* lab, int p) {
  int q = ((volatile const int*)lab)[p];
  while (q != p) { p = q; q = ((volatile const int*)lab)[p]; }
  return p;
}
__device__ __forceinline__ void uf_union(int* lab, int a, int b) {
  for (;;) {
    a = uf_find(lab, a);
    b = uf_find(lab, b);
    if (a == b) return;
    if (a < b) { const int t = a; a = b; b = t; }     // a > b: hang a under b
    const int old = atomicMin(lab + a, b);
    if (old == a) return;
    a = old;                                            // somebody else re-rooted a meanwhile: merge that root too
  }
}

template <typename TP>
__global__ void __launch_bounds__(OBS_THREADS, 2) k_find_obstacle(const TP* __restrict__ y_pred, int n_ch, int H, int W,
                                                               int* __restrict__ scratch_all, int* __restrict__ out,
                                                               unsigned* __restrict__ posbits_out) {
  extern __shared__ unsigned char sm[];   // [ceil(H*W/32)] words of positive bits, then [H*W] foreground flags
  __shared__ unsigned long long s_best;
  const long long item = blockIdx.x;
  const int tid = threadIdx.x, N = H * W;
  const TP* yp = y_pred + item * (long long)N * n_ch;
  unsigned* posbits = reinterpret_cast<unsigned*>(sm);   // two frames per SM fit this way: one loads while one clusters
  unsigned char* fg = sm + 4 * ((N + 31) / 32);
  int* lab = scratch_all + item * 5ll * N;
  int* minx = lab + N; int* maxx = minx + N; int* miny = maxx + N; int* maxy = miny + N;
  if (tid == 0) s_best = 0ull;
  for (int k = tid; k < (N + 31) / 32; k += OBS_THREADS) posbits[k] = 0u;
  __syncthreads();
  scan_positive_pixels<OBS_THREADS>(yp, N, n_ch, tid, [&](int p) { atomicOr(posbits + (p >> 5), 1u << (p & 31)); });
  __syncthreads();
  if (posbits_out)   // find_bbox_3d needs the same bits: hand them over instead of a second pass over the tensor
    for (int k = tid; k < (N + 31) / 32; k += OBS_THREADS) posbits_out[item * ((N + 31) / 32) + k] = posbits[k];
  for (int p = tid; p < N; p += OBS_THREADS) {
    const int r = p / W, c = p - r * W;
    int heat = 0;
    for (int y = max(r - 1, 2); y <= min(r + 2, H - 1); ++y)
      for (int x = max(c - 1, 2); x <= min(c + 2, W - 1); ++x) {
        const int q = y * W + x;
        heat += (posbits[q >> 5] >> (q & 31)) & 1u;
      }
    const bool f = heat > 2;                            // MIN_HEAT, predict.py:30,52
    fg[p] = f;
    if (f) {   // scratch is only ever read at foreground pixels
      lab[p] = p;
      minx[p] = W; maxx[p] = -1; miny[p] = H; maxy[p] = -1;
    }
  }
  __syncthreads();
  for (int p = tid; p < N; p += OBS_THREADS)
    if (fg[p]) {
      const int r = p / W, c = p - r * W;
      if (c > 0 && fg[p - 1]) uf_union(lab, p, p - 1);
      if (r > 0 && fg[p - W]) uf_union(lab, p, p - W);
    }
  __syncthreads();
  for (int p = tid; p < N; p += OBS_THREADS)
    if (fg[p]) {
      const int root = uf_find(lab, p), r = p / W, c = p - r * W;
      atomicMin(minx + root, c); atomicMax(maxx + root, c);
      atomicMin(miny + root, r); atomicMax(maxy + root, r);
    }
  __syncthreads();
  for (int p = tid; p < N; p += OBS_THREADS)
    if (fg[p] && ((volatile int*)lab)[p] == p) {
      const unsigned long long area = (unsigned long long)(maxx[p] - minx[p]) * (unsigned long long)(maxy[p] - miny[p]);
      if (area > 0) atomicMax(&s_best, (area << 32) | (unsigned)(0x7fffffff - p));
    }
  __syncthreads();
  if (tid != 0) return;
  int* o = out + 8 * item;   // found, centroid x, y, left, top, right, bottom, area
  const long long area = (long long)(s_best >> 32);
  if (area <= 100) {         // MIN_BBOX_AREA, predict.py:29,73
    for (int k = 0; k < 8; ++k) o[k] = 0;
    return;
  }
  const int root = 0x7fffffff - (int)(unsigned)(s_best & 0xffffffffu);
  const int l = minx[root] + 2, t = miny[root] + 2, r = maxx[root] - 2, b = maxy[root] - 2;   // predict.py:76
  o[0] = 1;
  o[1] = (int)((l + r) / 2.0);
  o[2] = (int)((t + b) / 2.0);
  o[3] = l; o[4] = t; o[5] = r; o[6] = b;
  o[7] = (int)area;
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

cudaError_t launch_find_bbox_3d(const void* distance, const void* height, int dtype, long long frame_stride, int pix_stride,
                                const void* y_pred, int pred_dtype, long long n, int H, int W, const int* bbox2d,
                                const double* centroid3d, double x_min, double y_min, double res_h, double res_v,
                                double* boxes, int* counts, long long cap, double* out_pred, double* out_box, int* n_boxes,
                                const unsigned* posbits, cudaStream_t st, long long* launches) {
  if (n <= 0) return cudaSuccess;
  const size_t smem = (size_t)W + H;
#define SFPE_BOX_LAUNCH(T, TP)                                                                                          \
  k_find_bbox_3d<T, TP><<<(unsigned)n, BOX_THREADS, smem, st>>>((const T*)distance, (const T*)height, frame_stride,       \
      pix_stride, (const TP*)y_pred, H, W, bbox2d, centroid3d, x_min, y_min, res_h, res_v, boxes, counts, cap, out_pred,  \
      out_box, n_boxes, posbits)
  if (dtype == SFPE_F32 && pred_dtype == SFPE_F32) SFPE_BOX_LAUNCH(float, float);
  else if (dtype == SFPE_F32) SFPE_BOX_LAUNCH(float, double);
  else if (pred_dtype == SFPE_F32) SFPE_BOX_LAUNCH(double, float);
  else SFPE_BOX_LAUNCH(double, double);
#undef SFPE_BOX_LAUNCH
  ++*launches;
  return cudaGetLastError();
}

cudaError_t launch_find_obstacle(const void* y_pred, int pred_dtype, int n_ch, long long n, int H, int W, int* scratch,
                                 int* out, unsigned* posbits_out, cudaStream_t st, long long* launches) {
  if (n <= 0) return cudaSuccess;
  const size_t smem = 4 * (((size_t)H * W + 31) / 32) + (size_t)H * W;
  cudaError_t e;
  if (pred_dtype == SFPE_F32) {
    e = cudaFuncSetAttribute(k_find_obstacle<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_find_obstacle<float><<<(unsigned)n, OBS_THREADS, smem, st>>>((const float*)y_pred, n_ch, H, W, scratch, out, posbits_out);
  } else {
    e = cudaFuncSetAttribute(k_find_obstacle<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_find_obstacle<double><<<(unsigned)n, OBS_THREADS, smem, st>>>((const double*)y_pred, n_ch, H, W, scratch, out, posbits_out);
  }
  ++*launches;
  return cudaGetLastError();
}

}  // namespace sfpe
