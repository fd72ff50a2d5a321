// Micro-benchmarks for the background zero fill of the fused kernel (run on the B200 box):
//   MB4  dedicated store warps (STG.128 stream) next to compute warps (FMA + LDS + LDG mix):
//        how much do the compute warps slow down, what store rate is reached?
//   MB5  TMA bulk stores (cp.async.bulk shared -> global) from a small zero buffer: bytes/clk/SM
//        vs chunk size and number of issuing threads; and the same next to the compute mix.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_zerofill ubench_zerofill.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("ERR %s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ void bulk_store(void* gdst, unsigned ssrc, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// mode bits: 1 = compute warps active, 2 = STG store warps active, 4 = TMA zero fill active
// warps [0, n_store) are store warps (mode&2) ; warp 0 lanes [0, n_tma) issue bulk stores (mode&4);
// the remaining warps do `iters` rounds of: 64x4 FMA, 4 LDS, 1 LDG.128 (coalesced, streaming), 1 smem atomicAdd
__global__ void __launch_bounds__(1024, 1) k_bg(int mode, int n_store, int n_tma, unsigned chunk, int iters,
                                                float4* zout, size_t zwords_per_cta, const float4* in, size_t in_words_per_cta,
                                                long long* cyc, long long* zbytes, float* sink) {
  extern __shared__ __align__(128) unsigned smem[];
  unsigned* sdata = smem + 4096;          // 16 KB zero buffer first, then 32 KB of data
  __shared__ volatile int done;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 4096 + 8192; i += 1024) smem[i] = i < 4096 ? 0u : i;
  if (tid == 0) done = 0;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  float4* zo = zout + (size_t)blockIdx.x * zwords_per_cta;
  const bool is_store = (mode & 2) && warp < n_store;
  const bool is_tma = (mode & 4) && warp == 0 && lane < n_tma && !is_store;
  const int first_compute = ((mode & 2) ? n_store : 0) > ((mode & 4) ? 1 : 0) ? n_store : ((mode & 4) ? 1 : 0);
  const long long t0 = clock64();
  if (is_store) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    size_t w = (size_t)warp * 32 + lane;
    long long nb = 0;
    const int fixed = (mode & 1) ? 0 : iters * 64;
    float4* p = zo + w;
    const size_t step = (size_t)n_store * 32;
    for (int k = 0; fixed ? k < fixed / 8 : !done; ++k) {
#pragma unroll
      for (int u = 0; u < 8; ++u) __stcs(p + u * step, z);
      p += 8 * step; nb += 128;
      if (p >= zo + zwords_per_cta - 8 * step) p = zo + w;
    }
    if (lane == 0) atomicAdd((unsigned long long*)&zbytes[blockIdx.x], (unsigned long long)nb * 32);
  } else if ((mode & 4) && warp == 0) {
    if (is_tma) {
      const unsigned zs = (unsigned)__cvta_generic_to_shared(smem);
      char* dst = (char*)zo + (size_t)lane * chunk;
      const size_t span = zwords_per_cta * 16;
      size_t off = 0;
      long long nb = 0;
      const int fixed = (mode & 1) ? 0 : iters * 16;
      for (int k = 0; fixed ? k < fixed : !done; ++k) {
        bulk_store(dst + (off & (span / 2 - 1)), zs, chunk);
        bulk_commit();
        bulk_wait_read<8>();
        off += (size_t)n_tma * chunk; nb += chunk;
      }
      bulk_wait<0>();
      atomicAdd((unsigned long long*)&zbytes[blockIdx.x], (unsigned long long)nb);
    }
  } else if (mode & 1) {
    float4* ip = (float4*)in + (size_t)blockIdx.x * in_words_per_cta;
    float a0 = tid, a1 = 1.f, a2 = 2.f, a3 = 3.f;
    size_t w = tid;
    for (int it = 0; it < iters; ++it) {
      const float4 v = __ldcg(ip + (w & (in_words_per_cta - 1))); w += 1024;
#pragma unroll 8
      for (int k = 0; k < 64; ++k) { a0 = fmaf(a0, 1.0001f, a1); a1 = fmaf(a1, 0.9999f, a2); a2 = fmaf(a2, 1.0002f, a3); a3 = fmaf(a3, 0.9998f, a0); }
      const unsigned h = (unsigned)(tid * 2654435761u + it * 40503u);
      a0 += __uint_as_float(sdata[h & 8191]) * 1e-30f + __uint_as_float(sdata[(h >> 5) & 8191]) * 1e-30f;
      atomicAdd(&sdata[(h >> 11) & 8191], 1u);
      a1 += v.x + v.y + v.z + v.w;
      ((unsigned*)(ip))[(w & (in_words_per_cta - 1))] = h;
    }
    if (a0 + a1 + a2 + a3 == 123.456f) sink[0] = a0;
    // the compute warps meet on a named barrier, then release the background warps
    const int ncomp = 1024 - first_compute * 32;
    asm volatile("bar.sync 1, %0;" ::"r"(ncomp) : "memory");
    if (tid == first_compute * 32) done = 1;
  }
  __syncthreads();
  const long long t1 = clock64();
  if (tid == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  int sms = 0; CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t zwords = (size_t)8 << 20, inwords = (size_t)2 << 20;   // 128 MB out, 32 MB in per CTA
  float4 *zout, *in; CHECK(cudaMalloc(&zout, sms * zwords * 16)); CHECK(cudaMalloc(&in, sms * inwords * 16));
  CHECK(cudaMemset(in, 0, sms * inwords * 16));
  long long *cyc, *zb; CHECK(cudaMalloc(&cyc, 1024 * 8)); CHECK(cudaMalloc(&zb, 1024 * 8));
  float* sink; CHECK(cudaMalloc(&sink, 64));
  const int SM = 200 * 1024;
  CHECK(cudaFuncSetAttribute(k_bg, cudaFuncAttributeMaxDynamicSharedMemorySize, SM));
  long long h[1024], hz[1024];
  auto run = [&](const char* tag, int grid, int mode, int n_store, int n_tma, unsigned chunk, int iters) {
    for (int rep = 0; rep < 2; ++rep) {
      CHECK(cudaMemset(zb, 0, 1024 * 8));
      k_bg<<<grid, 1024, SM>>>(mode, n_store, n_tma, chunk, iters, zout, zwords, in, inwords, cyc, zb, sink);
      CHECK(cudaDeviceSynchronize());
    }
    CHECK(cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost));
    CHECK(cudaMemcpy(hz, zb, grid * 8, cudaMemcpyDeviceToHost));
    double c = 0, b = 0; for (int i = 0; i < grid; ++i) { c += h[i]; b += hz[i]; } c /= grid; b /= grid;
    printf("%-34s grid %3d store_warps %2d tma_thr %2d chunk %5u: cycles/CTA %10.0f  cyc/iter %7.1f  fill B/clk/SM %6.2f\n", tag, grid,
           (mode & 2) ? n_store : 0, (mode & 4) ? n_tma : 0, chunk, c, c / iters, b / c);
  };
  const int IT = 2048;
  for (int grid : {16, 148}) {
    printf("== grid %d\n", grid);
    run("compute only (32 warps)", grid, 1, 0, 0, 0, IT);
    run("STG only, 1 store warp", grid, 2, 1, 0, 0, IT);
    run("STG only, 2 store warps", grid, 2, 2, 0, 0, IT);
    run("compute + 1 STG warp", grid, 3, 1, 0, 0, IT);
    run("compute + 2 STG warps", grid, 3, 2, 0, 0, IT);
    run("compute + 4 STG warps", grid, 3, 4, 0, 0, IT);
    for (int thr : {2, 4, 8, 16, 32})
      for (unsigned chunk : {256u, 512u, 1024u, 1536u, 2048u}) run("TMA only", grid, 4, 0, thr, chunk, IT);
    for (int thr : {8, 32})
      for (unsigned chunk : {512u, 1024u, 1536u}) run("compute + TMA", grid, 5, 0, thr, chunk, IT);
  }
  return 0;
}
