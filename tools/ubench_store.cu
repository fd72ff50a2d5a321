// Micro-benchmarks that size the fused kernel's store/zero-fill design (run on the B200 box):
//   MB1  streaming-store bandwidth per SM vs number of CTAs (1 CTA of 1024 threads per SM)
//   MB2  does a store stream overlap with ALU work of the same warps?  (burst size sweep)
//   MB3  latency of dependent HBM reads on 4 SMs while the other SMs stream stores
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_store ubench_store.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("ERR %s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

// ---- MB1 / MB2 ------------------------------------------------------------------------
// each thread: `iters` rounds of [alu FMAs] + [burst float4 streaming stores]
__global__ void __launch_bounds__(1024, 1) k_mix(float4* out, size_t words_per_cta, int iters, int alu, int burst,
                                                 long long* cyc, float* sink) {
  extern __shared__ unsigned smem[];   // occupy the SM: one CTA per SM
  float4* o = out + (size_t)blockIdx.x * words_per_cta;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  float a0 = threadIdx.x, a1 = 1.f, a2 = 2.f, a3 = 3.f;
  size_t w = threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll 8
    for (int k = 0; k < alu; ++k) { a0 = fmaf(a0, 1.0001f, a1); a1 = fmaf(a1, 0.9999f, a2); a2 = fmaf(a2, 1.0002f, a3); a3 = fmaf(a3, 0.9998f, a0); }
    for (int b = 0; b < burst; ++b) { __stcs(o + (w % words_per_cta), z); w += 1024; }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (a0 + a1 + a2 + a3 == 123.456f) sink[0] = a0;
}

// ---- MB3 ------------------------------------------------------------------------------
// blocks 0..3: one warp chases pointers through a big (cold) array; others stream stores
__global__ void __launch_bounds__(1024, 1) k_lat(const unsigned* chase, int hops, float4* out, size_t words_per_cta,
                                                 int store_iters, long long* cyc, unsigned* sink) {
  extern __shared__ unsigned smem[];
  if (blockIdx.x < 4) {
    if (threadIdx.x == 0) {
      unsigned p = blockIdx.x * 977u;
      const long long t0 = clock64();
      for (int h = 0; h < hops; ++h) p = __ldcg(chase + p);
      const long long t1 = clock64();
      cyc[blockIdx.x] = (t1 - t0) / hops;
      sink[blockIdx.x] = p;
    }
    return;
  }
  float4* o = out + (size_t)blockIdx.x * words_per_cta;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  size_t w = threadIdx.x;
  for (int it = 0; it < store_iters; ++it) { __stcs(o + (w % words_per_cta), z); w += 1024; }
}

int main() {
  int sms = 0; CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  int khz = 0; CHECK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
  printf("SMs %d, clock %d kHz\n", sms, khz);
  const size_t words_per_cta = (size_t)8 << 20;  // 128 MB per CTA region (float4 words: 8 Mi * 16 B)
  float4* out; CHECK(cudaMalloc(&out, (size_t)sms * words_per_cta * 16));
  long long* cyc; CHECK(cudaMalloc(&cyc, 1024 * 8));
  float* sink; CHECK(cudaMalloc(&sink, 64));
  const int SM = 200 * 1024;
  CHECK(cudaFuncSetAttribute(k_mix, cudaFuncAttributeMaxDynamicSharedMemorySize, SM));
  CHECK(cudaFuncSetAttribute(k_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, SM));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  long long h[1024];
  auto run = [&](int grid, int iters, int alu, int burst, const char* tag) {
    k_mix<<<grid, 1024, SM>>>(out, words_per_cta, 8, alu, burst, cyc, sink);  // warm
    CHECK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    k_mix<<<grid, 1024, SM>>>(out, words_per_cta, iters, alu, burst, cyc, sink);
    cudaEventRecord(e1);
    CHECK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    CHECK(cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost));
    double avg = 0; for (int i = 0; i < grid; ++i) avg += h[i]; avg /= grid;
    const double bytes = (double)grid * 1024 * 16 * iters * burst;
    printf("%-28s grid %3d alu %4d burst %3d: %8.3f ms  %7.1f GB/s  cycles/CTA %10.0f  B/clk/SM %6.2f  cyc/iter %8.1f\n", tag, grid,
           alu, burst, ms, bytes / ms / 1e6, avg, bytes / grid / avg, avg / iters);
  };
  printf("== MB1 store bandwidth vs active SMs\n");
  for (int grid : {1, 4, 8, 16, 37, 74, 111, 148}) run(grid, 4096, 0, 4, "stores only");
  printf("== MB2 overlap of stores with ALU work (148 CTAs)\n");
  run(148, 1024, 256, 0, "alu only (256 x4 fma)");
  run(148, 1024, 0, 8, "stores only (8)");
  run(148, 1024, 256, 8, "alu 256 + burst 8");
  run(148, 2048, 128, 4, "alu 128 + burst 4");
  run(148, 4096, 64, 2, "alu 64 + burst 2");
  run(148, 8192, 32, 1, "alu 32 + burst 1");
  run(148, 256, 1024, 32, "alu 1024 + burst 32");
  printf("== MB2b same on 16 CTAs (per-SM path, HBM not saturated)\n");
  run(16, 1024, 256, 0, "alu only");
  run(16, 1024, 0, 8, "stores only (8)");
  run(16, 1024, 256, 8, "alu 256 + burst 8");
  run(16, 8192, 32, 1, "alu 32 + burst 1");
  run(16, 256, 1024, 32, "alu 1024 + burst 32");
  run(16, 1024, 512, 8, "alu 512 + burst 8");
  printf("== MB3 dependent-read latency on 4 SMs vs store load on the rest\n");
  const size_t chase_n = (size_t)256 << 20;  // 1 GB of u32
  unsigned* chase; CHECK(cudaMalloc(&chase, chase_n * 4));
  {
    unsigned* hc = (unsigned*)malloc(chase_n * 4);
    unsigned long long s = 88172645463325252ull;
    for (size_t i = 0; i < chase_n; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; hc[i] = (unsigned)(s % chase_n); }
    CHECK(cudaMemcpy(chase, hc, chase_n * 4, cudaMemcpyHostToDevice));
    free(hc);
  }
  unsigned* usink; CHECK(cudaMalloc(&usink, 64));
  for (int grid : {4, 20, 41, 78, 148}) {
    for (int rep = 0; rep < 2; ++rep) {
      k_lat<<<grid, 1024, SM>>>(chase, 2000, out, words_per_cta, 1 << 16, cyc, usink);
      CHECK(cudaDeviceSynchronize());
    }
    CHECK(cudaMemcpy(h, cyc, 4 * 8, cudaMemcpyDeviceToHost));
    printf("store CTAs %3d: dependent HBM read latency %lld %lld %lld %lld cycles\n", grid - 4, h[0], h[1], h[2], h[3]);
  }
  return 0;
}
