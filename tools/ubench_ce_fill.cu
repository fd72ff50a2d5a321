// Can the zero fill of the BEV frames be taken off the SMs?  Measures, on one B200:
//  (1) cudaMemsetAsync / cudaMemcpyAsync(D2D from a zero buffer) throughput on an idle GPU;
//  (2) whether they make progress while a persistent kernel occupies every SM completely
//      (512 threads x 128 registers, 200 KB shared memory: no room for another CTA) -- i.e. whether
//      they run on the copy engines or as kernels; the persistent kernel polls a flag that is
//      written, stream-ordered, after the fill (bounded spin: it never hangs);
//  (3) the fill throughput next to a store-heavy persistent kernel.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_ce_fill tools/ubench_ce_fill.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

// occupies an SM completely; mode 0: spin on the flag; mode 1: also stream stores to `sink`
__global__ void __launch_bounds__(512, 1) k_occupy(volatile unsigned* flag, long long timeout_cycles, long long* waited,
                                                   int* seen, float4* sink, long long sink_words, int mode) {
  extern __shared__ unsigned smem[];
  float acc[112];   // force a full register file (>= 121 registers -> 128 allocated x 512 threads)
#pragma unroll
  for (int i = 0; i < 112; ++i) acc[i] = (float)(threadIdx.x + i);
  smem[threadIdx.x] = threadIdx.x;
  const long long t0 = clock64();
  int ok = 0;
  long long it = 0;
  while (clock64() - t0 < timeout_cycles) {
    if (*flag) { ok = 1; break; }
    if (mode == 1) {
      const float4 z = make_float4(acc[0], acc[1], acc[2], acc[3]);
      for (int u = 0; u < 16; ++u) {
        const long long w = (((long long)blockIdx.x * 512 + threadIdx.x) + (it * 16 + u) * (long long)gridDim.x * 512) % sink_words;
        __stcs(sink + w, z);
      }
      ++it;
    } else {
      __nanosleep(1000);
    }
#pragma unroll
    for (int i = 0; i < 112; ++i) acc[i] = acc[i] * 1.0001f + 0.5f;
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 112; ++i) s += acc[i];
  if (s == 1.2345f) smem[0] = 1;
  if (threadIdx.x == 0) { waited[blockIdx.x] = clock64() - t0; seen[blockIdx.x] = ok; }
}

int main() {
  const size_t bytes = 6ull << 30, zsrc_bytes = 64ull << 20;
  char *buf, *zsrc;
  float4* sink;
  unsigned *flag, *one;
  long long* waited;
  int* seen;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMalloc(&zsrc, zsrc_bytes));
  CK(cudaMemset(zsrc, 0, zsrc_bytes));
  const long long sink_words = (1ll << 30) / 16;
  CK(cudaMalloc(&sink, 1ull << 30));
  CK(cudaMalloc(&flag, 4)); CK(cudaMalloc(&one, 4));
  unsigned h1 = 1;
  CK(cudaMemcpy(one, &h1, 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&waited, 148 * 8)); CK(cudaMalloc(&seen, 148 * 4));
  cudaStream_t sa, sb;
  CK(cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&sb, cudaStreamNonBlocking));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float ms;
  // (1) idle throughput
  for (int rep = 0; rep < 2; ++rep) {
    CK(cudaEventRecord(e0, sb));
    CK(cudaMemsetAsync(buf, 0, bytes, sb));
    CK(cudaEventRecord(e1, sb));
    CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("idle   cudaMemsetAsync 6 GB: %.3f ms = %.0f GB/s\n", ms, bytes / ms / 1e6);
    CK(cudaEventRecord(e0, sb));
    for (size_t o = 0; o < bytes; o += zsrc_bytes) CK(cudaMemcpyAsync(buf + o, zsrc, zsrc_bytes, cudaMemcpyDeviceToDevice, sb));
    CK(cudaEventRecord(e1, sb));
    CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("idle   cudaMemcpyAsync D2D from a 64 MB zero buffer, 6 GB: %.3f ms = %.0f GB/s (written)\n", ms, bytes / ms / 1e6);
  }
  CK(cudaFuncSetAttribute(k_occupy, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  int nb = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_occupy, 512, 200 * 1024));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, k_occupy));
  printf("k_occupy: %d registers, %d CTA(s) per SM\n", fa.numRegs, nb);
  const long long timeout = 2000000000ll;   // ~1 s
  for (int mode = 0; mode < 2; ++mode)
    for (int kind = 0; kind < 2; ++kind) {
      CK(cudaMemset(flag, 0, 4));
      CK(cudaDeviceSynchronize());
      k_occupy<<<148, 512, 200 * 1024, sa>>>(flag, timeout, waited, seen, sink, sink_words, mode);
      CK(cudaEventRecord(e0, sb));
      if (kind == 0) CK(cudaMemsetAsync(buf, 0, bytes, sb));
      else for (size_t o = 0; o < bytes; o += zsrc_bytes) CK(cudaMemcpyAsync(buf + o, zsrc, zsrc_bytes, cudaMemcpyDeviceToDevice, sb));
      CK(cudaEventRecord(e1, sb));
      CK(cudaMemcpyAsync(flag, one, 4, cudaMemcpyDeviceToDevice, sb));
      CK(cudaDeviceSynchronize());
      CK(cudaEventElapsedTime(&ms, e0, e1));
      long long hw[148]; int hs[148];
      CK(cudaMemcpy(hw, waited, sizeof(hw), cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hs, seen, sizeof(hs), cudaMemcpyDeviceToHost));
      int nseen = 0; long long mx = 0;
      for (int i = 0; i < 148; ++i) { nseen += hs[i]; if (hw[i] > mx) mx = hw[i]; }
      printf("%s next to a full-SM persistent kernel (%s): fill %.3f ms = %.0f GB/s; %d/148 CTAs saw the flag, max wait %.3f ms\n",
             kind == 0 ? "cudaMemsetAsync" : "cudaMemcpyAsync D2D", mode == 0 ? "spinning" : "storing at full rate", ms, bytes / ms / 1e6,
             nseen, mx / 1.965e6);
    }
  return 0;
}
