// Micro-benchmarks that size the fused kernel's design (run on the B200 box):
//   remote DSMEM red/atom throughput, local smem atomics, global 64-bit atomicMin, store bandwidth.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_dsmem ubench_dsmem.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("ERR %s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ unsigned hash(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

__device__ __forceinline__ unsigned mapa(unsigned laddr, unsigned rank) {
  unsigned r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(laddr), "r"(rank)); return r;
}
__device__ __forceinline__ void red_add_cluster(unsigned raddr, unsigned v) {
  asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" ::"r"(raddr), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned atom_add_cluster(unsigned raddr, unsigned v) {
  unsigned o; asm volatile("atom.relaxed.cluster.shared::cluster.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(raddr), "r"(v) : "memory"); return o;
}

// mode 0: remote red (no return), 1: remote atom (return), 2: local smem atomicAdd (return), 3: local red
template <int MODE>
__global__ void __launch_bounds__(1024) k_dsmem(int words, int ops, unsigned* sink, long long* cyc) {
  extern __shared__ unsigned tile[];
  cg::cluster_group cl = cg::this_cluster();
  const unsigned csz = cl.num_blocks();
  for (int i = threadIdx.x; i < words; i += blockDim.x) tile[i] = 0;
  cl.sync();
  const unsigned base = (unsigned)__cvta_generic_to_shared(tile);
  unsigned acc = 0;
  unsigned h = hash(blockIdx.x * 1024 + threadIdx.x + 12345);
  const long long t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < ops; ++i) {
    h = hash(h + i);
    const unsigned w = h % (unsigned)words;
    const unsigned rk = (h >> 20) % csz;
    if (MODE == 0) red_add_cluster(mapa(base + 4 * w, rk), 1u << (8 * (h & 3)));
    if (MODE == 1) acc += atom_add_cluster(mapa(base + 4 * w, rk), 1u << (8 * (h & 3)));
    if (MODE == 2) acc += atomicAdd(&tile[w], 1u << (8 * (h & 3)));
    if (MODE == 3) atomicAdd(&tile[w], 1u << (8 * (h & 3)));
  }
  cl.sync();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  for (int i = threadIdx.x; i < words; i += blockDim.x) acc += tile[i];
  if (acc == 0xdeadbeef) sink[0] = acc;
}

// global 64-bit atomicMin with return on a small (L2 resident) per-cluster region
__global__ void __launch_bounds__(1024) k_gmin(unsigned long long* keys, int cells, int region_blocks, int ops, unsigned* sink, long long* cyc) {
  unsigned long long* my = keys + (size_t)(blockIdx.x / region_blocks) * cells;
  unsigned h = hash(blockIdx.x * 1024 + threadIdx.x + 999);
  unsigned acc = 0;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < ops; ++i) {
    h = hash(h + i);
    const unsigned long long old = atomicMin(&my[h % (unsigned)cells], ((unsigned long long)h << 32) | i);
    acc += (unsigned)(old >> 40);
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (acc == 0xdeadbeef) sink[0] = acc;
}

// streaming stores: each block writes `bytes_per_block` with float4 stores
__global__ void __launch_bounds__(1024) k_store(float4* out, size_t vec_per_block, int reps) {
  for (int r = 0; r < reps; ++r) {
    float4* o = out + ((size_t)blockIdx.x * reps + r) * vec_per_block;
    for (size_t i = threadIdx.x; i < vec_per_block; i += blockDim.x) o[i] = make_float4(1.f, 2.f, 3.f, (float)i);
  }
}

template <int MODE>
void run_dsmem(const char* name, int csz, int nclusters, int words, int ops, unsigned* sink, long long* d_cyc) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(csz * nclusters); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = (size_t)words * 4;
  cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csz; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CHECK(cudaFuncSetAttribute(k_dsmem<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, words * 4));
  if (csz > 8) CHECK(cudaFuncSetAttribute(k_dsmem<MODE>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  int maxc = 0;
  cudaError_t oe = cudaOccupancyMaxActiveClusters(&maxc, k_dsmem<MODE>, &cfg);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int it = 0; it < 2; ++it) {
    cudaEventRecord(e0);
    CHECK(cudaLaunchKernelEx(&cfg, k_dsmem<MODE>, words, ops, sink, d_cyc));
    cudaEventRecord(e1);
    CHECK(cudaDeviceSynchronize());
  }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[1024]; CHECK(cudaMemcpy(h, d_cyc, sizeof(long long) * csz * nclusters, cudaMemcpyDeviceToHost));
  long long mx = 0; double avg = 0; for (int i = 0; i < csz * nclusters; ++i) { mx = h[i] > mx ? h[i] : mx; avg += h[i]; } avg /= csz * nclusters;
  printf("%-28s csz=%d clusters=%d (max active %d, occ err %d) words=%d: %.3f ms, cycles/CTA avg %.0f max %lld -> %.2f cyc/op/SM (ops/CTA=%d)\n",
         name, csz, nclusters, maxc, (int)oe, words, ms, avg, mx, avg / (1024.0 * ops), 1024 * ops);
}

int main() {
  unsigned* sink; long long* d_cyc; CHECK(cudaMalloc(&sink, 4)); CHECK(cudaMalloc(&d_cyc, 8 * 4096));
  cudaDeviceProp pr; CHECK(cudaGetDeviceProperties(&pr, 0));
  printf("device %s SMs %d smem/block optin %zu\n", pr.name, pr.multiProcessorCount, pr.sharedMemPerBlockOptin);
  const int words = 45000;  // 180 KB tile
  const int ops = 64;
  for (int csz : {8, 4, 2}) {
    const int ncl = csz == 8 ? 16 : (csz == 4 ? 32 : 74);
    run_dsmem<0>("remote red.add (no ret)", csz, ncl, words, ops, sink, d_cyc);
    run_dsmem<1>("remote atom.add (ret)", csz, ncl, words, ops, sink, d_cyc);
  }
  run_dsmem<0>("remote red.add csz16", 16, 8, words, ops, sink, d_cyc);
  run_dsmem<2>("local atomicAdd (ret)", 1, 148, words, ops, sink, d_cyc);
  run_dsmem<3>("local red", 1, 148, words, ops, sink, d_cyc);
  run_dsmem<0>("red 18 clusters of 8?", 8, 18, words, ops, sink, d_cyc);
  {
    const int cells = 57632, nblk = 128, region = 8;
    unsigned long long* keys; CHECK(cudaMalloc(&keys, (size_t)cells * 8 * (nblk / region)));
    CHECK(cudaMemset(keys, 0xff, (size_t)cells * 8 * (nblk / region)));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int it = 0; it < 2; ++it) { cudaEventRecord(e0); k_gmin<<<nblk, 1024>>>(keys, cells, region, ops, sink, d_cyc); cudaEventRecord(e1); CHECK(cudaDeviceSynchronize()); }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[256]; CHECK(cudaMemcpy(h, d_cyc, 8 * nblk, cudaMemcpyDeviceToHost));
    double avg = 0; for (int i = 0; i < nblk; ++i) avg += h[i]; avg /= nblk;
    printf("global atomicMin u64 (ret), %d blocks: %.3f ms, %.2f cyc/op/SM, %.1f Gop/s\n", nblk, ms, avg / (1024.0 * ops), nblk * 1024.0 * ops / ms / 1e6);
  }
  {
    const size_t per_block = 720 * 1024 / 16;  // 720 KB in float4
    const int reps = 64;
    for (int nblk : {128, 148, 296}) {
      float4* out; size_t bytes = (size_t)nblk * reps * per_block * 16; CHECK(cudaMalloc(&out, bytes));
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      float best = 1e9;
      for (int it = 0; it < 3; ++it) { cudaEventRecord(e0); k_store<<<nblk, 1024>>>(out, per_block, reps); cudaEventRecord(e1); CHECK(cudaDeviceSynchronize()); float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best; }
      printf("store stream %d blocks x 1024 thr: %.1f GB in %.3f ms = %.0f GB/s\n", nblk, bytes / 1e9, best, bytes / best / 1e6);
      cudaFree(out);
    }
  }
  return 0;
}
