// How should 148 CTAs, each holding a full [28 features][256 bins] partial histogram in shared memory, add it into ONE global int64
// histogram?  Measures (B200): (a) every CTA issues one 64-bit global atomic per bin on the same 7168 addresses (what the
// histogram kernel's flush does); (b) the same into R replicas (contention / R); (c) clusters of C CTAs reduce their partials
// through distributed shared memory first, so one CTA per cluster issues the atomics; (d) plain stores into per-CTA slabs + a
// second kernel that sums the slabs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o flush_probe flush_probe.cu && ./flush_probe
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <vector>

namespace cg = cooperative_groups;
constexpr int kBins = 256, kF = 28, kThreads = 1024, kEntries = kF * kBins;      // 7168
constexpr int kSmemWords = 4 * kBins * 32;                                       // the kernel's 4 x 32 KB arrays

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ long long SumEntry(const unsigned int* hsm, int i) {
  // entry i = (array k, bin, lane q): sum of the 4 row sub-histograms, as the real flush does
  const int q = i & 7, bin = (i >> 3) & 255, k = i >> 11;
  long long s = 0;
  #pragma unroll
  for (int rr = 0; rr < 4; ++rr) s += hsm[k * (kBins * 32) + bin * 32 + ((rr + bin) & 3) * 8 + q];
  return s;
}

template <int R>
__global__ void __launch_bounds__(kThreads, 1) k_atomics(unsigned long long* work) {
  extern __shared__ unsigned int hsm[];
  for (int i = threadIdx.x; i < kSmemWords; i += kThreads) hsm[i] = 1u;
  __syncthreads();
  unsigned long long* w = work + static_cast<size_t>(blockIdx.x % R) * 8192;
  for (int i = threadIdx.x; i < 8192; i += kThreads) {
    const int f = ((i >> 11) * 8) + (i & 7);
    if (f < kF) atomicAdd(w + f * kBins + ((i >> 3) & 255), static_cast<unsigned long long>(SumEntry(hsm, i)));
  }
}

__global__ void __launch_bounds__(kThreads, 1) k_cluster(unsigned long long* work) {
  extern __shared__ unsigned int hsm[];
  cg::cluster_group cl = cg::this_cluster();
  for (int i = threadIdx.x; i < kSmemWords; i += kThreads) hsm[i] = 1u;
  cl.sync();
  const int C = cl.num_blocks(), r = cl.block_rank();
  // CTA r owns entries [r * 8192 / C, (r + 1) * 8192 / C) of every CTA of the cluster
  const int lo = r * (8192 / C), hi = lo + 8192 / C;
  for (int i = lo + threadIdx.x; i < hi; i += kThreads) {
    long long s = 0;
    for (int p = 0; p < C; ++p) {
      const unsigned int* peer = cl.map_shared_rank(hsm, p);
      s += SumEntry(peer, i);
    }
    const int f = ((i >> 11) * 8) + (i & 7);
    if (f < kF) atomicAdd(work + f * kBins + ((i >> 3) & 255), static_cast<unsigned long long>(s));
  }
  cl.sync();
}

__global__ void __launch_bounds__(kThreads, 1) k_slab_store(unsigned long long* slabs) {
  extern __shared__ unsigned int hsm[];
  for (int i = threadIdx.x; i < kSmemWords; i += kThreads) hsm[i] = 1u;
  __syncthreads();
  unsigned long long* w = slabs + static_cast<size_t>(blockIdx.x) * 8192;
  for (int i = threadIdx.x; i < 8192; i += kThreads) w[i] = static_cast<unsigned long long>(SumEntry(hsm, i));
}
__global__ void k_slab_sum(const unsigned long long* slabs, int n, unsigned long long* work) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 8192) return;
  unsigned long long s = 0;
  for (int c = 0; c < n; ++c) s += slabs[static_cast<size_t>(c) * 8192 + i];
  work[i] = s;
}
__global__ void __launch_bounds__(kThreads, 1) k_clear_only() {
  extern __shared__ unsigned int hsm[];
  for (int i = threadIdx.x; i < kSmemWords; i += kThreads) hsm[i] = 1u;
  __syncthreads();
  if (hsm[threadIdx.x] == 7u) printf("x");
}

template <typename F>
float TimeIt(F f, int reps = 20) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  f();
  cudaEventRecord(a);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  return ms * 1e3f / reps;
}

int main() {
  const int ctas = 148;
  const size_t smem = kSmemWords * 4;
  unsigned long long *work, *slabs;
  CK(cudaMalloc(&work, 8192 * 8 * 64));
  CK(cudaMalloc(&slabs, static_cast<size_t>(8192) * 8 * ctas));
  CK(cudaMemset(work, 0, 8192 * 8 * 64));
  CK(cudaFuncSetAttribute(k_atomics<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_atomics<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_atomics<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_atomics<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_atomics<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_slab_store, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_clear_only, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  printf("clear only (baseline)        : %7.2f us\n", TimeIt([&] { k_clear_only<<<ctas, kThreads, smem>>>(); }));
  printf("atomics, 1 copy              : %7.2f us\n", TimeIt([&] { k_atomics<1><<<ctas, kThreads, smem>>>(work); }));
  printf("atomics, 2 replicas          : %7.2f us\n", TimeIt([&] { k_atomics<2><<<ctas, kThreads, smem>>>(work); }));
  printf("atomics, 4 replicas          : %7.2f us\n", TimeIt([&] { k_atomics<4><<<ctas, kThreads, smem>>>(work); }));
  printf("atomics, 8 replicas          : %7.2f us\n", TimeIt([&] { k_atomics<8><<<ctas, kThreads, smem>>>(work); }));
  printf("atomics, 32 replicas         : %7.2f us\n", TimeIt([&] { k_atomics<32><<<ctas, kThreads, smem>>>(work); }));
  for (int ncta : {16, 37, 74}) {
    printf("atomics, 1 copy, %3d CTAs    : %7.2f us\n", ncta, TimeIt([&] { k_atomics<1><<<ncta, kThreads, smem>>>(work); }));
  }
  for (int C : {2, 4, 8}) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas / C * C); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int nclusters = 0;
    cudaOccupancyMaxActiveClusters(&nclusters, k_cluster, &cfg);
    float us = TimeIt([&] { cudaLaunchKernelEx(&cfg, k_cluster, work); });
    printf("cluster of %d (DSMEM) + atomics: %7.2f us   (grid %d CTAs, max active clusters %d, %s)\n", C, us, ctas / C * C, nclusters,
           cudaGetErrorString(cudaGetLastError()));
  }
  printf("slab stores                  : %7.2f us\n", TimeIt([&] { k_slab_store<<<ctas, kThreads, smem>>>(slabs); }));
  printf("slab stores + sum kernel     : %7.2f us\n", TimeIt([&] { k_slab_store<<<ctas, kThreads, smem>>>(slabs); k_slab_sum<<<8, 1024>>>(slabs, ctas, work); }));
  printf("sum kernel alone (8 CTAs)    : %7.2f us\n", TimeIt([&] { k_slab_sum<<<8, 1024>>>(slabs, ctas, work); }));
  CK(cudaDeviceSynchronize());
  return 0;
}
