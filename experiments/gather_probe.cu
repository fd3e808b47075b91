// Probe: DRAM bytes fetched per gathered 32-byte record for different load flavours, and the rate of returning 64-bit
// atomics on one address.  Run under `ncu --metrics dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum`.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <int MODE>
__global__ void k_gather(const uint8_t* __restrict__ rec, const int* __restrict__ idx, int n, unsigned* out) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  unsigned acc = 0;
  for (int g = warp * 32; g + 32 <= n; g += nwarps * 32) {
    const int my = idx[g + lane];
    if (MODE <= 3) {
      const int q = lane & 7, r = lane >> 3;
      #pragma unroll
      for (int s = 0; s < 8; ++s) {
        const int row = __shfl_sync(0xffffffffu, my, s * 4 + r);
        const unsigned* p = reinterpret_cast<const unsigned*>(rec + static_cast<size_t>(row) * 32) + q;
        unsigned v;
        if (MODE == 0) asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
        if (MODE == 1) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
        if (MODE == 2) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
        if (MODE == 3) asm volatile("ld.global.cv.u32 %0, [%1];" : "=r"(v) : "l"(p));
        acc += v;
      }
    } else if (MODE == 4) {            // 2 lanes x 16 B per row, 16 rows per step
      const int q = lane & 1, r = lane >> 1;
      #pragma unroll
      for (int s = 0; s < 2; ++s) {
        const int row = __shfl_sync(0xffffffffu, my, s * 16 + r);
        const uint4* p = reinterpret_cast<const uint4*>(rec + static_cast<size_t>(row) * 32) + q;
        uint4 v;
        asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
        acc += v.x + v.y + v.z + v.w;
      }
    } else if (MODE == 5) {            // 1 lane x 32 B per row (two 16 B loads), 32 rows per step
      const uint4* p = reinterpret_cast<const uint4*>(rec + static_cast<size_t>(my) * 32);
      uint4 a, b;
      asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(p));
      asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p + 1));
      acc += a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
    } else if (MODE == 6) {            // 1 lane x one 32 B load per row with an evict-first L2 policy
      const void* p = rec + static_cast<size_t>(my) * 32;
      unsigned v0, v1, v2, v3, v4, v5, v6, v7;
      asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3), "=r"(v4), "=r"(v5), "=r"(v6), "=r"(v7) : "l"(p));
      acc += v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
    }
  }
  if (acc == 0x12345678u) out[0] = acc;
}

__global__ void k_atom(unsigned long long* ctr, int per_warp, unsigned long long* sink) {
  unsigned long long a = 0;
  if ((threadIdx.x & 31) == 0)
    for (int i = 0; i < per_warp; ++i) a += atomicAdd(ctr, (1ULL << 32) | 1ULL);
  if (a == 1) sink[0] = a;
}
__global__ void k_atom_spread(unsigned long long* ctr, int per_warp, unsigned long long* sink) {
  unsigned long long a = 0;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if ((threadIdx.x & 31) == 0)
    for (int i = 0; i < per_warp; ++i) a += atomicAdd(ctr + (warp & 7) * 16, (1ULL << 32) | 1ULL);
  if (a == 1) sink[0] = a;
}

int main(int argc, char** argv) {
  const int N = 10000000;
  uint8_t* rec; CK(cudaMalloc(&rec, size_t(N) * 32)); CK(cudaMemset(rec, 1, size_t(N) * 32));
  unsigned* out; CK(cudaMalloc(&out, 64));
  unsigned long long* ctr; CK(cudaMalloc(&ctr, 4096)); CK(cudaMemset(ctr, 0, 4096));
  for (int dens : {2, 4, 16, 64}) {
    std::vector<int> h;
    unsigned s = 12345;
    for (int i = 0; i < N; ++i) { s = s * 1664525u + 1013904223u; if ((s >> 8) % dens == 0) h.push_back(i); }
    int n = static_cast<int>(h.size()) / 32 * 32;
    int* idx; CK(cudaMalloc(&idx, sizeof(int) * n)); CK(cudaMemcpy(idx, h.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
    printf("density 1/%d rows %d\n", dens, n);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
#define RUN(M) { CK(cudaMemset(rec, 1, size_t(N) * 32)); cudaEventRecord(e0); k_gather<M><<<148 * 2, 1024>>>(rec, idx, n, out); cudaEventRecord(e1); CK(cudaDeviceSynchronize()); \
                 float ms; cudaEventElapsedTime(&ms, e0, e1); printf("  mode %d: %.1f us  %.1f B/row-equivalent at 6.5TB/s\n", M, ms * 1e3, ms * 1e-3 * 6.5e12 / n); }
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
    CK(cudaFree(idx));
  }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0); k_atom<<<148 * 4, 256>>>(ctr, 16, ctr + 8); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("same-address returning u64 atomics: %d in %.1f us = %.2f ns each\n", 148 * 4 * 8 * 16, ms * 1e3, ms * 1e6 / (148 * 4 * 8 * 16));
    cudaEventRecord(e0); k_atom_spread<<<148 * 4, 256>>>(ctr + 32, 16, ctr + 8); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    cudaEventElapsedTime(&ms, e0, e1);
    printf("8-address returning u64 atomics: %d in %.1f us = %.2f ns each\n", 148 * 4 * 8 * 16, ms * 1e3, ms * 1e6 / (148 * 4 * 8 * 16));
  }
  return 0;
}
