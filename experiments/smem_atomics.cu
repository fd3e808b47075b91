// Microbenchmark: shared-memory update throughput on sm_100a for the histogram kernel design.
// Not part of the product; informs DESIGN.md (K4).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); return 1;}}while(0)

__device__ __forceinline__ uint32_t lcg(uint32_t& s){ s = s*1664525u + 1013904223u; return s>>24; }

// mode 0: red.shared.add.u32, conflict-free by construction: addr = bin*32+lane (one shared copy per CTA)
// mode 1: same but atomicAdd with return value used
// mode 2: warp-private copy, non-atomic LDS+IADD+STS
// mode 3: lanes = rows, same feature: addr = f*256+bin (random bank conflicts, CTA-shared copy)
// mode 4: float atomicAdd x2 (g,h) conflict-free layout
// mode 5: 64-bit atomicAdd conflict-free layout (bin*32+lane) of u64
// mode 6: mode 0 but bins skewed (3 distinct bins) -> still conflict-free by layout
// mode 7: mode 3 with skewed bins (3 distinct) -> heavy same-address conflicts
template<int MODE>
__global__ void k(uint32_t* out, long long* cyc, int iters){
  extern __shared__ uint32_t sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x>>5;
  int words = (MODE==2) ? 8192*nw : (MODE==5 ? 16384 : (MODE==4 ? 16384 : 8192));
  for(int i=threadIdx.x;i<words;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  uint32_t s = (blockIdx.x*blockDim.x+threadIdx.x)*2654435761u + 12345u;
  long long t0 = clock64();
  uint32_t acc=0;
  #pragma unroll 1
  for(int it=0; it<iters; ++it){
    #pragma unroll
    for(int u=0;u<16;++u){
      uint32_t bin = lcg(s);
      if(MODE==6||MODE==7) bin = bin % 3u;
      if(MODE==0||MODE==6){ atomicAdd(&sm[bin*32+lane], 0x00010001u); }
      else if(MODE==1){ acc += atomicAdd(&sm[bin*32+lane], 0x00010001u); }
      else if(MODE==2){ uint32_t* h = sm + warp*8192; h[bin*32+lane] += 0x00010001u; }
      else if(MODE==3||MODE==7){ atomicAdd(&sm[(u&31)*256+bin], 0x00010001u); }
      else if(MODE==4){ float* f=(float*)sm; atomicAdd(&f[bin*32+lane], 1.0f); atomicAdd(&f[8192+bin*32+lane], 0.5f);}
      else if(MODE==5){ unsigned long long* q=(unsigned long long*)sm; atomicAdd(&q[bin*32+lane], 0x0000000100000001ull);}
    }
  }
  long long t1 = clock64();
  __syncthreads();
  uint32_t v=acc; for(int i=threadIdx.x;i<words;i+=blockDim.x) v+=sm[i];
  out[blockIdx.x*blockDim.x+threadIdx.x]=v;
  if(threadIdx.x==0) cyc[blockIdx.x]=t1-t0;
}

// streaming read bandwidth reference: each thread reads uint4 grid-stride
__global__ void rd(const uint4* __restrict__ p, size_t n, uint32_t* out){
  uint32_t a=0; size_t i = blockIdx.x*(size_t)blockDim.x+threadIdx.x, st=(size_t)gridDim.x*blockDim.x;
  for(; i+3*st<n; i+=4*st){ uint4 x=p[i],y=p[i+st],z=p[i+2*st],w=p[i+3*st]; a+=x.x^y.y^z.z^w.w; }
  for(; i<n; i+=st){ uint4 x=p[i]; a+=x.x; }
  if(a==0x12345678u) out[0]=a;
}

template<int MODE> int run(const char* name, int threads, int ctas_per_sm, int nsm, uint32_t* out, long long* cyc){
  int nw=threads/32;
  size_t smem = (MODE==2)? (size_t)32768*nw : ((MODE==5||MODE==4)? 65536 : 32768);
  CK(cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = nsm*ctas_per_sm, iters=2000;
  k<MODE><<<grid,threads,smem>>>(out,cyc,10); CK(cudaDeviceSynchronize());
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); k<MODE><<<grid,threads,smem>>>(out,cyc,iters); cudaEventRecord(b); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms,a,b);
  long long hc[4096]; CK(cudaMemcpy(hc,cyc,sizeof(long long)*grid,cudaMemcpyDeviceToHost));
  double mc=0; for(int i=0;i<grid;i++) mc+= hc[i]; mc/=grid;
  double upd = (double)grid*threads*iters*16.0;
  double per_sm_cyc = (double)ctas_per_sm*threads*iters*16.0/mc;
  printf("%-34s thr=%4d cta/sm=%d  %.3f ms  %.2f Gupd/s  %.2f upd/cyc/SM (cyc/warp-instr=%.2f)\n", name, threads, ctas_per_sm, ms, upd/ms*1e-6, per_sm_cyc, 32.0/per_sm_cyc);
  return 0;
}

int main(){
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0));
  printf("%s SMs=%d smem/SM=%zu clock=%d kHz\n", p.name, p.multiProcessorCount, p.sharedMemPerMultiprocessor, p.clockRate);
  int nsm=p.multiProcessorCount;
  uint32_t* out; long long* cyc; CK(cudaMalloc(&out, 4096*1024*4)); CK(cudaMalloc(&cyc, 4096*8));
  for(int thr : {256, 512, 1024}){
    run<0>("red.shared u32 bank=lane", thr, 1, nsm, out, cyc);
  }
  run<0>("red.shared u32 bank=lane", 256, 4, nsm, out, cyc);
  run<0>("red.shared u32 bank=lane", 512, 4, nsm, out, cyc);
  run<1>("atom.shared u32 (ret) bank=lane", 512, 2, nsm, out, cyc);
  run<2>("warp-private LDS+STS", 128, 1, nsm, out, cyc);
  run<2>("warp-private LDS+STS", 192, 1, nsm, out, cyc);
  run<3>("red.shared lanes=rows random bins", 512, 2, nsm, out, cyc);
  run<4>("float atomicAdd x2 bank=lane", 512, 2, nsm, out, cyc);
  run<5>("u64 atomicAdd bank=lane", 512, 2, nsm, out, cyc);
  run<6>("red.shared bank=lane 3 bins", 512, 2, nsm, out, cyc);
  run<7>("red.shared lanes=rows 3 bins", 512, 2, nsm, out, cyc);
  // HBM read reference
  size_t bytes = (size_t)4<<30; uint4* buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf,1,bytes));
  for(int g : {nsm*4, nsm*8, nsm*16}){
    rd<<<g,512>>>(buf, bytes/16, out); CK(cudaDeviceSynchronize());
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a); for(int r=0;r<5;r++) rd<<<g,512>>>(buf, bytes/16, out); cudaEventRecord(b); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms,a,b);
    printf("read 4GiB grid=%d: %.1f GB/s\n", g, 5.0*bytes/ms*1e-6);
  }
  return 0;
}
