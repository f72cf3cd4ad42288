// Second micro-benchmark: how much of the shared-memory atomic cost is bank conflicts, and what do atomics into a cluster
// peer's shared memory (DSMEM) cost? Sizes the prize of a lane-per-bank ("transposed") histogram layout.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_atomics2 tools/ubench_atomics2.cu
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdint>
namespace cg = cooperative_groups;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

__device__ __forceinline__ uint32_t lcg(uint32_t& s){ s = s*1664525u + 1013904223u; return s >> 8; }

// ways = number of lanes that share one bank set: 1 -> every lane owns its bank (conflict free), 32 -> fully random banks.
// lanes are split in groups of `ways`; group g owns banks [g*ways, (g+1)*ways) and a lane picks one of them at random.
__global__ void __launch_bounds__(1024) k_ways(int iters, int rows, int ways, int natom, unsigned* out){
  extern __shared__ uint32_t sm[];
  for(int i=threadIdx.x;i<rows*32;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 12345u;
  const uint32_t lane = threadIdx.x & 31u, g0 = lane / ways * ways;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    uint32_t bank = g0 + (r >> 12) % ways;
    uint32_t row = r % (uint32_t)(rows - natom);
    for(int a=0;a<natom;a++) atomicAdd(&sm[(row+a)*32 + bank], r|1u);
  }
  __syncthreads();
  if(threadIdx.x==0) out[blockIdx.x]=sm[0];
}

// atomics into the shared memory of a random CTA of the cluster (own CTA included); lane-owned bank (ways=1) or random bank.
__global__ void __launch_bounds__(1024) k_dsmem(int iters, int rows, int ways, unsigned* out){
  extern __shared__ uint32_t sm[];
  cg::cluster_group cl = cg::this_cluster();
  for(int i=threadIdx.x;i<rows*32;i+=blockDim.x) sm[i]=0;
  cl.sync();
  const uint32_t nr = cl.num_blocks();
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 999u;
  const uint32_t lane = threadIdx.x & 31u, g0 = lane / ways * ways;
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm);
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    uint32_t bank = g0 + (r >> 12) % ways;
    uint32_t row = r % (uint32_t)rows;
    uint32_t dst = (r >> 17) % nr;
    uint32_t local = base + (row*32 + bank)*4, remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(dst));
    asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" :: "r"(remote), "r"(r|1u) : "memory");
  }
  cl.sync();
  if(threadIdx.x==0) out[blockIdx.x]=sm[0];
}

// Does the cost of ATOMS.ADD follow the number of active lanes? 5 atomics per iteration; each lane takes part in an iteration with
// probability pct/256 (drawn per lane and iteration, so the active lanes are scattered over the warp).
__global__ void __launch_bounds__(1024) k_lanes(int iters, int rows, int pct, unsigned* out){
  extern __shared__ uint32_t sm[];
  for(int i=threadIdx.x;i<rows*32;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 4711u;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    const bool on = (int)((r >> 3) & 255u) < pct;
    uint32_t row = (r >> 11) % (uint32_t)(rows - 5);
    uint32_t bank = r & 31u;
    if(on){
#pragma unroll
      for(int a=0;a<5;a++) atomicAdd(&sm[(row+a)*32 + bank], r|1u);
    }
  }
  __syncthreads();
  if(threadIdx.x==0) out[blockIdx.x]=sm[0];
}

// Clean version of the bank question: no division in the loop (masks only), 1024 rows x 32 banks, 5 atomics per iteration.
// mode 0: every lane owns its bank (conflict free); 1: random bank per lane; 2: random bank, all 5 atomics of an item in
// consecutive words (the histogram's [cnt, wY.lo, wY.hi, wYY.lo, wYY.hi] layout: bank = 5 * bin + k).
__global__ void __launch_bounds__(1024) k_clean(int iters, int mode, unsigned* out){
  extern __shared__ uint32_t sm[];
  for(int i=threadIdx.x;i<1029*32;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 31337u;
  const uint32_t lane = threadIdx.x & 31u;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    const uint32_t row = (r >> 5) & 1023u;
    if(mode == 6){  // the histogram's own sequence: red cnt, atom wY.lo / wYY.lo with the old value, red of hi + carry
      const uint32_t base = (((r >> 5) & 4095u) * 5u);
      atomicAdd(&sm[base], 1u);
      const uint32_t lo = r | 1u;
      const uint32_t o1 = atomicAdd(&sm[base + 1], lo);
      const uint32_t o2 = atomicAdd(&sm[base + 3], lo);
      atomicAdd(&sm[base + 2], 0xFFFFFFFFu + ((uint32_t)(o1 + lo) < lo ? 1u : 0u) + (r & 1u) + 1u);
      atomicAdd(&sm[base + 4], 0xFFFFFFFFu + ((uint32_t)(o2 + lo) < lo ? 1u : 0u) + (r & 2u) + 1u);
    } else if(mode == 2){
      const uint32_t base = (((r >> 5) & 4095u) * 5u);   // 4096 bins x 5 words
#pragma unroll
      for(int a=0;a<5;a++) atomicAdd(&sm[base + a], r|1u);
    } else {
      // modes 3, 4, 5: R = 2, 4, 8 interleaved copies -- a lane picks a random bank inside its own class (bank % R == lane % R)
      const uint32_t R = mode >= 3 ? (2u << (mode - 3)) : 1u;
      const uint32_t bank = mode == 0 ? lane : (mode >= 3 ? (((r & 31u) & ~(R - 1u)) | (lane & (R - 1u))) : (r & 31u));
#pragma unroll
      for(int a=0;a<5;a++) atomicAdd(&sm[(row+a)*32 + bank], r|1u);
    }
  }
  __syncthreads();
  if(threadIdx.x==0) out[blockIdx.x]=sm[0];
}

template<class F> float timeit(F f){
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms,a,b); return ms;
}

int main(){
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0));
  int sms=p.multiProcessorCount; printf("device %s SMs %d\n", p.name, sms);
  unsigned* out; CK(cudaMalloc(&out, 1<<20));
  const int iters=4096, rows=1280;     // 1280 rows x 32 banks x 4 B = 160 KB
  size_t smem=(size_t)rows*32*4;
  CK(cudaFuncSetAttribute(k_ways, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
  for(int natom : {1, 5}) for(int ways : {1, 2, 4, 8, 16, 32}){
    float ms=timeit([&]{k_ways<<<sms,1024,smem>>>(iters,rows,ways,natom,out);});
    double ops=(double)sms*1024*iters*natom;
    printf("smem ATOMS.ADD u32 x%d per item, %2d lanes share %2d banks : %8.3f ms  %7.1f Gops/s  %.2f lanes/clk/SM@1.9GHz\n", natom, ways, ways, ms, ops/ms*1e-6, ops/ms*1e-6/sms/1.9);
  }
  CK(cudaFuncSetAttribute(k_clean, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
  for(int mode : {0, 1, 2, 3, 4, 5, 6}){
    const char* nm[] = {"lane-owned bank (conflict free)", "random bank", "random bin, 5 consecutive words", "2 interleaved copies", "4 interleaved copies", "8 interleaved copies", "random bin, 2 of 5 return the old value"};
    float ms=timeit([&]{k_clean<<<sms,1024,1029*32*4>>>(iters,mode,out);});
    printf("clean ATOMS.ADD x5, %-32s: %8.3f ms  %.2f clk per warp instruction, %.1f lanes/clk/SM @1.9GHz\n", nm[mode], ms,
           ms*1e-3*1.9e9/(32.0*iters*5), 1024.0*iters*5/(ms*1e-3*1.9e9));
  }
  CK(cudaFuncSetAttribute(k_lanes, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
  for(int pct : {256, 128, 64, 32, 8}){
    float ms=timeit([&]{k_lanes<<<sms,1024,smem>>>(iters,rows,pct,out);});
    double winst=(double)sms*32*iters*5;   // warp-level atomic instructions issued (a warp skips one only if no lane is active)
    printf("smem ATOMS.ADD x5, %5.1f %% of the lanes active (scattered): %8.3f ms  %.2f clk per warp instruction @1.9GHz\n", pct/2.56, ms, ms*1e-3*1.9e9/(winst/sms/1.0)*1.0);
  }
  CK(cudaFuncSetAttribute(k_dsmem, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
  CK(cudaFuncSetAttribute(k_dsmem, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  for(int csz : {1, 2, 4, 8, 16}) for(int ways : {1, 32}){
    cudaLaunchConfig_t cfg = {};
    int ctas = sms / csz * csz;
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csz; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int maxc = 0; cudaOccupancyMaxActiveClusters(&maxc, k_dsmem, &cfg);
    if (maxc * csz < ctas) { ctas = maxc * csz; cfg.gridDim = dim3(ctas); }
    if (ctas == 0) { printf("dsmem cluster=%d: not launchable\n", csz); continue; }
    int its = 1024;
    cudaError_t err = cudaSuccess;
    float ms=timeit([&]{ cudaError_t e = cudaLaunchKernelEx(&cfg, k_dsmem, its, rows, ways, out); if (e != cudaSuccess) err = e; });
    if (err != cudaSuccess || cudaGetLastError() != cudaSuccess) { printf("dsmem cluster=%d: launch failed %s\n", csz, cudaGetErrorString(err)); continue; }
    double ops=(double)ctas*1024*its;
    printf("DSMEM red.add.u32 cluster=%2d ctas=%3d %s : %8.3f ms  %7.1f Gops/s  %.2f lanes/clk/SM@1.9GHz\n", csz, ctas, ways==1? "lane-owned bank":"random bank    ", ms, ops/ms*1e-6, ops/ms*1e-6/ctas/1.9);
  }
  return 0;
}
