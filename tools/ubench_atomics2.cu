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
