// Micro-benchmarks that size the histogram kernel's secondary ceiling on B200: shared-memory and L2 atomic throughput.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_atomics tools/ubench_atomics.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

__device__ __forceinline__ uint32_t lcg(uint32_t& s){ s = s*1664525u + 1013904223u; return s >> 8; }

template<int MODE> // 0: ATOMS.ADD no return, 1: with return (carry), 2: add64 emulation, 3: LDS+STS rmw (racy), 4: CAS64 atomicAdd(ull)
__global__ void __launch_bounds__(1024) k_smem(int iters, int words, int conflict_stride, unsigned* out){
  extern __shared__ uint32_t sm[];
  for(int i=threadIdx.x;i<words;i+=blockDim.x) sm[i]=0;
  __syncthreads();
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 12345u;
  uint32_t acc=0;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    uint32_t idx = (conflict_stride>0) ? ((r % (uint32_t)(words/conflict_stride))*conflict_stride) : (r % (uint32_t)words);
    if(MODE==0){ atomicAdd(&sm[idx], r|1u); }
    else if(MODE==1){ uint32_t old=atomicAdd(&sm[idx], r|1u); acc += old; }
    else if(MODE==2){ idx &= ~1u; uint32_t lo=r|1u; uint32_t old=atomicAdd(&sm[idx], lo); uint32_t c = ((uint32_t)(old+lo)<lo); uint32_t h=0xFFFFFFFFu+c; if(h) atomicAdd(&sm[idx+1], h); }
    else if(MODE==3){ uint32_t v=sm[idx]; sm[idx]=v+r; }
    else if(MODE==4){ idx &= ~1u; atomicAdd((unsigned long long*)&sm[idx], (unsigned long long)r); }
  }
  __syncthreads();
  if(threadIdx.x==0) out[blockIdx.x]=sm[0]+acc;
}

template<int MODE> // 0: RED.ADD.U64, 1: RED.ADD.U32, 2: atomicMax u64
__global__ void __launch_bounds__(256) k_gmem(int iters, size_t words64, unsigned long long* g){
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 777u;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    size_t idx = r % words64;
    if(MODE==0) atomicAdd(&g[idx], (unsigned long long)r);
    else if(MODE==1) atomicAdd(((unsigned*)g)+idx, r);
    else atomicMax(&g[idx], (unsigned long long)r);
  }
}

__global__ void __launch_bounds__(1024) k_match(int iters, int nkeys, unsigned* out){
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 99u;
  unsigned acc=0;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s) % nkeys;
    acc += __match_any_sync(0xffffffffu, r);
  }
  if(acc==0xdeadbeef) out[0]=acc;
}

// Does a warp shuffle compete with shared-memory atomics for the same pipe? 4 random ATOMS.ADD per iteration plus
// MODE 0: nothing, 1: six LDS.32 of a 32-entry table indexed by a random lane id, 2: six SHFL from a random lane.
template<int MODE>
__global__ void __launch_bounds__(1024) k_mix(int iters, int words, unsigned* out){
  extern __shared__ uint32_t sm[];
  for(int i=threadIdx.x;i<words+256;i+=blockDim.x) sm[i]=i;
  __syncthreads();
  uint32_t s = blockIdx.x*1315423911u + threadIdx.x*2654435761u + 4242u;
  uint32_t acc=0;
  uint32_t r0=threadIdx.x*7u, r1=threadIdx.x*11u, r2=threadIdx.x*13u, r3=threadIdx.x, r4=threadIdx.x*3u, r5=threadIdx.x*5u;
  for(int it=0; it<iters; it++){
    uint32_t r = lcg(s);
    uint32_t src = r & 31u;
    if(MODE==1){ const uint32_t* t = sm + words; acc += t[src] + t[32+src] + t[64+src] + t[96+src] + t[128+src] + t[160+src]; }
    if(MODE==2){ acc += __shfl_sync(0xffffffffu, r0, src) + __shfl_sync(0xffffffffu, r1, src) + __shfl_sync(0xffffffffu, r2, src)
                      + __shfl_sync(0xffffffffu, r3, src) + __shfl_sync(0xffffffffu, r4, src) + __shfl_sync(0xffffffffu, r5, src); }
    uint32_t idx = ((r>>5) + acc) % (uint32_t)(words-4);
    atomicAdd(&sm[idx], r|1u); atomicAdd(&sm[idx+1], r); atomicAdd(&sm[idx+2], r); atomicAdd(&sm[idx+3], r);
  }
  __syncthreads();
  if(threadIdx.x==0) out[blockIdx.x]=sm[0]+acc;
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
  const int iters=4096;
  const char* names[]={"ATOMS.ADD u32 (no return)","ATOMS.ADD u32 (return used)","add64 via 2x ATOMS (carry)","LDS+STS rmw (non-atomic)","atomicAdd u64 smem (CAS loop)"};
  int wordsv[]={1024*5, 8224*5, 48000};
  for(int wi=0; wi<3; wi++){
    int words=wordsv[wi]; size_t smem=(size_t)words*4;
    #define RUN(M) { CK(cudaFuncSetAttribute(k_smem<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024)); \
      for(int cs : {0, 32}) { for (int thr : {1024, 512}) { \
      int ctas = sms * ((smem<=100*1024 && thr<=1024) ? (thr==512?2:1) : 1); \
      float ms=timeit([&]{k_smem<M><<<ctas,thr,smem>>>(iters,words,cs,out);}); \
      double ops=(double)ctas*thr*iters; \
      printf("smem %-32s words=%6d %s thr=%4d ctas=%3d : %8.3f ms  %7.1f Gops/s  %.3f ops/clk/SM@1.9GHz\n", names[M], words, cs? "same-bank":"spread   ", thr, ctas, ms, ops/ms*1e-6, ops/ms*1e-6/sms/1.9); } } }
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4)
  }
  for(size_t bytes : {(size_t)64<<10, (size_t)6<<20, (size_t)64<<20}){
    unsigned long long* g; CK(cudaMalloc(&g, bytes)); CK(cudaMemset(g,0,bytes));
    size_t w=bytes/8;
    int ctas=sms*8;
    float ms0=timeit([&]{k_gmem<0><<<ctas,256>>>(1024,w,g);});
    float ms1=timeit([&]{k_gmem<1><<<ctas,256>>>(1024,w,g);});
    float ms2=timeit([&]{k_gmem<2><<<ctas,256>>>(1024,w,g);});
    double ops=(double)ctas*256*1024;
    printf("gmem footprint %6zu KB: RED.ADD.U64 %7.1f Gops/s | RED.ADD.U32 %7.1f Gops/s | ATOM.MAX.U64 %7.1f Gops/s\n", bytes>>10, ops/ms0*1e-6, ops/ms1*1e-6, ops/ms2*1e-6);
    cudaFree(g);
  }
  {
    int words=41120; size_t smem=(size_t)(words+256)*4;
    CK(cudaFuncSetAttribute(k_mix<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
    CK(cudaFuncSetAttribute(k_mix<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
    CK(cudaFuncSetAttribute(k_mix<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220*1024));
    float m0=timeit([&]{k_mix<0><<<sms,1024,smem>>>(2048,words,out);});
    float m1=timeit([&]{k_mix<1><<<sms,1024,smem>>>(2048,words,out);});
    float m2=timeit([&]{k_mix<2><<<sms,1024,smem>>>(2048,words,out);});
    printf("mix 4xATOMS per iter: alone %.3f ms | +6 LDS.32 (random lane of a 32-entry table) %.3f ms | +6 SHFL (random lane) %.3f ms\n", m0, m1, m2);
  }
  for(int nk : {4, 64, 1024}){
    float ms=timeit([&]{k_match<<<sms,1024>>>(2048,nk,out);});
    double ops=(double)sms*1024*2048;
    printf("match_any nkeys=%4d : %7.1f Glanes/s\n", nk, ops/ms*1e-6);
  }
  return 0;
}
