// Third micro-benchmark (round 2): candidate formulations of the histogram item loop, on real column-major data in HBM.
// Measures, per (row, column) item, what the accumulation scheme, the bin arithmetic, the record lookups and warp-level
// pre-aggregation of few-valued columns cost, so that the production kernel is re-formulated on evidence.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/ubench_hist3 tools/ubench_hist3.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);}}while(0)

__device__ __forceinline__ uint32_t hash32(uint32_t x){ x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// kind 0: bell-shaped continuous in [-4,4] (sum of 4 uniforms), 1: three values, 2: uniform
__global__ void gen_col(float* x, int64_t n, int kind, uint32_t seed){
  for(int64_t i = blockIdx.x*(int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x*blockDim.x){
    uint32_t h = hash32((uint32_t)i*2654435761u + seed);
    float v;
    if(kind == 1){ uint32_t r = h % 10u; v = r < 5 ? 0.0f : (r < 8 ? 1.09f : 2.17f); }
    else if(kind == 2){ v = (h >> 8) * (1.0f/16777216.0f) * 6.28f - 3.14f; }
    else { uint32_t h2 = hash32(h + 0x9e3779b9u); v = ((h & 0xFFFF) + (h >> 16) + (h2 & 0xFFFF) + (h2 >> 16)) * (1.0f/65536.0f) * 2.0f - 4.0f; }
    x[i] = v;
  }
}
__global__ void gen_y(float* y, uint16_t* code, int64_t n, int L){
  for(int64_t i = blockIdx.x*(int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x*blockDim.x){
    uint32_t h = hash32((uint32_t)i*40503u + 77u);
    y[i] = ((h >> 8) * (1.0f/16777216.0f)) * 1.4f - 0.7f;
    code[i] = (uint16_t)(hash32(h) % (uint32_t)L);
  }
}

__device__ __forceinline__ uint32_t smem_addr(const void* p){ return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void red_add(uint32_t a, uint32_t v){ asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t atom_add(uint32_t a, uint32_t v){ uint32_t o; asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(a), "r"(v) : "memory"); return o; }
__device__ __forceinline__ void red_add_hi(uint32_t a, uint32_t old, uint32_t lo, uint32_t hi){
  asm volatile("{\n .reg .u32 t, h;\n add.cc.u32 t, %1, %2;\n addc.u32 h, %3, 0;\n red.shared.add.u32 [%0], h;\n}" :: "r"(a), "r"(old), "r"(lo), "r"(hi) : "memory");
}
__device__ __forceinline__ uint32_t lds(uint32_t a){ uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint4 lds128(uint32_t a){ uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }

// one global atomic per CTA (148 K same-address atomics would cost ~70 us: not part of what is measured)
__device__ __forceinline__ void block_sum_out(unsigned long long acc, unsigned long long* out){
  __shared__ unsigned long long red[32];
  for(int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  __syncthreads();
  if((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if(threadIdx.x == 0){ unsigned long long t = 0; for(int i = 0; i < (int)(blockDim.x >> 5); i++) t += red[i]; atomicAdd(out, t); }
}

struct Args {
  const float* const* cols; int ncols; const float* y; const uint16_t* code; int64_t n_pad;
  float mn[16]; double st[16]; int nb; double sy, syy; int L; unsigned long long* out;
};

// SCHEME: 5 = [cnt, y.lo, y.hi, yy.lo, yy.hi] stride-5 records (round-1 layout); 4 = four planes [y.lo][y.hi][(yy<<18|1).lo][.hi];
//         3 = three planes [y32][yy32][cnt | carries] (32-bit quantities); 0 = no atomics (xor into a register)
// MATH:   0 = fp64 (reference formula), 1 = fp32 pre-filter with the fp64 fallback near bin edges
// AGG:    0 = every lane adds, 1 = ballot loop over the distinct bins of a warp (REDUX limbs, the first lane of each bin adds)
template<int SCHEME, int MATH, int AGG, int PF = 1>
__global__ void __launch_bounds__(1024, 1) root_kernel(Args a){
  extern __shared__ uint32_t sm[];
  const int nbp = a.nb + 1;
  const int wpc = SCHEME == 5 ? nbp * 5 : SCHEME == 4 ? nbp * 4 : nbp * 3;   // words per column
  for(int i = threadIdx.x; i < wpc * a.ncols + 32 * 8 + 32; i += blockDim.x) sm[i] = 0;
  __syncthreads();
  const uint32_t base = smem_addr(sm);
  const int lane = threadIdx.x & 31;
  uint32_t sinkx = 0;
  uint32_t* kmm = sm + wpc * a.ncols + 32 * 8;  // per column {~key(min), key(max)}
  const int64_t n_tiles = a.n_pad / 4096;
  for(int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x){
    const int64_t row = tile * 4096 + threadIdx.x * 4;
    const float4 yv = *reinterpret_cast<const float4*>(a.y + row);
    const float ys[4] = {yv.x, yv.y, yv.z, yv.w};
    long long qy[4], qyy[4];
#pragma unroll
    for(int j = 0; j < 4; j++){
      const double y = (double)ys[j];
      qy[j] = __double2ll_rn(__dmul_rn(y, a.sy));
      const long long q2 = __double2ll_rn(__dmul_rn(__dmul_rn(y, y), a.syy));
      qyy[j] = SCHEME == 4 ? ((q2 << 18) | 1LL) : q2;
    }
    float4 nxt = __ldcs(reinterpret_cast<const float4*>(a.cols[0] + row));
    float4 nxt2 = nxt;
    if(PF >= 2 && a.ncols > 1) nxt2 = __ldcs(reinterpret_cast<const float4*>(a.cols[1] + row));
#pragma unroll 1
    for(int c = 0; c < a.ncols; c++){
      const float4 cur = nxt;
      if(PF >= 2){
        nxt = nxt2;
        if(c + 2 < a.ncols) nxt2 = __ldcs(reinterpret_cast<const float4*>(a.cols[c + 2] + row));
      } else {
        if(c + 1 < a.ncols) nxt = __ldcs(reinterpret_cast<const float4*>(a.cols[c + 1] + row));
      }
      const float xs[4] = {cur.x, cur.y, cur.z, cur.w};
      const float mn = a.mn[c];
      const double st = a.st[c];
      const float st32 = (float)st * 2048.0f;
      int b[4];
#pragma unroll
      for(int j = 0; j < 4; j++){
        const float x = xs[j];
        int bb;
        if(MATH == 0){
          bb = __double2int_rz(__dmul_rn(__dsub_rn((double)x, (double)mn), st));
        } else {
          const int iv = __float2int_rz(__fmul_rn(__fsub_rn(x, mn), st32));
          bb = iv >> 11;
          if(((iv + 1) & 2047) <= 1) bb = __double2int_rz(__dmul_rn(__dsub_rn((double)x, (double)mn), st));
        }
        bb = max(0, min(bb, a.nb - 1));
        b[j] = (x != x) ? a.nb : bb;
      }
      {  // min / max of the column: one test per thread and column (4 rows), keys only move in the first tiles
        const float lo = fminf(fminf(xs[0], xs[1]), fminf(xs[2], xs[3])), hi = fmaxf(fmaxf(xs[0], xs[1]), fmaxf(xs[2], xs[3]));
        uint32_t ul = __float_as_uint(lo), uh = __float_as_uint(hi);
        const uint32_t kl = ~(ul ^ ((ul >> 31) ? 0xFFFFFFFFu : 0x80000000u)), kh = uh ^ ((uh >> 31) ? 0xFFFFFFFFu : 0x80000000u);
        if(kl > kmm[2 * c] || kh > kmm[2 * c + 1]){ atomicMax(&kmm[2 * c], kl); atomicMax(&kmm[2 * c + 1], kh); }
      }
      const uint32_t cb = base + (uint32_t)(c * wpc) * 4u;
      if(SCHEME == 0){
#pragma unroll
        for(int j = 0; j < 4; j++) sinkx ^= (uint32_t)b[j] * 2654435761u + (uint32_t)qy[j];
      } else if(AGG == 0){
        if(SCHEME == 5){
          uint32_t oy[4], oyy[4];
#pragma unroll
          for(int j = 0; j < 4; j++) red_add(cb + b[j] * 20, 1u);
#pragma unroll
          for(int j = 0; j < 4; j++) oy[j] = atom_add(cb + b[j] * 20 + 4, (uint32_t)qy[j]);
#pragma unroll
          for(int j = 0; j < 4; j++) oyy[j] = atom_add(cb + b[j] * 20 + 12, (uint32_t)qyy[j]);
#pragma unroll
          for(int j = 0; j < 4; j++){
            red_add_hi(cb + b[j] * 20 + 8, oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
            red_add_hi(cb + b[j] * 20 + 16, oyy[j], (uint32_t)qyy[j], (uint32_t)((unsigned long long)qyy[j] >> 32));
          }
        } else if(SCHEME == 4){
          uint32_t oy[4], oyy[4];
          const uint32_t ps = (uint32_t)nbp * 4u;
#pragma unroll
          for(int j = 0; j < 4; j++) oy[j] = atom_add(cb + b[j] * 4, (uint32_t)qy[j]);
#pragma unroll
          for(int j = 0; j < 4; j++) oyy[j] = atom_add(cb + 2 * ps + b[j] * 4, (uint32_t)qyy[j]);
#pragma unroll
          for(int j = 0; j < 4; j++){
            red_add_hi(cb + ps + b[j] * 4, oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
            red_add_hi(cb + 3 * ps + b[j] * 4, oyy[j], (uint32_t)qyy[j], (uint32_t)((unsigned long long)qyy[j] >> 32));
          }
        } else {
          uint32_t oy[4], oyy[4];
          const uint32_t ps = (uint32_t)nbp * 4u;
#pragma unroll
          for(int j = 0; j < 4; j++) oy[j] = atom_add(cb + b[j] * 4, (uint32_t)(qy[j] >> 8) + 0x20000000u);
#pragma unroll
          for(int j = 0; j < 4; j++) oyy[j] = atom_add(cb + ps + b[j] * 4, (uint32_t)(qyy[j] >> 8));
#pragma unroll
          for(int j = 0; j < 4; j++){
            const uint32_t ly = (uint32_t)(qy[j] >> 8) + 0x20000000u, lyy = (uint32_t)(qyy[j] >> 8);
            const uint32_t cy = (uint32_t)(oy[j] + ly) < ly ? 1u : 0u, cyy = (uint32_t)(oyy[j] + lyy) < lyy ? 1u : 0u;
            red_add(cb + 2 * ps + b[j] * 4, 1u + (cy << 18) + (cyy << 25));
          }
        }
      } else {
        // warp-aggregated: per distinct bin among the warp's 128 items, every thread first sums its own rows of that bin, then
        // four REDUX (26-bit limbs keep the sums inside 32 bits) give the warp's totals and one lane adds them
        const uint32_t ps = (uint32_t)nbp * 4u;
        unsigned todo = 0xFu;  // rows of this thread not yet added
        while(true){
          const unsigned has = __ballot_sync(0xFFFFFFFFu, todo != 0u);
          if(!has) break;
          const int leader = __ffs(has) - 1;
          const int jl = __ffs(todo) - 1;
          const int mine = jl == 0 ? b[0] : jl == 1 ? b[1] : jl == 2 ? b[2] : b[3];
          const int bl = __shfl_sync(0xFFFFFFFFu, mine, leader);
          long long sy = 0, sv = 0;
#pragma unroll
          for(int j = 0; j < 4; j++){
            const bool in = ((todo >> j) & 1u) && b[j] == bl;
            sy += in ? qy[j] : 0LL;
            sv += in ? qyy[j] : 0LL;
            todo &= in ? ~(1u << j) : ~0u;
          }
          const unsigned t_ylo = __reduce_add_sync(0xFFFFFFFFu, (unsigned)sy & 0x3FFFFFFu);
          const int t_yhi = __reduce_add_sync(0xFFFFFFFFu, (int)(sy >> 26));
          const unsigned t_vlo = __reduce_add_sync(0xFFFFFFFFu, (unsigned)sv & 0x3FFFFFFu);
          const int t_vhi = __reduce_add_sync(0xFFFFFFFFu, (int)(sv >> 26));
          if(lane == leader){
            const long long ty = ((long long)t_yhi << 26) + (long long)t_ylo;
            const long long tv = ((long long)t_vhi << 26) + (long long)t_vlo;
            const uint32_t o1 = atom_add(cb + bl * 4, (uint32_t)ty);
            const uint32_t o2 = atom_add(cb + 2 * ps + bl * 4, (uint32_t)tv);
            red_add_hi(cb + ps + bl * 4, o1, (uint32_t)ty, (uint32_t)((unsigned long long)ty >> 32));
            red_add_hi(cb + 3 * ps + bl * 4, o2, (uint32_t)tv, (uint32_t)((unsigned long long)tv >> 32));
          }
        }
      }
    }
  }
  __syncthreads();
  unsigned long long acc = sinkx;
  for(int c = threadIdx.x; c < 32; c += blockDim.x) acc += kmm[c];
  for(int i = threadIdx.x; i < wpc * a.ncols; i += blockDim.x) acc += sm[i];
  block_sum_out(acc, a.out);
}

// Levels below the root: one column per CTA, L leaves x (nb+1) bins; every item looks its (leaf) record up in shared memory.
// REC: 6 = six 32-bit loads from 7-word records (round-1), 4 = one 128-bit load of [place|nbin, min32, step32, binlo|binhi] from
//      12-word records (bin-range filter for the min/max keys; the rare slow path updates the keys), MATH as above.
template<int SCHEME, int REC, int MATH>
__global__ void __launch_bounds__(1024, 1) deep_kernel(Args a, int cols_per_pass){
  extern __shared__ uint32_t sm[];
  const int nbp = a.nb + 1, L = a.L;
  const int wpl = SCHEME == 5 ? nbp * 5 : nbp * 4;
  const int rstride = REC == 6 ? 7 : 12;
  uint32_t* T = sm;
  uint32_t* hist = sm + ((rstride * L + 3) & ~3);
  const int c = blockIdx.x % cols_per_pass, slice = blockIdx.x / cols_per_pass, nsl = gridDim.x / cols_per_pass;
  if(slice >= nsl) return;
  for(int i = threadIdx.x; i < wpl * L + 64; i += blockDim.x) hist[i] = 0;
  for(int l = threadIdx.x; l < L; l += blockDim.x){
    uint32_t* r = T + l * rstride;
    const double st = a.st[c];
    if(REC == 6){
      r[0] = (uint32_t)(l * wpl) | ((uint32_t)a.nb << 16); r[1] = __float_as_uint(a.mn[c]); r[2] = 0;
      r[3] = (uint32_t)__double2loint(st); r[4] = (uint32_t)__double2hiint(st); r[5] = 0; r[6] = 0;
    } else {
      r[0] = (uint32_t)(l * wpl) | ((uint32_t)a.nb << 16); r[1] = __float_as_uint(a.mn[c]); r[2] = __float_as_uint((float)st * 2048.0f);
      r[3] = 0xFFFF0000u;  // binlo = 0xFFFF, binhi = 0
      r[4] = 0; r[5] = 0; r[6] = (uint32_t)__double2loint(st); r[7] = (uint32_t)__double2hiint(st);
      r[8] = 0xFFFFu; r[9] = 0;  // lowest / highest bin seen (authoritative); r[3] is the snapshot the items test against
    }
  }
  __syncthreads();
  const uint32_t hist_a = smem_addr(hist), tab_a = smem_addr(T);
  const uint32_t ps = (uint32_t)nbp * 4u;  // plane stride inside one leaf block (SCHEME 4: [4 planes][nbp])
  const float* col = a.cols[c];
  const int64_t n_tiles = a.n_pad / 4096;
  for(int64_t tile = slice; tile < n_tiles; tile += nsl){
    const int64_t row = tile * 4096 + threadIdx.x * 4;
    const ushort4 cd = *reinterpret_cast<const ushort4*>(a.code + row);
    const int leaf[4] = {cd.x, cd.y, cd.z, cd.w};
    const float4 yv = *reinterpret_cast<const float4*>(a.y + row);
    const float4 xv = __ldcs(reinterpret_cast<const float4*>(col + row));
    const float ys[4] = {yv.x, yv.y, yv.z, yv.w};
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
    long long qy[4], qyy[4];
    uint32_t wd[4];
#pragma unroll
    for(int j = 0; j < 4; j++){
      const double y = (double)ys[j];
      qy[j] = __double2ll_rn(__dmul_rn(y, a.sy));
      const long long q2 = __double2ll_rn(__dmul_rn(__dmul_rn(y, y), a.syy));
      qyy[j] = SCHEME == 4 ? ((q2 << 18) | 1LL) : q2;
    }
#pragma unroll
    for(int j = 0; j < 4; j++){
      const float x = xs[j];
      const uint32_t rec = tab_a + (uint32_t)(leaf[j] * rstride) * 4u;
      int bb, nbin; uint32_t place;
      if(REC == 6){
        const uint32_t pn = lds(rec); place = pn & 0xFFFFu; nbin = (int)(pn >> 16);
        const double mn = (double)__uint_as_float(lds(rec + 4));
        const double st = __hiloint2double((int)lds(rec + 16), (int)lds(rec + 12));
        const uint32_t kmin = lds(rec + 20), kmax = lds(rec + 24);
        bb = max(0, min(__double2int_rz(__dmul_rn(__dsub_rn((double)x, mn), st)), nbin - 1));
        uint32_t u = __float_as_uint(x); const uint32_t key = u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u);
        if(x == x && (~key > kmin || key > kmax)){
          asm volatile("red.shared.max.u32 [%0], %1;" :: "r"(rec + 20), "r"(~key) : "memory");
          asm volatile("red.shared.max.u32 [%0], %1;" :: "r"(rec + 24), "r"(key) : "memory");
        }
      } else {
        const uint4 r = lds128(rec); place = r.x & 0xFFFFu; nbin = (int)(r.x >> 16);
        const float mn = __uint_as_float(r.y), st32 = __uint_as_float(r.z);
        if(MATH == 1){
          const int iv = __float2int_rz(__fmul_rn(__fsub_rn(x, mn), st32));
          bb = iv >> 11;
          if(((iv + 1) & 2047) <= 1) bb = __double2int_rz(__dmul_rn(__dsub_rn((double)x, (double)mn), __hiloint2double((int)lds(rec + 28), (int)lds(rec + 24))));
        } else {
          bb = __double2int_rz(__dmul_rn(__dsub_rn((double)x, (double)mn), __hiloint2double((int)lds(rec + 28), (int)lds(rec + 24))));
        }
        bb = max(0, min(bb, nbin - 1));
        const int blo = (int)(r.w >> 16), bhi = (int)(r.w & 0xFFFFu);
        if(x == x && (bb <= blo || bb >= bhi)){   // only values in (or beyond) the end bins seen so far can move the extremes
          uint32_t u = __float_as_uint(x); const uint32_t key = u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u);
          asm volatile("red.shared.max.u32 [%0], %1;" :: "r"(rec + 16), "r"(~key) : "memory");
          asm volatile("red.shared.max.u32 [%0], %1;" :: "r"(rec + 20), "r"(key) : "memory");
          atomicMin(reinterpret_cast<unsigned int*>(T + leaf[j] * rstride + 8), (unsigned)bb);
          atomicMax(reinterpret_cast<unsigned int*>(T + leaf[j] * rstride + 9), (unsigned)bb);
        }
      }
      bb = (x != x) ? nbin : bb;
      wd[j] = hist_a + (place + (uint32_t)(SCHEME == 5 ? bb * 5 : bb)) * 4u;
    }
    if(REC == 4){  // refresh the snapshots once per tile (a stale snapshot only sends more items down the slow path)
      if(threadIdx.x < L){ uint32_t* r = T + threadIdx.x * rstride; r[3] = (r[8] << 16) | r[9]; }
    }
    uint32_t oy[4], oyy[4];
    if(SCHEME == 5){
#pragma unroll
      for(int j = 0; j < 4; j++) red_add(wd[j], 1u);
#pragma unroll
      for(int j = 0; j < 4; j++) oy[j] = atom_add(wd[j] + 4, (uint32_t)qy[j]);
#pragma unroll
      for(int j = 0; j < 4; j++) oyy[j] = atom_add(wd[j] + 12, (uint32_t)qyy[j]);
#pragma unroll
      for(int j = 0; j < 4; j++){
        red_add_hi(wd[j] + 8, oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
        red_add_hi(wd[j] + 16, oyy[j], (uint32_t)qyy[j], (uint32_t)((unsigned long long)qyy[j] >> 32));
      }
    } else {
#pragma unroll
      for(int j = 0; j < 4; j++) oy[j] = atom_add(wd[j], (uint32_t)qy[j]);
#pragma unroll
      for(int j = 0; j < 4; j++) oyy[j] = atom_add(wd[j] + 2 * ps, (uint32_t)qyy[j]);
#pragma unroll
      for(int j = 0; j < 4; j++){
        red_add_hi(wd[j] + ps, oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
        red_add_hi(wd[j] + 3 * ps, oyy[j], (uint32_t)qyy[j], (uint32_t)((unsigned long long)qyy[j] >> 32));
      }
    }
  }
  __syncthreads();
  unsigned long long acc = 0;
  for(int i = threadIdx.x; i < wpl * L; i += blockDim.x) acc += hist[i];
  for(int i = threadIdx.x; i < rstride * L; i += blockDim.x) acc += T[i];
  block_sum_out(acc, a.out);
}

template<class F> float timeit(F f, int reps = 3){
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); CK(cudaDeviceSynchronize());
  cudaEventRecord(a); for(int i = 0; i < reps; i++) f(); cudaEventRecord(b); CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / reps;
}

int main(int argc, char** argv){
  const int64_t n = argc > 1 ? atoll(argv[1]) : 11000000;
  const int64_t n_pad = (n + 4095) / 4096 * 4096;
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount;
  printf("device %s SMs %d rows %lld\n", p.name, sms, (long long)n_pad);
  const int NC = 10;
  float* cols[NC]; const float** dcols; float* y; uint16_t* code; unsigned long long* out;
  // columns 0..7 continuous (bell), 8 uniform, 9 three-valued
  for(int c = 0; c < NC; c++){ CK(cudaMalloc(&cols[c], n_pad * 4)); gen_col<<<sms * 4, 256>>>(cols[c], n_pad, c == 9 ? 1 : c == 8 ? 2 : 0, 1000u + c); }
  CK(cudaMalloc(&dcols, NC * sizeof(float*))); CK(cudaMalloc(&y, n_pad * 4)); CK(cudaMalloc(&code, n_pad * 2)); CK(cudaMalloc(&out, 8));
  gen_y<<<sms * 4, 256>>>(y, code, n_pad, 32);
  CK(cudaDeviceSynchronize());
  Args a{}; a.y = y; a.code = code; a.n_pad = n_pad; a.out = out; a.L = 32;
  a.sy = ldexp(1.0, 62 - 24 + 0); a.syy = ldexp(1.0, 28);   // |y| < 1: 38-bit wY, 28-bit wYY (cnt folded below it)
  auto set_cols = [&](const int* idx, int k, int nb){
    const float* h[NC];
    for(int i = 0; i < k; i++){
      h[i] = cols[idx[i]];
      const float lo = idx[i] == 9 ? 0.0f : (idx[i] == 8 ? -3.15f : -4.0f), hi = idx[i] == 9 ? 2.18f : (idx[i] == 8 ? 3.15f : 4.0001f);
      a.mn[i] = lo; a.st[i] = nb / ((double)hi - (double)lo);
    }
    CK(cudaMemcpy(dcols, h, k * sizeof(float*), cudaMemcpyHostToDevice));
    a.cols = dcols; a.ncols = k; a.nb = nb;
  };
  const double clk = 1.9e9;
  auto report = [&](const char* name, float ms, double items){
    printf("%-78s %8.3f ms  %6.3f ns/Kitem  %6.2f SM-clk per warp-item (32 items) @1.9GHz  %7.1f Gitem/s\n", name, ms, ms * 1e6 / (items / 1e3),
           ms * 1e-3 * clk * sms / (items / 32.0), items / ms * 1e-6);
  };
  CK(cudaFuncSetAttribute(root_kernel<5,0,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<4,0,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<4,1,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<3,0,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<3,1,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<0,0,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<0,1,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<4,0,1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(root_kernel<4,1,1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  {
    const int idx[9] = {0, 1, 2, 3, 4, 5, 6, 7, 8};
    set_cols(idx, 9, 1024);
    const double items = (double)n_pad * 9;
    const size_t s5 = (size_t)(1025 * 5 * 9 + 320) * 4, s4 = (size_t)(1025 * 4 * 9 + 320) * 4, s3 = (size_t)(1025 * 3 * 9 + 320) * 4;
    printf("--- root level, 9 continuous f32 columns, 1024 bins, one 1024-thread CTA per SM\n");
    report("no atomics, fp64 bin arithmetic (loads + arithmetic floor)", timeit([&]{ root_kernel<0,0,0><<<sms, 1024, s3>>>(a); }), items);
    report("no atomics, fp32 pre-filter + fp64 fallback", timeit([&]{ root_kernel<0,1,0><<<sms, 1024, s3>>>(a); }), items);
    report("5 atomics [cnt,y.lo,y.hi,yy.lo,yy.hi] stride 5, fp64 (round 1)", timeit([&]{ root_kernel<5,0,0><<<sms, 1024, s5>>>(a); }), items);
    report("4 atomics, planes [y.lo][y.hi][yy<<18|1 .lo][.hi], fp64", timeit([&]{ root_kernel<4,0,0><<<sms, 1024, s4>>>(a); }), items);
    report("4 atomics, planes, fp32 pre-filter", timeit([&]{ root_kernel<4,1,0><<<sms, 1024, s4>>>(a); }), items);
    report("3 atomics [y32][yy32][cnt|carries], fp64", timeit([&]{ root_kernel<3,0,0><<<sms, 1024, s3>>>(a); }), items);
    report("3 atomics, fp32 pre-filter", timeit([&]{ root_kernel<3,1,0><<<sms, 1024, s3>>>(a); }), items);
    CK(cudaFuncSetAttribute(root_kernel<0,0,0,2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    CK(cudaFuncSetAttribute(root_kernel<4,0,0,2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    CK(cudaFuncSetAttribute(root_kernel<3,0,0,2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    report("no atomics, fp64, TWO columns prefetched", timeit([&]{ root_kernel<0,0,0,2><<<sms, 1024, s3>>>(a); }), items);
    report("4 atomics, planes, fp64, TWO columns prefetched", timeit([&]{ root_kernel<4,0,0,2><<<sms, 1024, s4>>>(a); }), items);
    report("3 atomics, fp64, TWO columns prefetched", timeit([&]{ root_kernel<3,0,0,2><<<sms, 1024, s3>>>(a); }), items);
    report("4 atomics, warp-aggregated (thread-local sums + REDUX) on CONTINUOUS columns (worst case)", timeit([&]{ root_kernel<4,1,1><<<sms, 1024, s4>>>(a); }, 1), items);
  }
  {
    const int idx[1] = {9};
    set_cols(idx, 1, 1024);
    const double items = (double)n_pad;
    const size_t s5 = (size_t)(1025 * 5 + 320) * 4;
    printf("--- root level, ONE three-valued column (p = .5/.3/.2), 1024 bins\n");
    report("5 atomics stride 5, fp64 (round 1, no copies)", timeit([&]{ root_kernel<5,0,0><<<sms, 1024, s5>>>(a); }), items);
    report("4 atomics planes, fp64", timeit([&]{ root_kernel<4,0,0><<<sms, 1024, s5>>>(a); }), items);
    report("4 atomics planes, warp-aggregated (thread-local sums + REDUX limbs), fp64", timeit([&]{ root_kernel<4,0,1><<<sms, 1024, s5>>>(a); }), items);
    report("4 atomics planes, warp-aggregated, fp32 pre-filter", timeit([&]{ root_kernel<4,1,1><<<sms, 1024, s5>>>(a); }), items);
    const int idx2[1] = {0};
    set_cols(idx2, 1, 1024);
    printf("--- root level, ONE continuous column, 1024 bins\n");
    report("5 atomics stride 5, fp64 (round 1)", timeit([&]{ root_kernel<5,0,0><<<sms, 1024, s5>>>(a); }), items);
    report("4 atomics planes, fp32 pre-filter", timeit([&]{ root_kernel<4,1,0><<<sms, 1024, s5>>>(a); }), items);
    report("no atomics, fp32 pre-filter", timeit([&]{ root_kernel<0,1,0><<<sms, 1024, s5>>>(a); }), items);
  }
  {
    const int idx[7] = {0, 1, 2, 3, 4, 5, 6};
    set_cols(idx, 7, 256);
    const double items = (double)n_pad * 7;
    printf("--- level 5: 32 leaves x 257 bins, one column per CTA (7 columns x 21 CTAs), records in shared memory\n");
    CK(cudaFuncSetAttribute(deep_kernel<5,6,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    CK(cudaFuncSetAttribute(deep_kernel<4,6,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    CK(cudaFuncSetAttribute(deep_kernel<4,4,0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    CK(cudaFuncSetAttribute(deep_kernel<4,4,1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    const size_t s5 = (size_t)(257 * 5 * 32 + 12 * 32 + 128) * 4, s4 = (size_t)(257 * 4 * 32 + 12 * 32 + 128) * 4;
    report("5 atomics, six 32-bit record loads, value test for min/max (round 1)", timeit([&]{ deep_kernel<5,6,0><<<147, 1024, s5>>>(a, 7); }), items);
    report("4 atomics, six 32-bit record loads", timeit([&]{ deep_kernel<4,6,0><<<147, 1024, s4>>>(a, 7); }), items);
    report("4 atomics, one 128-bit record load, bin-range filter for min/max, fp64", timeit([&]{ deep_kernel<4,4,0><<<147, 1024, s4>>>(a, 7); }), items);
    report("4 atomics, one 128-bit record load, bin-range filter, fp32 pre-filter", timeit([&]{ deep_kernel<4,4,1><<<147, 1024, s4>>>(a, 7); }), items);
    a.L = 4;
    gen_y<<<sms * 4, 256>>>(y, code, n_pad, 4);
    CK(cudaDeviceSynchronize());
    printf("--- level 2: 4 leaves x 257 bins (same kernels)\n");
    report("5 atomics, six 32-bit record loads (round 1)", timeit([&]{ deep_kernel<5,6,0><<<147, 1024, s5>>>(a, 7); }), items);
    report("4 atomics, one 128-bit record load, bin-range filter, fp32 pre-filter", timeit([&]{ deep_kernel<4,4,1><<<147, 1024, s4>>>(a, 7); }), items);
  }
  unsigned long long h; CK(cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost));
  printf("checksum %llu\n", h);
  return 0;
}
