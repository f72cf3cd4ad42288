// Cross-GPU merge of a level's accumulators over NVLink peer memory: replaces MRTask.reduce -> ScoreBuildHistogram.reduce ->
// DHistogram.add (T/ScoreBuildHistogram.java:74-88, T/DHistogram.java:434-447) for the row-sharded build.
//
// Every rank accumulates its level straight into an exchange buffer that its peers have mapped (CUDA IPC), laid out as
// [int64 sums | u64 min/max keys]. The merge is a two-shot all-reduce written as plain kernels over peer pointers:
//   barrier -> reduce-scatter (rank r pulls slice r of every peer's input, adds the sums / takes the max of the keys,
//   keeps the result in its own "out" half) -> barrier -> all-gather (every rank pulls every slice from its owner).
// One pass handles both reduction operators (NCCL needs two collectives), the payload is integer so the result does not
// depend on the order ranks are visited, and nothing is copied before or after: the histogram kernel writes the input in
// place and the finalize / split-search kernels read the result in place.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace h2ot {

constexpr int kXrMaxWorld = 16;

struct XrPeers {
  unsigned long long* buf[kXrMaxWorld];  // peer exchange buffers as mapped in THIS process ([in: cap][out: cap][flags])
};

// System-scope barrier between the ranks' GPUs: one CTA, thread p talks to peer p. flags live at the end of each rank's
// buffer: flags[src_rank] = last epoch src_rank has reached. A rank that waits longer than the timeout records an error
// instead of hanging the device.
__global__ void __launch_bounds__(32) xr_barrier_kernel(XrPeers peers, int world, int rank, size_t flag_off, unsigned long long epoch,
                                                        unsigned long long timeout_cycles, int* err) {
  const int p = threadIdx.x;
  if (p >= world) return;
  __threadfence_system();
  unsigned long long* remote = peers.buf[p] + flag_off + rank;
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(remote), "l"(epoch) : "memory");
  const unsigned long long* local = peers.buf[rank] + flag_off + p;
  const long long t0 = clock64();
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(local) : "memory");
    if (v >= epoch) break;
    if ((unsigned long long)(clock64() - t0) > timeout_cycles) {
      atomicExch(err, 1);
      break;
    }
    __nanosleep(200);
  }
}

// reduce-scatter: rank r owns words [r*n/world, (r+1)*n/world); words below n_sums are added, the rest are max-merged.
__global__ void __launch_bounds__(256) xr_reduce_scatter_kernel(XrPeers peers, int world, int rank, size_t cap, long long n_sums, long long n) {
  const long long lo = n * rank / world, hi = n * (rank + 1) / world;
  unsigned long long* out = peers.buf[rank] + cap;
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long acc = 0ULL;
    if (i < n_sums) {
      for (int p = 0; p < world; p++) acc += __ldcv(peers.buf[p] + i);
    } else {
      for (int p = 0; p < world; p++) {
        const unsigned long long v = __ldcv(peers.buf[p] + i);
        acc = v > acc ? v : acc;
      }
    }
    out[i] = acc;
  }
}

// all-gather: every word comes from the "out" half of the rank that owns its slice, into this rank's input half.
__global__ void __launch_bounds__(256) xr_all_gather_kernel(XrPeers peers, int world, int rank, size_t cap, long long n) {
  unsigned long long* in = peers.buf[rank];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    // owner o: largest o with n*o/world <= i
    int o = (int)(((i + 1) * world - 1) / n);
    while (n * o / world > i) o--;
    while (o + 1 < world && n * (o + 1) / world <= i) o++;
    in[i] = __ldcv(peers.buf[o] + cap + i);
  }
}

// The whole exchange as one cooperative launch (grid <= co-resident CTAs): barrier, reduce-scatter, barrier, all-gather, with
// grid-wide syncs between the phases. Saves three launches and their gaps per level over the four-kernel sequence.
__device__ __forceinline__ void xr_signal_wait(const XrPeers& peers, int world, int rank, size_t flag_off, unsigned long long epoch,
                                               unsigned long long timeout_cycles, int* err) {
  const int p = threadIdx.x;
  if (p >= world) return;
  __threadfence_system();
  unsigned long long* remote = peers.buf[p] + flag_off + rank;
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(remote), "l"(epoch) : "memory");
  const unsigned long long* local = peers.buf[rank] + flag_off + p;
  const long long t0 = clock64();
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(local) : "memory");
    if (v >= epoch) break;
    if ((unsigned long long)(clock64() - t0) > timeout_cycles) {
      atomicExch(err, 1);
      break;
    }
    __nanosleep(100);
  }
}

// (Remote loads take microseconds: every loop below issues all the loads of an iteration before it uses any of them -- one round
// trip per iteration instead of one per peer.)
__global__ void __launch_bounds__(256) xr_fused_kernel(XrPeers peers, int world, int rank, size_t cap, size_t flag_off, long long n_sums,
                                                       long long n, unsigned long long epoch, unsigned long long timeout_cycles, int* err) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  __shared__ long long s_lo[kXrMaxWorld + 1];
  if (threadIdx.x <= world) s_lo[threadIdx.x] = n * threadIdx.x / world;
  if (blockIdx.x == 0) xr_signal_wait(peers, world, rank, flag_off, epoch, timeout_cycles, err);
  grid.sync();
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
  {
    const long long lo = s_lo[rank], hi = s_lo[rank + 1];
    unsigned long long* out = peers.buf[rank] + cap;
    for (long long i = lo + tid; i < hi; i += stride) {
      unsigned long long v[kXrMaxWorld];
#pragma unroll
      for (int p = 0; p < kXrMaxWorld; p++) v[p] = p < world ? __ldcv(peers.buf[p] + i) : 0ULL;
      unsigned long long acc = 0ULL;
      if (i < n_sums) {
#pragma unroll
        for (int p = 0; p < kXrMaxWorld; p++) acc += v[p];
      } else {
#pragma unroll
        for (int p = 0; p < kXrMaxWorld; p++) acc = v[p] > acc ? v[p] : acc;
      }
      out[i] = acc;
    }
  }
  __threadfence_system();
  grid.sync();
  if (blockIdx.x == 0) xr_signal_wait(peers, world, rank, flag_off, epoch + 1, timeout_cycles, err);
  grid.sync();
  {
    unsigned long long* in = peers.buf[rank];
    constexpr int U = 4;
    for (long long i0 = tid; i0 < n; i0 += U * stride) {
      unsigned long long v[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const long long i = i0 + u * stride;
        if (i < n) {
          int o = 0;
#pragma unroll
          for (int q = 1; q < kXrMaxWorld; q++) o += (q < world && i >= s_lo[q]) ? 1 : 0;
          v[u] = __ldcv(peers.buf[o] + cap + i);
        }
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        const long long i = i0 + u * stride;
        if (i < n) in[i] = v[u];
      }
    }
  }
}


// ---- The exchange without barriers or fences: every 64-bit word travels as a 16-byte line {lo, flag, hi, flag} written with ONE
// store, flag = the exchange's epoch (the two 8-byte halves of an aligned 16-byte store are each written atomically, so a reader
// that sees both flags has the whole word). Ranks PUSH: (A) rank r stores word i of its input into the owner o(i)'s reduce-scatter
// area, slot [r][i - lo(o)]; (B) the owner polls the W - 1 lines of each of its words, reduces, keeps the result and stores it into
// every peer's all-gather area, slot [i]; (C) everybody polls the words it does not own. No grid-wide sync, no system-scope fence,
// a plain launch; latency = two NVLink store flights plus the polls. Twice the bytes on the wire, which at 0.7-5.5 MB per level is
// cheaper than the three barriers of the pull version (tools/xr_bench.py).
// Buffers alternate by epoch parity: a rank can only be one exchange ahead of a peer (it needs the peer's lines to finish one), so
// a line is never overwritten before it was read. A stale line carries an older epoch and is never mistaken for the current one.
struct __align__(16) XrLine {
  uint32_t d0, f0, d1, f1;
};
__device__ __forceinline__ void xr_line_store(XrLine* dst, unsigned long long v, uint32_t flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"((uint32_t)v), "r"(flag), "r"((uint32_t)(v >> 32)), "r"(flag)
               : "memory");
}
__device__ __forceinline__ bool xr_line_load(const XrLine* src, uint32_t flag, unsigned long long& v) {
  uint32_t d0, f0, d1, f1;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(d0), "=r"(f0), "=r"(d1), "=r"(f1) : "l"(src) : "memory");
  v = ((unsigned long long)d1 << 32) | d0;
  return f0 == flag && f1 == flag;
}

// rs_off / ag_off: word offsets of the two line areas inside every rank's buffer; half: words per parity of an area.
// The grid must be co-resident (the pollers of one rank wait for the pushers of another): <= 2 CTAs per SM.
__global__ void __launch_bounds__(256) xr_ll_kernel(XrPeers peers, int world, int rank, size_t rs_off, size_t ag_off, size_t half,
                                                    long long n_sums, long long n, uint32_t epoch, unsigned long long timeout_cycles, int* err) {
  __shared__ long long s_lo[kXrMaxWorld + 1];
  if (threadIdx.x <= world) s_lo[threadIdx.x] = n * threadIdx.x / world;
  __syncthreads();
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
  const size_t par = (size_t)(epoch & 1u) * half;
  const long long sl = (n + world - 1) / world;  // lines per sender in a reduce-scatter area
  unsigned long long* in = peers.buf[rank];
  const long long t0 = clock64();
  bool dead = false;
  // ---- A: push the words other ranks own
  for (long long i = tid; i < n; i += stride) {
    int o = 0;
#pragma unroll
    for (int q = 1; q < kXrMaxWorld; q++) o += (q < world && i >= s_lo[q]) ? 1 : 0;
    if (o == rank) continue;
    XrLine* dst = reinterpret_cast<XrLine*>(peers.buf[o] + rs_off + par) + (size_t)rank * sl + (size_t)(i - s_lo[o]);
    xr_line_store(dst, in[i], epoch);
  }
  // ---- B: reduce the words this rank owns, publish them
  {
    const long long lo = s_lo[rank], hi = s_lo[rank + 1];
    const XrLine* rs = reinterpret_cast<const XrLine*>(in + rs_off + par);
    for (long long i = lo + tid; i < hi && !dead; i += stride) {
      unsigned long long acc = in[i];
      const bool is_sum = i < n_sums;
      unsigned todo = ((1u << world) - 1u) & ~(1u << rank);
      while (todo) {
#pragma unroll
        for (int p = 0; p < kXrMaxWorld; p++) {
          if (!((todo >> p) & 1u)) continue;
          unsigned long long v;
          if (xr_line_load(rs + (size_t)p * sl + (size_t)(i - lo), epoch, v)) {
            acc = is_sum ? acc + v : (v > acc ? v : acc);
            todo &= ~(1u << p);
          }
        }
        if (todo && (unsigned long long)(clock64() - t0) > timeout_cycles) {
          atomicExch(err, 1);
          dead = true;
          break;
        }
      }
      in[i] = acc;
#pragma unroll
      for (int p = 0; p < kXrMaxWorld; p++)
        if (p < world && p != rank) xr_line_store(reinterpret_cast<XrLine*>(peers.buf[p] + ag_off + par) + (size_t)i, acc, epoch);
    }
  }
  // ---- C: collect the words the others own
  {
    const XrLine* ag = reinterpret_cast<const XrLine*>(in + ag_off + par);
    const long long lo = s_lo[rank], hi = s_lo[rank + 1];
    constexpr int U = 4;
    for (long long i0 = tid; i0 < n && !dead; i0 += U * stride) {
      unsigned todo = 0u;
#pragma unroll
      for (int u = 0; u < U; u++) {
        const long long i = i0 + u * stride;
        if (i < n && (i < lo || i >= hi)) todo |= 1u << u;
      }
      while (todo) {
#pragma unroll
        for (int u = 0; u < U; u++) {
          if (!((todo >> u) & 1u)) continue;
          const long long i = i0 + u * stride;
          unsigned long long v;
          if (xr_line_load(ag + (size_t)i, epoch, v)) {
            in[i] = v;
            todo &= ~(1u << u);
          }
        }
        if (todo && (unsigned long long)(clock64() - t0) > timeout_cycles) {
          atomicExch(err, 1);
          dead = true;
          break;
        }
      }
    }
  }
}

}  // namespace h2ot
