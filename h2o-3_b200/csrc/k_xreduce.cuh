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

__global__ void __launch_bounds__(256) xr_fused_kernel(XrPeers peers, int world, int rank, size_t cap, size_t flag_off, long long n_sums,
                                                       long long n, unsigned long long epoch, unsigned long long timeout_cycles, int* err) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  if (blockIdx.x == 0) xr_signal_wait(peers, world, rank, flag_off, epoch, timeout_cycles, err);
  grid.sync();
  {
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
  __threadfence_system();
  grid.sync();
  if (blockIdx.x == 0) xr_signal_wait(peers, world, rank, flag_off, epoch + 1, timeout_cycles, err);
  grid.sync();
  {
    unsigned long long* in = peers.buf[rank];
    for (int o = 0; o < world; o++) {
      const long long lo = n * o / world, hi = n * (o + 1) / world;
      const unsigned long long* src = peers.buf[o] + cap;
      for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) in[i] = __ldcv(src + i);
    }
  }
}

}  // namespace h2ot
