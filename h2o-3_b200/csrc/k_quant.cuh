// Per-leaf quantiles of the residuals y - (f + offset): what Quantile.StratifiedQuantilesTask computes for GBM's laplace / quantile
// leaf fitting (T/gbm/GBM.java:729-751, ComputeDiff :1132-1147; h2o-core/src/main/java/hex/quantile/Quantile.java:118-175) and, with
// one stratum and no tree, getInitialValueQuantile (:355-381). The reference homes in on the two order statistics around the
// fractional row prob * (n - 1) by refining a 1024-bin histogram until the bins that hold them are constant (Histo.findQuantile /
// refinePass, Quantile.java:267-325) and then interpolates (linearInterpolate, :344-350); the order statistics themselves do not
// depend on how they are found, so here the rows are sorted (stratum, value) once and the two values are read off.
#pragma once
#include "common.cuh"

namespace h2ot {

struct QuantKeysArgs {
  int64_t n, n_pad;
  int32_t leaf_start, n_strata, use_tree;
  const double* resp;
  const double* w;
  const double* off;
  const double* F;      // Tree_k (use_tree) or nullptr
  const int32_t* nids;  // rows' leaf ids after GammaPass (use_tree) or nullptr: every row is stratum 0
  uint64_t* keys;       // [n_pad] order-preserving key of the row's value; all-ones for rows that take no part
  uint32_t* strata;     // [n_pad] stratum, n_strata for rows that take no part
  unsigned long long* counts;  // [n_strata + 1]
};

__global__ void __launch_bounds__(256) quant_keys_kernel(QuantKeysArgs a) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n_pad; row += (int64_t)gridDim.x * blockDim.x) {
    uint32_t s = (uint32_t)a.n_strata;
    uint64_t key = ~0ULL;
    if (row < a.n) {
      const double y = a.resp[row];
      const double wv = a.w ? a.w[row] : 1.0;
      int st = 0;
      if (a.use_tree) {
        const int nid = a.nids[row];
        st = nid - a.leaf_start;
      }
      if (!isnan(y) && wv != 0.0 && !isnan(wv) && st >= 0 && st < a.n_strata) {
        const double d = a.use_tree ? __dsub_rn(y, __dadd_rn(a.F[row], a.off ? a.off[row] : 0.0)) : (a.off ? __dsub_rn(y, a.off[row]) : y);
        if (!isnan(d)) {
          s = (uint32_t)st;
          key = dkey(d);
        }
      }
    }
    a.keys[row] = key;
    a.strata[row] = s;
    // rows of a warp mostly share a few strata: one atomic per distinct stratum in the warp
    const unsigned peers = __match_any_sync(__activemask(), s);
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&a.counts[s], (unsigned long long)__popc(peers));
  }
}

__global__ void quant_gather_kernel(const uint64_t* sorted_keys, const long long* idx, int n, uint64_t* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = idx[i] >= 0 ? sorted_keys[idx[i]] : 0ULL;
}

}  // namespace h2ot
