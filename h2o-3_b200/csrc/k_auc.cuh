// AUC2.AUCBuilder on the device (h2o-core/src/main/java/hex/AUC2.java:356-581): the reference's streaming 400-bin histogram of
// (prediction, actual) pairs, whose merge rule defines the AUC the model reports (north star: "final logloss / AUC within 1e-5").
// The builder is sequential by construction -- every row is inserted in sorted position and the two neighbouring bins whose
// merge costs the least squared error are fused when the histogram is full -- so, like the reference (one builder per Chunk,
// AUC_Impl.map, :343-350), the rows are cut into chunks of consecutive rows and every chunk gets its own builder: here one WARP
// per chunk, its bins in shared memory, the 32 lanes sharing the linear scans (find_smallest_impl) and the array shifts of an
// insertion. The chunks' builders are folded on the host in row order with AUCBuilder.reduce (host_gbm.cpp), as
// MRTask.reduce does for the reference. fp64 throughout, same operation order as the Java (the TU is compiled with -fmad=false).
#pragma once
#include "common.cuh"

namespace h2ot {

constexpr int kAucBins = 400;                    // AUC2.NBINS (:36)
constexpr int kAucCap = 2 * kAucBins;            // the builder's arrays are twice the bin count (:366-369)
constexpr int kAucWarpsPerCta = 4;               // 4 arrays x 800 doubles x 4 warps = 102 KB of shared memory per CTA

struct AucArgs {
  int64_t n, chunk_rows;
  int32_t n_chunks, family, nclass;
  const double* resp;
  const double* w;     // nullptr: unit weights
  const double* off;   // nullptr: none
  const double* F;     // Tree_0 (bernoulli: the link-space score)
  double* out;         // [n_chunks][1 + 4 * kAucBins]: n, then ths / tps / fps / sqe of the chunk's builder
};

struct AucWarp {  // one builder, arrays in shared memory, scalars warp-uniform in registers
  double* ths;
  double* sqe;
  double* tps;
  double* fps;
  int n, ssx;
};

__device__ __forceinline__ double auc_k(const AucWarp& b, int i) { return __dadd_rn(b.tps[i], b.fps[i]); }
// compute_delta_error (:568-578): thresholds closer than a float ulp count as equal
__device__ __forceinline__ double auc_delta_error(double ths1, double n1, double ths0, double n0) {
  const double delta = (double)((float)ths1 - (float)ths0);
  if (delta == 0) return 0.0;
  return __ddiv_rn(__dmul_rn(__dmul_rn(__dmul_rn(delta, delta), n0), n1), __dadd_rn(n0, n1));
}
// combine_centers (:474-481)
__device__ __forceinline__ double auc_combine(double ths1, double n1, double ths0, double n0) {
  const double c = __ddiv_rn(__dadd_rn(__dmul_rn(ths0, n0), __dmul_rn(ths1, n1)), __dadd_rn(n0, n1));
  if (isnan(c) || isinf(c)) return __ddiv_rn(__dadd_rn(ths0, ths1), 2.0);
  return c;
}
// Double.compare order (Arrays.binarySearch): -0.0 < 0.0, NaN never gets here
__device__ __forceinline__ int auc_jcompare(double a, double b) {
  if (a < b) return -1;
  if (a > b) return 1;
  const long long x = __double_as_longlong(a), y = __double_as_longlong(b);
  return x == y ? 0 : (x < y ? -1 : 1);
}

// find_smallest_impl (:523-554), the linear scan shared by the warp: the first pair with a zero error increase wins outright,
// otherwise the first pair with the smallest squared error after the merge, otherwise (nothing comparable: infinities / NaNs)
// the first pair of closest centres.
__device__ __forceinline__ int auc_find_smallest_impl(const AucWarp& b, int lane) {
  if (b.n == 1) return 0;
  int first_zero = 0x7fffffff, best_i = 0x7fffffff;
  double best = DBL_MAX;
  for (int i = lane; i < b.n - 1; i += 32) {
    const double derr = auc_delta_error(b.ths[i + 1], auc_k(b, i + 1), b.ths[i], auc_k(b, i));
    if (derr == 0) { first_zero = min(first_zero, i); continue; }
    const double s = __dadd_rn(__dadd_rn(b.sqe[i], b.sqe[i + 1]), derr);
    if (s < best) { best = s; best_i = i; }  // strict: the earliest index of a lane's own candidates stays
  }
  // The sequential scan returns at the first zero, and only pairs BEFORE it took part in the minimum -- but a zero always wins
  // ("return i"), so only the position of the first zero matters.
  first_zero = __reduce_min_sync(0xFFFFFFFFu, first_zero);
  if (first_zero != 0x7fffffff) return first_zero;
  for (int o = 16; o > 0; o >>= 1) {
    const double os = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    const int oi = __shfl_xor_sync(0xFFFFFFFFu, best_i, o);
    if (os < best || (os == best && oi < best_i)) { best = os; best_i = oi; }
  }
  if (best_i != 0x7fffffff) return best_i;
  int mi = 0x7fffffff;
  double md = DBL_MAX;
  bool any = false;
  for (int i = lane; i < b.n - 1; i += 32) {
    const double d = __dsub_rn(b.ths[i + 1], b.ths[i]);
    if (!any || d < md) { md = d; mi = i; any = true; }
  }
  // sequential rule: minI = 0 with ths[1] - ths[0], then strictly smaller distances replace it
  for (int o = 16; o > 0; o >>= 1) {
    const double od = __shfl_xor_sync(0xFFFFFFFFu, md, o);
    const int oi = __shfl_xor_sync(0xFFFFFFFFu, mi, o);
    if (oi != 0x7fffffff && (mi == 0x7fffffff || od < md || (od == md && oi < mi))) { md = od; mi = oi; }
  }
  return mi == 0x7fffffff ? 0 : mi;
}
__device__ __forceinline__ int auc_find_smallest(AucWarp& b, int lane) {
  if (b.ssx == -1) b.ssx = auc_find_smallest_impl(b, lane);
  return b.ssx;
}

// arrays[from .. n) move one slot up (insertion) or arrays[from + 1 .. n) one slot down (removal); the warp walks the range in
// the direction that never overwrites an element before it was read
__device__ __forceinline__ void auc_shift_up(AucWarp& b, int from, int lane) {
  for (int base = b.n - 1; base >= from; base -= 32) {
    const int i = base - lane;
    double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    if (i >= from) { t0 = b.ths[i]; t1 = b.sqe[i]; t2 = b.tps[i]; t3 = b.fps[i]; }
    __syncwarp();
    if (i >= from) { b.ths[i + 1] = t0; b.sqe[i + 1] = t1; b.tps[i + 1] = t2; b.fps[i + 1] = t3; }
    __syncwarp();
  }
}
__device__ __forceinline__ void auc_shift_down(AucWarp& b, int to, int lane) {  // element to + 1 lands on to, ...
  for (int base = to + 1; base < b.n; base += 32) {
    const int i = base + lane;
    double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    if (i < b.n) { t0 = b.ths[i]; t1 = b.sqe[i]; t2 = b.tps[i]; t3 = b.fps[i]; }
    __syncwarp();
    if (i < b.n) { b.ths[i - 1] = t0; b.sqe[i - 1] = t1; b.tps[i - 1] = t2; b.fps[i - 1] = t3; }
    __syncwarp();
  }
}

// mergeOneBin (:483-505)
__device__ __forceinline__ void auc_merge_one_bin(AucWarp& b, int lane) {
  const int ssx = auc_find_smallest(b, lane);
  if (lane == 0) {
    const double k0 = auc_k(b, ssx), k1 = auc_k(b, ssx + 1);
    b.sqe[ssx] = __dadd_rn(__dadd_rn(b.sqe[ssx], b.sqe[ssx + 1]), auc_delta_error(b.ths[ssx + 1], k1, b.ths[ssx], k0));
    b.ths[ssx] = auc_combine(b.ths[ssx], k0, b.ths[ssx + 1], k1);
    b.tps[ssx] = __dadd_rn(b.tps[ssx], b.tps[ssx + 1]);
    b.fps[ssx] = __dadd_rn(b.fps[ssx], b.fps[ssx + 1]);
  }
  __syncwarp();
  auc_shift_down(b, ssx + 1, lane);
  b.n--;
  b.ssx = -1;
}

// perRow (:384-446). Every lane runs the scalar logic on the same values (shared-memory reads are broadcasts); lane 0 writes.
__device__ __forceinline__ void auc_per_row(AucWarp& b, double pred, int act, double w, int lane) {
  int low = 0, high = b.n - 1, idx = 0;
  bool found = false;
  while (low <= high) {
    const int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
    const int c = auc_jcompare(b.ths[mid], pred);
    if (c < 0) low = mid + 1;
    else if (c > 0) high = mid - 1;
    else { idx = mid; found = true; break; }
  }
  if (found) {
    if (lane == 0) {
      if (act == 0) b.fps[idx] = __dadd_rn(b.fps[idx], w);
      else b.tps[idx] = __dadd_rn(b.tps[idx], w);
    }
    __syncwarp();
    b.ssx = -1;
    return;
  }
  idx = low;
  if (b.n == kAucBins && idx > 0 && idx < b.n && b.ths[idx - 1] != b.ths[idx]) {
    const int ssx = auc_find_smallest(b, lane);
    const double dssx = __dadd_rn(__dadd_rn(b.sqe[ssx], b.sqe[ssx + 1]), auc_delta_error(b.ths[ssx + 1], auc_k(b, ssx + 1), b.ths[ssx], auc_k(b, ssx)));
    const double d0 = __dadd_rn(b.sqe[idx - 1], auc_delta_error(pred, w, b.ths[idx - 1], auc_k(b, idx - 1)));
    const double d1 = __dadd_rn(b.sqe[idx], auc_delta_error(b.ths[idx], auc_k(b, idx), pred, w));
    if (d0 < dssx || d1 < dssx) {
      if (d0 <= d1) idx--;
      if (ssx == idx - 1 || ssx == idx) b.ssx = -1;
      const double kk = auc_k(b, idx);
      __syncwarp();
      if (lane == 0) {
        if (act == 0) b.fps[idx] = __dadd_rn(b.fps[idx], w);
        else b.tps[idx] = __dadd_rn(b.tps[idx], w);
        b.sqe[idx] = __dadd_rn(b.sqe[idx], auc_delta_error(pred, w, b.ths[idx], kk));
        b.ths[idx] = auc_combine(b.ths[idx], kk, pred, w);
      }
      __syncwarp();
      return;
    }
  }
  if (idx == 0 || idx == b.n || idx == b.ssx) b.ssx = -1;
  else if (idx < b.ssx) b.ssx++;
  __syncwarp();
  auc_shift_up(b, idx, lane);
  if (lane == 0) {
    b.ths[idx] = pred;
    b.sqe[idx] = 0.0;
    b.tps[idx] = act == 0 ? 0.0 : w;
    b.fps[idx] = act == 0 ? w : 0.0;
  }
  __syncwarp();
  b.n++;
  if (b.n > kAucBins) auc_merge_one_bin(b, lane);
}

// One warp per chunk of consecutive rows. Bernoulli only (AUC2 is the binomial metric): pred = p1 = linkInv(f) as in
// score_metrics_kernel; rows without a response, with a response other than 0 / 1, or with weight 0 / NaN are skipped
// (ModelMetricsBinomial.MetricBuilderBinomial.perRow, h2o-core/src/main/java/hex/ModelMetricsBinomial.java).
__global__ void __launch_bounds__(32 * kAucWarpsPerCta) auc_build_kernel(AucArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunk = blockIdx.x * kAucWarpsPerCta + warp;
  if (chunk >= a.n_chunks) return;
  double* base = reinterpret_cast<double*>(smem_raw) + (size_t)warp * 4 * (kAucCap + 1);
  AucWarp b;
  b.ths = base;
  b.sqe = base + (kAucCap + 1);
  b.tps = base + 2 * (kAucCap + 1);
  b.fps = base + 3 * (kAucCap + 1);
  b.n = 0;
  b.ssx = -1;
  const int64_t r0 = (int64_t)chunk * a.chunk_rows, r1 = min(a.n, r0 + a.chunk_rows);
  for (int64_t rb = r0; rb < r1; rb += 32) {
    // the lanes fetch 32 rows at once, then the warp inserts them one after the other in row order
    const int64_t row = rb + lane;
    double pred = 0, wv = 0;
    int act = -1;
    if (row < r1) {
      const double y = a.resp[row];
      wv = a.w ? a.w[row] : 1.0;
      if (!isnan(y) && !(wv == 0.0 || isnan(wv))) {
        const int iact = (int)y;
        if (iact == 0 || iact == 1) {
          const double f = a.F[row] + (a.off ? a.off[row] : 0.0);
          pred = 1.0 / (1.0 + fmin(1e19, exp(-f)));
          act = isnan(pred) ? -1 : iact;
        }
      }
    }
    const unsigned valid = __ballot_sync(0xFFFFFFFFu, act >= 0);
    for (unsigned m = valid; m; m &= m - 1) {
      const int src = __ffs(m) - 1;
      const double p = __shfl_sync(0xFFFFFFFFu, pred, src);
      const double ww = __shfl_sync(0xFFFFFFFFu, wv, src);
      const int ac = __shfl_sync(0xFFFFFFFFu, act, src);
      auc_per_row(b, p, ac, ww, lane);
    }
  }
  __syncwarp();
  double* o = a.out + (size_t)chunk * (1 + 4 * kAucBins);
  if (lane == 0) o[0] = (double)b.n;
  for (int i = lane; i < kAucBins; i += 32) {
    const bool in = i < b.n;
    o[1 + i] = in ? b.ths[i] : 0.0;
    o[1 + kAucBins + i] = in ? b.tps[i] : 0.0;
    o[1 + 2 * kAucBins + i] = in ? b.fps[i] : 0.0;
    o[1 + 3 * kAucBins + i] = in ? b.sqe[i] : 0.0;
  }
}

}  // namespace h2ot
