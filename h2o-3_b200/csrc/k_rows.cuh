// Per-tree row passes and frame ingest kernels.
//  - rollup / encode : W/fvec/RollupStats.java:88-200 (min/max/isInt/naCnt) + lossless narrow storage
//  - init            : GBM.initializeModelSpecifics / MarkDecidedRows, T/gbm/GBM.java:282-349
//  - residuals       : GBM.ComputePredAndRes, T/gbm/GBM.java:1032-1081 (+ score1static :1486-1513, log_rescale G/GenModel.java:671-681)
//  - sample          : T/Sample.java:22-36 on PCG (W/util/RandomUtils.java:50-119)
//  - gamma           : GBM.GammaPass, T/gbm/GBM.java:1286-1407
//  - add_tree        : GBM.AddTreeContributions, T/gbm/GBM.java:1410-1470
#pragma once
#include "common.cuh"
#include "k_level.cuh"

namespace h2ot {

// ------------------------------------------------------------------------------------------ ingest
struct RollupDev {
  unsigned long long min_key;  // ~dkey(min) max-merged ; 0 == none
  unsigned long long max_key;  // dkey(max)
  unsigned long long nacnt;
  unsigned int not_int;        // != 0 if some value is not integral
  unsigned int not_f32;        // != 0 if some value is not fp32-exact
};

__global__ void __launch_bounds__(256) rollup_kernel(const double* v, int64_t n, RollupDev* out) {
  unsigned long long mn = 0, mx = 0, na = 0;
  unsigned int ni = 0, nf = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double d = v[i];
    if (isnan(d)) { na++; continue; }
    const unsigned long long k = dkey(d);
    if (~k > mn) mn = ~k;
    if (k > mx) mx = k;
    // RollupStats._isInt: ((long)d) == d  (Java saturating cast)
    const long long l = __double2ll_rz(d);
    if ((double)l != d) ni = 1;
    if ((double)(float)d != d) nf = 1;
  }
  // warp reduce then one atomic per warp
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long t;
    t = __shfl_xor_sync(0xffffffffu, mn, o); if (t > mn) mn = t;
    t = __shfl_xor_sync(0xffffffffu, mx, o); if (t > mx) mx = t;
    na += __shfl_xor_sync(0xffffffffu, na, o);
    ni |= __shfl_xor_sync(0xffffffffu, ni, o);
    nf |= __shfl_xor_sync(0xffffffffu, nf, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (mn) atomicMax(&out->min_key, mn);
    if (mx) atomicMax(&out->max_key, mx);
    if (na) atomicAdd(&out->nacnt, na);
    if (ni) atomicOr(&out->not_int, 1u);
    if (nf) atomicOr(&out->not_f32, 1u);
  }
}

// n_pad >= n : pad rows are written as NA
__global__ void __launch_bounds__(256) encode_kernel(const double* v, int64_t n, int64_t n_pad, int storage, double base, void* out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
    const double d = i < n ? v[i] : __longlong_as_double(0x7ff8000000000000LL);
    switch (storage) {
      case H2OT_STORE_F32: reinterpret_cast<float*>(out)[i] = (float)d; break;
      case H2OT_STORE_F64: reinterpret_cast<double*>(out)[i] = d; break;
      case H2OT_STORE_U8: reinterpret_cast<unsigned char*>(out)[i] = isnan(d) ? 255 : (unsigned char)(int)(d - base); break;
      default: reinterpret_cast<unsigned short*>(out)[i] = isnan(d) ? 65535 : (unsigned short)(int)(d - base); break;
    }
  }
}

// ------------------------------------------------------------------------------------------ init
struct InitArgs {
  int64_t n, n_pad;
  int32_t K;
  const double* resp;
  const double* w;
  int32_t* const* nids;  // [K]
  unsigned long long* skipped;
};
__global__ void __launch_bounds__(256) init_rows_kernel(InitArgs a) {
  unsigned long long cnt = 0;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n_pad; row += (int64_t)gridDim.x * blockDim.x) {
    int v = H2OT_DECIDED_ROW;
    if (row < a.n) {
      const double wv = a.w ? a.w[row] : 1.0;
      const bool skip = (wv == 0.0) || isnan(a.resp[row]);
      v = skip ? H2OT_DECIDED_ROW : 0;
      if (skip) cnt++;
    }
    for (int k = 0; k < a.K; k++) a.nids[k][row] = v;
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(a.skipped, cnt);
}

// ------------------------------------------------------------------------------------------ initial value
// SharedTree.InitialValue (T/SharedTree.java:1138-1185): sum w (y - offset) and sum w over rows with a response and w != 0,
// plus what DRF's set-up needs: weight per class and max |(float) y|. Block partials are written out and summed by the host in
// block order, so the result does not depend on scheduling.
__device__ __forceinline__ double exp_cap(double x) { return fmin(1e19, exp(x)); }  // LogExpUtil.exp (H/DistributionFactory.java:534)

struct InitStatsArgs {
  int64_t n;
  int32_t nclass, family;
  double tweedie_power;
  const double* resp;
  const double* w;
  const double* off;
  double* partial;         // [gridDim.x][2] : num, den
  double* class_w;         // [nclass] (nclass > 1), atomically accumulated
  unsigned int* ymax_bits; // float bits of max |(float) y|
};
constexpr int kInitStatsClasses = 1024;  // classes with block-private sums (power of two); nclass above that is rejected by the host
__global__ void __launch_bounds__(256) init_stats_kernel(InitStatsArgs a) {
  __shared__ double s_cls[kInitStatsClasses];
  if (a.class_w)
    for (int i = threadIdx.x; i < kInitStatsClasses; i += blockDim.x) s_cls[i] = 0.0;
  __syncthreads();
  double num = 0, den = 0;
  float ymax = 0.f;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (int64_t)gridDim.x * blockDim.x) {
    const double y = a.resp[row];
    if (isnan(y)) continue;
    const double wv = a.w ? a.w[row] : 1.0;
    ymax = fmaxf(ymax, fabsf((float)y));
    if (a.class_w && wv != 0.0) {  // class prior (DRF only): block-private sums, flushed once per block
      const int c = (int)y;
      if (c >= 0 && c < a.nclass) atomicAdd(&s_cls[c & (kInitStatsClasses - 1)], wv);
    }
    if (wv == 0.0) continue;
    const double o = a.off ? a.off[row] : 0.0;
    // Distribution.initFNum / initFDenom (H/DistributionFactory.java)
    if (a.family == H2OT_FAMILY_POISSON) {
      num += __dmul_rn(wv, y);
      den += __dmul_rn(wv, exp_cap(o));
    } else if (a.family == H2OT_FAMILY_GAMMA) {
      num += __dmul_rn(__dmul_rn(wv, y), exp_cap(-o));
      den += wv;
    } else if (a.family == H2OT_FAMILY_TWEEDIE) {
      num += __dmul_rn(__dmul_rn(wv, y), exp_cap(__dmul_rn(o, 1.0 - a.tweedie_power)));
      den += __dmul_rn(wv, exp_cap(__dmul_rn(o, 2.0 - a.tweedie_power)));
    } else {
      num += __dmul_rn(wv, __dsub_rn(y, o));
      den += wv;
    }
  }
  __shared__ double s_num[256], s_den[256];
  s_num[threadIdx.x] = num;
  s_den[threadIdx.x] = den;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) {
      s_num[threadIdx.x] += s_num[threadIdx.x + o];
      s_den[threadIdx.x] += s_den[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    a.partial[2 * blockIdx.x] = s_num[0];
    a.partial[2 * blockIdx.x + 1] = s_den[0];
  }
  for (int o = 16; o > 0; o >>= 1) ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
  if ((threadIdx.x & 31) == 0 && ymax > 0.f) atomicMax(a.ymax_bits, __float_as_uint(ymax));
  if (a.class_w)
    for (int i = threadIdx.x; i < a.nclass && i < kInitStatsClasses; i += blockDim.x)
      if (s_cls[i] != 0.0) atomicAdd(&a.class_w[i], s_cls[i]);
}

// ------------------------------------------------------------------------------------------ residuals
struct ResidArgs {
  int64_t n;
  int32_t K, nclass, family;
  const double* resp;
  const double* w;
  const double* off;
  const double* const* F;  // [K] Tree_k (fp64: holds the double init_f until a row's first update, float valued after)
  float* const* work;      // [K]
  unsigned int* absmax;   // [K + 2] float bits of max |work| per class-tree, then max |gammaNum / w| and max |gammaDenom / w|
  double tweedie_power;
};

__device__ __forceinline__ double tree_val(const ResidArgs& a, int k, int64_t row) {
  return a.F[k][row];
}

__global__ void __launch_bounds__(256) residuals_kernel(ResidArgs a) {
  float amaxk = 0.f, gnum = 0.f, gden = 0.f;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (int64_t)gridDim.x * blockDim.x) {
    const double wv = a.w ? a.w[row] : 1.0;
    if (wv == 0.0) continue;
    const double y = a.resp[row];
    if (isnan(y)) continue;
    const double f = tree_val(a, 0, row) + (a.off ? a.off[row] : 0.0);
    if (a.family == H2OT_FAMILY_MULTINOMIAL) {
      // score1static: fs[1] = f ; fs[k+1] = Tree_k ; log_rescale: max, exp(fs - max), sum in k order
      double maxval = f;
      for (int k = 1; k < a.nclass; k++) maxval = fmax(maxval, tree_val(a, k, row));
      double dsum = 0;
      for (int k = 0; k < a.nclass; k++) dsum += exp((k == 0 ? f : tree_val(a, k, row)) - maxval);
      const int yi = (int)y;
      for (int k = 0; k < a.nclass; k++) {
        const double e = exp((k == 0 ? f : tree_val(a, k, row)) - maxval);
        float r;
        if (isinf(dsum)) r = (float)((double)(yi == k ? 1.0f : 0.0f) - (double)(isinf(e) ? 1.0f : 0.0f));
        else r = (float)((double)(yi == k ? 1.0f : 0.0f) - (double)(float)(e / dsum));
        a.work[k][row] = r;
        amaxk = fmaxf(amaxk, fabsf(r));
      }
    } else {
      double r;
      // per-row magnitude of the GammaPass terms (upper bounds, rounded up): they scale the fixed-point leaf sums
      double tn = 0, td = 1.0;
      if (a.family == H2OT_FAMILY_BERNOULLI) r = y - 1.0 / (1.0 + fmin(1e19, exp(-f)));  // LogitFunction.linkInv
      else if (a.family == H2OT_FAMILY_POISSON) {
        const double ef = exp_cap(f);
        r = y - ef;
        tn = fabs(y);
        td = fabs(y) + fabs((double)(float)r) + 1.0;  // gammaDenom = w (y - z), z the float-rounded residual
      } else if (a.family == H2OT_FAMILY_GAMMA) {
        r = __dsub_rn(__dmul_rn(y, exp_cap(-f)), 1.0);
      } else if (a.family == H2OT_FAMILY_TWEEDIE) {
        const double e1 = exp_cap(__dmul_rn(f, 1.0 - a.tweedie_power)), e2 = exp_cap(__dmul_rn(f, 2.0 - a.tweedie_power));
        r = __dsub_rn(__dmul_rn(y, e1), e2);
        tn = fabs(__dmul_rn(y, e1));
        td = e2;
      } else if (a.family == H2OT_FAMILY_LAPLACE) r = f > y ? -0.5 : 0.5;                                     // H/DistributionFactory.java:418-420
      else if (a.family == H2OT_FAMILY_QUANTILE) r = y > f ? 0.5 * a.tweedie_power : 0.5 * (a.tweedie_power - 1.0);  // :433 (alpha)
      else r = y - f;
      const float rf = (float)r;
      a.work[0][row] = rf;
      amaxk = fmaxf(amaxk, fabsf(rf));
      if (a.family == H2OT_FAMILY_GAMMA) tn = (double)fabsf(rf) + 1.0;
      else if (a.family != H2OT_FAMILY_POISSON && a.family != H2OT_FAMILY_TWEEDIE) tn = (double)fabsf(rf);
      gnum = fmaxf(gnum, __double2float_ru(tn));
      gden = fmaxf(gden, __double2float_ru(td));
    }
  }
  if (a.family == H2OT_FAMILY_MULTINOMIAL) { gnum = 1.f; gden = 1.f; }
  for (int o = 16; o > 0; o >>= 1) {
    gnum = fmaxf(gnum, __shfl_xor_sync(0xffffffffu, gnum, o));
    gden = fmaxf(gden, __shfl_xor_sync(0xffffffffu, gden, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (gnum > 0.f) atomicMax(&a.absmax[a.K], __float_as_uint(gnum));
    if (gden > 0.f) atomicMax(&a.absmax[a.K + 1], __float_as_uint(gden));
  }
  for (int o = 16; o > 0; o >>= 1) amaxk = fmaxf(amaxk, __shfl_xor_sync(0xffffffffu, amaxk, o));
  // one bound shared by all K class-trees (max over k) keeps the kernel simple
  if ((threadIdx.x & 31) == 0 && amaxk > 0.f)
    for (int k = 0; k < a.K; k++) atomicMax(&a.absmax[k], __float_as_uint(amaxk));
}

// ------------------------------------------------------------------------------------------ sample
__device__ __forceinline__ uint32_t pcg_next(unsigned long long& state, unsigned long long inc) {
  const unsigned long long old = state;
  state = old * 6364136223846793005ULL + inc;
  const uint32_t xorshifted = (uint32_t)(((old >> 18) ^ old) >> 27);
  const uint32_t rot = (uint32_t)(old >> 59);
  return (xorshifted >> rot) | (xorshifted << ((-rot) & 31));
}
// rate_per_class: Sample._rate_per_class (T/Sample.java:14, 30), indexed by the response class; nullptr == null
__global__ void __launch_bounds__(256) sample_kernel(int64_t n, int64_t row_offset, long long tree_seed, float rate,
                                                     double rate_d, const double* rate_per_class, int n_rates, const double* resp, int32_t* nids) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
    bool skip = isnan(resp[row]);
    if (!skip) {
      if (rate_per_class) {
        const long long cls = (long long)resp[row];  // (int) ys.at8(row)
        rate_d = (cls >= 0 && cls < n_rates) ? rate_per_class[cls] : 0.0;
      }
      // rand.setSeed(_tree._seed + row + nids.start()) ; nextFloat() = next(24) / (float)(1 << 24)
      const unsigned long long inc = 3ULL;  // (seq << 1) | 1 with seq == 1
      unsigned long long state = 0ULL;
      pcg_next(state, inc);
      state += (unsigned long long)tree_seed + (unsigned long long)(row + row_offset);
      pcg_next(state, inc);
      const uint32_t hi = pcg_next(state, inc);
      const uint32_t lo = pcg_next(state, inc);
      (void)hi;
      const int bits24 = (int)(lo & 0xFFFFFFu);  // next(24) keeps the low 24 bits of nextLong()
      const float u = (float)bits24 / (float)(1 << 24);
      skip = (double)u >= rate_d;  // float compared with the double rate
    }
    if (skip) nids[row] = H2OT_OUT_OF_BAG;
  }
}

// ------------------------------------------------------------------------------------------ gamma
struct GammaArgs {
  int64_t n, n_pad;
  int32_t family, leaf_start, n_leaf_nodes, rows_per_cta;
  const double* resp;
  const double* w;
  const float* work;
  int32_t* nids;
  const h2ot_decided_node* nodes;
  const uint8_t* bitsets;
  const ColRef* cols;
  long long* acc;  // [n_leaf_nodes][4] : num.hi, num.lo, den.hi, den.lo (fixed point, split at bit 24)
  double s_num, s_den;
  const double* F;    // Tree_k (tweedie needs f = Tree_k + offset)
  const double* off;
  double tweedie_power;
};
constexpr int kGammaSmemLeaves = 2048;
constexpr int kGammaSplitBits = 24;

// One row of GammaPass: final descent (writes the leaf nid) and the row's contribution to its leaf's sums.
__device__ __forceinline__ void gamma_row(const GammaArgs& a, int64_t row, int nid, double y, double wv, float work, int& nid_out, int& idx,
                                          long long& qn, long long& qd) {
  idx = -1;
  nid_out = nid;
  if (wv == 0.0 || isnan(y)) return;
  const bool oob = nid <= H2OT_OUT_OF_BAG;
  if (oob) nid = -nid + H2OT_OUT_OF_BAG;
  if (nid < 0) return;
  h2ot_decided_node dn = a.nodes[nid];
  if (dn.col < 0) {
    if (dn.pid < 0) return;  // the root found no split: the tree is a stump, no leaf to fit
    dn = a.nodes[dn.pid];
  }
  const int leafnid = child_node_id(dn, a.cols, a.bitsets, row);
  nid_out = leafnid;
  if (oob) return;
  const int i = leafnid - a.leaf_start;
  if (i < 0 || i >= a.n_leaf_nodes) return;
  const double z = (double)work;
  double num = __dmul_rn(wv, z);
  double den;
  if (a.family == H2OT_FAMILY_GAUSSIAN) den = wv;
  else if (a.family == H2OT_FAMILY_BERNOULLI) {
    const double ff = __dsub_rn(y, z);
    den = __dmul_rn(__dmul_rn(wv, ff), __dsub_rn(1.0, ff));
  } else if (a.family == H2OT_FAMILY_POISSON) {  // gammaNum = w y, gammaDenom = w (y - z)
    num = __dmul_rn(wv, y);
    den = __dmul_rn(wv, __dsub_rn(y, z));
  } else if (a.family == H2OT_FAMILY_GAMMA) {  // w (z + 1), w
    num = __dmul_rn(wv, __dadd_rn(z, 1.0));
    den = wv;
  } else if (a.family == H2OT_FAMILY_TWEEDIE) {  // w y exp(f (1 - p)), w exp(f (2 - p))
    const double f = a.F[row] + (a.off ? a.off[row] : 0.0);
    num = __dmul_rn(__dmul_rn(wv, y), exp_cap(__dmul_rn(f, 1.0 - a.tweedie_power)));
    den = __dmul_rn(wv, exp_cap(__dmul_rn(f, 2.0 - a.tweedie_power)));
  } else {
    const double az = fabs(z);
    den = __dmul_rn(wv, __dmul_rn(az, __dsub_rn(1.0, az)));
  }
  qn = __double2ll_rn(__dmul_rn(num, a.s_num));
  qd = __double2ll_rn(__dmul_rn(den, a.s_den));
  idx = i;
}

// Four consecutive rows per thread (independent load chains in flight); rows_per_cta is a multiple of 1024.
__global__ void __launch_bounds__(256) gamma_kernel(GammaArgs a) {
  __shared__ uint32_t s_acc[kGammaSmemLeaves * 4];
  const bool use_s = a.n_leaf_nodes <= kGammaSmemLeaves;
  if (use_s)
    for (int i = threadIdx.x; i < a.n_leaf_nodes * 4; i += blockDim.x) s_acc[i] = 0;
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * a.rows_per_cta;
  const int64_t hi = min(a.n_pad, lo + (int64_t)a.rows_per_cta);
  for (int64_t row0 = lo + (int64_t)threadIdx.x * 4; row0 < hi; row0 += (int64_t)blockDim.x * 4) {
    const int4 nv = *reinterpret_cast<const int4*>(a.nids + row0);
    const float4 wk = *reinterpret_cast<const float4*>(a.work + row0);
    const double2 y01 = *reinterpret_cast<const double2*>(a.resp + row0), y23 = *reinterpret_cast<const double2*>(a.resp + row0 + 2);
    const int nid[4] = {nv.x, nv.y, nv.z, nv.w};
    const float wrk[4] = {wk.x, wk.y, wk.z, wk.w};
    const double ys[4] = {y01.x, y01.y, y23.x, y23.y};
    int out[4], idx[4];
    long long qn[4], qd[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const bool in = row0 + j < a.n;
      const double wv = !in ? 0.0 : (a.w ? a.w[row0 + j] : 1.0);
      gamma_row(a, row0 + j, nid[j], ys[j], wv, wrk[j], out[j], idx[j], qn[j], qd[j]);
    }
    if (out[0] != nid[0] || out[1] != nid[1] || out[2] != nid[2] || out[3] != nid[3])
      *reinterpret_cast<int4*>(a.nids + row0) = make_int4(out[0], out[1], out[2], out[3]);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (idx[j] < 0) continue;
      if (use_s) {
        smem_add64(&s_acc[idx[j] * 4], qn[j]);
        smem_add64(&s_acc[idx[j] * 4 + 2], qd[j]);
      } else {
        unsigned long long* g = reinterpret_cast<unsigned long long*>(a.acc + (size_t)idx[j] * 4);
        atomicAdd(g, (unsigned long long)(qn[j] >> kGammaSplitBits));
        atomicAdd(g + 1, (unsigned long long)(qn[j] & ((1LL << kGammaSplitBits) - 1)));
        atomicAdd(g + 2, (unsigned long long)(qd[j] >> kGammaSplitBits));
        atomicAdd(g + 3, (unsigned long long)(qd[j] & ((1LL << kGammaSplitBits) - 1)));
      }
    }
  }
  __syncthreads();
  if (use_s)
    for (int i = threadIdx.x; i < a.n_leaf_nodes * 2; i += blockDim.x) {
      const long long v = (long long)(((unsigned long long)s_acc[i * 2 + 1] << 32) | s_acc[i * 2]);
      if (v == 0) continue;
      unsigned long long* g = reinterpret_cast<unsigned long long*>(a.acc + (size_t)(i >> 1) * 4 + (i & 1) * 2);
      atomicAdd(g, (unsigned long long)(v >> kGammaSplitBits));
      atomicAdd(g + 1, (unsigned long long)(v & ((1LL << kGammaSplitBits) - 1)));
    }
}

// ------------------------------------------------------------------------------------------ add tree
struct AddTreeArgs {
  int64_t n;
  int32_t n_nodes;
  const double* resp;
  const double* w;
  int32_t* nids;
  double* F;
  const float* node_pred;
  const double* node_factor;  // nullptr, or per node 1 + nextGaussian * pred_noise_bandwidth (T/gbm/GBM.java:1458-1463)
};
__global__ void __launch_bounds__(256) add_tree_kernel(AddTreeArgs a) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (int64_t)gridDim.x * blockDim.x) {
    const int nid = a.nids[row];
    int nv = 0;  // FRESH
    if (nid < 0) {
      const double wv = a.w ? a.w[row] : 1.0;
      if (wv == 0.0 || isnan(a.resp[row])) nv = H2OT_DECIDED_ROW;
      a.nids[row] = nv;
      continue;
    }
    a.nids[row] = nv;
    const float pred = nid < a.n_nodes ? a.node_pred[nid] : 0.f;
    const double contrib = (a.node_factor && nid < a.n_nodes) ? __dmul_rn(a.node_factor[nid], (double)pred) : (double)pred;
    a.F[row] = (double)(float)__dadd_rn(a.F[row], contrib);  // float-rounded every tree (q3)
  }
}

// ------------------------------------------------------------------------------------------ score a finished tree
// SharedTreeMojoModel.scoreTree on the node table: root to leaf per row, prediction added in double.
struct ScoreTreeArgs {
  int64_t n;
  int32_t n_nodes, leaf_start;
  const h2ot_decided_node* nodes;
  const uint8_t* bitsets;
  const ColRef* cols;
  double* F;
  const float* node_pred;
};
__global__ void __launch_bounds__(256) score_tree_kernel(ScoreTreeArgs a) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (int64_t)gridDim.x * blockDim.x) {
    int nid = 0;
    for (int depth = 0; depth < 64 && nid >= 0 && nid < a.leaf_start && nid < a.n_nodes; depth++) {
      const h2ot_decided_node dn = a.nodes[nid];
      if (dn.col < 0) break;  // a root that never split is the leaf
      nid = child_node_id(dn, a.cols, a.bitsets, row);
    }
    const float pred = (nid >= 0 && nid < a.n_nodes) ? a.node_pred[nid] : 0.f;
    a.F[row] = __dadd_rn(a.F[row], (double)pred);
  }
}

// ------------------------------------------------------------------------------------------ training metrics
// hex.tree.Score (T/Score.java:57-141) for GBM on the resident Tree_k columns: score2 / score1 (T/SharedTree.java:782-789,
// T/gbm/GBM.java:1486-1513) then ModelMetrics{Regression,Binomial,Multinomial}.MetricBuilder.perRow — the squared-error and
// log-loss sums (H/ModelMetricsBinomial.java perRow, MathUtils.logloss). Block partials in block order: deterministic.
struct ScoreArgs2 {
  int64_t n;
  int32_t K, nclass, family, pad;
  const double* resp;
  const double* w;
  const double* off;
  const double* const* F;
  double* partial;  // [gridDim.x][4] : count, wcount, sumsqe, logloss
};
__global__ void __launch_bounds__(256) score_metrics_kernel(ScoreArgs2 a) {
  double cnt = 0, wc = 0, sse = 0, ll = 0;
  const double kMaxLL = 34.538776394910684;  // -log(1e-15)
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (int64_t)gridDim.x * blockDim.x) {
    const double y = a.resp[row];
    if (isnan(y)) continue;
    const double wv = a.w ? a.w[row] : 1.0;
    if (wv == 0.0 || isnan(wv)) continue;
    const double f = a.F[0][row] + (a.off ? a.off[row] : 0.0);
    double err;
    if (a.family == H2OT_FAMILY_GAUSSIAN || a.family == H2OT_FAMILY_LAPLACE || a.family == H2OT_FAMILY_QUANTILE) {  // identity link
      err = y - f;
    } else if (a.family >= H2OT_FAMILY_POISSON) {
      err = y - exp_cap(f);  // ModelMetricsRegression on the response scale: ds[0] = linkInv(f)
    } else {
      const int iact = (int)y;
      double pact;
      if (a.family == H2OT_FAMILY_BERNOULLI) {
        if (iact != 0 && iact != 1) continue;
        const double p1 = 1.0 / (1.0 + fmin(1e19, exp(-f)));
        pact = iact == 1 ? p1 : 1.0 - p1;
      } else {
        double maxval = f;
        for (int k = 1; k < a.nclass; k++) maxval = fmax(maxval, a.F[k][row]);
        double dsum = 0, eact = 0;
        for (int k = 0; k < a.nclass; k++) {
          const double e = exp((k == 0 ? f : a.F[k][row]) - maxval);
          dsum += e;
          if (k == iact) eact = e;
        }
        pact = (iact >= 0 && iact < a.nclass) ? eact / dsum : 0.0;
      }
      err = 1.0 - pact;
      ll += wv * fmin(kMaxLL, -log(1.0 - err));
    }
    sse += wv * err * err;
    cnt += 1.0;
    wc += wv;
  }
  __shared__ double sh[4][256];
  sh[0][threadIdx.x] = cnt; sh[1][threadIdx.x] = wc; sh[2][threadIdx.x] = sse; sh[3][threadIdx.x] = ll;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
      for (int q = 0; q < 4; q++) sh[q][threadIdx.x] += sh[q][threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x < 4) a.partial[4 * blockIdx.x + threadIdx.x] = sh[threadIdx.x][0];
}

// ------------------------------------------------------------------------------------------ DRF
// DRF.initializeModelSpecifics (T/drf/DRF.java:139-153): Work_k <- 1 where the response is class k (classification),
// Work_0 <- (float) response (regression); Tree_k <- 0, NIDs_k <- 0 (pad rows DECIDED), OUT_BAG_TREES <- 0.
struct DrfInitArgs {
  int64_t n, n_pad;
  int32_t K, nclass;
  const double* resp;
  double* const* F;
  float* const* work;
  int32_t* const* nids;
  double* oobt;
};
__global__ void __launch_bounds__(256) drf_init_kernel(DrfInitArgs a) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n_pad; row += (int64_t)gridDim.x * blockDim.x) {
    const bool in = row < a.n;
    const double y = in ? a.resp[row] : __longlong_as_double(0x7ff8000000000000LL);
    for (int k = 0; k < a.K; k++) {
      float wv = 0.f;
      if (!isnan(y)) wv = a.nclass > 1 ? ((int)y == k ? 1.f : 0.f) : (float)y;
      a.work[k][row] = wv;
      a.F[k][row] = 0.0;
      a.nids[k][row] = in ? 0 : H2OT_DECIDED_ROW;
    }
    a.oobt[row] = 0.0;
  }
}

// DRF.CollectPreds (T/drf/DRF.java:255-330): rows the tree held out of bag receive its vote; every row's nid is reset.
struct CollectArgs {
  int64_t n;
  int32_t n_nodes, root_is_leaf, update_oobt, pad;
  const double* w;
  int32_t* nids;
  double* F;
  double* oobt;
  const float* node_pred;
  const h2ot_decided_node* nodes;
  const uint8_t* bitsets;
  const ColRef* cols;
};
__global__ void __launch_bounds__(256) drf_collect_kernel(CollectArgs a) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (int64_t)gridDim.x * blockDim.x) {
    const double wv = a.w ? a.w[row] : 1.0;
    int nid = a.nids[row];
    const bool oob = nid <= H2OT_OUT_OF_BAG;
    if (wv != 0.0 && oob) {
      nid = -nid + H2OT_OUT_OF_BAG;
      int leafnid = 0;
      if (!a.root_is_leaf) {
        h2ot_decided_node dn = a.nodes[nid];
        if (dn.col < 0) dn = a.nodes[dn.pid];  // bottomed out or unable to decide: the parent's decision
        leafnid = child_node_id(dn, a.cols, a.bitsets, row);
      }
      const float pred = (leafnid >= 0 && leafnid < a.n_nodes) ? a.node_pred[leafnid] : 0.f;
      a.F[row] = (double)(float)__dadd_rn(a.F[row], (double)pred);
    }
    a.nids[row] = 0;
    if (a.update_oobt && oob) a.oobt[row] = __dadd_rn(a.oobt[row], wv);
  }
}

}  // namespace h2ot
