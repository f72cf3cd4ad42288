// K3: split search. One CTA per (leaf, col) restates DTree.findBestSplitPoint (T/DTree.java:984-1487,
// constraint == 0, no bounds) + Split.splatNumeric (:253-291) + fillBitSet (:1907-1941); a second kernel
// restates DecidedNode.bestCol (:616-656) per leaf.
//
// Bit-exactness notes:
//  * the cumulative sums are done sequentially in bin order by single threads (fp64 adds do not commute
//    bit-for-bit), six scans run concurrently on six threads;
//  * the candidate scan is done in parallel. The reference's sequential accept rule
//      (se < best) || (se == best && |b - mid| < |best_b - mid|)
//    selects, among all valid candidates, the lexicographic minimum of (se, |b - mid|, visit order), which is
//    what the block reduction computes; the min_rows `continue` / `break` asymmetry (:1170-1171) becomes a
//    block-wide "first break index";
//  * all products/sums use explicit round-to-nearest intrinsics, so no FMA contraction can change a bit.
#pragma once
#include "common.cuh"
#include "k_level.cuh"

namespace h2ot {

struct SplitArgs {
  int32_t L, C;
  const DescDev* descs;
  const double* vals;
  const double* min2;
  const double* maxIn;
  const uint8_t* has_vals;
  double min_rows, min_split_improvement;
  h2ot_split* cand;  // [L*C]
  const double* sp;  // the level's split-point pool or nullptr
  int32_t fused;     // 1: the pair's finalize (fixed point -> _vals, min2 / maxIn / wNA, has_vals) runs here first, fin holds its arguments
  FinalizeArgs fin;
};

constexpr int kSplitThreads = 128;

struct Cand {
  double se;
  int dist;
  int ord;  // 2*b + opt ; -1 for the initial (NAvsREST / none) state
};
__device__ __forceinline__ bool cand_less(const Cand& x, const Cand& y) {
  // NaN se never wins (the reference's `<` / `==` tests are false for NaN)
  if (isnan(x.se)) return false;
  if (isnan(y.se)) return true;
  if (x.se < y.se) return true;
  if (x.se > y.se) return false;
  if (x.dist < y.dist) return true;
  if (x.dist > y.dist) return false;
  return x.ord < y.ord;
}

__device__ __forceinline__ float ulpf_dev(float f) {
  f = fabsf(f);
  if (isinf(f) || isnan(f)) return f;
  if (f == FLT_MAX) return ldexpf(1.0f, 104);
  return __uint_as_float(__float_as_uint(f) + 1) - f;
}
// W/util/MathUtils.java:381-389
__device__ __forceinline__ bool eq_one_small_ulp(float a, float b) {
  if (isnan(a) && isnan(b)) return true;
  float small_ulp = fminf(ulpf_dev(a), ulpf_dev(b));
  return fabsf(a - b) <= small_ulp;
}

__device__ __forceinline__ double se_of(double wyy, double wy, double w) {  // wYY - wY*wY/w, clamped at 0
  double s = __dsub_rn(wyy, __ddiv_rn(__dmul_rn(wy, wy), w));
  return s < 0 ? 0.0 : s;
}

__global__ void __launch_bounds__(kSplitThreads) find_split_kernel(SplitArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int p = blockIdx.x;
  const DescDev d = a.descs[p];
  h2ot_split* out = a.cand + p;
  const int tid = threadIdx.x;
  if (tid == 0) {
    out->col = -1;
    out->splat = __int_as_float(0x7fc00000);
    out->bs_nbits = out->bs_bitoff = out->bs_nbytes = 0;
  }
  if (d.nbin < 0) {
    if (a.fused) finalize_pair(a.fin, p, nullptr);
    return;
  }
  if (!a.fused && !a.has_vals[p]) return;
  const int nbins = d.nbin;
  const int n1 = nbins + 1;
  const bool is_cat_sorted = (d.flags & H2OT_DESC_ISCAT) && d.step == 1.0;
  // shared layout (doubles): orig[3*n1] | v[3*nbins] (sorted copy, cats only) | wlo,wYlo,wYYlo,whi,wYhi,wYYhi [6*n1] | avgs[n1] | idxs (int)[n1]
  double* orig = reinterpret_cast<double*>(smem_raw);
  double* sorted = orig + 3 * n1;
  double* wlo = sorted + (is_cat_sorted ? 3 * nbins : 0);
  double* wYlo = wlo + n1;
  double* wYYlo = wYlo + n1;
  double* whi = wYYlo + n1;
  double* wYhi = whi + n1;
  double* wYYhi = wYhi + n1;
  double* avgs = wYYhi + n1;
  int* idxs = reinterpret_cast<int*>(avgs + (is_cat_sorted ? n1 : 0));
  __shared__ int s_break;
  __shared__ Cand s_c[kSplitThreads];

  if (a.fused) {
    if (!finalize_pair(a.fin, p, orig)) return;
  } else {
    const double* gv = a.vals + d.voff;
    const int n3 = 3 * n1;
    for (int i0 = tid; i0 < n3; i0 += 8 * kSplitThreads) {  // eight loads in flight per thread
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; u++) x[u] = i0 + u * kSplitThreads < n3 ? gv[i0 + u * kSplitThreads] : 0.0;
#pragma unroll
      for (int u = 0; u < 8; u++)
        if (i0 + u * kSplitThreads < n3) orig[i0 + u * kSplitThreads] = x[u];
    }
  }
  if (tid == 0) s_break = 0x7fffffff;
  __syncthreads();
  const double wNA = orig[3 * nbins], wYNA = orig[3 * nbins + 1], wYYNA = orig[3 * nbins + 2];
  // bestCol filter (T/DTree.java:625,633): actNBins() <= 1
  if (nbins + (wNA > 0 ? 1 : 0) <= 1) return;

  const double* v = orig;
  if (is_cat_sorted) {  // :1006-1033 — stable ascending sort of bins by mean response
    for (int i = tid; i < n1; i += blockDim.x) {
      avgs[i] = i < nbins ? (orig[3 * i] == 0 ? -DBL_MAX : __ddiv_rn(orig[3 * i + 1], orig[3 * i])) : DBL_MAX;
    }
    __syncthreads();
    for (int i = tid; i < n1; i += blockDim.x) {  // rank by counting == stable sort position
      const double ai = avgs[i];
      int r = 0;
      for (int j = 0; j < n1; j++) {
        const double aj = avgs[j];
        r += (aj < ai || (aj == ai && j < i)) ? 1 : 0;
      }
      idxs[r] = i;
    }
    __syncthreads();
    for (int i = tid; i < nbins; i += blockDim.x) {
      const int id = idxs[i];
      sorted[3 * i] = orig[3 * id];
      sorted[3 * i + 1] = orig[3 * id + 1];
      sorted[3 * i + 2] = orig[3 * id + 2];
    }
    __syncthreads();
    v = sorted;
  }

  // :1036-1067 and :1090-1121 — six sequential scans on six threads. The additions are a dependent chain (the reference's order is
  // the result), but the loads and stores around them need not be: eight values are fetched before, and stored after, eight adds
  // (left to the compiler, every load waited for the previous store -- source and destination could alias -- and an element cost
  // ~100 clocks instead of one fp64 add).
  if (tid < 3) {
    double* dst = tid == 0 ? wlo : (tid == 1 ? wYlo : wYYlo);
    double s = 0.0;
    dst[0] = 0.0;
    int b = 1;
    for (; b + 7 <= nbins; b += 8) {
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; u++) x[u] = v[3 * (b - 1 + u) + tid];
#pragma unroll
      for (int u = 0; u < 8; u++) {
        s = __dadd_rn(s, x[u]);
        x[u] = s;
      }
#pragma unroll
      for (int u = 0; u < 8; u++) dst[b + u] = x[u];
    }
    for (; b <= nbins; b++) {
      s = __dadd_rn(s, v[3 * (b - 1) + tid]);
      dst[b] = s;
    }
  } else if (tid >= 32 && tid < 35) {
    const int j = tid - 32;
    double* dst = j == 0 ? whi : (j == 1 ? wYhi : wYYhi);
    double s = 0.0;
    dst[nbins] = 0.0;
    int b = nbins - 1;
    for (; b - 7 >= 0; b -= 8) {
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; u++) x[u] = v[3 * (b - u) + j];
#pragma unroll
      for (int u = 0; u < 8; u++) {
        s = __dadd_rn(s, x[u]);
        x[u] = s;
      }
#pragma unroll
      for (int u = 0; u < 8; u++) dst[b - u] = x[u];
    }
    for (; b >= 0; b--) {
      s = __dadd_rn(s, v[3 * b + j]);
      dst[b] = s;
    }
  }
  __syncthreads();

  const double min_rows = a.min_rows;
  const double tot = __dadd_rn(wlo[nbins], wNA);
  if (tot < 2 * min_rows) return;  // :1071
  {
    const double syy = __dadd_rn(wYYlo[nbins], wYYNA), sy = __dadd_rn(wYlo[nbins], wYNA);
    const double var = __dsub_rn(__dmul_rn(syy, tot), __dmul_rn(sy, sy));
    if (((float)var) == 0.0f) return;  // :1081
  }
  double seNonNA = __dsub_rn(wYYhi[0], __ddiv_rn(__dmul_rn(wYhi[0], wYhi[0]), whi[0]));  // :1128
  if (seNonNA < 0) seNonNA = 0;
  const int mid = nbins >> 1;

  // first break index (:1169-1171)
  for (int b = 1 + tid; b <= nbins - 1; b += blockDim.x) {
    if (v[3 * b] == 0) continue;
    if (__dadd_rn(wlo[b], wNA) < min_rows) continue;
    if (__dadd_rn(whi[b], wNA) < min_rows) atomicMin(&s_break, b);
  }
  __syncthreads();
  const int bstop = s_break;

  // initial state (:1123-1159)
  Cand best;
  best.se = __dadd_rn(DBL_MAX, DBL_MAX);
  best.dist = mid;  // |0 - mid|
  best.ord = -1;
  double seNA = 0;
  const bool na_vs_rest = wNA >= min_rows;
  if (na_vs_rest) {
    seNA = __dsub_rn(wYYNA, __ddiv_rn(__dmul_rn(wYNA, wYNA), wNA));
    if (seNA < 0) seNA = 0;
    best.se = __dadd_rn(seNonNA, seNA);
  }
  // candidates (:1168-1348)
  for (int b = 1 + tid; b <= nbins - 1 && b < bstop; b += blockDim.x) {
    if (v[3 * b] == 0) continue;
    if (__dadd_rn(wlo[b], wNA) < min_rows) continue;
    const int dist = abs(b - mid);
    if (wNA == 0) {
      Cand c;
      c.se = __dadd_rn(se_of(wYYlo[b], wYlo[b], wlo[b]), se_of(wYYhi[b], wYhi[b], whi[b]));
      c.dist = dist;
      c.ord = 2 * b;
      if (cand_less(c, best)) best = c;
    } else {
      {  // NAs left
        const double selo = se_of(__dadd_rn(wYYlo[b], wYYNA), __dadd_rn(wYlo[b], wYNA), __dadd_rn(wlo[b], wNA));
        const double sehi = se_of(wYYhi[b], wYhi[b], whi[b]);
        if (__dadd_rn(wlo[b], wNA) >= min_rows && whi[b] >= min_rows) {
          Cand c;
          c.se = __dadd_rn(selo, sehi);
          c.dist = dist;
          c.ord = 2 * b;
          if (cand_less(c, best)) best = c;
        }
      }
      {  // NAs right
        const double selo = se_of(wYYlo[b], wYlo[b], wlo[b]);
        const double sehi = se_of(__dadd_rn(wYYhi[b], wYYNA), __dadd_rn(wYhi[b], wYNA), __dadd_rn(whi[b], wNA));
        if (wlo[b] >= min_rows && __dadd_rn(whi[b], wNA) >= min_rows) {
          Cand c;
          c.se = __dadd_rn(selo, sehi);
          c.dist = dist;
          c.ord = 2 * b + 1;
          if (cand_less(c, best)) best = c;
        }
      }
    }
  }
  s_c[tid] = best;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (tid < s) {
      if (cand_less(s_c[tid + s], s_c[tid])) s_c[tid] = s_c[tid + s];
    }
    __syncthreads();
  }
  if (tid != 0) return;

  // ---- thread 0: materialise the winner and run the post checks (:1350-1487)
  const Cand win = s_c[0];
  int bestb = 0, nasplit = H2OT_NA_NONE;
  double best_seL = DBL_MAX, best_seR = DBL_MAX, seBefore = seNonNA;
  double nLeft = 0, nRight = 0, predLeft = 0, predRight = 0, tree_p0 = 0, tree_p1 = 0;
  if (na_vs_rest) {  // :1140-1159
    const double sy = __dadd_rn(wYhi[0], wYNA);
    const double seAll = __dsub_rn(__dadd_rn(wYYhi[0], wYYNA), __ddiv_rn(__dmul_rn(sy, sy), __dadd_rn(whi[0], wNA)));
    best_seL = seNonNA;
    best_seR = seNA;
    nasplit = H2OT_NA_VS_REST;
    seBefore = seAll;
    nLeft = whi[0];
    predLeft = wYhi[0];
    nRight = wNA;
    predRight = wYNA;
    tree_p0 = __ddiv_rn(predLeft, nLeft);
    tree_p1 = __ddiv_rn(predRight, nRight);
  }
  if (win.ord >= 0) {
    const int b = win.ord >> 1;
    bestb = b;
    if (wNA == 0) {  // :1180-1230
      best_seL = se_of(wYYlo[b], wYlo[b], wlo[b]);
      best_seR = se_of(wYYhi[b], wYhi[b], whi[b]);
      nLeft = wlo[b];
      nRight = whi[b];
      predLeft = wYlo[b];
      predRight = wYhi[b];
      tree_p0 = __ddiv_rn(wYlo[b], wlo[b]);
      tree_p1 = __ddiv_rn(wYhi[b], whi[b]);
    } else if ((win.ord & 1) == 0) {  // NAs left :1233-1287
      best_seL = se_of(__dadd_rn(wYYlo[b], wYYNA), __dadd_rn(wYlo[b], wYNA), __dadd_rn(wlo[b], wNA));
      best_seR = se_of(wYYhi[b], wYhi[b], whi[b]);
      nLeft = __dadd_rn(wlo[b], wNA);
      nRight = whi[b];
      predLeft = __dadd_rn(wYlo[b], wYNA);
      predRight = wYhi[b];
      nasplit = H2OT_NA_LEFT;
      tree_p0 = __ddiv_rn(__dadd_rn(wYlo[b], wYNA), __dadd_rn(wlo[b], wNA));
      tree_p1 = __ddiv_rn(wYhi[b], whi[b]);
    } else {  // NAs right :1290-1346
      best_seL = se_of(wYYlo[b], wYlo[b], wlo[b]);
      best_seR = se_of(__dadd_rn(wYYhi[b], wYYNA), __dadd_rn(wYhi[b], wYNA), __dadd_rn(whi[b], wNA));
      nLeft = wlo[b];
      nRight = __dadd_rn(whi[b], wNA);
      predLeft = wYlo[b];
      predRight = __dadd_rn(wYhi[b], wYNA);
      nasplit = H2OT_NA_RIGHT;
      tree_p0 = __ddiv_rn(wYlo[b], wlo[b]);
      tree_p1 = __ddiv_rn(__dadd_rn(wYhi[b], wYNA), __dadd_rn(whi[b], wNA));
    }
  }
  if (bestb == 0 && nasplit == H2OT_NA_NONE) return;                                                     // :1350
  const double thresh = __dmul_rn(seBefore, __dsub_rn(1.0, a.min_split_improvement));
  if (!(__dadd_rn(best_seL, best_seR) < thresh)) return;                                                 // :1356, :1452
  if (eq_one_small_ulp((float)__ddiv_rn(predLeft, nLeft), (float)__ddiv_rn(predRight, nRight))) return;  // :1363
  if (nLeft < min_rows || nRight < min_rows) return;                                                     // :1368
  const double node_p0 = __ddiv_rn(predLeft, nLeft), node_p1 = __ddiv_rn(predRight, nRight);
  if (eq_one_small_ulp((float)tree_p0, (float)tree_p1)) return;                                          // :1457

  int equal = 0;
  if (is_cat_sorted && nasplit != H2OT_NA_VS_REST) {  // :1465-1473, IcedBitSet ctor W/util/IcedBitSet.java:27-34
    int off = __double2int_rz(d.min);
    int nbits = nbins, bitoff = off;
    if (bitoff + nbits <= 32) { bitoff = 0; nbits = 32; }
    const int nbytes = ((nbits - 1) >> 3) + 1;
    if (nbytes > H2OT_MAX_BITSET_BYTES) return;  // guarded on the host (nbins_cats <= 4096)
    uint8_t* bs = out->bs;
    for (int i = 0; i < nbytes; i++) bs[i] = 0;
#define BS_SET(ix) bs[((ix) - bitoff) >> 3] |= (uint8_t)(1u << (((ix) - bitoff) & 7))
#define BS_CLR(ix) bs[((ix) - bitoff) >> 3] &= (uint8_t) ~(1u << (((ix) - bitoff) & 7))
#define BS_HAS(ix) ((bs[((ix) - bitoff) >> 3] >> (((ix) - bitoff) & 7)) & 1)
    for (int i = bestb; i < nbins; i++) BS_SET(idxs[i] + off);  // fillBitSet :1907-1941
    int wentRight = 0, wentLeft = 0;
    for (int i = 0; i < nbins; i++)
      if (orig[3 * i] > 0) {
        if (BS_HAS(i + off)) wentRight++;
        else wentLeft++;
      }
    const bool shouldGoLeft = wentLeft >= wentRight;
    for (int i = 0; i < nbins; i++)
      if (orig[3 * i] == 0) {
        if (BS_HAS(i + off) && shouldGoLeft) BS_CLR(i + off);
        if (!BS_HAS(i + off) && !shouldGoLeft) BS_SET(i + off);
      }
    int card = 0;
    for (int i = 0; i < nbytes; i++) card += __popc((unsigned)bs[i]);
#undef BS_SET
#undef BS_CLR
#undef BS_HAS
    if (card == 0 || card == nbits) return;
    equal = (bitoff + nbits <= 32) ? 2 : 3;
    out->bs_nbits = nbits;
    out->bs_bitoff = bitoff;
    out->bs_nbytes = nbytes;
  }
  if (nasplit == H2OT_NA_NONE) nasplit = nLeft > nRight ? H2OT_LEFT : H2OT_RIGHT;  // :1476-1478

  float splat = -1.0f;  // Split.splat :240-242
  if (nasplit != H2OT_NA_VS_REST && equal == 0) {  // splatNumeric :253-291 (on the ORIGINAL bin order)
    int x = bestb - 1;
    while (x >= 0 && orig[3 * x] == 0) x--;
    int n = bestb;
    while (n < nbins && orig[3 * n] == 0) n++;
    // binAt (T/DHistogram.java:308-312): through the node's split points when it has them
    const double* pts = (d.flags & H2OT_DESC_SPLITPTS) ? a.sp + d.sp_off : nullptr;
    const bool absolute = (d.flags & H2OT_DESC_ABSOLUTE) != 0;
    double lo = bin_at_pts(d.min, d.step, x + 1, pts, absolute);
    double hi = bin_at_pts(d.min, d.step, n, pts, absolute);
    if (d.flags & H2OT_DESC_ISINT) {
      lo = d.step == 1 ? __dsub_rn(lo, 1.0) : floor(lo);
      hi = d.step == 1 ? hi : ceil(hi);
    }
    splat = (float)__ddiv_rn(__dadd_rn(lo, hi), 2.0);
    if ((d.flags & H2OT_DESC_CHECKFLOAT) && lo != hi && (lo > (double)splat || hi < (double)splat))
      splat = __int_as_float(0x7fc00000);
  }
  out->col = p % a.C;
  out->bin = bestb;
  out->nasplit = nasplit;
  out->equal = equal;
  out->splat = splat;
  out->se = seBefore;
  out->se0 = best_seL;
  out->se1 = best_seR;
  out->n0 = nLeft;
  out->n1 = nRight;
  out->p0 = node_p0;
  out->p1 = node_p1;
  out->tree_p0 = tree_p0;
  out->tree_p1 = tree_p1;
}

// DecidedNode.bestCol (T/DTree.java:616-656): first strictly smaller se() in _scoreCols order. One warp per leaf: lanes take
// columns, a shuffle reduction keeps the lexicographic minimum of (se, position in _scoreCols), the warp copies the winner.
// (one full warp per leaf; best_col_kernel below, or the first warp of next_level_kernel when the levels run back to back)
__device__ __forceinline__ void best_col_warp(int l, int C, const DescDev* descs, const h2ot_split* cand, h2ot_split* o) {  // o: the leaf's record
  const int lane = threadIdx.x & 31;
  int bc = -1, border = 0x7fffffff;
  double bse = 0;
  for (int c = lane; c < C; c += 32) {
    const h2ot_split& s = cand[(size_t)l * C + c];
    if (s.col < 0) continue;
    const DescDev& d = descs[(size_t)l * C + c];
    if ((d.flags & H2OT_DESC_CHECKFLOAT) && isnan(s.splat)) continue;  // :644-651
    const double se = __dadd_rn(s.se0, s.se1);
    const int order = (int)(int16_t)(uint16_t)d.pad;  // position in the leaf's _scoreCols
    if (bc < 0 || se < bse || (se == bse && order < border)) {
      bc = c;
      bse = se;
      border = order;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
    const int oo = __shfl_xor_sync(0xffffffffu, border, o);
    const double os = __shfl_xor_sync(0xffffffffu, bse, o);
    if (oc >= 0 && (bc < 0 || os < bse || (os == bse && oo < border))) {
      bc = oc;
      bse = os;
      border = oo;
    }
  }
  if (bc < 0) {
    if (lane == 0) {
      o->col = -1;
      o->splat = __int_as_float(0x7fc00000);
    }
    return;
  }
  // copy header + statistics (104 bytes) and the used part of the bitset, word-wise
  const uint32_t* src = reinterpret_cast<const uint32_t*>(cand + (size_t)l * C + bc);
  uint32_t* dst = reinterpret_cast<uint32_t*>(o);
  const int nbytes = cand[(size_t)l * C + bc].bs_nbytes;
  const int nwords = 26 + (nbytes + 3) / 4;  // offsetof(h2ot_split, bs) == 104
  for (int i = lane; i < nwords; i += 32) dst[i] = src[i];
}
__global__ void __launch_bounds__(32) best_col_kernel(int L, int C, const DescDev* descs, const h2ot_split* cand, h2ot_split* best_out) {
  best_col_warp(blockIdx.x, C, descs, cand, best_out + blockIdx.x);
}

}  // namespace h2ot
