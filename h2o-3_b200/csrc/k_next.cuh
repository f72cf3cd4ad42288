// The step between two levels of a tree, on the device: ScoreBuildOneTree.onCompletion's DecidedNode constructor
// (h2o-algos/src/main/java/hex/tree/DTree.java:703-726) with Split.nextLevelHistos (:312-417) and the DHistogram constructor
// rule (h2o-algos/src/main/java/hex/tree/DHistogram.java:190-269, find_maxEx :453-459) -- what host_gbm.cpp's build_layer_k does
// between two h2ot_level_step calls, without the round trip to the host that bounds a level on small shards.
//
// Node ids are STATIC while a tree is built this way: the leaves of level d are the slots 2^d - 1 .. 2^(d+1) - 2 of a complete
// binary tree (child w of slot s at level d is slot 2 s + w of level d + 1), so every launch parameter of every level is known
// before the tree starts and the levels can be enqueued back to back. A slot without a node has no rows and no histograms; the
// host renumbers the nodes in creation order (H2O's numbering) when it mirrors the finished tree.
// Only builds without per-node column sampling and with uniform-adaptive histograms come here (UndecidedNode.scoreCols == null).
#pragma once
#include "common.cuh"
#include "k_level.cuh"
#include "k_rows.cuh"
#include "k_split.cuh"

namespace h2ot {

struct NextArgs {
  int32_t C, Lp;               // columns; slots of the parent level (2^(d-1))
  int32_t parent_base;         // static id of the parent level's first slot (Lp - 1)
  int32_t last_level;          // the children only need to exist (the tree stops here): no descriptors are written
  int32_t nbins, nbins_cats;
  double min_rows;
  const DescDev* pdescs;       // parent level [Lp][C]
  const double* pmaxex;        // parent level _maxEx [Lp][C]
  const h2ot_split* best;      // parent level [Lp]
  const h2ot_split* cand;      // nullptr, or the parent level's candidates [Lp][C]: DecidedNode.bestCol runs here first and fills best_out
  h2ot_split* best_out;        // == best when cand is set
  const double* min2;          // [Lp * C]
  const double* maxIn;
  const double* wna;
  const uint8_t* has_vals;
  DescDev* ndescs;             // next level [2 Lp][C]: layout fields (gbin, voff, smem_off) pre-filled, the rest written here
  double* nmaxex;              // next level _maxEx [2 Lp][C]
  const int32_t* nb_cap;       // per column: the largest nbin the next level's layout has room for
  h2ot_decided_node* nodes;    // node table, static ids
  uint8_t* bitsets;            // pool: H2OT_MAX_BITSET_BYTES per static id
  uint8_t* exists;             // [static id]: the slot holds a node
  h2ot_split* tree_splits;     // [static id]: the node's split, kept for the host's mirror of the tree
  int32_t* err;                // set to 1 if a descriptor does not fit the layout (cannot happen unless the caps are wrong)
  const double* conc_src;      // step behind the root level only: the root's per-column bin concentration [C] ...
  double* conc_dst;            // ... copied into the tree's output block (the host's planner reads it after the tree)
};

// Start of a tree: node table all -1 (col = pid = -1), only the root exists, no error, no concentration measured.
__global__ void __launch_bounds__(256) tree_init_kernel(uint32_t* node_words, int n_node_words, uint8_t* exists, int S, int32_t* err, double* conc,
                                                        int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_node_words) node_words[i] = 0xFFFFFFFFu;
  if (i < S) exists[i] = i == 0 ? 1 : 0;
  if (i < C) conc[i] = 0.0;
  if (i == 0) *err = 0;
}

__device__ __forceinline__ double next_ulp(double d) {  // Math.ulp(double)
  if (isnan(d)) return d;
  d = fabs(d);
  if (isinf(d)) return d;
  if (d == DBL_MAX) return ldexp(1.0, 971);
  return __longlong_as_double(__double_as_longlong(d) + 1) - d;
}
__device__ __forceinline__ bool next_within_one_small_ulp(double a, double b) {  // water/util/MathUtils.java:391-397
  if (isnan(a) && isnan(b)) return true;
  return fabs(a - b) <= fmin(next_ulp(a), next_ulp(b));
}
__device__ __forceinline__ long long next_d2l(double d) {  // Java (long) cast
  if (isnan(d)) return 0;
  if (d >= 9223372036854775807.0) return 0x7FFFFFFFFFFFFFFFLL;
  if (d <= -9223372036854775808.0) return (long long)0x8000000000000000ULL;
  return (long long)d;
}

// One CTA per parent slot.
__global__ void __launch_bounds__(128) next_level_kernel(NextArgs a) {
  const int s = blockIdx.x;
  const int pid = a.parent_base + s;                 // static id of the parent node
  const int child_base = 2 * a.Lp - 1 + 2 * s;       // static ids of its two children
  __shared__ int s_cnt[2];
  __shared__ h2ot_split sp_s;
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  if (a.conc_src && s == 0)
    for (int c = threadIdx.x; c < a.C; c += blockDim.x) a.conc_dst[c] = a.conc_src[c];
  if (a.cand) {
    // DecidedNode.bestCol (T/DTree.java:616-656), otherwise a kernel of its own between the split search and this one: the first
    // warp picks the leaf's column and copies the winner straight into shared memory; the record is then published for the host
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sp_s);
    for (int i = threadIdx.x; i < (int)(sizeof(h2ot_split) / 4); i += blockDim.x) dst[i] = 0u;
    __syncthreads();
    if (threadIdx.x < 32) best_col_warp(s, a.C, a.pdescs, a.cand, &sp_s);
    __syncthreads();
    uint32_t* pub = reinterpret_cast<uint32_t*>(a.best_out + s);
    for (int i = threadIdx.x; i < (int)(sizeof(h2ot_split) / 4); i += blockDim.x) pub[i] = dst[i];
  } else {  // the split record, word-wise into shared memory
    const uint32_t* src = reinterpret_cast<const uint32_t*>(a.best + s);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sp_s);
    for (int i = threadIdx.x; i < (int)(sizeof(h2ot_split) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const bool node = a.exists[pid] != 0;
  const bool split = node && sp_s.col >= 0;
  // ---- Split.nextLevelHistos for both ways, one (way, column) per thread
  for (int t = threadIdx.x; t < 2 * a.C; t += blockDim.x) {
    const int way = t / a.C, j = t - way * a.C;
    DescDev* nd = a.last_level ? nullptr : a.ndescs + (size_t)(2 * s + way) * a.C + j;
    bool present = false;
    double cmin = 0, cmax_ex = 0, cstep = 0;
    int cnbin = -1;
    uint32_t cflags = 0;
    if (split) {
      const double n = way == 0 ? sp_s.n0 : sp_s.n1;
      const double se = way == 0 ? sp_s.se0 : sp_s.se1;
      const DescDev h = a.pdescs[(size_t)s * a.C + j];
      if (!(n < a.min_rows) && !(se <= 1e-30) && h.nbin >= 0) {
        const size_t pi = (size_t)s * a.C + j;
        const bool has_vals = a.has_vals[pi] != 0;
        const int is_int = (h.flags & H2OT_DESC_ISCAT) ? 2 : ((h.flags & H2OT_DESC_ISINT) ? 1 : 0);
        const int adj_nbins = max(h.nbin >> 1, a.nbins);
        double mn, mx;
        bool go = true;
        if (!has_vals || sp_s.equal > 1) {
          mn = h.min;
          mx = a.pmaxex[pi];
        } else {
          mn = a.min2[pi];
          const double max_in = a.maxIn[pi];
          if (max_in == mn) go = false;
          double u = next_ulp(max_in);  // find_maxEx
          if (is_int > 0 && 1 > u) u = 1;
          const double res = max_in + u;
          mx = isinf(res) ? max_in : res;
        }
        if (go && sp_s.nasplit == H2OT_NA_VS_REST && way == 1) go = false;
        if (go && sp_s.col == j) {
          if (sp_s.equal == 0) {
            double spl = (double)sp_s.splat;
            if (is_int > 0) spl = (double)(float)ceil(spl);
            if (sp_s.nasplit != H2OT_NA_VS_REST) {
              if (way == 0) mx = spl;
              else mn = spl;
            }
          } else if (sp_s.equal == 1) {
            if (way == 1) go = false;
          }
        }
        if (go && mn > mx) go = false;
        if (go && next_within_one_small_ulp(mn, mx)) go = false;
        if (go) {
          const double range = mx - mn;
          if (range < 1e-280 && isinf((double)adj_nbins / range)) go = false;
        }
        if (go && is_int > 0 && !(mn + 1 < mx)) go = false;
        if (go) {  // DHistogram constructor (make_hist)
          const bool has_na_bin = has_vals ? a.wna[pi] > 0 : (h.flags & H2OT_DESC_INITNA) != 0;
          const bool has_nas = ((sp_s.nasplit == H2OT_NA_LEFT && way == 0) || (sp_s.nasplit == H2OT_NA_RIGHT && way == 1)) && has_na_bin;
          int xbins = is_int == 2 ? a.nbins_cats : adj_nbins;
          double step;
          bool ok = true;
          if (is_int > 0 && mx - mn <= xbins) {
            xbins = (int)(unsigned short)(next_d2l(mx) - next_d2l(mn));  // (char) cast
            step = 1.0;
          } else {
            step = __ddiv_rn((double)xbins, __dsub_rn(mx, mn));
            if (step <= 0 || isinf(step) || isnan(step)) ok = false;
          }
          if (ok) {
            present = true;
            cmin = mn;
            cmax_ex = mx;
            cstep = step;
            cnbin = (int)(unsigned short)xbins;
            cflags = H2OT_DESC_ACTIVE | (is_int > 0 ? H2OT_DESC_ISINT : 0) | (is_int == 2 ? H2OT_DESC_ISCAT : 0) | (h.flags & H2OT_DESC_INTOPT) |
                     (has_nas ? H2OT_DESC_INITNA : 0);
          }
        }
      }
    }
    if (present) atomicAdd(&s_cnt[way], 1);
    if (nd) {
      if (present && (cnbin < 1 || cnbin > a.nb_cap[j])) {
        atomicExch(a.err, 1);
        present = false;
      }
      nd->nbin = present ? cnbin : -1;
      nd->min = cmin;
      nd->step = cstep;
      nd->flags = present ? cflags : 0u;
      nd->pad = (uint32_t)j;  // _scoreCols == null: bestCol's order is the column index
      a.nmaxex[(size_t)(2 * s + way) * a.C + j] = cmax_ex;
    }
  }
  __syncthreads();
  // ---- the DecidedNode itself, its children's existence, the mirror record
  const bool c0 = split && s_cnt[0] > 0, c1 = split && s_cnt[1] > 0;
  if (!a.last_level && !(c0 && c1)) {
    // a child that does not exist must not keep descriptors written for the other columns... none were: present == false for
    // every column of a way whose count is zero. Nothing to undo.
  }
  if (threadIdx.x == 0) {
    a.exists[child_base] = c0 ? 1 : 0;
    a.exists[child_base + 1] = c1 ? 1 : 0;
    h2ot_decided_node dn;
    dn.nid = pid;
    dn.pid = pid == 0 ? -1 : (pid - 1) >> 1;
    dn.col = split ? sp_s.col : -1;
    dn.equal = split ? sp_s.equal : 0;
    dn.nasplit = split ? sp_s.nasplit : H2OT_NA_NONE;
    dn.splat = split ? sp_s.splat : __int_as_float(0x7fc00000);
    dn.child[0] = c0 ? child_base : -1;
    dn.child[1] = c1 ? child_base + 1 : -1;
    dn.bs_off = pid * H2OT_MAX_BITSET_BYTES;
    dn.bs_nbits = split ? sp_s.bs_nbits : 0;
    dn.bs_bitoff = split ? sp_s.bs_bitoff : 0;
    dn.pad = 0;
    if (node) a.nodes[pid] = dn;
    // DTree.UndecidedNode.do_not_split (:504-511): a node that found no split is cut off from its parent, whose rows then stay
    // with the parent's decision
    if (node && !split && pid > 0) a.nodes[(pid - 1) >> 1].child[(pid - 1) & 1] = -1;
  }
  if (node) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(a.tree_splits + pid);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&sp_s);
    for (int i = threadIdx.x; i < (int)(sizeof(h2ot_split) / 4); i += blockDim.x) dst[i] = src[i];
    if (split && sp_s.equal >= 2)
      for (int i = threadIdx.x; i < sp_s.bs_nbytes; i += blockDim.x) a.bitsets[(size_t)pid * H2OT_MAX_BITSET_BYTES + i] = sp_s.bs[i];
  }
}

// After the last level: every child pointer that does not lead to a split node becomes a LeafNode (DTree leaf insertion,
// T/gbm/GBM.java:655-679), numbered by its static slot: leaf id = S + slot, S = 2^(D+1) - 1. A child is a leaf when it never existed
// or found no split (-1, see above) or sits on level D, which is never searched (slot >= nS = 2^D - 1).
__global__ void __launch_bounds__(128) leafify_kernel(h2ot_decided_node* nodes, const uint8_t* exists, int nS, int S) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nS || !exists[p]) return;
  h2ot_decided_node dn = nodes[p];
  if (dn.col < 0) return;
#pragma unroll
  for (int w = 0; w < 2; w++)
    if (dn.child[w] < 0 || dn.child[w] >= nS) nodes[p].child[w] = S + 2 * p + 1 + w;
}

// GammaPass's sums to leaf values: fitBestConstants for gaussian / bernoulli / multinomial (T/gbm/GBM.java:813-847), the same
// double arithmetic as host_gbm.cpp. pred[0 .. S) = 0 (rows that stayed on a node that is no leaf add nothing), pred[S + slot] = value.
struct LeafValueArgs {
  int32_t S;
  const long long* acc;  // [S][4]: num.hi, num.lo, den.hi, den.lo (gamma_kernel's split fixed point)
  int32_t e_num, e_den;
  double rate;           // learn_rate (annealed) x (K - 1) / K for multinomial
  int32_t multinomial;
  double max_abs;
  float* pred;           // [2 S]
  float* pred_out;       // [S] (nullable): the leaf values again, inside the tree's output block
};
__global__ void __launch_bounds__(128) leaf_value_kernel(LeafValueArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * a.S) return;
  if (i < a.S) {
    a.pred[i] = 0.f;
    return;
  }
  const long long* q = a.acc + (size_t)(i - a.S) * 4;
  const double num = __dadd_rn(ldexp((double)q[0], kGammaSplitBits - a.e_num), ldexp((double)q[1], -a.e_num));
  const double den = __dadd_rn(ldexp((double)q[2], kGammaSplitBits - a.e_den), ldexp((double)q[3], -a.e_den));
  const double gamma = den == 0 ? 0.0 : __ddiv_rn(num, den);
  double gf = __dmul_rn(a.rate, gamma);
  if (a.multinomial) {
    if (gf > 1e4) gf = (double)1e4f;
    else if (gf < -1e4) gf = (double)-1e4f;
  }
  if (isnan(gf)) gf = 0;
  else if (isinf(gf)) gf = (gf > 0 ? 1.0 : -1.0) * (double)1e4f;
  if (gf > a.max_abs) gf = a.max_abs;
  if (gf < -a.max_abs) gf = -a.max_abs;
  a.pred[i] = (float)gf;
  if (a.pred_out) a.pred_out[i - a.S] = (float)gf;
}

}  // namespace h2ot
