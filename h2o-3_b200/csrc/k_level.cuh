// Per-level kernels: score_decide (row -> leaf), histogram accumulation, finalize.
// Replaces ScoreBuildHistogram2.setupLocal phases 1+2 and postGlobal
// (T/ScoreBuildHistogram2.java:139-168, 192-310, 412-466; T/DHistogram.java:609-704, 734-740).
#pragma once
#include "common.cuh"

namespace h2ot {

// Device-side histogram descriptor for one (leaf, col): host desc + placement.
struct DescDev {
  double min, step;
  int32_t nbin;      // -1: inactive
  int32_t smem_off;  // word offset of bin 0 inside the CTA's shared histogram (smem mode)
  int64_t gbin;      // global bin offset (3 int64 accumulators per bin)
  int64_t voff;      // offset (doubles) of the block in the packed vals output
  uint32_t flags;    // H2OT_DESC_*
  uint32_t pad;
};
struct GroupDev {
  int32_t c_begin, c_end;  // column range [c_begin, c_end)
  int32_t smem_words;      // histogram words needed by this group
  int32_t mode;            // 0: shared-memory privatised, 1: global atomics
};

// ---------------------------------------------------------------------------------------------
// K2a: score_decide. One thread per row. T/ScoreBuildHistogram2.java:139-168 (+ :201-211 at the root),
// DecidedNode.getChildNodeID T/DTree.java:728-763.
// ---------------------------------------------------------------------------------------------
struct ScoreArgs {
  int64_t n, n_pad;
  int32_t leaf, n_leaves;
  int32_t* nids;
  uint16_t* code;
  const h2ot_decided_node* nodes;  // indexed by nid
  const uint8_t* bitsets;
  const ColRef* cols;
  unsigned long long* leaf_rows;  // [n_leaves]
};

__device__ __forceinline__ int child_node_id(const h2ot_decided_node& dn, const ColRef* cols, const uint8_t* bitsets,
                                             int64_t row) {
  double d = load_value(cols[dn.col], row);
  int bin = -1;
  bool isNA = isnan(d);
  if (!isNA) {
    if (dn.nasplit == H2OT_NA_VS_REST) bin = 0;
    else if (dn.equal == 0) bin = (d >= (double)dn.splat) ? 1 : 0;  // double vs float (q4)
    else {
      int b = __double2int_rz(d);
      int idx = b - dn.bs_bitoff;
      if (idx >= 0 && idx < dn.bs_nbits) bin = (bitsets[dn.bs_off + (idx >> 3)] >> (idx & 7)) & 1;
      else isNA = true;  // out-of-range level follows the NA direction (q5)
    }
  }
  if (isNA) bin = (dn.nasplit == H2OT_NA_LEFT || dn.nasplit == H2OT_LEFT) ? 0 : 1;
  return dn.child[bin];
}

constexpr int kScoreSmemLeaves = 4096;

__global__ void __launch_bounds__(256) score_decide_kernel(ScoreArgs a) {
  __shared__ unsigned int s_cnt[kScoreSmemLeaves];
  const bool use_s = a.n_leaves <= kScoreSmemLeaves;
  if (use_s)
    for (int i = threadIdx.x; i < a.n_leaves; i += blockDim.x) s_cnt[i] = 0;
  __syncthreads();
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < a.n_pad; row += (int64_t)gridDim.x * blockDim.x) {
    int res = -1;
    if (row < a.n) {
      int nid = a.nids[row];
      if (a.leaf == 0) {
        res = (nid == H2OT_DECIDED_ROW) ? -1 : 0;  // OOB rows are histogrammed at the root (q1)
      } else if (nid == H2OT_DECIDED_ROW) {
        res = -1;
      } else {
        bool oob = nid <= H2OT_OUT_OF_BAG;
        if (oob) nid = -nid + H2OT_OUT_OF_BAG;
        h2ot_decided_node dn = a.nodes[nid];
        bool done = false;
        if (dn.col < 0) {          // leftover non-split
          if (dn.pid < 0) { res = nid - a.leaf; done = true; }
          else {
            nid = dn.pid;
            int xnid = oob ? (-nid + H2OT_OUT_OF_BAG) : nid;
            a.nids[row] = xnid;
            dn = a.nodes[nid];
          }
        }
        if (!done) {
          nid = child_node_id(dn, a.cols, a.bitsets, row);
          if (nid != H2OT_DECIDED_ROW) {
            if (oob) nid = -nid + H2OT_OUT_OF_BAG;
            a.nids[row] = nid;
          }
          res = nid - a.leaf;
        }
      }
    }
    a.code[row] = (res >= 0 && res < a.n_leaves) ? (uint16_t)res : (uint16_t)kSkipCode;
    if (res >= 0 && res < a.n_leaves) {
      if (use_s) atomicAdd(&s_cnt[res], 1u);
      else atomicAdd(&a.leaf_rows[res], 1ULL);
    }
  }
  __syncthreads();
  if (use_s)
    for (int i = threadIdx.x; i < a.n_leaves; i += blockDim.x)
      if (s_cnt[i]) atomicAdd(&a.leaf_rows[i], (unsigned long long)s_cnt[i]);
}

// ---------------------------------------------------------------------------------------------
// K2b: histogram accumulation
// ---------------------------------------------------------------------------------------------
struct HistArgs {
  int64_t n_pad;
  int32_t L, C, n_groups, n_slices;
  const uint16_t* code;
  const float* work;
  const double* w;  // nullptr: unit weights
  const ColRef* cols;
  const DescDev* descs;  // [L][C]
  const GroupDev* groups;
  long long* acc;              // [total_bins][3] fixed point
  unsigned long long* minmax;  // [L][C][2] : {~dkey(min2), dkey(maxIn)}, 0 == untouched
  double sy, syy, sw;          // 2^k scales of wY, wYY, w
};

struct __align__(16) DescS {  // shared-memory copy, 32 B
  double min, step;
  int32_t nbin;  // -1 inactive
  int32_t soff;
  uint32_t flags;
  uint32_t pad;
};
struct __align__(16) MinMaxS {
  unsigned long long mn, mx;
};

constexpr int kHistThreads = 1024;
constexpr int kRowsPerThread = 4;
constexpr int kTileRows = kHistThreads * kRowsPerThread;  // == kRowPad

// Loads 4 consecutive rows of one column as doubles (NaN == NA). 128-bit / 64-bit / 32-bit vector loads.
__device__ __forceinline__ void load4(const ColRef& c, int64_t row, double (&x)[4]) {
  const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
  switch (c.storage) {
    case H2OT_STORE_F32: {
      float4 v = __ldcs(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(c.data) + row));
      x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
      break;
    }
    case H2OT_STORE_F64: {
      const double2* p = reinterpret_cast<const double2*>(reinterpret_cast<const double*>(c.data) + row);
      double2 a = __ldcs(p), b = __ldcs(p + 1);
      x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y;
      break;
    }
    case H2OT_STORE_U8: {
      uchar4 v = __ldcs(reinterpret_cast<const uchar4*>(reinterpret_cast<const unsigned char*>(c.data) + row));
      x[0] = v.x == 255 ? kNaN : c.base + (double)v.x;
      x[1] = v.y == 255 ? kNaN : c.base + (double)v.y;
      x[2] = v.z == 255 ? kNaN : c.base + (double)v.z;
      x[3] = v.w == 255 ? kNaN : c.base + (double)v.w;
      break;
    }
    default: {
      ushort4 v = __ldcs(reinterpret_cast<const ushort4*>(reinterpret_cast<const unsigned short*>(c.data) + row));
      x[0] = v.x == 65535 ? kNaN : c.base + (double)v.x;
      x[1] = v.y == 65535 ? kNaN : c.base + (double)v.y;
      x[2] = v.z == 65535 ? kNaN : c.base + (double)v.z;
      x[3] = v.w == 65535 ? kNaN : c.base + (double)v.w;
      break;
    }
  }
}

// Shared-memory privatised histograms. One CTA = (column group, row-tile stride). Per bin WPB 32-bit words:
//   unit weights (WPB 5): [cnt, wY.lo, wY.hi, wYY.lo, wYY.hi]
//   weighted     (WPB 7): [w.lo, w.hi, wY.lo, wY.hi, wYY.lo, wYY.hi, pad]
template <int WPB, bool HAS_W>
__global__ void __launch_bounds__(kHistThreads, 1) hist_smem_kernel(HistArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = blockIdx.x % a.n_groups;
  const int slice = blockIdx.x / a.n_groups;
  const GroupDev grp = a.groups[g];
  const int ncg = grp.c_end - grp.c_begin;
  const int npairs = ncg * a.L;
  DescS* descS = reinterpret_cast<DescS*>(smem_raw);
  MinMaxS* mmS = reinterpret_cast<MinMaxS*>(descS + npairs);
  ColRef* colS = reinterpret_cast<ColRef*>(mmS + npairs);
  uint32_t* hist = reinterpret_cast<uint32_t*>(colS + ncg);

  for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
    const int gc = p / a.L, l = p - gc * a.L;
    const DescDev d = a.descs[(size_t)l * a.C + grp.c_begin + gc];
    DescS s;
    s.min = d.min; s.step = d.step; s.nbin = d.nbin; s.soff = d.smem_off; s.flags = d.flags; s.pad = 0;
    descS[p] = s;
    mmS[p].mn = 0ULL; mmS[p].mx = 0ULL;
  }
  for (int i = threadIdx.x; i < ncg; i += blockDim.x) colS[i] = a.cols[grp.c_begin + i];
  for (int i = threadIdx.x; i < grp.smem_words; i += blockDim.x) hist[i] = 0u;
  __syncthreads();

  const int64_t n_tiles = a.n_pad / kTileRows;
  for (int64_t tile = slice; tile < n_tiles; tile += a.n_slices) {
    const int64_t row = tile * kTileRows + (int64_t)threadIdx.x * kRowsPerThread;
    const ushort4 cd = *reinterpret_cast<const ushort4*>(a.code + row);
    int leaf[4] = {cd.x, cd.y, cd.z, cd.w};
    if (leaf[0] == kSkipCode && leaf[1] == kSkipCode && leaf[2] == kSkipCode && leaf[3] == kSkipCode) continue;
    const float4 yv = *reinterpret_cast<const float4*>(a.work + row);
    const float ys[4] = {yv.x, yv.y, yv.z, yv.w};
    long long qy[4], qyy[4], qw[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      double y = (double)ys[j];
      double wy, wyy;
      if (HAS_W) {
        double wv = a.w[row + j];
        if (wv == 0.0) leaf[j] = kSkipCode;  // updateHisto skips w == 0 before min/max tracking (q7)
        wy = __dmul_rn(wv, y);
        wyy = __dmul_rn(wy, y);
        qw[j] = __double2ll_rn(__dmul_rn(wv, a.sw));
      } else {
        wy = y;
        wyy = __dmul_rn(y, y);
        qw[j] = 1;
      }
      qy[j] = __double2ll_rn(__dmul_rn(wy, a.sy));
      qyy[j] = __double2ll_rn(__dmul_rn(wyy, a.syy));
    }
    for (int gc = 0; gc < ncg; gc++) {
      double x[4];
      load4(colS[gc], row, x);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        if (leaf[j] == kSkipCode) continue;
        const int p = gc * a.L + leaf[j];
        const DescS d = descS[p];
        if (d.nbin < 0) continue;
        int b;
        if (isnan(x[j])) b = d.nbin;
        else {
          b = bin_of(x[j], d.min, d.step, d.nbin, (d.flags & H2OT_DESC_CHECKFLOAT) != 0);
          const unsigned long long k = dkey(x[j]);
          const MinMaxS m = mmS[p];
          if (~k > m.mn) atomicMax(&mmS[p].mn, ~k);
          if (k > m.mx) atomicMax(&mmS[p].mx, k);
        }
        uint32_t* wd = hist + d.soff + b * WPB;
        if (HAS_W) {
          smem_add64(wd, qw[j]);
          smem_add64(wd + 2, qy[j]);
          smem_add64(wd + 4, qyy[j]);
        } else {
          atomicAdd(wd, 1u);
          smem_add64(wd + 1, qy[j]);
          smem_add64(wd + 3, qyy[j]);
        }
      }
    }
  }
  __syncthreads();

  // deterministic merge: integer adds into the global fixed-point accumulators
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int p = warp; p < npairs; p += nwarps) {
    const DescS d = descS[p];
    if (d.nbin < 0) continue;
    const int gc = p / a.L, l = p - gc * a.L;
    const DescDev dd = a.descs[(size_t)l * a.C + grp.c_begin + gc];
    for (int b = lane; b <= d.nbin; b += 32) {
      const uint32_t* wd = hist + d.soff + b * WPB;
      long long vw, vy, vyy;
      if (HAS_W) {
        vw = (long long)(((unsigned long long)wd[1] << 32) | wd[0]);
        vy = (long long)(((unsigned long long)wd[3] << 32) | wd[2]);
        vyy = (long long)(((unsigned long long)wd[5] << 32) | wd[4]);
      } else {
        vw = (long long)wd[0];
        vy = (long long)(((unsigned long long)wd[2] << 32) | wd[1]);
        vyy = (long long)(((unsigned long long)wd[4] << 32) | wd[3]);
      }
      unsigned long long* dst = reinterpret_cast<unsigned long long*>(a.acc + (dd.gbin + b) * 3);
      if (vw) atomicAdd(dst, (unsigned long long)vw);
      if (vy) atomicAdd(dst + 1, (unsigned long long)vy);
      if (vyy) atomicAdd(dst + 2, (unsigned long long)vyy);
    }
    if (lane == 0) {
      const MinMaxS m = mmS[p];
      unsigned long long* mm = a.minmax + ((size_t)l * a.C + grp.c_begin + gc) * 2;
      if (m.mn) atomicMax(mm, m.mn);
      if (m.mx) atomicMax(mm + 1, m.mx);
    }
  }
}

// Global (L2) atomics variant: any shape (used when one column's leaves x bins exceed shared memory, and as a
// cross-check of the shared-memory kernel in the tests). One thread = 4 rows, loops over the group's columns.
template <bool HAS_W>
__global__ void __launch_bounds__(256) hist_global_kernel(HistArgs a, int c_begin, int c_end) {
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < a.n_pad / 4; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = q * 4;
    const ushort4 cd = *reinterpret_cast<const ushort4*>(a.code + row);
    int leaf[4] = {cd.x, cd.y, cd.z, cd.w};
    if (leaf[0] == kSkipCode && leaf[1] == kSkipCode && leaf[2] == kSkipCode && leaf[3] == kSkipCode) continue;
    const float4 yv = *reinterpret_cast<const float4*>(a.work + row);
    const float ys[4] = {yv.x, yv.y, yv.z, yv.w};
    long long qy[4], qyy[4], qw[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      double y = (double)ys[j];
      double wy, wyy;
      if (HAS_W) {
        double wv = a.w[row + j];
        if (wv == 0.0) leaf[j] = kSkipCode;
        wy = __dmul_rn(wv, y);
        wyy = __dmul_rn(wy, y);
        qw[j] = __double2ll_rn(__dmul_rn(wv, a.sw));
      } else {
        wy = y;
        wyy = __dmul_rn(y, y);
        qw[j] = 1;
      }
      qy[j] = __double2ll_rn(__dmul_rn(wy, a.sy));
      qyy[j] = __double2ll_rn(__dmul_rn(wyy, a.syy));
    }
    for (int c = c_begin; c < c_end; c++) {
      double x[4];
      load4(a.cols[c], row, x);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        if (leaf[j] == kSkipCode) continue;
        const DescDev d = a.descs[(size_t)leaf[j] * a.C + c];
        if (d.nbin < 0) continue;
        int b;
        if (isnan(x[j])) b = d.nbin;
        else {
          b = bin_of(x[j], d.min, d.step, d.nbin, (d.flags & H2OT_DESC_CHECKFLOAT) != 0);
          const unsigned long long k = dkey(x[j]);
          unsigned long long* mm = a.minmax + ((size_t)leaf[j] * a.C + c) * 2;
          if (~k > mm[0]) atomicMax(mm, ~k);
          if (k > mm[1]) atomicMax(mm + 1, k);
        }
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(a.acc + (d.gbin + b) * 3);
        atomicAdd(dst, (unsigned long long)qw[j]);
        if (qy[j]) atomicAdd(dst + 1, (unsigned long long)qy[j]);
        if (qyy[j]) atomicAdd(dst + 2, (unsigned long long)qyy[j]);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// finalize == postGlobal: fixed point -> fp64 and reducePrecision (T/DHistogram.java:734-740)
// ---------------------------------------------------------------------------------------------
struct FinalizeArgs {
  int32_t L, C;
  const DescDev* descs;
  const long long* acc;
  const unsigned long long* minmax;
  const unsigned long long* leaf_rows;
  double* vals;        // packed
  double* min2;        // [L*C]
  double* maxIn;       // [L*C]
  uint8_t* has_vals;   // [L*C]
  double* wna;         // [L*C] weight of the NA bucket (DHistogram.wNA, T/DHistogram.java:84)
  double inv_sy, inv_syy, inv_sw;
};

__global__ void __launch_bounds__(128) finalize_kernel(FinalizeArgs a) {
  const int p = blockIdx.x;  // (l, c)
  const int l = p / a.C;
  const DescDev d = a.descs[p];
  const bool has = d.nbin >= 0 && a.leaf_rows[l] > 0;
  if (threadIdx.x == 0) {
    a.has_vals[p] = has ? 1 : 0;
    const unsigned long long mn = a.minmax[(size_t)p * 2], mx = a.minmax[(size_t)p * 2 + 1];
    // untouched sentinels: DBL_MAX / -DBL_MAX (T/DHistogram.java:218-219); the int-optimised path
    // leaves (double)Integer.MAX_VALUE / MIN_VALUE behind once it has run (T/DHistogram.java:677-703)
    const bool intopt = has && (d.flags & H2OT_DESC_INTOPT);
    a.min2[p] = mn ? dkey_inv(~mn) : (intopt ? 2147483647.0 : DBL_MAX);
    a.maxIn[p] = mx ? dkey_inv(mx) : (intopt ? -2147483648.0 : -DBL_MAX);
    a.wna[p] = has ? __dmul_rn((double)a.acc[(d.gbin + d.nbin) * 3], a.inv_sw) : 0.0;
  }
  if (!has) return;
  for (int b = threadIdx.x; b <= d.nbin; b += blockDim.x) {
    const long long* s = a.acc + (d.gbin + b) * 3;
    double* v = a.vals + d.voff + (size_t)b * 3;
    v[0] = __dmul_rn((double)s[0], a.inv_sw);
    v[1] = (double)(float)__dmul_rn((double)s[1], a.inv_sy);
    v[2] = (double)(float)__dmul_rn((double)s[2], a.inv_syy);
  }
}

// per-row bin codes of one column under one descriptor (parity hook for DHistogram.bin)
__global__ void bin_codes_kernel(int64_t n, ColRef col, double mn, double step, int nbin, int checkFloat, uint16_t* out) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
    double x = load_value(col, row);
    out[row] = (uint16_t)(isnan(x) ? nbin : bin_of(x, mn, step, nbin, checkFloat != 0));
  }
}

}  // namespace h2ot
