// Per-level kernels: score_decide (row -> leaf), histogram accumulation, finalize.
// Replaces ScoreBuildHistogram2.setupLocal phases 1+2 and postGlobal
// (T/ScoreBuildHistogram2.java:139-168, 192-310, 412-466; T/DHistogram.java:609-704, 734-740).
#pragma once
#include <type_traits>

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
  uint32_t pad;      // low 16 bits: h2ot_hist_desc::order
  int32_t sp_off;    // H2OT_DESC_SPLITPTS: first split point in the level's pool (HistArgs::sp)
  int32_t sp_n;      // number of split points
  int32_t zero_pos;  // _zeroSplitPntPos (Arrays.binarySearch(_splitPts, 0.0), negative if absent)
  int32_t pad2;
};
// DHistogram.bin of a non-NA value under one descriptor, any histogram type
__device__ __forceinline__ int bin_desc(const DescDev& d, const double* sp, double x) {
  if (d.flags & H2OT_DESC_SPLITPTS)
    return bin_splitpts(x, d.min, d.step, d.nbin, sp + d.sp_off, d.sp_n, d.zero_pos, (d.flags & H2OT_DESC_ABSOLUTE) != 0,
                        (d.flags & H2OT_DESC_CHECKFLOAT) != 0);
  return bin_of(x, d.min, d.step, d.nbin, (d.flags & H2OT_DESC_CHECKFLOAT) != 0);
}
struct GroupDev {
  int32_t cols_off, ncols;  // the group's columns: group_cols[cols_off .. cols_off + ncols) (all of one storage type)
  int32_t smem_words;       // histogram words needed by this group
  int32_t mode;             // 0: shared-memory privatised, 1: global atomics
  int32_t storage;          // H2OT_STORE_* shared by the group's columns
  int32_t n_slices;         // CTAs working on this group (proportional to its estimated cost)
  int32_t cta_begin;        // first blockIdx.x of the group inside its launch
  int32_t l_begin, l_count; // leaves [l_begin, l_begin + l_count) handled by the group: all of them unless one column's
                            // histograms exceed shared memory, then the column is cut into leaf ranges (one group each)
  int32_t pad;
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

// One row of score_decide: returns the leaf-relative id (or -1) and updates nid in place.
__device__ __forceinline__ int score_row(const ScoreArgs& a, int64_t row, int& nid_io) {
  int nid = nid_io;
  if (a.leaf == 0) return (nid == H2OT_DECIDED_ROW) ? -1 : 0;  // OOB rows are histogrammed at the root (q1)
  if (nid == H2OT_DECIDED_ROW) return -1;
  const bool oob = nid <= H2OT_OUT_OF_BAG;
  if (oob) nid = -nid + H2OT_OUT_OF_BAG;
  h2ot_decided_node dn = a.nodes[nid];
  if (dn.col < 0) {  // leftover non-split
    if (dn.pid < 0) return nid - a.leaf;
    nid = dn.pid;
    nid_io = oob ? (-nid + H2OT_OUT_OF_BAG) : nid;
    dn = a.nodes[nid];
  }
  nid = child_node_id(dn, a.cols, a.bitsets, row);
  if (nid != H2OT_DECIDED_ROW) {
    if (oob) nid = -nid + H2OT_OUT_OF_BAG;  // stays OOB-encoded (negative): skipped below the root
    nid_io = nid;
  }
  return nid - a.leaf;
}

// Four consecutive rows per thread: 128-bit loads / stores of the nid column, one 64-bit store of the four leaf codes.
__global__ void __launch_bounds__(256) score_decide_kernel(ScoreArgs a) {
  __shared__ unsigned int s_cnt[kScoreSmemLeaves];
  const bool use_s = a.n_leaves <= kScoreSmemLeaves;
  if (use_s)
    for (int i = threadIdx.x; i < a.n_leaves; i += blockDim.x) s_cnt[i] = 0;
  __syncthreads();
  const int64_t nq = a.n_pad / 4;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row0 = q * 4;
    int4 nv = *reinterpret_cast<const int4*>(a.nids + row0);
    int nid[4] = {nv.x, nv.y, nv.z, nv.w};
    const int old[4] = {nv.x, nv.y, nv.z, nv.w};
    unsigned short code[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int res = -1;
      if (row0 + j < a.n) res = score_row(a, row0 + j, nid[j]);
      const bool in = res >= 0 && res < a.n_leaves;
      code[j] = in ? (unsigned short)res : (unsigned short)kSkipCode;
      if (in) {
        if (use_s) atomicAdd(&s_cnt[res], 1u);
        else atomicAdd(&a.leaf_rows[res], 1ULL);
      }
    }
    *reinterpret_cast<ushort4*>(a.code + row0) = make_ushort4(code[0], code[1], code[2], code[3]);
    if (nid[0] != old[0] || nid[1] != old[1] || nid[2] != old[2] || nid[3] != old[3])
      *reinterpret_cast<int4*>(a.nids + row0) = make_int4(nid[0], nid[1], nid[2], nid[3]);
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
  int32_t group_begin, pad0;
  const int32_t* group_cols;
  const uint16_t* code;
  const float* work;
  const double* w;  // nullptr: unit weights
  const ColRef* cols;
  const DescDev* descs;  // [L][C]
  const GroupDev* groups;
  long long* acc;              // [total_bins][3] fixed point
  unsigned long long* minmax;  // [L][C][2] : {~dkey(min2), dkey(maxIn)}, 0 == untouched
  double sy, syy, sw;          // 2^k scales of wY, wYY, w
  const double* sp;            // the level's split-point pool (descriptors with H2OT_DESC_SPLITPTS), else nullptr
};

// Shared-memory descriptor table of one CTA: one record of kTStride 32-bit words per (group column, leaf) pair
// p = gc * L + leaf. A warp's lanes sit in different leaves once the tree is a few levels deep; the odd stride puts the
// same word of up to 32 different records into 32 different banks (one wavefront per lookup), whereas 16/32-byte aligned
// records would serialise on the few bank-quads they start in.
enum { kTPlace = 0, kTMinLo, kTMinHi, kTStepLo, kTStepHi, kTKmin, kTKmax, kTStride = 7 };
//   kTPlace: bits 0..15 word offset of bin 0 in the CTA's histogram (0xFFFF: no histogram for this pair), bits 16..31 nbin
//   kTMinLo / kTMinHi: _min as fp64 halves -- or, for a column all of whose minima are fp32-exact (always the case for
//                      f32-stored columns: minima are observed values or float split points), the float bits in kTMinLo alone
//   kTKmin / kTKmax: ~key32 / key32 of the smallest / largest value seen (0: none) -- F32 / U8 / U16 storage
struct __align__(16) MinMaxS {  // 64-bit keys, only for F64-stored columns
  unsigned long long mn, mx;
};
struct Double4 {
  double2 a, b;
};
constexpr uint32_t kInactive = 0xFFFFu;  // place field of a pair without a histogram

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

// DHistogram.bin with the float-split correction (DRF), kept out of line: inlined it pushes the histogram loop past 64
// registers and the spills go through the same L1 pipe as the atomics
__device__ __noinline__ int bin_checkfloat(double x, double mn, double step, int nbin) { return bin_of(x, mn, step, nbin, true); }
// Same result without the two fp64 divisions for values that sit well inside their bin. The correction compares x with the
// FLOAT-rounded edges of its bin, float(min + i/step): an edge moves by at most 2^-24 |edge| (+ fp64 noise) when rounded, i.e. by
// less than g = 2^-22 (|min| step + nbin) bin widths, so a value whose fractional position is inside (g, 1 - g) keeps the
// truncated index. Everything else (edges, +-Inf, below-min) takes the exact path. Not applied to the last bin, where the
// reference skips the correction as well (idx + 1 < length fails).
__device__ __forceinline__ int bin_checkfloat_fast(double x, double mn, double step, int nbin) {
  const double pos = __dmul_rn(__dsub_rn(x, mn), step);
  const int idx = __double2int_rz(pos);
  if (idx >= nbin - 1) return nbin - 1;
  // x == min exactly (every row holding the node's smallest value of a few-valued column): bin 0 cannot be left downwards, and
  // upwards only if float(min + 1/step) <= min, i.e. a bin narrower than half a float ulp of min
  if (pos == 0.0 && step * fabs(mn) < 4194304.0 && step < 1e37) return 0;
  const double frac = pos - (double)idx;
  const double g = 2.384185791015625e-07 * (fabs(mn) * step + (double)nbin);  // 2^-22
  if (frac > g && frac < 1.0 - g) return idx;
  return bin_checkfloat(x, mn, step, nbin);
}

// ---- shared-memory access by 32-bit address, predicated in PTX: the compiler wraps every conditional atomic in a
// BSSY / BRA / BSYNC sequence, which tripled the instruction count of the inner loop.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void red_add_u32(uint32_t addr, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atom_add_u32(uint32_t addr, uint32_t v) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(addr), "r"(v) : "memory");
  return old;
}
// [addr] += hi + carry(old + lo). Unconditional: the high halves are non-zero for nearly every row (negative values, values
// past 2^32), and ptxas turns a predicated shared-memory atomic into a BSSY / BRA / BSYNC block that costs more than the add.
__device__ __forceinline__ void red_add_hi(uint32_t addr, uint32_t old, uint32_t lo, uint32_t hi) {
  asm volatile("{\n .reg .u32 t, h;\n add.cc.u32 t, %1, %2;\n addc.u32 h, %3, 0;\n red.shared.add.u32 [%0], h;\n}" ::"r"(addr), "r"(old),
               "r"(lo), "r"(hi)
               : "memory");
}
// if (on && v > cur) [addr] = max([addr], v)
__device__ __forceinline__ void red_max_if_gt(uint32_t addr, uint32_t v, uint32_t cur, uint32_t on) {
  asm volatile(
      "{\n .reg .pred q;\n setp.gt.u32 q, %1, %2;\n setp.ne.and.u32 q, %3, 0, q;\n @q red.shared.max.u32 [%0], %1;\n}" ::"r"(addr), "r"(v),
      "r"(cur), "r"(on) : "memory");
}

// R (row, col) items of one thread, staged so that independent work overlaps: record lookups, fp64 bin arithmetic, then
// the atomics grouped by word (the carries out of the low words are consumed last). Rows that take no part (skip code,
// no histogram for the pair) are steered to a per-lane sink bin instead of branching around the atomics.
// x: values as doubles; na: NA mask; key: order-preserving 32-bit keys (unused when KEY64).
struct RecRegs {  // one descriptor record in registers (root level: the same record serves every row of a column)
  uint32_t place, kmin, kmax;
  int nbin;
  double mn, st;
};
template <bool MN32>
__device__ __forceinline__ RecRegs load_rec(uint32_t rec) {
  RecRegs r;
  const uint32_t pn = lds_u32(rec + kTPlace * 4);
  r.place = pn & 0xFFFFu;
  r.nbin = (int)(pn >> 16);
  if (MN32) r.mn = (double)__uint_as_float(lds_u32(rec + kTMinLo * 4));
  else r.mn = __hiloint2double((int)lds_u32(rec + kTMinHi * 4), (int)lds_u32(rec + kTMinLo * 4));
  r.st = __hiloint2double((int)lds_u32(rec + kTStepHi * 4), (int)lds_u32(rec + kTStepLo * 4));
  r.kmin = lds_u32(rec + kTKmin * 4);
  r.kmax = lds_u32(rec + kTKmax * 4);
  return r;
}

template <int WPB, bool HAS_W, bool KEY64, bool CHECKFLOAT, bool ONE_LEAF, bool MN32, bool AGG, int J0, int R>
__device__ __forceinline__ void hist_items(uint32_t hist_a, uint32_t tab_a, uint32_t sink_a, MinMaxS* mm64, int pbase, int replica, const RecRegs& shared_rec,
                                           const int (&leaf)[4],
                                           const double (&x)[4], const bool (&na)[4], const uint32_t (&key)[4],
                                           const long long (&qw)[4], const long long (&qy)[4], const long long (&qyy)[4]) {
  uint32_t wd[4];
#pragma unroll
  for (int j = J0; j < J0 + R; j++) {
    const bool row_on = leaf[j] != kSkipCode;
    const int p = pbase + ((row_on && !ONE_LEAF) ? leaf[j] : 0);
    const uint32_t rec = tab_a + (uint32_t)p * (kTStride * 4);
    RecRegs rr;
    if (ONE_LEAF) rr = shared_rec;
    else rr = load_rec<MN32>(rec);
    const uint32_t place = rr.place;
    const uint32_t on = (row_on && place != kInactive) ? 1u : 0u;
    const int nbin = rr.nbin;
    const double mn = rr.mn, st = rr.st;
    int b = nbin;
    if (CHECKFLOAT) {
      // the float-split correction costs two fp64 divisions; with per-node mtries most (leaf, column) pairs have no histogram,
      // so it is only evaluated for rows that take part (a divergent branch is cheaper than dividing for nothing)
      // (the sign bit of the record's step carries the descriptor's H2OT_DESC_CHECKFLOAT flag)
      if (on && !na[j]) {
        const double sa = fabs(st);
        b = st < 0 ? bin_checkfloat_fast(x[j], mn, sa, nbin) : max(0, min(__double2int_rz(__dmul_rn(__dsub_rn(x[j], mn), sa)), nbin - 1));
      }
    } else {
      b = max(0, min(__double2int_rz(__dmul_rn(__dsub_rn(x[j], mn), st)), nbin - 1));  // +-Inf saturate onto the end bins
      b = na[j] ? nbin : b;
    }
    if (KEY64) {
      if (on && !na[j]) {
        const unsigned long long k = dkey(x[j]);
        const MinMaxS m = mm64[p];
        if (~k > m.mn) atomicMax(&mm64[p].mn, ~k);
        if (k > m.mx) atomicMax(&mm64[p].mx, k);
      }
    } else {
      // min / max keys: a new extreme is rare once a tile or two have gone by, so ONE branch guards both updates
      if (on && !na[j] && (~key[j] > rr.kmin || key[j] > rr.kmax)) {
        asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(rec + kTKmin * 4), "r"(~key[j]) : "memory");
        asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(rec + kTKmax * 4), "r"(key[j]) : "memory");
      }
    }
    // replica r of a pair sits r * (nbin + 1) bins after replica 0 (replica == 0 unless the planner replicated the column)
    wd[j] = on ? hist_a + (place + (uint32_t)((replica * (nbin + 1) + b) * WPB)) * 4u : sink_a;
  }
  constexpr uint32_t OW = 0, OY = (HAS_W ? 2 : 1) * 4, OYY = (HAS_W ? 4 : 3) * 4;
  if (AGG && !HAS_W) {
    // Few-valued column: the warp's rows pile up on a handful of bins, and same-address atomics serialise (a 3-valued column
    // cost 3.5x a continuous one). Per distinct bin among the warp's R x 32 items: every thread sums its own rows of that bin,
    // five REDUX give the warp's totals (24-bit limbs keep each of them inside 32 bits for addends below 2^46, which the
    // planner guarantees before it flags a column: frames of 2^18 rows or more), and one lane adds them.
    const unsigned lane = threadIdx.x & 31u;
    unsigned todo = 0u;
#pragma unroll
    for (int j = J0; j < J0 + R; j++) todo |= (wd[j] != sink_a ? 1u : 0u) << j;
    while (true) {
      const unsigned has = __ballot_sync(0xFFFFFFFFu, todo != 0u);
      if (!has) break;
      const int leader = __ffs(has) - 1;
      uint32_t mine = 0u;
#pragma unroll
      for (int j = J0 + R - 1; j >= J0; j--)
        if ((todo >> j) & 1u) mine = wd[j];
      const uint32_t wl = __shfl_sync(0xFFFFFFFFu, mine, leader);
      long long sy = 0, sv = 0;
      unsigned cnt = 0u;
#pragma unroll
      for (int j = J0; j < J0 + R; j++) {
        const bool in = ((todo >> j) & 1u) && wd[j] == wl;
        sy += in ? qy[j] : 0LL;
        sv += in ? qyy[j] : 0LL;
        cnt += in ? 1u : 0u;
        todo &= in ? ~(1u << j) : ~0u;
      }
      const unsigned t_ylo = __reduce_add_sync(0xFFFFFFFFu, (unsigned)sy & 0xFFFFFFu);
      const int t_yhi = __reduce_add_sync(0xFFFFFFFFu, (int)(sy >> 24));
      const unsigned t_vlo = __reduce_add_sync(0xFFFFFFFFu, (unsigned)sv & 0xFFFFFFu);
      const unsigned t_vhi = __reduce_add_sync(0xFFFFFFFFu, (unsigned)((unsigned long long)sv >> 24));
      const unsigned t_cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
      if ((int)lane == leader) {
        const long long ty = ((long long)t_yhi << 24) + (long long)t_ylo;
        const unsigned long long tv = ((unsigned long long)t_vhi << 24) + (unsigned long long)t_vlo;
        red_add_u32(wl + OW, t_cnt);
        const uint32_t o1 = atom_add_u32(wl + OY, (uint32_t)ty);
        const uint32_t o2 = atom_add_u32(wl + OYY, (uint32_t)tv);
        red_add_hi(wl + OY + 4, o1, (uint32_t)ty, (uint32_t)((unsigned long long)ty >> 32));
        red_add_hi(wl + OYY + 4, o2, (uint32_t)tv, (uint32_t)(tv >> 32));
      }
    }
    return;
  }
  // low words first (their return values carry into the high words), then the high words
  uint32_t oy[4], oyy[4], ow[4];
#pragma unroll
  for (int j = J0; j < J0 + R; j++) {
    if (HAS_W) ow[j] = atom_add_u32(wd[j] + OW, (uint32_t)qw[j]);
    else red_add_u32(wd[j] + OW, 1u);
  }
#pragma unroll
  for (int j = J0; j < J0 + R; j++) oy[j] = atom_add_u32(wd[j] + OY, (uint32_t)qy[j]);
#pragma unroll
  for (int j = J0; j < J0 + R; j++) oyy[j] = atom_add_u32(wd[j] + OYY, (uint32_t)qyy[j]);
#pragma unroll
  for (int j = J0; j < J0 + R; j++) {
    if (HAS_W) red_add_hi(wd[j] + OW + 4, ow[j], (uint32_t)qw[j], (uint32_t)((unsigned long long)qw[j] >> 32));
    red_add_hi(wd[j] + OY + 4, oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
    red_add_hi(wd[j] + OYY + 4, oyy[j], (uint32_t)qyy[j], (uint32_t)((unsigned long long)qyy[j] >> 32));
  }
}

#ifndef H2OT_ITEMS_IN_FLIGHT
#define H2OT_ITEMS_IN_FLIGHT 2
#endif
constexpr int kItemsInFlight = H2OT_ITEMS_IN_FLIGHT;  // rows of a thread whose lookups / atomics are interleaved (2 or 4)

// Shared-memory privatised histograms. One CTA = (column group, row-tile stride). Per bin WPB 32-bit words:
//   unit weights (WPB 5): [cnt, wY.lo, wY.hi, wYY.lo, wYY.hi]
//   weighted     (WPB 7): [w.lo, w.hi, wY.lo, wY.hi, wYY.lo, wYY.hi, pad]
// followed by 32 sink bins (one per lane) that swallow the adds of rows taking no part.
template <int WPB, bool HAS_W, int STORAGE, bool CHECKFLOAT, bool ONE_LEAF>
__global__ void __launch_bounds__(kHistThreads, 1) hist_smem_kernel(HistArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr bool KEY64 = STORAGE == H2OT_STORE_F64;
  int g = a.group_begin;
  while (g + 1 < a.group_begin + a.n_groups && (int)blockIdx.x >= a.groups[g + 1].cta_begin) g++;
  const GroupDev grp = a.groups[g];
  const int slice = (int)blockIdx.x - grp.cta_begin;
  const int ncg = grp.ncols;
  const int LG = grp.l_count;  // leaves of this group
  const int npairs = ncg * LG;
  MinMaxS* mmS = reinterpret_cast<MinMaxS*>(smem_raw);
  ColRef* colS = reinterpret_cast<ColRef*>(mmS + (KEY64 ? npairs : 0));
  int32_t* cidxS = reinterpret_cast<int32_t*>(colS + ncg);
  int32_t* mn32S = cidxS + ncg;  // per group column: 1 when every active minimum of the column is fp32-exact
  int32_t* repS = mn32S + ncg;   // per group column: histogram replicas (1, 2 or 4; lanes pick one by lane id)
  uint32_t* T = reinterpret_cast<uint32_t*>(repS + ncg);
  uint32_t* hist = T + (size_t)kTStride * npairs;
  uint32_t* sink = hist + grp.smem_words;

  for (int i = threadIdx.x; i < ncg; i += blockDim.x) {
    const int cv = a.group_cols[grp.cols_off + i];  // column | log2(replicas) << 24 | aggregate << 29
    const int c = cv & 0xFFFFFF;
    cidxS[i] = c;
    repS[i] = (1 << ((cv >> 24) & 31)) | (((cv >> 29) & 1) << 8);  // replicas | 0x100: aggregate inside the warp
    mn32S[i] = 1;
    colS[i] = a.cols[c];
  }
  __syncthreads();
  for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
    const int gc = p / LG, l = grp.l_begin + (p - gc * LG);
    const DescDev d = a.descs[(size_t)l * a.C + (a.group_cols[grp.cols_off + gc] & 0xFFFFFF)];
    uint32_t* r = T + (size_t)p * kTStride;
    r[kTMinLo] = (uint32_t)__double2loint(d.min);
    r[kTMinHi] = (uint32_t)__double2hiint(d.min);
    r[kTStepLo] = (uint32_t)__double2loint(d.step);
    r[kTStepHi] = (uint32_t)__double2hiint(d.step) | ((CHECKFLOAT && (d.flags & H2OT_DESC_CHECKFLOAT)) ? 0x80000000u : 0u);
    r[kTPlace] = (d.nbin < 0 ? kInactive : (uint32_t)d.smem_off) | ((uint32_t)max(d.nbin, 1) << 16);
    r[kTKmin] = 0u;
    r[kTKmax] = 0u;
    if (KEY64) { mmS[p].mn = 0ULL; mmS[p].mx = 0ULL; }
    if (d.nbin >= 0 && (double)(float)d.min != d.min) mn32S[gc] = 0;  // benign race: every writer stores 0
  }
  __syncthreads();
  for (int p = threadIdx.x; p < npairs; p += blockDim.x) {  // columns with fp32-exact minima keep the float bits in kTMinLo
    const int gc = p / LG;
    if (mn32S[gc]) {
      uint32_t* r = T + (size_t)p * kTStride;
      r[kTMinLo] = __float_as_uint((float)__hiloint2double((int)r[kTMinHi], (int)r[kTMinLo]));
    }
  }
  for (int i = threadIdx.x; i < grp.smem_words + 32 * WPB; i += blockDim.x) hist[i] = 0u;
  __syncthreads();
  const uint32_t hist_a = smem_addr(hist), tab_a = smem_addr(T);
  const uint32_t sink_a = smem_addr(sink) + (threadIdx.x & 31) * (WPB * 4);

  using RawT = typename std::conditional<STORAGE == H2OT_STORE_F32, uint4,
               typename std::conditional<STORAGE == H2OT_STORE_U16, uint2,
               typename std::conditional<STORAGE == H2OT_STORE_U8, unsigned int, Double4>::type>::type>::type;
  constexpr int kElem = STORAGE == H2OT_STORE_F32 ? 4 : STORAGE == H2OT_STORE_F64 ? 8 : STORAGE == H2OT_STORE_U8 ? 1 : 2;

  const int64_t n_tiles = a.n_pad / kTileRows;
  for (int64_t tile = slice; tile < n_tiles; tile += grp.n_slices) {
    const int64_t row = tile * kTileRows + (int64_t)threadIdx.x * kRowsPerThread;
    const ushort4 cd = __ldg(reinterpret_cast<const ushort4*>(a.code + row));
    int leaf[4] = {cd.x, cd.y, cd.z, cd.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {  // group-relative leaf; rows of other leaf ranges take no part here
      const int rel = leaf[j] - grp.l_begin;
      leaf[j] = (leaf[j] != kSkipCode && (unsigned)rel < (unsigned)LG) ? rel : kSkipCode;
    }
    // a warp none of whose rows takes part skips the tile; the decision is warp-uniform because the item loop of a few-valued
    // column uses warp-wide votes and reductions (a lane on its own must not leave)
    if (__all_sync(0xFFFFFFFFu, leaf[0] == kSkipCode && leaf[1] == kSkipCode && leaf[2] == kSkipCode && leaf[3] == kSkipCode)) continue;
    const float4 yv = __ldg(reinterpret_cast<const float4*>(a.work + row));
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
    // software pipeline over the group's columns: the next column's 4 values are in flight while this one is binned
    auto fetch = [&](int gc) -> RawT {
      const char* base = reinterpret_cast<const char*>(colS[gc].data) + row * kElem;
      if constexpr (STORAGE == H2OT_STORE_F64) {
        Double4 r;
        r.a = __ldcs(reinterpret_cast<const double2*>(base));
        r.b = __ldcs(reinterpret_cast<const double2*>(base) + 1);
        return r;
      } else {
        return __ldcs(reinterpret_cast<const RawT*>(base));
      }
    };
    RawT nxt = fetch(0);
    for (int gc = 0; gc < ncg; gc++) {
      const RawT cur = nxt;
      if (gc + 1 < ncg) nxt = fetch(gc + 1);
      const int pbase = gc * LG;
      double x[4];
      bool na[4];
      uint32_t key[4] = {0u, 0u, 0u, 0u};
      if constexpr (STORAGE == H2OT_STORE_F32) {
        const float xf[4] = {__uint_as_float(cur.x), __uint_as_float(cur.y), __uint_as_float(cur.z), __uint_as_float(cur.w)};
#pragma unroll
        for (int j = 0; j < 4; j++) { x[j] = (double)xf[j]; na[j] = xf[j] != xf[j]; key[j] = fkey(xf[j]); }
      } else if constexpr (STORAGE == H2OT_STORE_U8) {
        const double cb = colS[gc].base;
        const unsigned xc[4] = {cur & 255u, (cur >> 8) & 255u, (cur >> 16) & 255u, cur >> 24};
#pragma unroll
        for (int j = 0; j < 4; j++) { x[j] = cb + (double)xc[j]; na[j] = xc[j] == 255u; key[j] = xc[j] + 1u; }
      } else if constexpr (STORAGE == H2OT_STORE_U16) {
        const double cb = colS[gc].base;
        const unsigned xc[4] = {cur.x & 65535u, cur.x >> 16, cur.y & 65535u, cur.y >> 16};
#pragma unroll
        for (int j = 0; j < 4; j++) { x[j] = cb + (double)xc[j]; na[j] = xc[j] == 65535u; key[j] = xc[j] + 1u; }
      } else {
        x[0] = cur.a.x; x[1] = cur.a.y; x[2] = cur.b.x; x[3] = cur.b.y;
#pragma unroll
        for (int j = 0; j < 4; j++) na[j] = isnan(x[j]);
      }
      RecRegs rec0{};
      const int rs = repS[gc];
      const int replica = (int)(threadIdx.x & 31) & ((rs & 0xFF) - 1);
      constexpr bool kCanAgg = !HAS_W && !CHECKFLOAT;  // (the planner never flags a column of a weighted or float-split build)
      const bool agg = kCanAgg && (rs & 0x100) != 0;   // uniform per column
#define H2OT_ITEMS(MN32_, AGG_)                                                                                                         \
  do {                                                                                                                                  \
    if (ONE_LEAF) rec0 = load_rec<MN32_>(tab_a + (uint32_t)pbase * (kTStride * 4)); /* root level: one record per column */            \
    hist_items<WPB, HAS_W, KEY64, CHECKFLOAT, ONE_LEAF, MN32_, AGG_, 0, kItemsInFlight>(hist_a, tab_a, sink_a, mmS, pbase, replica, rec0, \
                                                                                        leaf, x, na, key, qw, qy, qyy);                 \
    if (kItemsInFlight < 4)                                                                                                             \
      hist_items<WPB, HAS_W, KEY64, CHECKFLOAT, ONE_LEAF, MN32_, AGG_, 2, 2>(hist_a, tab_a, sink_a, mmS, pbase, replica, rec0, leaf, x,   \
                                                                             na, key, qw, qy, qyy);                                     \
  } while (0)
      if (mn32S[gc]) {  // uniform per column
        if (agg) H2OT_ITEMS(true, kCanAgg);
        else H2OT_ITEMS(true, false);
      } else {
        if (agg) H2OT_ITEMS(false, kCanAgg);
        else H2OT_ITEMS(false, false);
      }
#undef H2OT_ITEMS
    }
  }
  __syncthreads();

  // deterministic merge: integer adds into the global fixed-point accumulators
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int p = warp; p < npairs; p += nwarps) {
    const uint32_t* r = T + (size_t)p * kTStride;
    const uint32_t place = r[kTPlace] & 0xFFFFu;
    if (place == kInactive) continue;
    const int nbin = (int)(r[kTPlace] >> 16);
    const int gc = p / LG, l = grp.l_begin + (p - gc * LG);
    const size_t lc = (size_t)l * a.C + cidxS[gc];
    const DescDev dd = a.descs[lc];
    const int nrep = repS[gc] & 0xFF;
    for (int b = lane; b <= nbin; b += 32) {
      long long vw = 0, vy = 0, vyy = 0;
      for (int r = 0; r < nrep; r++) {
        const uint32_t* wd = hist + place + (r * (nbin + 1) + b) * WPB;
        if (HAS_W) {
          vw += (long long)(((unsigned long long)wd[1] << 32) | wd[0]);
          vy += (long long)(((unsigned long long)wd[3] << 32) | wd[2]);
          vyy += (long long)(((unsigned long long)wd[5] << 32) | wd[4]);
        } else {
          vw += (long long)wd[0];
          vy += (long long)(((unsigned long long)wd[2] << 32) | wd[1]);
          vyy += (long long)(((unsigned long long)wd[4] << 32) | wd[3]);
        }
      }
      unsigned long long* dst = reinterpret_cast<unsigned long long*>(a.acc + (dd.gbin + b) * 3);
      if (vw) atomicAdd(dst, (unsigned long long)vw);
      if (vy) atomicAdd(dst + 1, (unsigned long long)vy);
      if (vyy) atomicAdd(dst + 2, (unsigned long long)vyy);
    }
    if (lane == 0) {
      // block-local keys -> the global order-preserving 64-bit form {~dkey(min), dkey(max)} (rank independent)
      unsigned long long gmn = 0ULL, gmx = 0ULL;
      if (KEY64) {
        gmn = mmS[p].mn;
        gmx = mmS[p].mx;
      } else {
        const double cb = colS[gc].base;
        const uint32_t kmin = r[kTKmin], kmax = r[kTKmax];
        if (kmin) {
          const uint32_t k = ~kmin;
          gmn = ~dkey(STORAGE == H2OT_STORE_F32 ? (double)fkey_inv(k) : cb + (double)(k - 1u));
        }
        if (kmax) gmx = dkey(STORAGE == H2OT_STORE_F32 ? (double)fkey_inv(kmax) : cb + (double)(kmax - 1u));
      }
      unsigned long long* mm = a.minmax + lc * 2;
      if (gmn) atomicMax(mm, gmn);
      if (gmx) atomicMax(mm + 1, gmx);
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
          b = bin_desc(d, a.sp, x[j]);
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
  double* conc;        // [L*C] sum_b (w_b / sum w)^2 : chance that two rows share a bin (cost model of the atomics), nullable
  double inv_sy, inv_syy, inv_sw;
};

// One (leaf, column) pair, by a CTA of 128 threads. smem_vals (nullable): the pair's finalised values are also left there
// ([3 (nbin + 1)] doubles; the split search fused behind this reads them without the round trip through global memory).
// Returns whether the pair has values (uniform over the CTA).
__device__ __forceinline__ bool finalize_pair(const FinalizeArgs& a, int p, double* smem_vals) {
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
  if (!has) {
    if (a.conc && threadIdx.x == 0) a.conc[p] = 0.0;
    return false;
  }
  double sw = 0, sww = 0;
  for (int b = threadIdx.x; b <= d.nbin; b += blockDim.x) {
    const long long* s = a.acc + (d.gbin + b) * 3;
    const long long s0 = s[0], s1 = s[1], s2 = s[2];
    double* v = a.vals + d.voff + (size_t)b * 3;
    const double v0 = __dmul_rn((double)s0, a.inv_sw);
    const double v1 = (double)(float)__dmul_rn((double)s1, a.inv_sy);
    const double v2 = (double)(float)__dmul_rn((double)s2, a.inv_syy);
    v[0] = v0;
    v[1] = v1;
    v[2] = v2;
    if (smem_vals) {
      smem_vals[3 * b] = v0;
      smem_vals[3 * b + 1] = v1;
      smem_vals[3 * b + 2] = v2;
    }
    sw += (double)s0;
    sww += (double)s0 * (double)s0;
  }
  if (a.conc) {  // not part of the reference: feeds the planner's per-column cost estimate
    __shared__ double r_w[128], r_ww[128];
    r_w[threadIdx.x] = sw;
    r_ww[threadIdx.x] = sww;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
      if ((int)threadIdx.x < o) {
        r_w[threadIdx.x] += r_w[threadIdx.x + o];
        r_ww[threadIdx.x] += r_ww[threadIdx.x + o];
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) a.conc[p] = r_w[0] > 0 ? r_ww[0] / (r_w[0] * r_w[0]) : 0.0;
  }
  return true;
}

__global__ void __launch_bounds__(128) finalize_kernel(FinalizeArgs a) { finalize_pair(a, blockIdx.x, nullptr); }

// per-row bin codes of one column under one descriptor (parity hook for DHistogram.bin)
__global__ void bin_codes_kernel(int64_t n, ColRef col, DescDev d, const double* sp, uint16_t* out) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
    double x = load_value(col, row);
    out[row] = (uint16_t)(isnan(x) ? d.nbin : bin_desc(d, sp, x));
  }
}

}  // namespace h2ot
