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
// p = gc * L + leaf. Records are 48 bytes apart: a 128-bit lookup by 32 lanes sitting in 32 different leaves touches banks
// 12 i .. 12 i + 3 (mod 32), and 8 consecutive records cover the 32 banks exactly once, so the lookup takes the minimal four
// wavefronts (one when the lanes share a few leaves, as on the first levels: identical addresses are broadcast).
//   hot quad (one ld.shared.v4 per item):
//     kTPlace : bits 0..15 word offset of the pair's accumulators in the CTA's histogram (0xFFFF: none), bits 16..31 nbin
//     kTMin32 : _min as float bits            } fp32 pre-filter of DHistogram.bin: trunc((x - min32) * st32) with st32 = step * 2^11;
//     kTSt32  : float(step) * 2048, or NaN    } the 11 fraction bits tell whether the fp64 result could differ (then the exact
//               when the pair is not eligible } path runs). NaN sends every item of the pair down the exact path.
//     kTSnap  : [lowest bin seen : 16 | highest bin seen : 16], a (possibly stale) copy of kTBinLo / kTBinHi: only items whose bin
//               is at or beyond these can move the pair's min / max, everything else skips the key updates
//   cold words: kTKmin / kTKmax ~key32 / key32 of the smallest / largest value (0: none; F32 / U8 / U16 storage),
//     kTStepLo / kTStepHi the fp64 step (sign bit: the descriptor's H2OT_DESC_CHECKFLOAT), kTBinLo / kTBinHi the bin range seen,
//     kTMinLo / kTMinHi the fp64 _min
enum { kTPlace = 0, kTMin32, kTSt32, kTSnap, kTKmin, kTKmax, kTStepLo, kTStepHi, kTBinLo, kTBinHi, kTMinLo, kTMinHi, kTStride = 12 };
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

// Accumulator planes of one (leaf, column) pair with nbin + 1 bins (the last one is the NA bucket), each plane nbin + 1 words:
//   unit weights (kPlanesUnit):  [wY.lo][wY.hi][wYY.lo][cnt << 19 | wYY.hi]
//   weighted (kPlanesW):         [w.lo][w.hi][wY.lo][wY.hi][wYY.lo][wYY.hi]
// int64 fixed point kept as 32-bit halves (shared memory has native atomics for 32-bit integers only; 64-bit and fp adds are CAS
// loops). Plane-major, so that the banks of one atomic instruction follow the bins (uniform over the 32 banks) whatever the
// number of planes. With unit weights the row count shares a word with the high half of wYY -- four atomics per item instead
// of five: the count field holds 13 bits and the wYY.hi field 19; an add that leaves the count at kSpillAt or more empties the
// word into the global accumulators (atom.exch, then two global atomics; rare: a bin has to take 2048 rows of one CTA first).
// Every thread has at most four unchecked adds outstanding, so between the add that reaches kSpillAt and the exchange that
// follows it at most 4096 more rows can land: count < 2048 + 4096 < 2^13, wYY.hi < 6144 * 65 < 2^19 for 38-bit addends.
constexpr int kPlanesUnit = 4, kPlanesW = 6;
constexpr uint32_t kCntShift = 19, kHiMask = (1u << kCntShift) - 1u, kSpillAt = 2048u << kCntShift;

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

// ---- shared-memory access by 32-bit address, in PTX: the compiler wraps every conditional atomic in a
// BSSY / BRA / BSYNC sequence, which tripled the instruction count of the inner loop.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
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

// DHistogram.bin in fp64 from the record's cold words (pairs that are not eligible for the fp32 pre-filter, and the few items
// of eligible pairs that sit within 2^-11 of a bin edge). Out of line: it runs for a fraction of a percent of the items.
template <bool CHECKFLOAT>
__device__ __noinline__ int bin_exact_rec(uint32_t rec, double x, int nbin) {
  const double mn = __hiloint2double((int)lds_u32(rec + kTMinHi * 4), (int)lds_u32(rec + kTMinLo * 4));
  const double st = __hiloint2double((int)lds_u32(rec + kTStepHi * 4), (int)lds_u32(rec + kTStepLo * 4));
  if (CHECKFLOAT && st < 0) return bin_checkfloat_fast(x, mn, -st, nbin);
  return max(0, min(__double2int_rz(__dmul_rn(__dsub_rn(x, mn), fabs(st))), nbin - 1));  // +-Inf saturate onto the end bins
}

// An item whose bin is at or beyond the pair's bin range seen so far: it may hold a new smallest / largest value. Rare once a
// tile has gone by (the end bins of a node's range hold a fraction of a percent of its rows).
__device__ __noinline__ void minmax_slow32(uint32_t rec, uint32_t key, int bin) {
  asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(rec + kTKmin * 4), "r"(~key) : "memory");
  asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(rec + kTKmax * 4), "r"(key) : "memory");
  asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(rec + kTBinLo * 4), "r"((uint32_t)bin) : "memory");
  asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(rec + kTBinHi * 4), "r"((uint32_t)bin) : "memory");
}
__device__ __noinline__ void minmax_slow64(uint32_t rec, MinMaxS* mm, double x, int bin) {
  const unsigned long long k = dkey(x);
  atomicMax(&mm->mn, ~k);
  atomicMax(&mm->mx, k);
  asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(rec + kTBinLo * 4), "r"((uint32_t)bin) : "memory");
  asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(rec + kTBinHi * 4), "r"((uint32_t)bin) : "memory");
}

// The count | wYY.hi word of one bin has reached kSpillAt rows: empty it into the global accumulators. w3: the word's address,
// bin0_w3: the address of bin 0's word in the same plane; gbin: the pair's first global bin.
__device__ __noinline__ void spill_cnt_word(uint32_t w3, uint32_t bin0_w3, long long* acc, long long gbin) {
  uint32_t g;
  asm volatile("atom.shared.exch.b32 %0, [%1], %2;" : "=r"(g) : "r"(w3), "r"(0u) : "memory");
  if (g) {
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(acc + (gbin + (long long)((w3 - bin0_w3) >> 2)) * 3);
    atomicAdd(dst, (unsigned long long)(g >> kCntShift));
    atomicAdd(dst + 2, (unsigned long long)(g & kHiMask) << 32);
  }
}

#ifndef H2OT_ITEMS_IN_FLIGHT
#define H2OT_ITEMS_IN_FLIGHT 2
#endif
constexpr int kItemsInFlight = H2OT_ITEMS_IN_FLIGHT;  // rows of a thread whose lookups / atomics are interleaved (2 or 4)

// What one thread knows about the column it is working on (uniform over the CTA)
struct ColCtx {
  uint32_t hist_a, tab_a, sink_a;  // shared-memory addresses: histogram words, record table, this lane's sink word
  int pbase;                       // first record of the column: gc * LG
  int replica;                     // this lane's copy of the column's bins
  int c;                           // frame column
  int l_begin;                     // first leaf of the group
  double xbase;                    // U8 / U16 storage: value = xbase + code (the items carry the codes); 0 otherwise
  bool agg;                        // few-valued column: aggregate the warp's rows per distinct bin before touching shared memory
};

// R (row, col) items of one thread, staged so that independent work overlaps: record lookups, bin arithmetic, then the atomics
// grouped by plane (the carries out of the low words are consumed last). Rows that take no part (skip code, no histogram for
// the pair) add into a per-lane sink word instead of branching around the atomics.
template <bool HAS_W, bool KEY64, bool CHECKFLOAT, bool ONE_LEAF, int J0, int R, class XT>
__device__ __forceinline__ void hist_items(const HistArgs& a, const ColCtx& cx, MinMaxS* mm64, const uint4& rec0, const int (&leaf)[4],
                                           const XT (&x)[4], const uint32_t (&key)[4], const bool (&na)[4],
                                           const long long (&qw)[4], const long long (&qy)[4], const long long (&qyy)[4]) {
  constexpr int NPL = HAS_W ? kPlanesW : kPlanesUnit;
  uint32_t wd[4], ps[4], on4[4];
#pragma unroll
  for (int j = J0; j < J0 + R; j++) {
    const bool row_on = leaf[j] != kSkipCode;
    const int p = cx.pbase + ((row_on && !ONE_LEAF) ? leaf[j] : 0);
    const uint32_t rec = cx.tab_a + (uint32_t)p * (kTStride * 4);
    const uint4 q = ONE_LEAF ? rec0 : lds_v4(rec);
    const uint32_t place = q.x & 0xFFFFu;
    const int nbin = (int)(q.x >> 16);
    const bool on = row_on && place != kInactive;
    int bb;
    if (CHECKFLOAT || KEY64) {
      // fp64 all the way (DRF's float-split correction, f64-stored columns); the correction costs two fp64 divisions near bin
      // edges, so it is only evaluated for rows that take part
      bb = 0;
      if (on && !na[j]) {
        const double mn = __hiloint2double((int)lds_u32(rec + kTMinHi * 4), (int)lds_u32(rec + kTMinLo * 4));
        const double st = __hiloint2double((int)lds_u32(rec + kTStepHi * 4), (int)lds_u32(rec + kTStepLo * 4));
        const double xd = KEY64 ? (double)x[j] : cx.xbase + (double)x[j];
        bb = (CHECKFLOAT && st < 0) ? bin_checkfloat_fast(xd, mn, -st, nbin)
                                    : max(0, min(__double2int_rz(__dmul_rn(__dsub_rn(xd, mn), fabs(st))), nbin - 1));
      }
    } else {
      // fp32 pre-filter: x (a float value, or a U8 / U16 code) and min32 (= _min, or _min - base for codes) are exact floats, so
      // (x - min32) * st32 is within 3 * 2^-24 relative of the real product, i.e. within nbin * 2^11 * 1.8e-7 < 1 of it in units of
      // 2^-11 bins (eligible pairs have nbin <= 2048). Only if the 11 fraction bits are all 0 or all 1 can the fp64 truncation
      // of the reference differ: those items take the exact path.
      const int iv = __float2int_rz(__fmul_rn(__fsub_rn((float)x[j], __uint_as_float(q.y)), __uint_as_float(q.z)));
      bb = iv >> 11;
      if (on && !na[j] && ((iv + 1) & 2047) <= 1) bb = bin_exact_rec<false>(rec, cx.xbase + (double)x[j], nbin);
      bb = max(0, min(bb, nbin - 1));
    }
    // min / max of the pair: only values in (or beyond) the end bins seen so far can move them
    if (on && !na[j] && (bb <= (int)(q.w >> 16) || bb >= (int)(q.w & 0xFFFFu))) {
      if (KEY64) minmax_slow64(rec, mm64 + p, (double)x[j], bb);
      else minmax_slow32(rec, key[j], bb);
    }
    const int b = na[j] ? nbin : bb;
    // replica r of a pair sits r * NPL planes after replica 0 (replica == 0 unless the planner replicated the column)
    ps[j] = on ? (uint32_t)(nbin + 1) * 4u : 128u;
    wd[j] = on ? cx.hist_a + (place + (uint32_t)(cx.replica * NPL * (nbin + 1) + b)) * 4u : cx.sink_a;
    on4[j] = on ? 1u : 0u;
  }
  if (HAS_W) {
    uint32_t ow[4], oy[4], oyy[4];
#pragma unroll
    for (int j = J0; j < J0 + R; j++) ow[j] = atom_add_u32(wd[j], (uint32_t)qw[j]);
#pragma unroll
    for (int j = J0; j < J0 + R; j++) oy[j] = atom_add_u32(wd[j] + 2 * ps[j], (uint32_t)qy[j]);
#pragma unroll
    for (int j = J0; j < J0 + R; j++) oyy[j] = atom_add_u32(wd[j] + 4 * ps[j], (uint32_t)qyy[j]);
#pragma unroll
    for (int j = J0; j < J0 + R; j++) {
      red_add_hi(wd[j] + ps[j], ow[j], (uint32_t)qw[j], (uint32_t)((unsigned long long)qw[j] >> 32));
      red_add_hi(wd[j] + 3 * ps[j], oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
      red_add_hi(wd[j] + 5 * ps[j], oyy[j], (uint32_t)qyy[j], (uint32_t)((unsigned long long)qyy[j] >> 32));
    }
  } else if (!cx.agg) {
    uint32_t oy[4], oyy[4];
#pragma unroll
    for (int j = J0; j < J0 + R; j++) oy[j] = atom_add_u32(wd[j], (uint32_t)qy[j]);
#pragma unroll
    for (int j = J0; j < J0 + R; j++) oyy[j] = atom_add_u32(wd[j] + 2 * ps[j], (uint32_t)qyy[j]);
#pragma unroll
    for (int j = J0; j < J0 + R; j++) red_add_hi(wd[j] + ps[j], oy[j], (uint32_t)qy[j], (uint32_t)((unsigned long long)qy[j] >> 32));
#pragma unroll
    for (int j = J0; j < J0 + R; j++) {
      const uint32_t lo = (uint32_t)qyy[j];
      const uint32_t h = (uint32_t)((unsigned long long)qyy[j] >> 32) + ((uint32_t)(oyy[j] + lo) < lo ? 1u : 0u) + (1u << kCntShift);
      const uint32_t w3 = wd[j] + 3 * ps[j];
      const uint32_t o3 = atom_add_u32(w3, h);
      if (on4[j] && o3 + h >= kSpillAt) {
        const int l = cx.l_begin + (ONE_LEAF ? 0 : leaf[j]);
        const uint32_t pn = lds_u32(cx.tab_a + (uint32_t)(cx.pbase + (ONE_LEAF ? 0 : leaf[j])) * (kTStride * 4));
        const uint32_t bin0 = cx.hist_a + ((pn & 0xFFFFu) + (uint32_t)((cx.replica * NPL + 3) * (int)((pn >> 16) + 1))) * 4u;
        spill_cnt_word(w3, bin0, a.acc, a.descs[(size_t)l * a.C + cx.c].gbin);
      }
    }
  } else {
    // Few-valued column: the warp's rows pile up on a handful of words, and same-address atomics serialise (a 3-valued column
    // cost 3.5x a continuous one). Per distinct word among the warp's R x 32 items: every thread sums its own rows of that word,
    // five REDUX give the warp's totals (26-bit limbs keep each of them inside 32 bits), and one lane adds them.
    const unsigned lane = threadIdx.x & 31u;
    unsigned todo = 0u;
#pragma unroll
    for (int j = J0; j < J0 + R; j++) todo |= on4[j] << j;
    while (true) {
      const unsigned has = __ballot_sync(0xFFFFFFFFu, todo != 0u);
      if (!has) break;
      const int leader = __ffs(has) - 1;
      uint32_t mine = 0u, mine_ps = 0u;
      int mine_leaf = 0;  // (selected in the unrolled loop: a run-time index would push leaf[] into local memory)
#pragma unroll
      for (int j = J0 + R - 1; j >= J0; j--)
        if ((todo >> j) & 1u) { mine = wd[j]; mine_ps = ps[j]; mine_leaf = leaf[j]; }
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
      const unsigned t_ylo = __reduce_add_sync(0xFFFFFFFFu, (unsigned)sy & 0x3FFFFFFu);
      const int t_yhi = __reduce_add_sync(0xFFFFFFFFu, (int)(sy >> 26));
      const unsigned t_vlo = __reduce_add_sync(0xFFFFFFFFu, (unsigned)sv & 0x3FFFFFFu);
      const unsigned t_vhi = __reduce_add_sync(0xFFFFFFFFu, (unsigned)((unsigned long long)sv >> 26));
      const unsigned t_cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
      if ((int)lane == leader) {
        const long long ty = ((long long)t_yhi << 26) + (long long)t_ylo;
        const unsigned long long tv = ((unsigned long long)t_vhi << 26) + (unsigned long long)t_vlo;
        const uint32_t o1 = atom_add_u32(wl, (uint32_t)ty);
        const uint32_t o2 = atom_add_u32(wl + 2 * mine_ps, (uint32_t)tv);
        red_add_hi(wl + mine_ps, o1, (uint32_t)ty, (uint32_t)((unsigned long long)ty >> 32));
        const uint32_t lo = (uint32_t)tv;
        const uint32_t h = (uint32_t)(tv >> 32) + ((uint32_t)(o2 + lo) < lo ? 1u : 0u) + (t_cnt << kCntShift);
        const uint32_t w3 = wl + 3 * mine_ps;
        const uint32_t o3 = atom_add_u32(w3, h);
        if (o3 + h >= kSpillAt) {
          const int lf = ONE_LEAF ? 0 : mine_leaf;
          const uint32_t pn = lds_u32(cx.tab_a + (uint32_t)(cx.pbase + lf) * (kTStride * 4));
          const uint32_t bin0 = cx.hist_a + ((pn & 0xFFFFu) + (uint32_t)((cx.replica * NPL + 3) * (int)((pn >> 16) + 1))) * 4u;
          spill_cnt_word(w3, bin0, a.acc, a.descs[(size_t)(cx.l_begin + lf) * a.C + cx.c].gbin);
        }
      }
    }
  }
}

// Shared-memory privatised histograms. One CTA = (column group, row-tile stride); see the accumulator planes above. After the
// histogram words come 32 x 6 sink words (lane-owned, six planes 128 bytes apart) that swallow the adds of rows taking no part.
template <bool HAS_W, int STORAGE, bool CHECKFLOAT, bool ONE_LEAF>
__global__ void __launch_bounds__(kHistThreads, 1) hist_smem_kernel(HistArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr bool KEY64 = STORAGE == H2OT_STORE_F64;
  constexpr int NPL = HAS_W ? kPlanesW : kPlanesUnit;
  int g = a.group_begin;
  while (g + 1 < a.group_begin + a.n_groups && (int)blockIdx.x >= a.groups[g + 1].cta_begin) g++;
  const GroupDev grp = a.groups[g];
  const int slice = (int)blockIdx.x - grp.cta_begin;
  const int ncg = grp.ncols;
  const int LG = grp.l_count;  // leaves of this group
  const int npairs = ncg * LG;
  uint32_t* T = reinterpret_cast<uint32_t*>(smem_raw);  // 16-byte aligned records first
  MinMaxS* mmS = reinterpret_cast<MinMaxS*>(T + (size_t)kTStride * npairs);
  ColRef* colS = reinterpret_cast<ColRef*>(mmS + (KEY64 ? npairs : 0));
  int32_t* cidxS = reinterpret_cast<int32_t*>(colS + ncg);
  int32_t* repS = cidxS + ncg;  // per group column: histogram replicas (lanes pick one by lane id) | 0x100 when aggregated
  uint32_t* hist = reinterpret_cast<uint32_t*>(repS + ncg);
  uint32_t* sink = hist + grp.smem_words;

  for (int i = threadIdx.x; i < ncg; i += blockDim.x) {
    const int cv = a.group_cols[grp.cols_off + i];  // column | log2(replicas) << 24 | aggregate << 29
    const int c = cv & 0xFFFFFF;
    cidxS[i] = c;
    repS[i] = (1 << ((cv >> 24) & 31)) | (((cv >> 29) & 1) << 8);
    colS[i] = a.cols[c];
  }
  for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
    const int gc = p / LG, l = grp.l_begin + (p - gc * LG);
    const int c = a.group_cols[grp.cols_off + gc] & 0xFFFFFF;
    const DescDev d = a.descs[(size_t)l * a.C + c];
    const ColRef cr = a.cols[c];
    uint32_t* r = T + (size_t)p * kTStride;
    const bool chk = CHECKFLOAT && (d.flags & H2OT_DESC_CHECKFLOAT);
    r[kTMinLo] = (uint32_t)__double2loint(d.min);
    r[kTMinHi] = (uint32_t)__double2hiint(d.min);
    r[kTStepLo] = (uint32_t)__double2loint(d.step);
    r[kTStepHi] = (uint32_t)__double2hiint(d.step) | (chk ? 0x80000000u : 0u);
    r[kTPlace] = (d.nbin < 0 ? kInactive : (uint32_t)d.smem_off) | ((uint32_t)max(d.nbin, 1) << 16);
    r[kTKmin] = 0u;
    r[kTKmax] = 0u;
    r[kTBinLo] = 0xFFFFu;
    r[kTBinHi] = 0u;
    r[kTSnap] = 0xFFFF0000u;
    // fp32 pre-filter: the pair's min (relative to the column's base for U8 / U16 codes, which are exact floats themselves)
    // must be an exact float, the scaled step a normal float, nbin <= 2048 (error bound of the pre-filter) and the offset
    // below 2^20 bins (so that the fp64 rounding of min - base stays far inside that bound)
    const bool coded = STORAGE == H2OT_STORE_U8 || STORAGE == H2OT_STORE_U16;
    const double mrel = coded ? d.min - cr.base : d.min;
    const float mn32 = (float)mrel, st32 = (float)d.step;
    const bool ok = !KEY64 && !chk && d.nbin >= 1 && d.nbin <= 2048 && (double)mn32 == mrel && d.step > 1e-30 && d.step < 1e30 &&
                    (!coded || fabs(mrel) * d.step < 1048576.0);
    r[kTMin32] = __float_as_uint(mn32);
    r[kTSt32] = ok ? __float_as_uint(st32 * 2048.0f) : 0x7FC00000u;
    if (KEY64) { mmS[p].mn = 0ULL; mmS[p].mx = 0ULL; }
  }
  for (int i = threadIdx.x; i < grp.smem_words + 32 * kPlanesW; i += blockDim.x) hist[i] = 0u;
  __syncthreads();
  ColCtx cx;
  cx.hist_a = smem_addr(hist);
  cx.tab_a = smem_addr(T);
  cx.sink_a = smem_addr(sink) + (threadIdx.x & 31) * 4;
  cx.l_begin = grp.l_begin;

  using RawT = typename std::conditional<STORAGE == H2OT_STORE_F32, uint4,
               typename std::conditional<STORAGE == H2OT_STORE_U16, uint2,
               typename std::conditional<STORAGE == H2OT_STORE_U8, unsigned int, Double4>::type>::type>::type;
  using XT = typename std::conditional<KEY64, double, float>::type;
  constexpr int kElem = STORAGE == H2OT_STORE_F32 ? 4 : STORAGE == H2OT_STORE_F64 ? 8 : STORAGE == H2OT_STORE_U8 ? 1 : 2;

  const int64_t n_tiles = a.n_pad / kTileRows;
  for (int64_t tile = slice; tile < n_tiles; tile += grp.n_slices) {
    // refresh the bin-range snapshots the items test against (a stale one only sends more items down the slow path)
    for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
      uint32_t* r = T + (size_t)p * kTStride;
      r[kTSnap] = (r[kTBinLo] << 16) | r[kTBinHi];
    }
    const int64_t row = tile * kTileRows + (int64_t)threadIdx.x * kRowsPerThread;
    const ushort4 cd = __ldg(reinterpret_cast<const ushort4*>(a.code + row));
    int leaf[4] = {cd.x, cd.y, cd.z, cd.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {  // group-relative leaf; rows of other leaf ranges take no part here
      const int rel = leaf[j] - grp.l_begin;
      leaf[j] = (leaf[j] != kSkipCode && (unsigned)rel < (unsigned)LG) ? rel : kSkipCode;
    }
    // a warp none of whose rows takes part skips the tile; the decision is warp-uniform because the item loop below uses
    // warp-wide votes and reductions (a lane on its own must not leave)
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
      // the count | wYY.hi packing needs addends below 2^38; the scale guarantees it for |y| within the declared bound
      if (!HAS_W) qyy[j] = min(qyy[j], (1LL << 38) - 1);
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
      cx.pbase = gc * LG;
      cx.c = cidxS[gc];
      cx.xbase = (STORAGE == H2OT_STORE_U8 || STORAGE == H2OT_STORE_U16) ? colS[gc].base : 0.0;
      XT x[4];
      bool na[4];
      uint32_t key[4] = {0u, 0u, 0u, 0u};
      if constexpr (STORAGE == H2OT_STORE_F32) {
        x[0] = __uint_as_float(cur.x); x[1] = __uint_as_float(cur.y); x[2] = __uint_as_float(cur.z); x[3] = __uint_as_float(cur.w);
#pragma unroll
        for (int j = 0; j < 4; j++) { na[j] = x[j] != x[j]; key[j] = fkey(x[j]); }
      } else if constexpr (STORAGE == H2OT_STORE_U8) {
        const unsigned xc[4] = {cur & 255u, (cur >> 8) & 255u, (cur >> 16) & 255u, cur >> 24};
#pragma unroll
        for (int j = 0; j < 4; j++) { x[j] = (float)xc[j]; na[j] = xc[j] == 255u; key[j] = xc[j] + 1u; }
      } else if constexpr (STORAGE == H2OT_STORE_U16) {
        const unsigned xc[4] = {cur.x & 65535u, cur.x >> 16, cur.y & 65535u, cur.y >> 16};
#pragma unroll
        for (int j = 0; j < 4; j++) { x[j] = (float)xc[j]; na[j] = xc[j] == 65535u; key[j] = xc[j] + 1u; }
      } else {
        x[0] = cur.a.x; x[1] = cur.a.y; x[2] = cur.b.x; x[3] = cur.b.y;
#pragma unroll
        for (int j = 0; j < 4; j++) na[j] = isnan(x[j]);
      }
      const int rs = repS[gc];
      cx.replica = (int)(threadIdx.x & 31) & ((rs & 0xFF) - 1);
      cx.agg = !HAS_W && (rs & 0x100) != 0;
      uint4 rec0 = make_uint4(0u, 0u, 0u, 0u);
      if (ONE_LEAF) rec0 = lds_v4(cx.tab_a + (uint32_t)cx.pbase * (kTStride * 4));  // root level: one record per column
      hist_items<HAS_W, KEY64, CHECKFLOAT, ONE_LEAF, 0, kItemsInFlight>(a, cx, mmS, rec0, leaf, x, key, na, qw, qy, qyy);
      if (kItemsInFlight < 4) hist_items<HAS_W, KEY64, CHECKFLOAT, ONE_LEAF, 2, 2>(a, cx, mmS, rec0, leaf, x, key, na, qw, qy, qyy);
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
    const int pl = nbin + 1;
    for (int b = lane; b <= nbin; b += 32) {
      long long vw = 0, vy = 0, vyy = 0;
      for (int rr = 0; rr < nrep; rr++) {
        const uint32_t* wd = hist + place + rr * NPL * pl + b;
        if (HAS_W) {
          vw += (long long)(((unsigned long long)wd[pl] << 32) | wd[0]);
          vy += (long long)(((unsigned long long)wd[3 * pl] << 32) | wd[2 * pl]);
          vyy += (long long)(((unsigned long long)wd[5 * pl] << 32) | wd[4 * pl]);
        } else {
          vy += (long long)(((unsigned long long)wd[pl] << 32) | wd[0]);
          vw += (long long)(wd[3 * pl] >> kCntShift);
          vyy += (long long)(((unsigned long long)(wd[3 * pl] & kHiMask) << 32) | wd[2 * pl]);
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
  if (!has) {
    if (a.conc && threadIdx.x == 0) a.conc[p] = 0.0;
    return;
  }
  double sw = 0, sww = 0;
  for (int b = threadIdx.x; b <= d.nbin; b += blockDim.x) {
    const long long* s = a.acc + (d.gbin + b) * 3;
    double* v = a.vals + d.voff + (size_t)b * 3;
    v[0] = __dmul_rn((double)s[0], a.inv_sw);
    v[1] = (double)(float)__dmul_rn((double)s[1], a.inv_sy);
    v[2] = (double)(float)__dmul_rn((double)s[2], a.inv_syy);
    sw += (double)s[0];
    sww += (double)s[0] * (double)s[0];
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
}

// per-row bin codes of one column under one descriptor (parity hook for DHistogram.bin)
__global__ void bin_codes_kernel(int64_t n, ColRef col, DescDev d, const double* sp, uint16_t* out) {
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
    double x = load_value(col, row);
    out[row] = (uint16_t)(isnan(x) ? d.nbin : bin_desc(d, sp, x));
  }
}

}  // namespace h2ot
