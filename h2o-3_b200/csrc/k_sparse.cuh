// Histogram accumulation for SPARSE levels: per-node column sampling (DRF mtries, GBM col_sample_rate) gives every leaf its
// own few `_scoreCols` (T/DTree.java:443-499; q8: T/ScoreBuildHistogram2.java:421-422), so most (leaf, column) pairs have no
// histogram. The column-major kernel (k_level.cuh) would test rows x all columns; here a CTA owns a RANGE OF LEAVES, scans
// the u16 leaf codes of its rows and, for a row of one of its leaves, gathers only that leaf's columns. Same bins, same
// fixed-point sums, same keys as hist_smem_kernel -- integer adds commute, so both kernels give identical bits.
// (T/DHistogram.java:609-658 updateHisto, :272-306 bin.)
#pragma once
#include "k_level.cuh"

namespace h2ot {

// one record per (leaf of the range, slot): the leaf's j-th active column. 11 words: odd stride, so the same word of 32
// different records sits in 32 different banks.
enum { kSPlace = 0, kSMinLo, kSMinHi, kSStepLo, kSStepHi, kSKmin, kSKmax, kSCol, kSSpOff, kSSpN, kSZero, kSStride = 11 };
//   kSSpOff: first split point of the pair in HistArgs::sp, 0xFFFFFFFF when the pair bins uniformly; kSSpN: their number, bit 31 =
//   absolute points; kSZero: _zeroSplitPntPos. Only read by the SPLITPTS instantiations.
constexpr int kSparseQueueWords = (kHistThreads / 32) * (64 + 32);  // compaction queues of the CTA's warps
constexpr int kSparseColsSmem = 512;  // ColRef table is staged in shared memory up to this many frame columns

template <int WPB, bool HAS_W, bool CHECKFLOAT, bool SPLITPTS>
__global__ void __launch_bounds__(kHistThreads, 1) hist_sparse_kernel(HistArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int g = a.group_begin;
  while (g + 1 < a.group_begin + a.n_groups && (int)blockIdx.x >= a.groups[g + 1].cta_begin) g++;
  const GroupDev grp = a.groups[g];
  const int slice = (int)blockIdx.x - grp.cta_begin;
  const int LG = grp.l_count, M = grp.ncols;  // leaves of the range, slots per leaf
  const int npairs = LG * M;
  const bool any64 = grp.storage != 0;  // some active column is stored as f64: 64-bit min/max keys are kept beside the records
  const bool cols_in_smem = a.C <= kSparseColsSmem;
  MinMaxS* mmS = reinterpret_cast<MinMaxS*>(smem_raw);
  ColRef* colS = reinterpret_cast<ColRef*>(mmS + (any64 ? npairs : 0));
  uint32_t* T = reinterpret_cast<uint32_t*>(colS + (cols_in_smem ? a.C : 0));
  uint32_t* queue = T + (size_t)kSStride * npairs;  // per warp: 64 row ids (u32) then, after all of them, 64 leaf ids (u16)
  uint32_t* hist = queue + kSparseQueueWords;
  const ColRef* colP = cols_in_smem ? colS : a.cols;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;

  if (cols_in_smem)
    for (int i = threadIdx.x; i < a.C; i += blockDim.x) colS[i] = a.cols[i];
  for (int ll = warp; ll < LG; ll += nwarps) {  // compact each leaf's active columns into its slots
    const size_t l = (size_t)(grp.l_begin + ll);
    int cnt = 0;
    for (int c0 = 0; c0 < a.C; c0 += 32) {
      const int c = c0 + lane;
      DescDev d{};
      d.nbin = -1;
      if (c < a.C) d = a.descs[l * a.C + c];
      const bool act = d.nbin >= 0;
      const unsigned m = __ballot_sync(0xffffffffu, act);
      if (act) {
        const int p = ll * M + cnt + __popc(m & ((1u << lane) - 1u));
        uint32_t* r = T + (size_t)p * kSStride;
        r[kSPlace] = (uint32_t)d.smem_off | ((uint32_t)d.nbin << 16);
        r[kSMinLo] = (uint32_t)__double2loint(d.min);
        r[kSMinHi] = (uint32_t)__double2hiint(d.min);
        r[kSStepLo] = (uint32_t)__double2loint(d.step);
        r[kSStepHi] = (uint32_t)__double2hiint(d.step) | ((CHECKFLOAT && (d.flags & H2OT_DESC_CHECKFLOAT)) ? 0x80000000u : 0u);
        r[kSKmin] = 0u;
        r[kSKmax] = 0u;
        r[kSCol] = (uint32_t)c;
        if (SPLITPTS) {
          const bool sp = (d.flags & H2OT_DESC_SPLITPTS) != 0;
          r[kSSpOff] = sp ? (uint32_t)d.sp_off : 0xFFFFFFFFu;
          r[kSSpN] = (uint32_t)d.sp_n | ((d.flags & H2OT_DESC_ABSOLUTE) ? 0x80000000u : 0u);
          r[kSZero] = (uint32_t)d.zero_pos;
        }
        if (any64) { mmS[p].mn = 0ULL; mmS[p].mx = 0ULL; }
      }
      cnt += __popc(m);
    }
    for (int j = cnt + lane; j < M; j += 32) T[(size_t)(ll * M + j) * kSStride + kSPlace] = kInactive | (1u << 16);
  }
  for (int i = threadIdx.x; i < grp.smem_words; i += blockDim.x) hist[i] = 0u;
  __syncthreads();
  const uint32_t hist_a = smem_addr(hist), tab_a = smem_addr(T);

  // One (row, leaf) item: gather the leaf's columns for that row and accumulate. Runs with any subset of lanes.
  auto process = [&](const int64_t row, const int relleaf) {
    const double y = (double)__ldg(a.work + row);
    double wy, wyy;
    long long qw = 1;
    if (HAS_W) {
      const double wv = a.w[row];
      if (wv == 0.0) return;  // updateHisto skips w == 0 before min/max tracking (q7)
      wy = __dmul_rn(wv, y);
      wyy = __dmul_rn(wy, y);
      qw = __double2ll_rn(__dmul_rn(wv, a.sw));
    } else {
      wy = y;
      wyy = __dmul_rn(y, y);
    }
    const long long qy = __double2ll_rn(__dmul_rn(wy, a.sy));
    const long long qyy = __double2ll_rn(__dmul_rn(wyy, a.syy));
    const int pbase = relleaf * M;
#pragma unroll 1
    for (int s = 0; s < M; s++) {
      const int p = pbase + s;
      const uint32_t rec = tab_a + (uint32_t)p * (kSStride * 4);
      const uint32_t pn = lds_u32(rec + kSPlace * 4);
      const uint32_t place = pn & 0xFFFFu;
      if (place == kInactive) break;  // slots are compacted
      const int nbin = (int)(pn >> 16);
      const double mn = __hiloint2double((int)lds_u32(rec + kSMinHi * 4), (int)lds_u32(rec + kSMinLo * 4));
      const double st = __hiloint2double((int)lds_u32(rec + kSStepHi * 4), (int)lds_u32(rec + kSStepLo * 4));
      const ColRef cr = colP[lds_u32(rec + kSCol * 4)];
      double x;
      bool na;
      uint32_t key = 0u;
      switch (cr.storage) {
        case H2OT_STORE_F32: {
          const float xf = __ldg(reinterpret_cast<const float*>(cr.data) + row);
          x = (double)xf; na = xf != xf; key = fkey(xf);
          break;
        }
        case H2OT_STORE_F64: {
          x = __ldg(reinterpret_cast<const double*>(cr.data) + row);
          na = isnan(x);
          break;
        }
        case H2OT_STORE_U8: {
          const unsigned v = __ldg(reinterpret_cast<const unsigned char*>(cr.data) + row);
          x = cr.base + (double)v; na = v == 255u; key = v + 1u;
          break;
        }
        default: {
          const unsigned v = __ldg(reinterpret_cast<const unsigned short*>(cr.data) + row);
          x = cr.base + (double)v; na = v == 65535u; key = v + 1u;
          break;
        }
      }
      int b = nbin;
      if (!na) {
        const double sa = CHECKFLOAT ? fabs(st) : st;  // sign bit of the record's step == the descriptor's H2OT_DESC_CHECKFLOAT
        uint32_t spo = 0xFFFFFFFFu;
        if (SPLITPTS) spo = lds_u32(rec + kSSpOff * 4);
        if (SPLITPTS && spo != 0xFFFFFFFFu) {
          const uint32_t spn = lds_u32(rec + kSSpN * 4);
          b = bin_splitpts(x, mn, sa, nbin, a.sp + spo, (int)(spn & 0x7FFFFFFFu), (int)lds_u32(rec + kSZero * 4), (spn >> 31) != 0,
                           CHECKFLOAT && st < 0);
        } else if (CHECKFLOAT && st < 0) b = bin_checkfloat_fast(x, mn, sa, nbin);
        else b = max(0, min(__double2int_rz(__dmul_rn(__dsub_rn(x, mn), sa)), nbin - 1));
        if (cr.storage == H2OT_STORE_F64) {
          const unsigned long long k = dkey(x);
          const MinMaxS m = mmS[p];
          if (~k > m.mn) atomicMax(&mmS[p].mn, ~k);
          if (k > m.mx) atomicMax(&mmS[p].mx, k);
        } else {
          red_max_if_gt(rec + kSKmin * 4, ~key, lds_u32(rec + kSKmin * 4), 1u);
          red_max_if_gt(rec + kSKmax * 4, key, lds_u32(rec + kSKmax * 4), 1u);
        }
      }
      const uint32_t wd = hist_a + (place + (uint32_t)(b * WPB)) * 4u;
      constexpr uint32_t OW = 0, OY = (HAS_W ? 2 : 1) * 4, OYY = (HAS_W ? 4 : 3) * 4;
      uint32_t ow = 0;
      if (HAS_W) ow = atom_add_u32(wd + OW, (uint32_t)qw);
      else red_add_u32(wd + OW, 1u);
      const uint32_t oy = atom_add_u32(wd + OY, (uint32_t)qy);
      const uint32_t oyy = atom_add_u32(wd + OYY, (uint32_t)qyy);
      if (HAS_W) red_add_hi(wd + OW + 4, ow, (uint32_t)qw, (uint32_t)((unsigned long long)qw >> 32));
      red_add_hi(wd + OY + 4, oy, (uint32_t)qy, (uint32_t)((unsigned long long)qy >> 32));
      red_add_hi(wd + OYY + 4, oyy, (uint32_t)qyy, (uint32_t)((unsigned long long)qyy >> 32));
    }
  };

  // Rows of this leaf range are a fraction of the rows scanned (row sampling, other ranges): each warp compacts them into
  // a 64-entry queue and works them off 32 at a time, so that the item loop runs with full warps.
  uint32_t* qrow = queue + warp * 64;
  uint16_t* qrel = reinterpret_cast<uint16_t*>(queue + nwarps * 64) + warp * 64;
  const unsigned lt = (1u << lane) - 1u;
  int qn = 0;  // warp-uniform
  const int64_t n_tiles = a.n_pad / kTileRows;
  for (int64_t tile = slice; tile < n_tiles; tile += grp.n_slices) {
    const int64_t row0 = tile * kTileRows + (int64_t)threadIdx.x * kRowsPerThread;
    const ushort4 cd = __ldg(reinterpret_cast<const ushort4*>(a.code + row0));
    const int codes[4] = {cd.x, cd.y, cd.z, cd.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int r = codes[j] - grp.l_begin;
      const bool act = codes[j] != kSkipCode && (unsigned)r < (unsigned)LG;
      const unsigned m = __ballot_sync(0xffffffffu, act);
      if (m == 0u) continue;
      if (act) {
        const int q = qn + __popc(m & lt);
        qrow[q] = (uint32_t)(row0 + j);
        qrel[q] = (uint16_t)r;
      }
      qn += __popc(m);
      __syncwarp();
      if (qn >= 32) {
        const uint32_t prow = qrow[lane];
        const int prel = qrel[lane];
        const int rest = qn - 32;
        uint32_t trow = 0;
        uint16_t trel = 0;
        if (lane < rest) { trow = qrow[32 + lane]; trel = qrel[32 + lane]; }
        __syncwarp();
        if (lane < rest) { qrow[lane] = trow; qrel[lane] = trel; }
        __syncwarp();
        qn = rest;
        process((int64_t)prow, prel);
      }
    }
  }
  if (lane < qn) process((int64_t)qrow[lane], (int)qrel[lane]);
  __syncthreads();

  // deterministic merge into the global fixed-point accumulators (same layout as hist_smem_kernel's)
  for (int p = warp; p < npairs; p += nwarps) {
    const uint32_t* r = T + (size_t)p * kSStride;
    const uint32_t place = r[kSPlace] & 0xFFFFu;
    if (place == kInactive) continue;
    const int nbin = (int)(r[kSPlace] >> 16);
    const int c = (int)r[kSCol];
    const size_t lc = (size_t)(grp.l_begin + p / M) * a.C + c;
    const DescDev dd = a.descs[lc];
    for (int b = lane; b <= nbin; b += 32) {
      const uint32_t* wd = hist + place + b * WPB;
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
      const ColRef cr = colP[c];
      unsigned long long gmn = 0ULL, gmx = 0ULL;
      if (cr.storage == H2OT_STORE_F64) {
        gmn = mmS[p].mn;
        gmx = mmS[p].mx;
      } else {
        const uint32_t kmin = r[kSKmin], kmax = r[kSKmax];
        if (kmin) {
          const uint32_t k = ~kmin;
          gmn = ~dkey(cr.storage == H2OT_STORE_F32 ? (double)fkey_inv(k) : cr.base + (double)(k - 1u));
        }
        if (kmax) gmx = dkey(cr.storage == H2OT_STORE_F32 ? (double)fkey_inv(kmax) : cr.base + (double)(kmax - 1u));
      }
      unsigned long long* mm = a.minmax + lc * 2;
      if (gmn) atomicMax(mm, gmn);
      if (gmx) atomicMax(mm + 1, gmx);
    }
  }
}

}  // namespace h2ot
