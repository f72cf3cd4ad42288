// Shared device/host helpers for the h2ot engine (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <string>

#include "../../include/h2ot.h"

namespace h2ot {

// ---------------------------------------------------------------------------------------------
// error plumbing: nothing throws across the C ABI
// ---------------------------------------------------------------------------------------------
void set_error(const std::string& msg);
#define H2OT_FAIL(code, msg)  \
  do {                        \
    ::h2ot::set_error(msg);   \
    return (code);            \
  } while (0)
#define CUDA_TRY(expr)                                                                                      \
  do {                                                                                                      \
    cudaError_t _e = (expr);                                                                                \
    if (_e != cudaSuccess) {                                                                                \
      ::h2ot::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " (" + __FILE__ + ":" + \
                        std::to_string(__LINE__) + ")");                                                    \
      return H2OT_ERR_CUDA;                                                                                 \
    }                                                                                                       \
  } while (0)

constexpr int kSkipCode = 0xFFFF;  // leaf code of a row that takes no part in this level
constexpr int kRowPad = 4096;      // row arrays are padded to a multiple of this (vector loads need no tail path)

// ---------------------------------------------------------------------------------------------
// column references (column-major, one narrow array per column, resident in HBM)
// ---------------------------------------------------------------------------------------------
struct ColRef {
  const void* data;
  double base;      // value = base + code for U8 / U16 storage
  int32_t storage;  // H2OT_STORE_*
  int32_t pad;
};

// value of one row as a double (NaN == NA) — used by the scalar paths (score / gamma / debug)
__device__ __forceinline__ double load_value(const ColRef& c, int64_t row) {
  switch (c.storage) {
    case H2OT_STORE_F32: return (double)__ldg(reinterpret_cast<const float*>(c.data) + row);
    case H2OT_STORE_F64: return __ldg(reinterpret_cast<const double*>(c.data) + row);
    case H2OT_STORE_U8: {
      unsigned v = __ldg(reinterpret_cast<const unsigned char*>(c.data) + row);
      return v == 255u ? __longlong_as_double(0x7ff8000000000000LL) : c.base + (double)v;
    }
    default: {
      unsigned v = __ldg(reinterpret_cast<const unsigned short*>(c.data) + row);
      return v == 65535u ? __longlong_as_double(0x7ff8000000000000LL) : c.base + (double)v;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// order-preserving integer keys for min/max tracking (0 is reserved for "no value yet")
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t dkey(double d) {
  union { double d; uint64_t u; } x;
  x.d = d;
  return x.u ^ ((x.u >> 63) ? 0xFFFFFFFFFFFFFFFFULL : 0x8000000000000000ULL);
}
__host__ __device__ __forceinline__ double dkey_inv(uint64_t k) {
  union { double d; uint64_t u; } x;
  x.u = k ^ ((k >> 63) ? 0x8000000000000000ULL : 0xFFFFFFFFFFFFFFFFULL);
  return x.d;
}
__device__ __forceinline__ uint32_t fkey(float f) {
  uint32_t u = __float_as_uint(f);
  return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(uint32_t k) {
  return __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xFFFFFFFFu));
}

// ---------------------------------------------------------------------------------------------
// DHistogram.bin (T/DHistogram.java:272-306) for the UniformAdaptive layout. fp64, no FMA contraction
// (sub then mul cannot fuse, and the TU is built with -fmad=false). (int) cast == Java's saturating d2i.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int bin_of(double x, double mn, double step, int nbin, bool checkFloat) {
  if (isinf(x)) return x < 0 ? 0 : nbin - 1;
  double pos = __dmul_rn(__dsub_rn(x, mn), step);
  int idx = __double2int_rz(pos);  // saturating, NaN -> 0 : same as the JVM
  if (checkFloat) {                // _splitPtsFloat correction, :289-301 ; binAt(i) = _min + i/_step (:308-312)
    if (idx + 1 < nbin) {
      float splitAt = (float)__dadd_rn(mn, __ddiv_rn((double)(idx + 1), step));
      if (x >= (double)splitAt) idx++;
      if (idx > 0) {
        float cur = (float)__dadd_rn(mn, __ddiv_rn((double)idx, step));
        if (!(x >= (double)cur)) idx--;
      }
    }
  }
  if (idx == nbin) idx--;
  // the reference asserts 0 <= idx < nbin; clamp so that a violated precondition cannot corrupt memory
  idx = max(0, min(idx, nbin - 1));
  return idx;
}

// ---------------------------------------------------------------------------------------------
// DHistogram.bin / binAt with split points (histogram_type Random / QuantilesGlobal / UniformRobust):
// T/DHistogram.java:281-286, 308-312. pts: the node's _splitPts (n of them), zero_pos: _zeroSplitPntPos
// (Arrays.binarySearch(_splitPts, 0.0), negative when absent), absolute: _absoluteSplitPts.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double bin_at_pts(double mn, double step, int b, const double* pts, bool absolute) {
  if (absolute) return pts[b];
  return __dadd_rn(mn, __ddiv_rn(pts ? pts[b] : (double)b, step));
}
// java.util.Arrays.binarySearch for a key that is neither NaN nor a zero (those never get here): same probe sequence, so
// the same index among duplicates
__device__ __forceinline__ int jbinary_search(const double* a, int n, double key) {
  int low = 0, high = n - 1;
  while (low <= high) {
    const int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
    const double v = a[mid];
    if (v < key) low = mid + 1;
    else if (v > key) high = mid - 1;
    else return mid;
  }
  return -(low + 1);
}
__device__ __forceinline__ int bin_splitpts(double x, double mn, double step, int nbin, const double* pts, int n, int zero_pos,
                                            bool absolute, bool checkFloat) {
  if (isinf(x)) return x < 0 ? 0 : nbin - 1;
  const double pos = absolute ? x : __dmul_rn(__dsub_rn(x, mn), step);
  int idx = pos == 0.0 ? zero_pos : jbinary_search(pts, n, pos);
  if (idx < 0) idx = -idx - 2;
  if (checkFloat && idx + 1 < nbin) {  // :287-301 with _splitPtsFloat[i] = (float) binAt(i)
    const float splitAt = (float)bin_at_pts(mn, step, idx + 1, pts, absolute);
    if (x >= (double)splitAt) idx++;
    if (idx > 0) {
      const float cur = (float)bin_at_pts(mn, step, idx, pts, absolute);
      if (!(x >= (double)cur)) idx--;
    }
  }
  if (idx == nbin) idx--;
  return max(0, min(idx, nbin - 1));  // the reference asserts the range; clamp so that a violated precondition cannot corrupt memory
}

// 64-bit signed accumulate into two adjacent 32-bit shared-memory words (lo, hi). Shared memory has
// native atomics only for 32-bit integers on sm_100a (64-bit / fp adds compile to CAS loops), so the
// carry is propagated by hand; integer adds commute, so the result is order independent.
__device__ __forceinline__ void smem_add64(uint32_t* w, long long v) {
  uint32_t lo = (uint32_t)v, hi = (uint32_t)((unsigned long long)v >> 32);
  uint32_t carry = 0;
  if (lo) {
    uint32_t old = atomicAdd(w, lo);
    carry = ((uint32_t)(old + lo) < lo) ? 1u : 0u;
  }
  uint32_t h = hi + carry;
  if (h) atomicAdd(w + 1, h);
}

__host__ __device__ __forceinline__ int ceil_log2_u64(uint64_t v) {
  int b = 0;
  while (b < 63 && (1ULL << b) < v) b++;
  return b;
}

}  // namespace h2ot
