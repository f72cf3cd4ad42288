// Chunk codecs on the device: Chunk.atd_impl of every numeric Chunk subclass (W/fvec/C*Chunk.java), so that the host can
// hand over Chunk._mem instead of running getDoubles on the CPU. One thread per row (dense codecs) or per
// stored element (sparse codecs). Byte order is little endian (UnsafeUtils on x86).
#pragma once
#include "common.cuh"

namespace h2ot {

struct ChunkArgs {
  const uint8_t* mem;  // device copy of Chunk._mem
  double* dst;         // first output row
  int64_t n;           // rows of the chunk
  int32_t codec;
  // CSChunk header (W/fvec/CSChunk.java:52-62): value = decimal ? (bias + x) / scale : (bias + x) * scale
  int64_t bias;
  double scale;
  int32_t decimal;
  // CBS: bits per value; CUD: number of uniques; CX*: id / value sizes, stored elements, NA-sparse flag
  int32_t bpv, nuniq, id_sz, val_sz, na_sparse;
  int64_t n_elems;
  double con;     // C0D / C0L constant
  int32_t* err;   // set to 1 when a sparse id is out of range
};

template <typename T>
__device__ __forceinline__ T ld_le(const uint8_t* p) {  // unaligned little-endian load
  unsigned long long v = 0;
#pragma unroll
  for (int i = 0; i < (int)sizeof(T); i++) v |= (unsigned long long)p[i] << (8 * i);
  if (sizeof(T) == 8) return (T)v;
  if (sizeof(T) == 4) return (T)(unsigned int)v;
  if (sizeof(T) == 2) return (T)(unsigned short)v;
  return (T)(unsigned char)v;
}

__device__ __forceinline__ double cs_value(const ChunkArgs& a, long long x) {  // CSChunk.getD :52-54
  const double m = (double)(a.bias + x);
  return a.decimal ? __ddiv_rn(m, a.scale) : __dmul_rn(m, a.scale);
}

constexpr int kCSOff = 16;  // CSChunk._OFF = 8 + 4 + 4

__global__ void __launch_bounds__(256) chunk_decode_kernel(ChunkArgs a) {
  const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * blockDim.x) {
    double v;
    switch (a.codec) {
      case H2OT_CODEC_C0D:
      case H2OT_CODEC_C0L: v = a.con; break;
      case H2OT_CODEC_C1: {
        const unsigned x = a.mem[i];
        v = x == 0xFFu ? kNaN : (double)x;
        break;
      }
      case H2OT_CODEC_C1N: v = (double)a.mem[i]; break;
      case H2OT_CODEC_C1S: {
        const unsigned x = a.mem[kCSOff + i];
        v = x == 0xFFu ? kNaN : cs_value(a, (long long)x);
        break;
      }
      case H2OT_CODEC_C2: {
        const short x = reinterpret_cast<const short*>(a.mem)[i];
        v = x == (short)-32768 ? kNaN : (double)x;
        break;
      }
      case H2OT_CODEC_C2S: {
        const short x = reinterpret_cast<const short*>(a.mem + kCSOff)[i];
        v = x == (short)-32768 ? kNaN : cs_value(a, (long long)x);
        break;
      }
      case H2OT_CODEC_C4: {
        const int x = reinterpret_cast<const int*>(a.mem)[i];
        v = x == INT32_MIN ? kNaN : (double)x;
        break;
      }
      case H2OT_CODEC_C4S: {
        const int x = reinterpret_cast<const int*>(a.mem + kCSOff)[i];
        v = x == INT32_MIN ? kNaN : cs_value(a, (long long)x);
        break;
      }
      case H2OT_CODEC_C4F: {
        const float x = reinterpret_cast<const float*>(a.mem)[i];
        v = x != x ? kNaN : (double)x;
        break;
      }
      case H2OT_CODEC_C8: {
        const long long x = reinterpret_cast<const long long*>(a.mem)[i];
        v = x == INT64_MIN ? kNaN : (double)x;
        break;
      }
      case H2OT_CODEC_C8D: v = reinterpret_cast<const double*>(a.mem)[i]; break;
      case H2OT_CODEC_CBS: {  // CBSChunk.read :63-68, 133-135
        const int64_t bit = i * a.bpv;
        const int b = (int)(signed char)a.mem[2 + (bit >> 3)];
        const int x = (b >> (int)(bit & 7)) & (1 | a.bpv);
        v = x == 2 ? kNaN : (double)x;
        break;
      }
      case H2OT_CODEC_CUD: {  // CUDChunk.atd_impl :44-47
        const int which = (int)(signed char)a.mem[8 + ((int64_t)a.nuniq << 3) + i] + 128;
        v = __longlong_as_double(ld_le<long long>(a.mem + 8 + ((int64_t)which << 3)));
        break;
      }
      default:  // CXI / CXF: rows without a stored element (CXIChunk.atd_impl :111-116)
        v = a.na_sparse ? kNaN : 0.0;
        break;
    }
    a.dst[i] = v;
  }
}

// stored elements of a sparse chunk, after chunk_decode_kernel filled the default (CXIChunk.getVal :36-48, CXFChunk.getVal :12-18)
__global__ void __launch_bounds__(256) chunk_scatter_kernel(ChunkArgs a) {
  const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
  const int elem = a.id_sz + a.val_sz;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < a.n_elems; k += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t* p = a.mem + 8 + k * elem;
    const int64_t id = a.id_sz == 4 ? (int64_t)ld_le<int>(p) : (int64_t)ld_le<unsigned short>(p);
    if (id < 0 || id >= a.n) {
      *a.err = 1;
      continue;
    }
    double v;
    if (a.codec == H2OT_CODEC_CXF) {
      v = a.val_sz == 4 ? (double)__uint_as_float(ld_le<unsigned int>(p + 4)) : __longlong_as_double(ld_le<long long>(p + 4));
      if (v != v) v = kNaN;
    } else {
      switch (a.val_sz) {
        case 0: v = 1.0; break;
        case 2: { const short x = ld_le<short>(p + a.id_sz); v = x == (short)-32768 ? kNaN : (double)x; break; }
        case 4: { const int x = ld_le<int>(p + a.id_sz); v = x == INT32_MIN ? kNaN : (double)x; break; }
        default: { const long long x = ld_le<long long>(p + a.id_sz); v = x == INT64_MIN ? kNaN : (double)x; break; }
      }
    }
    a.dst[id] = v;
  }
}

// decode of a finalized column back to doubles (h2ot_debug_get_column)
__global__ void __launch_bounds__(256) column_values_kernel(ColRef c, int64_t row0, int64_t n, double* out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = load_value(c, row0 + i);
}

}  // namespace h2ot
