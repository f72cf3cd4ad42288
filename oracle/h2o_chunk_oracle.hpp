// TEST INFRASTRUCTURE ONLY: CPU restatement of H2O's numeric chunk codecs -- the byte layouts that
// NewChunk.compress writes and the values Chunk.atd_impl reads back. Nothing under h2o-3_b200/ may include this file.
// Pinned by the reference's own unit-test vectors (tests/test_oracle_chunks.py: C1ChunkTest, C2ChunkTest, C4ChunkTest,
// C1SChunkTest.test_inflate_impl / test_precision, C2SChunkTest, C4SChunkTest, CBSChunkTest): the encoders below restate the
// writers, the decoders the readers, and the round trip must give the values those tests assert -- for test_precision that is
// Double.parseDouble("<mantissa>e<exponent>"), a truth independent of both.
// Paths: W = h2o-core/src/main/java/water.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <unordered_map>
#include <vector>

namespace orc_chunk {

enum Codec { C0D = 0, C0L, C1, C1N, C1S, C2, C2S, C4, C4S, C4F, C8, C8D, CBS, CUD, CXI, CXF };

template <typename T>
inline T get(const uint8_t* m, int64_t off) {  // W/util/UnsafeUtils.java:9-14 (native order == little endian on x86)
  T v;
  std::memcpy(&v, m + off, sizeof(T));
  return v;
}
template <typename T>
inline void put(std::vector<uint8_t>& m, int64_t off, T v) {
  std::memcpy(m.data() + off, &v, sizeof(T));
}

// W/util/PrettyPrint.java:80-130
inline double pow10_int(int exp) {
  static const double powers10[] = {0.0000000001, 0.000000001, 0.00000001, 0.0000001, 0.000001, 0.00001, 0.0001, 0.001, 0.01, 0.1, 1.0,
                                    10.0, 100.0, 1000.0, 10000.0, 100000.0, 1000000.0, 10000000.0, 100000000.0, 1000000000.0,
                                    10000000000.0};
  return (exp >= -10 && exp <= 10) ? powers10[exp + 10] : std::pow(10.0, exp);
}
inline double pow10(int64_t m, int e) { return e < 0 ? (double)m / pow10_int(-e) : (double)m * pow10_int(e); }
inline int64_t pow10i(int exp) {
  int64_t r = 1;
  if (exp > -1 && exp < 19) {
    for (int i = 0; i < exp; i++) r *= 10;
    return r;
  }
  return (int64_t)std::pow(10.0, exp);
}

constexpr int kCSOff = 8 + 4 + 4;  // W/fvec/CSChunk.java:18
constexpr int64_t kNAS[4] = {0xFF, INT16_MIN, INT32_MIN, INT64_MIN};  // W/fvec/NewChunk.java:1435

// ---------------------------------------------------------------------------------------------- readers (Chunk.atd_impl)
struct CSHeader {  // W/fvec/CSChunk.java:57-64
  int64_t bias;
  double scale;
  bool decimal;
  int sz_log;
};
inline CSHeader cs_header(const uint8_t* m) {
  CSHeader h;
  h.bias = get<int64_t>(m, 0);
  const int x = get<int32_t>(m, 8);
  const int sz = get<int32_t>(m, 12);
  h.decimal = sz < 0;
  h.sz_log = sz < 0 ? -sz - 1 : sz;
  h.scale = pow10(1, h.decimal ? -x : x);
  return h;
}
inline double cs_getD(const CSHeader& h, int64_t x, int64_t na) {  // :52-54
  const double nan = std::numeric_limits<double>::quiet_NaN();
  return x == na ? nan : (h.decimal ? (double)(h.bias + x) / h.scale : (double)(h.bias + x) * h.scale);
}

// number of rows as initFromBytes derives it; -1: malformed
inline int64_t rows_of(int codec, const uint8_t* m, int64_t len) {
  switch (codec) {
    case C0D: case C0L: return len == 12 ? get<int32_t>(m, 8) : -1;           // W/fvec/C0DChunk.java:12-21
    case C1: case C1N: return len;                                             // W/fvec/C1Chunk.java:10
    case C1S: return len >= kCSOff ? len - kCSOff : -1;                        // CSChunk.java:59
    case C2: return len >> 1;                                                  // W/fvec/C2Chunk.java:12
    case C2S: return len >= kCSOff ? (len - kCSOff) >> 1 : -1;
    case C4: case C4F: return len >> 2;                                        // W/fvec/C4Chunk.java:10
    case C4S: return len >= kCSOff ? (len - kCSOff) >> 2 : -1;
    case C8: case C8D: return len >> 3;                                        // W/fvec/C8Chunk.java:10
    case CBS: return len >= 2 ? ((len - 2) * 8 - (int8_t)m[0]) / (int8_t)m[1] : -1;  // W/fvec/CBSChunk.java:155-160
    case CUD: case CXI: case CXF: return len >= 8 ? get<int32_t>(m, 0) : -1;   // W/fvec/CUDChunk.java:93, CXIChunk.java:60
    default: return -1;
  }
}

inline double atd(int codec, const uint8_t* m, int64_t len, int64_t i) {
  const double nan = std::numeric_limits<double>::quiet_NaN();
  switch (codec) {
    case C0D: return get<double>(m, 0);                                              // C0DChunk.java:28
    case C0L: return (double)get<int64_t>(m, 0);                                     // C0LChunk.java:20
    case C1: { const int64_t r = 0xFF & m[i]; return r == 0xFF ? nan : (double)r; }  // C1Chunk.java:20-23
    case C1N: return (double)(0xFF & m[i]);                                          // C1NChunk.java:11
    case C1S: return cs_getD(cs_header(m), 0xFF & m[kCSOff + i], 0xFF);              // C1SChunk.java:31-33
    case C2: { const int r = get<int16_t>(m, i << 1); return r == INT16_MIN ? nan : (double)r; }   // C2Chunk.java:18-21
    case C2S: return cs_getD(cs_header(m), get<int16_t>(m, kCSOff + 2 * i), INT16_MIN);             // C2SChunk.java:32-37
    case C4: { const int64_t r = get<int32_t>(m, i << 2); return r == INT32_MIN ? nan : (double)r; }  // C4Chunk.java:16-19
    case C4S: return cs_getD(cs_header(m), get<int32_t>(m, kCSOff + 4 * i), INT32_MIN);             // C4SChunk.java:24-30
    case C4F: { const float r = get<float>(m, i << 2); return std::isnan(r) ? nan : (double)r; }     // C4FChunk.java:15-18
    case C8: { const int64_t r = get<int64_t>(m, i << 3); return r == INT64_MIN ? nan : (double)r; }  // C8Chunk.java:16-19
    case C8D: return get<double>(m, i << 3);                                                          // C8DChunk.java:16
    case CBS: {                                                                      // CBSChunk.java:41-44, 63-68, 133-135
      const int bpv = (int8_t)m[1];
      const int64_t bix = 2 + ((i * bpv) >> 3);
      const int off = (int)((bpv * i) & 7);
      const int8_t b = (int8_t)(((int)(int8_t)m[bix] >> off) & (1 | bpv));
      return b == 2 ? nan : (double)b;
    }
    case CUD: {                                                                      // CUDChunk.java:44-47
      const int nu = get<int32_t>(m, 4);
      const int which = (int)(int8_t)m[8 + ((int64_t)nu << 3) + i] + 128;
      return get<double>(m, 8 + ((int64_t)which << 3));
    }
    case CXI: case CXF: {                                                            // CXIChunk.java:58-116, CXFChunk.java:12-36
      const int id_sz = m[4] & 0xFF, val_sz = m[5] & 0xFF, elem = id_sz + val_sz;
      const bool is_na = (0xFF & m[6]) == 1;
      const int64_t nnz = (len - 8) / elem;
      int64_t lb = 0, ub = nnz;  // findOffset's binary search over the stored ids
      while (lb < ub) {
        const int64_t mid = lb + ((ub - lb) >> 1);
        const int64_t x = 8 + mid * elem;
        const int64_t id = id_sz == 4 ? get<int32_t>(m, x) : (0xFFFF & get<int16_t>(m, x));
        if (id == i) {
          if (codec == CXF) return elem == 8 ? (double)get<float>(m, x + 4) : get<double>(m, x + 4);
          int64_t v;
          switch (val_sz) {
            case 0: v = 1; break;
            case 2: v = get<int16_t>(m, x + 2); break;
            case 4: v = get<int32_t>(m, x + 4); break;
            default: v = get<int64_t>(m, x + 4); break;
          }
          const int64_t na = val_sz == 2 ? (int64_t)INT16_MIN : val_sz == 4 ? (int64_t)INT32_MIN : val_sz == 8 ? INT64_MIN : -1;
          return (val_sz != 0 && v == na) ? nan : (double)v;
        }
        if (id < i) lb = mid + 1;
        else ub = mid;
      }
      return is_na ? nan : 0.0;
    }
    default: return nan;
  }
}

// ---------------------------------------------------------------------------------------------- writers (NewChunk.compress)
// bufX with bias 0 / scale 0: C1 / C1N / C2 / C4 / C8 (W/fvec/NewChunk.java:1482-1534)
inline std::vector<uint8_t> enc_int(int log, const std::vector<int64_t>& v, const std::vector<uint8_t>& na) {
  std::vector<uint8_t> m(v.size() << log);
  for (size_t i = 0; i < v.size(); i++) {
    const int64_t le = na[i] ? kNAS[log] : v[i];
    switch (log) {
      case 0: m[i] = (uint8_t)le; break;
      case 1: put<int16_t>(m, i << 1, (int16_t)le); break;
      case 2: put<int32_t>(m, i << 2, (int32_t)le); break;
      default: put<int64_t>(m, i << 3, le); break;
    }
  }
  return m;
}

// the fixed-point ("fpoint") branch of NewChunk.compress (:1395-1409) + bufX + the C1S/C2S/C4SChunk constructors
// (W/fvec/C1SChunk.java:8-22, CSChunk.java:24-39). values are mantissa * 10^exponent; log picks the width (0, 1, 2).
inline std::vector<uint8_t> enc_scaled(int log, std::vector<int64_t> man, std::vector<int32_t> exp, const std::vector<uint8_t>& na) {
  const size_t n = man.size();
  for (size_t i = 0; i < n; i++) {  // NewChunk.addNum (:615, :624-628): canonical zero, trailing decimal zeros folded into the exponent
    if (man[i] == 0) exp[i] = 0;
    while (exp[i] < 0 && (man[i] / 10) * 10 == man[i]) {
      man[i] /= 10;
      exp[i]++;
    }
  }
  int xmin = INT32_MAX;  // NewChunk.compress :1284-1312: smallest exponent over the non-zero values, all trailing zeros stripped
  for (size_t i = 0; i < n; i++) {
    if (na[i] || man[i] == 0) continue;
    int64_t l = man[i];
    int x = exp[i];
    while ((l / 10) * 10 == l) {
      l /= 10;
      x++;
    }
    xmin = std::min(xmin, x);
  }
  if (xmin == INT32_MAX) xmin = 0;
  std::vector<int64_t> le(n, 0);
  int64_t lemin = INT64_MAX;
  for (size_t i = 0; i < n; i++) {
    if (na[i]) continue;
    const int x = exp[i] - xmin;
    le[i] = x >= 0 ? man[i] * pow10i(x) : man[i] / pow10i(-x);
    lemin = std::min(lemin, le[i]);
  }
  if (lemin == INT64_MAX) lemin = 0;
  const int64_t bias = log == 0 ? lemin : log == 1 ? 32767 + lemin : 2147483647LL + lemin;  // :1398-1407
  std::vector<uint8_t> m(kCSOff + (n << log));
  for (size_t i = 0; i < n; i++) {
    const int64_t s = na[i] ? kNAS[log] : le[i] - bias;
    switch (log) {
      case 0: m[kCSOff + i] = (uint8_t)s; break;
      case 1: put<int16_t>(m, kCSOff + (i << 1), (int16_t)s); break;
      default: put<int32_t>(m, kCSOff + (i << 2), (int32_t)s); break;
    }
  }
  put<int64_t>(m, 0, bias);
  put<int32_t>(m, 8, xmin);
  put<int32_t>(m, 12, log);
  if (xmin < 0) {  // precision check of the constructor: fall back to division by 10^-scale when the product is off
    const double scale = pow10(1, xmin), div = pow10(1, -xmin);
    for (size_t i = 0; i < n; i++) {
      if (na[i]) continue;
      const int64_t x = log == 0 ? (int64_t)(0xFF & m[kCSOff + i]) : log == 1 ? (int64_t)get<int16_t>(m.data(), kCSOff + (i << 1))
                                                                               : (int64_t)get<int32_t>(m.data(), kCSOff + (i << 2));
      if ((double)(bias + x) * scale != (double)(x + bias) / div) {
        put<int32_t>(m, 12, -log - 1);  // setDecimal :34-38
        break;
      }
    }
  }
  return m;
}

inline std::vector<uint8_t> enc_c0(bool is_long, double dcon, int64_t lcon, int32_t len) {  // C0DChunk.java:14-21, C0LChunk.java:11-15
  std::vector<uint8_t> m(12);
  if (is_long) put<int64_t>(m, 0, lcon);
  else put<double>(m, 0, dcon);
  put<int32_t>(m, 8, len);
  return m;
}

// NewChunk.bufB + CBSChunk(len, bpv) (W/fvec/NewChunk.java:1593-1601, CBSChunk.java:28-38, 57-61, 137-139)
inline std::vector<uint8_t> enc_cbs(const std::vector<uint8_t>& v, const std::vector<uint8_t>& na, int bpv) {
  const int64_t len = (int64_t)v.size();
  const int gap = (int)((8 - ((len * bpv) & 7)) & 7);
  const int64_t clen = 2 + (len >> (3 - bpv + 1)) + (gap == 0 ? 0 : 1);
  std::vector<uint8_t> m((size_t)clen, 0);
  m[0] = (uint8_t)gap;
  m[1] = (uint8_t)bpv;
  for (int64_t i = 0; i < len; i++) {
    const int val = na[i] ? 2 : (v[i] ? 1 : 0);
    if (!val) continue;
    const int64_t bix = 2 + ((i * bpv) >> 3);
    const int off = (int)((bpv * i) & 7);
    m[bix] = (uint8_t)(m[bix] | (val << off));
  }
  return m;
}

// CUDChunk constructor (W/fvec/CUDChunk.java:21-38). The reference numbers the uniques in HashMap iteration order; any
// order decodes to the same values, this one uses first-seen order.
inline std::vector<uint8_t> enc_cud(const std::vector<double>& v) {
  std::unordered_map<int64_t, int> idx;
  std::vector<int64_t> uniq;
  for (double d : v) {
    int64_t bits;
    std::memcpy(&bits, &d, 8);
    if (!idx.count(bits)) {
      idx[bits] = (int)uniq.size();
      uniq.push_back(bits);
    }
  }
  const int nu = (int)uniq.size();
  std::vector<uint8_t> m(8 + ((size_t)nu << 3) + v.size());
  put<int32_t>(m, 0, (int32_t)v.size());
  put<int32_t>(m, 4, nu);
  for (int j = 0; j < nu; j++) put<int64_t>(m, 8 + ((int64_t)j << 3), uniq[j]);
  for (size_t i = 0; i < v.size(); i++) {
    int64_t bits;
    std::memcpy(&bits, &v[i], 8);
    m[8 + ((size_t)nu << 3) + i] = (uint8_t)(int8_t)(idx[bits] - 128);
  }
  return m;
}

// NewChunk.bufS (W/fvec/NewChunk.java:1437-1459): sparse integers
inline std::vector<uint8_t> enc_cxi(int32_t len, int id_sz, int val_sz, bool na_sparse, const std::vector<int32_t>& ids,
                                    const std::vector<int64_t>& vals, const std::vector<uint8_t>& na) {
  const int elem = id_sz + val_sz;
  std::vector<uint8_t> m(8 + ids.size() * elem, 0);
  put<int32_t>(m, 0, len);
  m[4] = (uint8_t)id_sz;
  m[5] = (uint8_t)val_sz;
  m[6] = na_sparse ? 1 : 0;
  const int64_t NA = val_sz == 2 ? (int64_t)INT16_MIN : val_sz == 4 ? (int64_t)INT32_MIN : val_sz == 8 ? INT64_MIN : -1;
  for (size_t i = 0; i < ids.size(); i++) {
    const int64_t o = 8 + (int64_t)i * elem;
    if (id_sz == 2) put<int16_t>(m, o, (int16_t)ids[i]);
    else put<int32_t>(m, o, ids[i]);
    const int64_t val = na[i] ? NA : vals[i];
    switch (val_sz) {
      case 0: break;
      case 2: put<int16_t>(m, o + id_sz, (int16_t)val); break;
      case 4: put<int32_t>(m, o + id_sz, (int32_t)val); break;
      default: put<int64_t>(m, o + id_sz, val); break;
    }
  }
  return m;
}

// NewChunk.bufD (:1462-1480): sparse floats / doubles
inline std::vector<uint8_t> enc_cxf(int32_t len, int valsz, bool na_sparse, const std::vector<int32_t>& ids,
                                    const std::vector<double>& vals) {
  const int elem = valsz + 4;
  std::vector<uint8_t> m(8 + ids.size() * elem, 0);
  put<int32_t>(m, 0, len);
  m[4] = 4;
  m[5] = (uint8_t)valsz;
  m[6] = na_sparse ? 1 : 0;
  for (size_t i = 0; i < ids.size(); i++) {
    const int64_t o = 8 + (int64_t)i * elem;
    put<int32_t>(m, o, ids[i]);
    if (valsz == 4) put<float>(m, o + 4, (float)vals[i]);
    else put<double>(m, o + 4, vals[i]);
  }
  return m;
}

}  // namespace orc_chunk
