// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
//
// CPU restatement (C++17) of the reference's hex.tree histogram hot path, used ONLY as the
// parity checker (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference leg).
// Nothing under h2o-3_b200/ includes, links or calls this file.
//
// Parity status: PINNED BY SOURCE + by the reference's own fixture-free unit tests
// (tests/test_oracle_golden.py replays hex/tree/DTreeTest, DTreeSplitTest, DHistogramTest,
// HistogramTest, GBMTest.testPubDev6356, SharedTreeTest.checkNAPredictor vectors). The reference is
// Java and no JVM exists in this image, so the oracle cannot be pinned against reference OUTPUT
// beyond those vectors (see DESIGN.md "Oracle").
//
// Every function cites the reference file:line it follows (paths relative to /root/reference):
//   T/  = h2o-algos/src/main/java/hex/tree/
//   W/  = h2o-core/src/main/java/water/
//   H/  = h2o-core/src/main/java/hex/
//   G/  = h2o-genmodel/src/main/java/hex/genmodel/
//
// Build with: g++ -std=c++17 -O2 -ffp-contract=off  (no -ffast-math; IEEE fp64 semantics needed).
#pragma once
#include <algorithm>
#include <atomic>
#include <cassert>
#include <chrono>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace h2o_oracle {

// ----------------------------------------------------------------------------------------------
// Java numeric semantics helpers
// ----------------------------------------------------------------------------------------------
// Java (int) cast of a double: NaN -> 0, saturating, truncation toward zero (JLS 5.1.3).
inline int32_t j_d2i(double d) {
  if (std::isnan(d)) return 0;
  if (d >= 2147483647.0) return INT32_MAX;
  if (d <= -2147483648.0) return INT32_MIN;
  return (int32_t)d;
}
inline int64_t j_d2l(double d) {
  if (std::isnan(d)) return 0;
  if (d >= 9223372036854775807.0) return INT64_MAX;
  if (d <= -9223372036854775808.0) return INT64_MIN;
  return (int64_t)d;
}
// Java Math.ulp(double)
inline double j_ulp(double d) {
  if (std::isnan(d)) return d;
  d = std::fabs(d);
  if (std::isinf(d)) return d;
  if (d == DBL_MAX) return std::ldexp(1.0, 971);
  return std::nextafter(d, INFINITY) - d;
}
inline float j_ulpf(float f) {
  if (std::isnan(f)) return f;
  f = std::fabs(f);
  if (std::isinf(f)) return f;
  if (f == FLT_MAX) return std::ldexp(1.0f, 104);
  return std::nextafterf(f, INFINITY) - f;
}
// W/util/MathUtils.java:381-397
inline bool equalsWithinOneSmallUlp(float a, float b) {
  if (std::isnan(a) && std::isnan(b)) return true;
  float ulp_a = j_ulpf(a), ulp_b = j_ulpf(b);
  float small_ulp = std::fmin(ulp_a, ulp_b);
  float absdiff = std::fabs(a - b);
  return absdiff <= small_ulp;
}
inline bool equalsWithinOneSmallUlp(double a, double b) {
  if (std::isnan(a) && std::isnan(b)) return true;
  double ulp_a = j_ulp(a), ulp_b = j_ulp(b);
  double small_ulp = std::fmin(ulp_a, ulp_b);
  double absdiff = std::fabs(a - b);
  return absdiff <= small_ulp;
}

// ----------------------------------------------------------------------------------------------
// RNGs
// ----------------------------------------------------------------------------------------------
// W/util/RandomUtils.java:50-119 (PCGRNG: PCG-XSH-RR 64/32) layered under java.util.Random's
// nextFloat (next(24)/2^24) / nextDouble; getRNG(seed) == PCGRNG(seed, 1) (RandomUtils.java:16-21).
struct PCGRNG {
  uint64_t _state = 0, _inc = 0;
  explicit PCGRNG(int64_t seed, int64_t seq = 1) {
    // super(0) -> Random ctor calls the overridden setSeed(0) while _inc is still 0 => _state == 0
    _state = 0;
    _inc = ((uint64_t)seq << 1) | 1u;
    nextInt();
    _state += (uint64_t)seed;
    nextInt();
  }
  void setSeed(int64_t seed) {  // RandomUtils.java:67-72
    _state = 0;
    nextInt();
    _state += (uint64_t)seed;
    nextInt();
  }
  int32_t nextInt() {  // :77-83
    uint64_t oldstate = _state;
    _state = oldstate * 6364136223846793005ULL + _inc;
    uint32_t xorshifted = (uint32_t)(((oldstate >> 18) ^ oldstate) >> 27);
    uint32_t rot = (uint32_t)(oldstate >> 59);
    return (int32_t)((xorshifted >> rot) | (xorshifted << ((-rot) & 31)));
  }
  int64_t nextLong() {  // :84
    uint64_t hi = (uint64_t)(uint32_t)nextInt();
    uint64_t lo = (uint64_t)(uint32_t)nextInt();
    return (int64_t)((hi << 32) | lo);
  }
  int32_t next(int bits) {  // :86
    int64_t nextseed = nextLong();
    return (int32_t)(nextseed & ((1LL << bits) - 1));
  }
  int32_t nextInt(int32_t bound) {  // :88-118
    uint64_t threshold = (uint64_t)((-(int64_t)bound % (int64_t)bound)) & 0xFFFFFFFFULL;
    for (;;) {
      uint64_t r = (uint64_t)(uint32_t)nextInt();
      if (r >= threshold) return (int32_t)(r % (uint64_t)bound);
    }
  }
  float nextFloat() { return next(24) / (float)(1 << 24); }  // java.util.Random.nextFloat
  double nextDouble() {                                        // java.util.Random.nextDouble
    return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53;
  }
  // java.util.Random.nextGaussian (polar method) right after a setSeed, which drops a cached second value: only the first of
  // the pair is ever used by AddTreeContributions (T/gbm/GBM.java:1460-1463). StrictMath.log / sqrt: libm's.
  double nextGaussianAfterSeed() {
    double v1, v2, s;
    do {
      v1 = 2 * nextDouble() - 1;
      v2 = 2 * nextDouble() - 1;
      s = v1 * v1 + v2 * v2;
    } while (s >= 1 || s == 0);
    const double multiplier = std::sqrt(-2 * std::log(s) / s);
    return v1 * multiplier;
  }
};

// java.util.Random (LCG) — only needed to replay TT/DHistogramTest.java:173-206 inputs.
struct JavaRandom {
  uint64_t seed;
  bool haveNext = false;
  double nextNext = 0;
  explicit JavaRandom(int64_t s) { seed = ((uint64_t)s ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
  int32_t next(int bits) {
    seed = (seed * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
    return (int32_t)((int64_t)seed >> (48 - bits));
  }
  int32_t nextInt(int32_t bound) {
    int32_t r = next(31);
    int32_t m = bound - 1;
    if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
    for (int32_t u = r; u - (r = u % bound) + m < 0; u = next(31)) {
    }
    return r;
  }
  double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
  double nextGaussian() {
    if (haveNext) { haveNext = false; return nextNext; }
    double v1, v2, s;
    do {
      v1 = 2 * nextDouble() - 1;
      v2 = 2 * nextDouble() - 1;
      s = v1 * v1 + v2 * v2;
    } while (s >= 1 || s == 0);
    double multiplier = std::sqrt(-2 * std::log(s) / s);
    nextNext = v2 * multiplier;
    haveNext = true;
    return v1 * multiplier;
  }
};

// ----------------------------------------------------------------------------------------------
// IcedBitSet — W/util/IcedBitSet.java:20-111
// ----------------------------------------------------------------------------------------------
struct IcedBitSet {
  std::vector<uint8_t> _val;
  int _nbits = 0, _bitoff = 0;
  IcedBitSet() = default;
  IcedBitSet(int nbits, int bitoff) {  // :27-34
    if (bitoff + nbits <= 32) { bitoff = 0; nbits = 32; }
    _val.assign(bytes(nbits), 0);
    _nbits = nbits;
    _bitoff = bitoff;
  }
  static int bytes(int nbits) { return ((nbits - 1) >> 3) + 1; }
  bool isInRange(int idx) const { idx -= _bitoff; return idx >= 0 && idx < _nbits; }
  bool contains(int idx) const { idx -= _bitoff; return (_val[idx >> 3] & (uint8_t)(1u << (idx & 7))) != 0; }
  void set(int idx) { idx -= _bitoff; _val[idx >> 3] |= (uint8_t)(1u << (idx & 7)); }
  void clear(int idx) { idx -= _bitoff; _val[idx >> 3] &= (uint8_t) ~(1u << (idx & 7)); }
  int cardinality() const {
    int n = 0;
    for (uint8_t b : _val) n += __builtin_popcount(b);
    return n;
  }
  int size() const { return _nbits; }
  int numBytes() const { return bytes(_nbits); }
  int max() const { return _bitoff + _nbits; }
};

// ----------------------------------------------------------------------------------------------
// Parameters (T/SharedTreeModel.java:51-98, T/gbm/GBMModel.java:24-41 defaults)
// ----------------------------------------------------------------------------------------------
enum Family { GAUSSIAN = 0, BERNOULLI = 1, MULTINOMIAL = 2, POISSON = 3, GAMMA = 4, TWEEDIE = 5, LAPLACE = 6, QUANTILE = 7 };
enum NASplitDir { NA_None = 0, NAvsREST = 1, NALeft = 2, NARight = 3, Left = 4, Right = 5 };  // T/DHistogram.java:152-168

struct Params {
  int ntrees = 50;
  int max_depth = 5;
  double min_rows = 10;
  int nbins = 20;
  int nbins_cats = 1024;
  int nbins_top_level = 1 << 10;
  double min_split_improvement = 1e-5;
  double sample_rate = 1.0;
  std::vector<double> sample_rate_per_class;  // empty == null (T/SharedTreeModel.java:59)
  double pred_noise_bandwidth = 0;            // T/gbm/GBMModel.java:28
  double col_sample_rate = 1.0;
  double col_sample_rate_per_tree = 1.0;
  double col_sample_rate_change_per_level = 1.0;
  double learn_rate = 0.1;
  double learn_rate_annealing = 1.0;
  double max_abs_leafnode_pred = DBL_MAX;
  int64_t seed = 1234;
  int family = BERNOULLI;
  int nclass = 1;  // 1 regression, 2 binomial, >2 multinomial
  bool drf = false;
  int mtries = -1;  // DRF only
  int nthreads = 1; // CPU threads for the SBH2 restatement (baseline timing)
  double tweedie_power = 1.5;  // H/Model.java Parameters._tweedie_power; for family QUANTILE this slot holds _quantile_alpha
  int histogram_type = 0;      // HistoType: 0 UniformAdaptive (== AUTO), 1 Random (T/SharedTreeModel.java HistogramType)
};

// ----------------------------------------------------------------------------------------------
// DHistogram — T/DHistogram.java
// ----------------------------------------------------------------------------------------------
// java.util.Arrays.binarySearch(double[], double): <, > first, then Double.doubleToLongBits order (-0.0 < 0.0, NaN last)
inline int j_binarySearch(const std::vector<double>& a, double key) {
  int low = 0, high = (int)a.size() - 1;
  while (low <= high) {
    const int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
    const double midVal = a[mid];
    if (midVal < key) low = mid + 1;
    else if (midVal > key) high = mid - 1;
    else {
      auto bits = [](double d) { int64_t b; if (std::isnan(d)) return (int64_t)0x7ff8000000000000LL; memcpy(&b, &d, 8); return b; };
      const int64_t midBits = bits(midVal), keyBits = bits(key);
      if (midBits == keyBits) return mid;
      else if (midBits < keyBits) low = mid + 1;
      else high = mid - 1;
    }
  }
  return -(low + 1);
}
// W/util/ArrayUtils.java:1985-2002
inline std::vector<double> limitToRange(const std::vector<double>& pts, double min, double maxEx) {
  int start = j_binarySearch(pts, min);
  if (start < 0) start = -start - 1;
  if (start == (int)pts.size()) start--;
  if (pts[start] > min && start > 0) start--;
  int end = j_binarySearch(pts, maxEx);
  if (end < 0) end = -end - 1;
  if (end < start) end = start;
  return std::vector<double>(pts.begin() + start, pts.begin() + end);
}
// W/util/ArrayUtils.java:1955-1983 (used by GlobalQuantilesCalc and GuidedSplitPoints before the points reach DHistogram)
inline std::vector<double> makeUniqueAndLimitToRange(const std::vector<double>& splitPoints, double min, double maxEx) {
  double last = splitPoints[0];
  std::vector<double> out;
  for (size_t i = 0; i < splitPoints.size(); ++i) {
    const double pos = splitPoints[i];
    if (pos >= min && out.empty()) {
      out.push_back(min);
      if (pos > min) out.push_back(pos);
      last = pos;
    } else if (pos > maxEx) {
      break;
    } else if (pos > min && pos < maxEx && (i == 0 || pos != last)) {
      out.push_back(pos);
      last = pos;
    }
  }
  if (out.empty()) return std::vector<double>{min};
  return out;
}
// W/util/ArrayUtils.java:1933-1952
inline std::vector<double> padUniformly(const std::vector<double>& orig, int newLength) {
  const int origLength = (int)orig.size();
  if (newLength <= origLength || origLength <= 1) return orig;
  const int extraPoints = newLength - origLength;
  const int extraPointsPerBin = extraPoints / (origLength - 1);
  std::vector<double> res((size_t)newLength);
  int pos = 0;
  const int rem = extraPoints - extraPointsPerBin * (origLength - 1);
  for (int i = 0; i < origLength - 1; ++i) {
    const double startPos = orig[i];
    const double delta = orig[i + 1] - startPos;
    const int ext = extraPointsPerBin + (i < rem ? 1 : 0);
    res[pos++] = startPos;
    for (int j = 0; j < ext; ++j) res[pos++] = startPos + (j + 0.5) / ext * delta;
  }
  res[pos] = orig[origLength - 1];
  return res;
}

// SharedTreeModel.SharedTreeParameters.HistogramType (T/SharedTreeModel.java); RoundRobin resolves to one of the others
enum HistoType { HT_UniformAdaptive = 0, HT_Random = 1, HT_QuantilesGlobal = 2, HT_UniformRobust = 3 };

struct DHistogram {
  // split points (T/DHistogram.java:56-66): Random -> relative positions in [0, nbin); QuantilesGlobal / custom -> absolute values
  int _histoType = HT_UniformAdaptive;
  std::vector<double> _splitPts;          // empty == Java null
  std::vector<double> _globalSplitPts;    // what DKV.getGet(_globalSplitPointsKey).splitPts would hold (empty == none)
  bool _globalCanRefine = true;
  std::vector<double> _customSplitPoints;
  bool _absoluteSplitPts = false;
  int _zeroSplitPntPos = 0;
  uint8_t _isInt = 0;  // 0 float, 1 int, 2 categorical
  bool _intOpt = false;
  uint16_t _nbin = 0;  // Java char
  double _step = 0;
  double _min = 0, _maxEx = 0;
  int32_t _minInt = 0;
  bool _initNA = false;
  double _minSplitImprovement = 0;
  bool _checkFloatSplits = false;
  int64_t _seed = 0;
  std::vector<double> _vals;  // empty == Java null
  bool _has_vals = false;
  static constexpr int _vals_dim = 3;
  double _min2 = DBL_MAX, _maxIn = -DBL_MAX;
  std::vector<float> _splitPtsFloat;
  static constexpr int32_t INT_NA = INT32_MIN;  // :52

  DHistogram() = default;
  // ctor :190-269 (UniformAdaptive only; constraints/uplift/custom split points are out of scope)
  DHistogram(int nbins, int nbins_cats, uint8_t isInt, double min, double maxEx, bool intOpt, bool initNA,
             double minSplitImprovement, int64_t seed, bool checkFloatSplits, int histoType = 0) {
    if (!(maxEx > min)) throw std::runtime_error("DHistogram: maxEx <= min");
    _histoType = histoType;
    _isInt = isInt;
    _min = min;
    _maxEx = maxEx;
    _min2 = DBL_MAX;
    _maxIn = -DBL_MAX;
    _initNA = initNA;
    _intOpt = intOpt;
    _minInt = j_d2i(min);
    _minSplitImprovement = minSplitImprovement;
    _seed = seed;
    int xbins = isInt == 2 ? nbins_cats : nbins;
    if (isInt > 0 && maxEx - min <= xbins) {
      xbins = (uint16_t)(j_d2l(maxEx) - j_d2l(min));  // (char) cast
      _step = 1.0f;
    } else {
      _step = xbins / (maxEx - min);
      if (_step <= 0 || std::isinf(_step) || std::isnan(_step)) throw std::range_error("StepOutOfRange");
    }
    _nbin = (uint16_t)xbins;
    _checkFloatSplits = checkFloatSplits;
  }

  double w(int i) const { return _vals[_vals_dim * i + 0]; }
  double wY(int i) const { return _vals[_vals_dim * i + 1]; }
  double wYY(int i) const { return _vals[_vals_dim * i + 2]; }
  double wNA() const { return _vals[_vals_dim * _nbin + 0]; }
  double wYNA() const { return _vals[_vals_dim * _nbin + 1]; }
  double wYYNA() const { return _vals[_vals_dim * _nbin + 2]; }
  int nbins() const { return _nbin; }
  bool hasNABin() const { return !_has_vals ? _initNA : wNA() > 0; }  // :335-340
  int actNBins() const { return nbins() + (hasNABin() ? 1 : 0); }    // :317-319
  double bins(int b) const { return w(b); }

  // :308-312
  double binAt(int b) const {
    if (_absoluteSplitPts) return _splitPts[b];
    return _min + (_splitPts.empty() ? b : _splitPts[b]) / _step;
  }
  // :742-749
  static std::vector<double> makeRandomSplitPoints(int nbin, PCGRNG& rng) {
    std::vector<double> splitPts((size_t)nbin);
    splitPts[0] = 0;
    for (int i = 1; i < nbin; i++) splitPts[i] = (double)(rng.nextFloat() * (float)nbin);  // float * int in float
    std::sort(splitPts.begin(), splitPts.end());
    return splitPts;
  }

  // :343-417
  void init() {
    assert(!_has_vals);
    if (_histoType == HT_Random) {  // every node makes the same split points (:347-352)
      const double d = ((_step + 0.324) * _min + 8.3425) + 89.342 * _maxEx;
      int64_t bits;
      memcpy(&bits, &d, 8);
      const int64_t seed = (int64_t)((uint64_t)bits + (uint64_t)0xDECAF * (uint64_t)_nbin + (uint64_t)0xC0FFEE * (uint64_t)_isInt + (uint64_t)_seed);
      PCGRNG rng(seed);
      _splitPts = makeRandomSplitPoints(_nbin, rng);
    } else if (_histoType == HT_QuantilesGlobal) {  // :353-376
      if (!_globalSplitPts.empty()) {
        _splitPts = limitToRange(_globalSplitPts, _min, _maxEx);
        if (_globalCanRefine && _splitPts.size() > 1 && (int)_splitPts.size() < _nbin) _splitPts = padUniformly(_splitPts, _nbin);
        if (_splitPts.size() <= 1) {
          _splitPts.clear();
          _histoType = HT_UniformAdaptive;
        } else {
          _absoluteSplitPts = true;
          _nbin = (uint16_t)_splitPts.size();
        }
      }
    } else if (_histoType == HT_UniformRobust) {  // :378-385, 419-430
      if (!_customSplitPoints.empty()) {
        _splitPts = limitToRange(_customSplitPoints, _min, _maxEx);
        if (_splitPts.size() <= 1) {
          _splitPts.clear();
          _histoType = HT_UniformAdaptive;
        } else {
          _absoluteSplitPts = true;
          _nbin = (uint16_t)(_splitPts.size() - 1);
        }
      } else {
        _histoType = HT_UniformAdaptive;
      }
    }
    if (!_splitPts.empty()) {  // :389-401 canonical zero, cached position of zero
      _zeroSplitPntPos = j_binarySearch(_splitPts, 0.0);
      if (_zeroSplitPntPos < 0) {
        const int nzPos = j_binarySearch(_splitPts, -0.0);
        if (nzPos >= 0) {
          _splitPts[nzPos] = 0.0;
          _zeroSplitPntPos = nzPos;
        }
      }
    }
    _vals.assign((size_t)_vals_dim * _nbin + _vals_dim, 0.0);
    _has_vals = true;
    if (_checkFloatSplits) {
      _splitPtsFloat.resize(_nbin);
      for (int i = 0; i < _nbin; i++) _splitPtsFloat[i] = (float)binAt(i);
    }
  }

  // :272-306
  int bin(double col_data) const {
    if (std::isnan(col_data)) return _nbin;
    if (std::isinf(col_data)) return col_data < 0 ? 0 : _nbin - 1;
    double pos = _absoluteSplitPts ? col_data : ((col_data - _min) * _step);
    int idx1;
    if (!_splitPts.empty()) {
      idx1 = pos == 0.0 ? _zeroSplitPntPos : j_binarySearch(_splitPts, pos);
      if (idx1 < 0) idx1 = -idx1 - 2;
    } else {
      idx1 = j_d2i(pos);
    }
    if (!_splitPtsFloat.empty()) {
      if (idx1 + 1 < (int)_splitPtsFloat.size()) {
        float splitAt = _splitPtsFloat[idx1 + 1];
        if (col_data >= splitAt) idx1++;
        if (idx1 > 0) {
          if (!(col_data >= _splitPtsFloat[idx1])) idx1--;
        }
      }
    }
    if (idx1 == _nbin) idx1--;
    return idx1;
  }

  int nonEmptyBins() const {  // :323-333
    if (!_has_vals) return 0;
    int nonEmpty = 0;
    for (size_t i = 0; i < _vals.size() - _vals_dim; i += _vals_dim)
      if (_vals[i] > 0) nonEmpty++;
    return nonEmpty;
  }

  // :609-658 (vals_dim == 3, no uplift)
  void updateHisto(const double* ws, const double* cs, const double* ys, const int* rows, int hi, int lo) {
    for (int r = lo; r < hi; ++r) {
      const int k = rows[r];
      const double weight = ws == nullptr ? 1 : ws[k];
      if (weight == 0) continue;
      const double col_data = cs[k];
      if (col_data < _min2) _min2 = col_data;
      if (col_data > _maxIn) _maxIn = col_data;
      const double y = ys[r];
      double wy = weight * y;
      double wyy = wy * y;
      int b = bin(col_data);
      const int s = _vals_dim * b;
      _vals[s + 0] += weight;
      _vals[s + 1] += wy;
      _vals[s + 2] += wyy;
    }
  }

  // :674-704
  void updateHistoInt(const double* ws, const int32_t* cs, const double* ys, const int* rows, int hi, int lo) {
    int min2_int = _min2 == DBL_MAX ? INT32_MAX : j_d2i(_min2);
    // NB: the reference compares against -Double.MIN_VALUE (sic, :678), which never equals the
    // initial -Double.MAX_VALUE, so the cast path is always taken: (int)(-DBL_MAX) == INT_MIN.
    int maxIn_int = _maxIn == -DBL_MIN ? INT32_MIN : j_d2i(_maxIn);
    for (int r = lo; r < hi; r++) {
      const int k = rows[r];
      const double weight = ws == nullptr ? 1 : ws[k];
      if (weight == 0) continue;
      const int col_data = cs[k];
      int b;
      if (col_data != INT_NA) {
        if (col_data < min2_int) min2_int = col_data;
        if (col_data > maxIn_int) maxIn_int = col_data;
        b = col_data - _minInt;
      } else
        b = _nbin;
      const double y = ys[r];
      double wy = weight * y;
      double wyy = wy * y;
      const int s = _vals_dim * b;
      _vals[s + 0] += weight;
      _vals[s + 1] += wy;
      _vals[s + 2] += wyy;
    }
    _min2 = min2_int;
    _maxIn = maxIn_int;
  }

  // :434-447
  void add(const DHistogram& dsh) {
    if (!dsh._has_vals) return;
    if (!_has_vals) {
      _vals = dsh._vals;
      _has_vals = true;
      if (_checkFloatSplits && _splitPtsFloat.empty()) _splitPtsFloat = dsh._splitPtsFloat;
    } else {
      for (size_t i = 0; i < _vals.size(); i++) _vals[i] += dsh._vals[i];
    }
    if (_min2 > dsh._min2) _min2 = dsh._min2;
    if (_maxIn < dsh._maxIn) _maxIn = dsh._maxIn;
  }

  double find_min() const { return _min2; }
  double find_maxIn() const { return _maxIn; }
  double find_maxEx() const { return find_maxEx(_maxIn, _isInt); }
  static double find_maxEx(double maxIn, int isInt) {  // :454-459
    double ulp = j_ulp(maxIn);
    if (isInt > 0 && 1 > ulp) ulp = 1;
    double res = maxIn + ulp;
    return std::isinf(res) ? maxIn : res;
  }

  // :734-740
  void reducePrecision() {
    if (!_has_vals) return;
    for (size_t i = 0; i < _vals.size(); i += _vals_dim) {
      _vals[i + 1] = (double)(float)_vals[i + 1];
      _vals[i + 2] = (double)(float)_vals[i + 2];
    }
  }
};
using HistPtr = std::shared_ptr<DHistogram>;  // nullptr == Java null entry in DHistogram[]

// ----------------------------------------------------------------------------------------------
// Split — T/DTree.java:182-417
// ----------------------------------------------------------------------------------------------
struct Split {
  int _col = -1, _bin = 0;
  int _nasplit = NA_None;
  std::optional<IcedBitSet> _bs;
  int8_t _equal = 0;
  double _se = 0, _se0 = 0, _se1 = 0, _n0 = 0, _n1 = 0, _p0 = 0, _p1 = 0, _tree_p0 = 0, _tree_p1 = 0;
  double se() const { return _se0 + _se1; }
  double pre_split_se() const { return _se; }
  bool isNumericSplit() const { return _nasplit != NAvsREST && (_equal == 0 || _equal == 1); }  // :244-246

  // :253-291
  float splatNumeric(const DHistogram& h) const {
    int x = _bin - 1;
    while (x >= 0 && h.bins(x) == 0) x--;
    int n = _bin;
    while (n < h.nbins() && h.bins(n) == 0) n++;
    double lo = h.binAt(x + 1);
    double hi = h.binAt(n);
    if (h._isInt > 0) {
      lo = h._step == 1 ? lo - 1 : std::floor(lo);
      hi = h._step == 1 ? hi : std::ceil(hi);
    }
    const float splitAt = (float)((lo + hi) / 2.0);
    if (h._checkFloatSplits && lo != hi && (lo > splitAt || hi < splitAt)) return NAN;
    return splitAt;
  }
  float splat(const std::vector<HistPtr>& hs) const {  // :240-242
    return isNumericSplit() ? splatNumeric(*hs[_col]) : -1.0f;
  }

  // :312-417 (no constraints / interaction constraints / UniformRobust)
  // Returns empty vector for Java null.
  std::vector<HistPtr> nextLevelHistos(const std::vector<HistPtr>& currentHistos, int way, double splat,
                                       const Params& parms, bool& isNull) const {
    isNull = true;
    double n = way == 0 ? _n0 : _n1;
    if (n < parms.min_rows) return {};
    double se = way == 0 ? _se0 : _se1;
    if (se <= 1e-30) return {};
    int cnt = 0;
    std::vector<HistPtr> nhists(currentHistos.size());
    for (size_t j = 0; j < currentHistos.size(); j++) {
      const HistPtr& h = currentHistos[j];
      if (!h) continue;
      const int adj_nbins = std::max(h->nbins() >> 1, parms.nbins);
      double min, maxEx;
      if (!h->_has_vals || _equal > 1) {
        min = h->_min;
        maxEx = h->_maxEx;
      } else {
        min = h->find_min();
        if (h->find_maxIn() == min) continue;
        maxEx = h->find_maxEx();
      }
      if (_nasplit == NAvsREST) {
        if (way == 1) continue;
      }
      if (_col == (int)j) {
        switch (_equal) {
          case 0: {
            if (_nasplit != NAvsREST) {
              if (h->_vals[h->_vals_dim * _bin] == 0) throw std::runtime_error("unimpl: empty split bin");
            }
            double split = splat;
            if (h->_isInt > 0) split = (float)std::ceil(split);
            if (_nasplit != NAvsREST) {
              if (way == 0) maxEx = split;
              else min = split;
            }
            break;
          }
          case 1:
            if (way == 1) continue;
            break;
          case 2:
          case 3:
            break;
          default:
            throw std::runtime_error("fail");
        }
      }
      if (min > maxEx) continue;
      if (equalsWithinOneSmallUlp(min, maxEx)) continue;
      if (std::isinf(adj_nbins / (maxEx - min))) continue;
      if (h->_isInt > 0 && !(min + 1 < maxEx)) continue;
      const bool hasNAs = ((_nasplit == NALeft && way == 0) || (_nasplit == NARight && way == 1)) && h->hasNABin();
      // DHistogram.make :504-512 ; seed h._seed*0xDECAF+(way+1)
      nhists[j] = std::make_shared<DHistogram>(adj_nbins, parms.nbins_cats, h->_isInt, min, maxEx, h->_intOpt, hasNAs,
                                               parms.min_split_improvement,
                                               (int64_t)((uint64_t)h->_seed * 0xDECAFULL + (uint64_t)(way + 1)),
                                               h->_checkFloatSplits, parms.histogram_type);
      cnt++;
    }
    if (cnt == 0) return {};
    isNull = false;
    return nhists;
  }
};

// W/util/ArrayUtils.java:1565-1596 — stable ascending sort of indices by value (insertion sort below the
// cutoff, TimSort (stable) above; both orderings equal a stable sort for non-NaN keys).
inline void sort_idxs(std::vector<int>& idxs, const std::vector<double>& values) {
  if ((int)idxs.size() < 500) {
    for (size_t i = 0; i < idxs.size(); i++)
      for (size_t j = i; j > 0 && values[idxs[j - 1]] > values[idxs[j]]; j--) std::swap(idxs[j], idxs[j - 1]);
  } else {
    std::stable_sort(idxs.begin(), idxs.end(), [&](int x, int y) { return values[x] < values[y]; });
  }
}

// T/DTree.java:1907-1941
inline int8_t fillBitSet(const DHistogram& hs, int off, const std::vector<int>& idxs, int best, int nbins, IcedBitSet& bs) {
  for (int i = best; i < nbins; i++) bs.set(idxs[i] + off);
  int nonEmptyThatWentRight = 0, nonEmptyThatWentLeft = 0;
  for (int i = 0; i < nbins; i++) {
    if (hs.w(i) > 0) {
      if (bs.contains(i + off)) nonEmptyThatWentRight++;
      else nonEmptyThatWentLeft++;
    }
  }
  bool shouldGoLeft = nonEmptyThatWentLeft >= nonEmptyThatWentRight;
  for (int i = 0; i < nbins; i++) {
    if (hs.w(i) == 0) {
      if (bs.contains(i + off) && shouldGoLeft) bs.clear(i + off);
      if (!bs.contains(i + off) && !shouldGoLeft) bs.set(i + off);
    }
  }
  if (bs.cardinality() == 0 || bs.cardinality() == bs.size()) return -1;
  return (int8_t)(bs.max() <= 32 ? 2 : 3);
}

// T/DTree.java:984-1487 with constraint==0, min=max=NaN, useBounds=false, dist=null (vals_dim==3).
inline std::optional<Split> findBestSplitPoint(const DHistogram& hs, int col, double min_rows) {
  if (!hs._has_vals) return std::nullopt;
  const int nbins = hs.nbins();
  const int vals_dim = hs._vals_dim;
  const double* vals = hs._vals.data();
  std::vector<double> sorted_vals;
  std::vector<int> idxs;
  bool have_idxs = false;
  if (hs._isInt == 2 && hs._step == 1) {  // :1006-1033
    idxs.resize(nbins + 1);
    for (int i = 0; i < nbins + 1; i++) idxs[i] = i;
    std::vector<double> avgs(nbins + 1);
    for (int i = 0; i < nbins; i++) avgs[i] = hs.w(i) == 0 ? -DBL_MAX : hs.wY(i) / hs.w(i);
    avgs[nbins] = DBL_MAX;
    sort_idxs(idxs, avgs);
    sorted_vals.assign((size_t)vals_dim * nbins, 0.0);
    for (int i = 0; i < nbins; i++) {
      int id = idxs[i];
      sorted_vals[vals_dim * i + 0] = hs._vals[vals_dim * id + 0];
      sorted_vals[vals_dim * i + 1] = hs._vals[vals_dim * id + 1];
      sorted_vals[vals_dim * i + 2] = hs._vals[vals_dim * id + 2];
    }
    vals = sorted_vals.data();
    have_idxs = true;
  }
  // :1036-1067
  std::vector<double> wlo(nbins + 1, 0.0), wYlo(nbins + 1, 0.0), wYYlo(nbins + 1, 0.0);
  for (int b = 1; b <= nbins; b++) {
    int id = vals_dim * (b - 1);
    double n0 = wlo[b - 1], n1 = vals[id + 0];
    if (n0 == 0 && n1 == 0) continue;
    double m0 = wYlo[b - 1], m1 = vals[id + 1];
    double s0 = wYYlo[b - 1], s1 = vals[id + 2];
    wlo[b] = n0 + n1;
    wYlo[b] = m0 + m1;
    wYYlo[b] = s0 + s1;
  }
  const double wNA = hs.wNA();
  double tot = wlo[nbins] + wNA;
  if (tot < 2 * min_rows) return std::nullopt;  // :1071
  double wYNA = hs.wYNA();
  double wYYNA = hs.wYYNA();
  double var = (wYYlo[nbins] + wYYNA) * tot - (wYlo[nbins] + wYNA) * (wYlo[nbins] + wYNA);
  if (((float)var) == 0.0f) return std::nullopt;  // :1081
  // :1090-1121
  std::vector<double> whi(nbins + 1, 0.0), wYhi(nbins + 1, 0.0), wYYhi(nbins + 1, 0.0);
  for (int b = nbins - 1; b >= 0; b--) {
    double n0 = whi[b + 1], n1 = vals[vals_dim * b];
    if (n0 == 0 && n1 == 0) continue;
    double m0 = wYhi[b + 1], m1 = vals[vals_dim * b + 1];
    double s0 = wYYhi[b + 1], s1 = vals[vals_dim * b + 2];
    whi[b] = n0 + n1;
    wYhi[b] = m0 + m1;
    wYYhi[b] = s0 + s1;
  }
  double best_seL = DBL_MAX, best_seR = DBL_MAX;
  int nasplit = NA_None;
  double seNonNA = wYYhi[0] - wYhi[0] * wYhi[0] / whi[0];  // :1128
  if (seNonNA < 0) seNonNA = 0;
  double seBefore = seNonNA;
  double nLeft = 0, nRight = 0, predLeft = 0, predRight = 0, tree_p0 = 0, tree_p1 = 0;
  if (wNA >= min_rows) {  // :1140-1159
    double seAll = (wYYhi[0] + wYYNA) - (wYhi[0] + wYNA) * (wYhi[0] + wYNA) / (whi[0] + wNA);
    double seNA = wYYNA - wYNA * wYNA / wNA;
    if (seNA < 0) seNA = 0;
    best_seL = seNonNA;
    best_seR = seNA;
    nasplit = NAvsREST;
    seBefore = seAll;
    nLeft = whi[0];
    predLeft = wYhi[0];
    nRight = wNA;
    predRight = wYNA;
    tree_p0 = predLeft / nLeft;
    tree_p1 = predRight / nRight;
  }
  int best = 0;
  int8_t equal = 0;
  for (int b = 1; b <= nbins - 1; b++) {  // :1168-1348
    if (vals[vals_dim * b] == 0) continue;
    if (wlo[b] + wNA < min_rows) continue;
    if (whi[b] + wNA < min_rows) break;
    if (wNA == 0) {
      double selo = wYYlo[b] - wYlo[b] * wYlo[b] / wlo[b];
      double sehi = wYYhi[b] - wYhi[b] * wYhi[b] / whi[b];
      if (selo < 0) selo = 0;
      if (sehi < 0) sehi = 0;
      if ((selo + sehi < best_seL + best_seR) ||
          (selo + sehi == best_seL + best_seR && std::abs(b - (nbins >> 1)) < std::abs(best - (nbins >> 1)))) {
        double tmpPredLeft = wYlo[b] / wlo[b];
        double tmpPredRight = wYhi[b] / whi[b];
        best_seL = selo;
        best_seR = sehi;
        best = b;
        nLeft = wlo[best];
        nRight = whi[best];
        predLeft = wYlo[best];
        predRight = wYhi[best];
        tree_p0 = tmpPredLeft;
        tree_p1 = tmpPredRight;
      }
    } else {
      {  // option 1: NAs left :1233-1287
        double selo = wYYlo[b] + wYYNA - (wYlo[b] + wYNA) * (wYlo[b] + wYNA) / (wlo[b] + wNA);
        double sehi = wYYhi[b] - wYhi[b] * wYhi[b] / whi[b];
        if (selo < 0) selo = 0;
        if (sehi < 0) sehi = 0;
        if ((selo + sehi < best_seL + best_seR) ||
            (selo + sehi == best_seL + best_seR && std::abs(b - (nbins >> 1)) < std::abs(best - (nbins >> 1)))) {
          if ((wlo[b] + wNA) >= min_rows && whi[b] >= min_rows) {
            double tmpPredLeft = (wYlo[b] + wYNA) / (wlo[b] + wNA);
            double tmpPredRight = wYhi[b] / whi[b];
            best_seL = selo;
            best_seR = sehi;
            best = b;
            nLeft = wlo[best] + wNA;
            nRight = whi[best];
            predLeft = wYlo[best] + wYNA;
            predRight = wYhi[best];
            nasplit = NALeft;
            tree_p0 = tmpPredLeft;
            tree_p1 = tmpPredRight;
          }
        }
      }
      {  // option 2: NAs right :1290-1346
        double selo = wYYlo[b] - wYlo[b] * wYlo[b] / wlo[b];
        double sehi = wYYhi[b] + wYYNA - (wYhi[b] + wYNA) * (wYhi[b] + wYNA) / (whi[b] + wNA);
        if (selo < 0) selo = 0;
        if (sehi < 0) sehi = 0;
        if ((selo + sehi < best_seL + best_seR) ||
            (selo + sehi == best_seL + best_seR && std::abs(b - (nbins >> 1)) < std::abs(best - (nbins >> 1)))) {
          if (wlo[b] >= min_rows && (whi[b] + wNA) >= min_rows) {
            double tmpPredLeft = wYlo[b] / wlo[b];
            double tmpPredRight = (wYhi[b] + wYNA) / (whi[b] + wNA);
            best_seL = selo;
            best_seR = sehi;
            best = b;
            nLeft = wlo[best];
            nRight = whi[best] + wNA;
            predLeft = wYlo[best];
            predRight = wYhi[best] + wYNA;
            nasplit = NARight;
            tree_p0 = tmpPredLeft;
            tree_p1 = tmpPredRight;
          }
        }
      }
    }
  }
  if (best == 0 && nasplit == NA_None) return std::nullopt;                                   // :1350
  if (!(best_seL + best_seR < seBefore * (1 - hs._minSplitImprovement))) return std::nullopt;   // :1356
  if (equalsWithinOneSmallUlp((float)(predLeft / nLeft), (float)(predRight / nRight))) return std::nullopt;  // :1363
  if (nLeft < min_rows || nRight < min_rows) return std::nullopt;                               // :1368
  const double node_p0 = predLeft / nLeft;
  const double node_p1 = predRight / nRight;
  if (!(best_seL + best_seR < seBefore * (1 - hs._minSplitImprovement))) return std::nullopt;   // :1452
  if (equalsWithinOneSmallUlp((float)tree_p0, (float)tree_p1)) return std::nullopt;             // :1457
  std::optional<IcedBitSet> bs;
  if (have_idxs && nasplit != NAvsREST) {  // :1465-1473
    const int off = j_d2i(hs._min);
    bs = IcedBitSet(nbins, off);
    equal = fillBitSet(hs, off, idxs, best, nbins, *bs);
    if (equal < 0) return std::nullopt;
  }
  if (nasplit == NA_None) nasplit = nLeft > nRight ? Left : Right;  // :1476-1478
  Split s;
  s._col = col;
  s._bin = best;
  s._nasplit = nasplit;
  s._bs = bs;
  s._equal = equal;
  s._se = seBefore;
  s._se0 = best_seL;
  s._se1 = best_seR;
  s._n0 = nLeft;
  s._n1 = nRight;
  s._p0 = node_p0;
  s._p1 = node_p1;
  s._tree_p0 = tree_p0;
  s._tree_p1 = tree_p1;
  return s;
}

// ----------------------------------------------------------------------------------------------
// DTree and nodes — T/DTree.java:36-133, 438-511, 587-763, 936-965
// ----------------------------------------------------------------------------------------------
constexpr int NO_PARENT = -1;
constexpr int DECIDED_ROW = -1;              // T/ScoreBuildHistogram.java:61
constexpr int OUT_OF_BAG = -2;               // :63
constexpr int UNDECIDED_CHILD_NODE_ID = -1;  // :65
constexpr int FRESH = 0;
inline bool isOOBRow(int nid) { return nid <= OUT_OF_BAG; }      // :69
inline bool isDecidedRow(int nid) { return nid == DECIDED_ROW; }  // :70
inline int oob2Nid(int oobNid) { return -oobNid + OUT_OF_BAG; }   // :71
inline int nid2Oob(int nid) { return -nid + OUT_OF_BAG; }         // :72

enum NodeKind { UNDECIDED = 0, DECIDED = 1, LEAF = 2 };

struct Node {
  int kind = UNDECIDED;
  int _pid = NO_PARENT, _nid = 0;
  // Undecided
  std::vector<HistPtr> _hs;
  bool _scoreColsNull = true;
  std::vector<int> _scoreCols;
  // Decided
  std::optional<Split> _split;
  float _splat = NAN;
  int _nids[2] = {-1, -1};
  // Leaf
  float _pred = 0;
};

struct DTree {
  int _ncols;
  int64_t _seed;
  std::vector<Node> _ns;
  int _leaves = 0, _depth = 0;
  int _mtrys, _mtrys_per_tree;
  PCGRNG _rand;
  std::vector<int> _cols;
  const Params* _parms;

  int len() const { return (int)_ns.size(); }
  Node& node(int i) { return _ns[i]; }
  const Node& node(int i) const { return _ns[i]; }

  // :56-58
  int actual_mtries() const {
    return std::min(std::max(1, j_d2i((double)_mtrys * std::pow(_parms->col_sample_rate_change_per_level, _depth))), _ncols);
  }
  // :60-86
  DTree(int ncols, int mtrys, int mtrys_per_tree, int64_t seed, const Params* parms)
      : _ncols(ncols), _seed(seed), _mtrys(mtrys), _mtrys_per_tree(mtrys_per_tree), _rand(seed), _parms(parms) {
    std::vector<int> activeCols(ncols);
    for (int i = 0; i < ncols; ++i) activeCols[i] = i;
    int len = ncols;
    if (mtrys_per_tree < ncols) {
      PCGRNG colSampleRNG((int64_t)((uint64_t)seed * 0xDA7AULL));
      for (int i = 0; i < mtrys_per_tree; i++) {
        if (len == 0) break;
        int idx2 = colSampleRNG.nextInt(len);
        int col = activeCols[idx2];
        activeCols[idx2] = activeCols[--len];
        activeCols[len] = col;
      }
      activeCols = std::vector<int>(activeCols.begin() + len, activeCols.end());
    }
    _cols = activeCols;
  }

  // UndecidedNode ctor :443-450 + scoreCols :462-499
  int newUndecided(int pid, std::vector<HistPtr> hs) {
    Node n;
    n.kind = UNDECIDED;
    n._pid = pid;
    n._nid = len();
    n._hs = std::move(hs);
    _ns.push_back(std::move(n));
    Node& u = _ns.back();
    scoreCols(u);
    return u._nid;
  }
  void scoreCols(Node& u) {
    const int hslen = (int)u._hs.size();
    if (actual_mtries() == hslen && _mtrys_per_tree == hslen) {
      u._scoreColsNull = true;
      return;
    }
    const std::vector<int>& activeCols = _cols;
    std::vector<int> cols(activeCols.size());
    int len = 0;
    for (size_t i = 0; i < activeCols.size(); i++) {
      int idx = activeCols[i];
      if (!u._hs[idx]) continue;
      cols[len++] = idx;
    }
    int choices = len;
    int mtries = actual_mtries();
    if (choices > 0) {
      for (int i = 0; i < mtries; i++) {
        if (len == 0) break;
        int idx2 = _rand.nextInt(len);
        int col = cols[idx2];
        cols[idx2] = cols[--len];
        cols[len] = col;
      }
    }
    u._scoreColsNull = false;
    u._scoreCols.assign(cols.begin() + len, cols.begin() + choices);
  }

  // UndecidedNode.doNotSplit :504-511
  void doNotSplit(int nid) {
    int pid = _ns[nid]._pid;
    if (pid == NO_PARENT) return;
    Node& dn = _ns[pid];
    for (int i = 0; i < 2; i++)
      if (dn._nids[i] == nid) { dn._nids[i] = UNDECIDED_CHILD_NODE_ID; return; }
    throw std::runtime_error("doNotSplit: fail");
  }

  // DecidedNode.bestCol :616-656
  std::optional<Split> bestCol(const Node& u, const std::vector<HistPtr>& hs) const {
    std::optional<Split> best;
    if (hs.empty()) return best;
    const int maxCols = u._scoreColsNull ? (int)hs.size() : (int)u._scoreCols.size();
    for (int i = 0; i < maxCols; i++) {
      int col = u._scoreColsNull ? i : u._scoreCols[i];
      if (!hs[col] || hs[col]->actNBins() <= 1) continue;
      std::optional<Split> s = findBestSplitPoint(*hs[col], col, _parms->min_rows);
      if (!s) continue;
      if (!best || s->se() < best->se()) {
        if (hs[s->_col]->_checkFloatSplits) {
          float splitAt = s->splat(hs);
          if (std::isnan(splitAt)) continue;
        }
        best = s;
      }
    }
    return best;
  }

  // DecidedNode(UndecidedNode n, DHistogram hs[]) :703-726 — replaces node n in place.
  void decide(int nid, const std::vector<HistPtr>& hs) {
    Node u = _ns[nid];  // copy of the undecided node (scoreCols needed by bestCol)
    Node dn;
    dn.kind = DECIDED;
    dn._pid = u._pid;
    dn._nid = u._nid;
    dn._split = bestCol(u, hs);
    if (!dn._split) {
      dn._splat = NAN;
      dn._nids[0] = dn._nids[1] = UNDECIDED_CHILD_NODE_ID;
      _ns[nid] = dn;
      return;
    }
    dn._splat = dn._split->splat(hs);
    _ns[nid] = dn;
    for (int way = 0; way < 2; way++) {
      bool isNull;
      std::vector<HistPtr> nhists = _ns[nid]._split->nextLevelHistos(hs, way, _ns[nid]._splat, *_parms, isNull);
      int child = isNull ? UNDECIDED_CHILD_NODE_ID : newUndecided(nid, std::move(nhists));
      _ns[nid]._nids[way] = child;
    }
  }

  int newLeaf(int pid) {  // LeafNode(tree,pid) :938
    Node n;
    n.kind = LEAF;
    n._pid = pid;
    n._nid = len();
    _ns.push_back(n);
    _leaves++;
    return _ns.back()._nid;
  }
  void newLeafAt(int pid, int nid) {  // LeafNode(tree,pid,nid) :939
    Node n;
    n.kind = LEAF;
    n._pid = pid;
    n._nid = nid;
    _ns[nid] = n;
    _leaves++;
  }
};

// ----------------------------------------------------------------------------------------------
// Frame (column store stand-in for Frame/Vec/Chunk read side; W/fvec/Chunk.java:785-794 getDoubles)
// ----------------------------------------------------------------------------------------------
struct Column {
  std::vector<double> v;  // NaN == NA
  bool isCat = false;
  int cardinality = 0;  // domain().length for categoricals
  // rollups (W/fvec/RollupStats.java:31-48,88-200)
  double min = NAN, max = NAN;
  int64_t nacnt = 0;
  bool isInt = true;
  void rollup() {
    min = DBL_MAX;
    max = -DBL_MAX;
    nacnt = 0;
    isInt = true;
    for (double d : v) {
      if (std::isnan(d)) { nacnt++; continue; }
      if (d < min) min = d;
      if (d > max) max = d;
      if ((double)j_d2l(d) != d) isInt = false;
    }
    if (nacnt == (int64_t)v.size()) { min = max = NAN; }
  }
  bool isConstNoNA() const { return min == max && nacnt == 0; }  // Vec.isConst(true) W/fvec/Vec.java:861-865
};

struct Frame {
  std::vector<Column> cols;
  std::vector<double> y;   // response (class index for classification); NaN == NA
  std::vector<double> w;   // empty == no weights column
  std::vector<double> off; // empty == no offset
  int64_t nrows() const { return (int64_t)y.size(); }
  int ncols() const { return (int)cols.size(); }
};

// T/DHistogram.java:526-554
inline bool useIntOpt(const Column& v, const Params& parms) {
  if (parms.histogram_type != HT_UniformAdaptive && parms.histogram_type != HT_UniformRobust) return false;  // :537-542
  if (v.isCat) return v.cardinality < parms.nbins_cats;
  if (v.isInt) {
    double min = v.min, max = v.max;
    double intLen = max - min;
    return intLen < parms.nbins && j_d2i(max) - j_d2i(min) == j_d2i(intLen);
  }
  return false;
}

// T/DHistogram.java:474-502
inline std::vector<HistPtr> initialHist(const Frame& fr, int nbins, int64_t seed, const Params& parms, bool checkFloatSplits) {
  std::vector<HistPtr> hs(fr.ncols());
  for (int c = 0; c < fr.ncols(); c++) {
    const Column& v = fr.cols[c];
    const double minIn = v.isCat ? 0 : std::fmax(v.min, -DBL_MAX);
    const double maxIn = v.isCat ? v.cardinality - 1 : std::fmin(v.max, DBL_MAX);
    const double maxEx = v.isCat ? v.cardinality : DHistogram::find_maxEx(maxIn, v.isInt ? 1 : 0);
    const int64_t vlen = (int64_t)v.v.size();
    try {
      uint8_t type = (uint8_t)(v.isCat ? 2 : (v.isInt ? 1 : 0));
      bool intOpt = useIntOpt(v, parms);
      if (v.nacnt == vlen || v.isConstNoNA()) hs[c] = nullptr;
      else
        hs[c] = std::make_shared<DHistogram>(nbins, parms.nbins_cats, type, minIn, maxEx, intOpt, v.nacnt > 0,
                                             parms.min_split_improvement, seed, checkFloatSplits, parms.histogram_type);
    } catch (const std::range_error&) {
      hs[c] = nullptr;
    }
  }
  return hs;
}

// ----------------------------------------------------------------------------------------------
// DecidedNode.getChildNodeID — T/DTree.java:728-763
// ----------------------------------------------------------------------------------------------
inline int getChildNodeID(const Node& dn, const Frame& fr, int64_t row) {
  const Split& sp = *dn._split;
  double d = fr.cols[sp._col].v[row];
  int bin = -1;
  bool isNA = std::isnan(d);
  if (!isNA) {
    if (sp._nasplit == NAvsREST) bin = 0;
    else if (sp._equal == 0) bin = d >= dn._splat ? 1 : 0;  // double vs float compare (q4)
    else if (sp._equal >= 2) {
      int b = j_d2i(d);
      if (sp._bs->isInRange(b)) bin = sp._bs->contains(b) ? 1 : 0;
      else isNA = true;
    }
  }
  if (isNA) {
    if (sp._nasplit == NALeft || sp._nasplit == Left) bin = 0;
    else if (sp._nasplit == NARight || sp._nasplit == Right || sp._nasplit == NAvsREST) bin = 1;
    else if (sp._nasplit == NA_None) bin = 1;
    else throw std::runtime_error("unimpl nasplit");
  }
  return dn._nids[bin];
}

// ----------------------------------------------------------------------------------------------
// ScoreBuildHistogram2 — T/ScoreBuildHistogram2.java:139-168 (score_decide), :192-270 (phase 1),
// :375-453 (phase 2), :455-466 (postGlobal); threads: RangeWorkAllocator-style fixed partition
// (:348-373) so the oracle is deterministic for a given thread count.
// ----------------------------------------------------------------------------------------------
struct LevelCapture {  // analogue of the histogram monitor hook (T/SharedTree.java:1253-1289)
  int tree = 0, k = 0, leaf = 0, nleaves = 0;
  // per (leaf-relative node n, column c): descriptor + filled histogram (after reducePrecision)
  struct Entry {
    bool present = false;  // DHistogram non-null and tracked by scoreCols
    bool has_vals = false;
    uint8_t isInt = 0;
    bool intOpt = false;
    int nbin = 0;
    double min = 0, maxEx = 0, step = 0, min2 = 0, maxIn = 0;
    std::vector<double> vals;
  };
  std::vector<Entry> entries;  // [nleaves*ncols]
  std::vector<int32_t> nids_after;  // nids column after score_decide
};

struct SBH2 {
  const Frame& fr;
  const Params& parms;
  DTree& tree;
  int leaf;
  std::vector<std::vector<HistPtr>>& hcs;  // [leaf-relative node][col]
  const std::vector<double>& work;         // ys (Work_k)
  std::vector<int32_t>& nids;
  int nthreads;
  int64_t chunk_rows;

  // score_decide :139-168 on rows [lo,hi); nnids -> res (leaf-relative ids)
  void score_decide(int64_t lo, int64_t hi, std::vector<int32_t>& res) {
    for (int64_t row = lo; row < hi; row++) {
      int nid = nids[row];
      if (isDecidedRow(nid)) { res[row] = nid - leaf; continue; }
      bool oob = isOOBRow(nid);
      if (oob) nid = oob2Nid(nid);
      const Node* dn = &tree.node(nid);
      if (!dn->_split) {
        if (dn->_pid == NO_PARENT) { res[row] = nid - leaf; continue; }
        nid = dn->_pid;
        int xnid = oob ? nid2Oob(nid) : nid;
        nids[row] = xnid;
        res[row] = xnid - leaf;
        dn = &tree.node(nid);
      }
      nid = getChildNodeID(*dn, fr, row);
      if (!isDecidedRow(nid)) {
        if (oob) nid = nid2Oob(nid);
        nids[row] = nid;
      }
      res[row] = nid - leaf;
    }
  }

  void run() {
    const int64_t N = fr.nrows();
    const int numLeafs = (int)hcs.size();
    const int ncols = fr.ncols();
    const int64_t nchunks = std::max<int64_t>(1, (N + chunk_rows - 1) / chunk_rows);
    std::vector<int32_t> nnids(N, 0);
    std::vector<std::vector<int>> nhs(nchunks), rss(nchunks);
    std::vector<std::vector<double>> yss(nchunks);
    // ---- phase 1 (:192-270), parallel over chunks
    auto phase1 = [&](int64_t c) {
      int64_t lo = c * chunk_rows, hi = std::min(N, lo + chunk_rows);
      if (leaf > 0) score_decide(lo, hi, nnids);
      else
        for (int64_t r = lo; r < hi; r++) nnids[r] = isDecidedRow(nids[r]) ? DECIDED_ROW : 0;  // :204-211 (q1)
      std::vector<int>& nh = nhs[c];
      nh.assign(numLeafs + 1, 0);
      for (int64_t r = lo; r < hi; r++)
        if (nnids[r] >= 0) nh[nnids[r] + 1]++;
      for (int i = 0; i < numLeafs; i++) nh[i + 1] += nh[i];
      std::vector<int>& rows = rss[c];
      rows.assign(hi - lo, 0);
      for (int64_t r = lo; r < hi; r++)
        if (nnids[r] >= 0) rows[nh[nnids[r]]++] = (int)(r - lo);
      // gather ys in leaf order (:254-262)
      std::vector<double>& ys = yss[c];
      ys.assign(hi - lo, 0.0);
      for (int n = 0; n < (int)nh.size(); n++) {
        const int l = (n == 0 ? 0 : nh[n - 1]);
        const int h = nh[n];
        for (int i = l; i < h; i++) ys[i] = work[lo + rows[i]];
      }
    };
    parallel_for(nchunks, phase1);
    // ---- phase 2 (:272-310, 375-453): per column, workers over fixed chunk ranges with private copies
    // active columns (:90-102)
    std::vector<char> active(ncols, 1);
    bool allNull = false;
    {
      std::vector<char> a(ncols, 0);
      for (int n = 0; n < numLeafs; n++) {
        const Node& u = tree.node(n + leaf);
        if (!u._scoreColsNull) {
          for (int c : u._scoreCols) a[c] = 1;
        } else { allNull = true; break; }
      }
      if (!allNull) active = a;
    }
    const int nw = std::max(1, nthreads);
    const int64_t rangeLen = (nchunks + nw - 1) / nw;
    auto computeChunk = [&](int col, int64_t id, std::vector<HistPtr>& lh, std::vector<int32_t>& ibuf) {
      const std::vector<int>& nh = nhs[id];
      const std::vector<int>& rs = rss[id];
      const int64_t base = id * chunk_rows;
      const int len = (int)rs.size();
      const double* cs = fr.cols[col].v.data() + base;
      const double* ws = fr.w.empty() ? nullptr : fr.w.data() + base;
      bool extracted = false;
      for (int n = 0; n < numLeafs; n++) {
        const Node& u = tree.node(n + leaf);
        if (u._scoreColsNull || std::find(u._scoreCols.begin(), u._scoreCols.end(), col) != u._scoreCols.end()) {
          HistPtr& h = lh[n];
          int hi = nh[n];
          int lo = (n == 0 ? 0 : nh[n - 1]);
          if (hi == lo || !h) continue;
          if (!h->_has_vals) h->init();
          if (h->_intOpt) {
            if (!extracted) {  // extractData :715-728 -> getIntegers with INT_NA
              ibuf.resize(len);
              for (int i = 0; i < len; i++) ibuf[i] = std::isnan(cs[i]) ? DHistogram::INT_NA : j_d2i(cs[i]);
              extracted = true;
            }
            h->updateHistoInt(ws, ibuf.data(), yss[id].data(), rs.data(), hi, lo);
          } else {
            h->updateHisto(ws, cs, yss[id].data(), rs.data(), hi, lo);
          }
        }
      }
    };
    std::vector<int> colsToDo;
    for (int c = 0; c < ncols; c++)
      if (active[c]) colsToDo.push_back(c);
    // tasks = (col, worker); each worker gets a private deep copy, merged in worker order (mergeHistos :313-321)
    struct Task { int col; int w; std::vector<HistPtr> lh; };
    std::vector<Task> tasks;
    for (int c : colsToDo)
      for (int w = 0; w < nw; w++) {
        if (w * rangeLen >= nchunks) break;
        Task t;
        t.col = c;
        t.w = w;
        t.lh.resize(numLeafs);
        for (int n = 0; n < numLeafs; n++)
          t.lh[n] = hcs[n][c] ? (w == 0 ? hcs[n][c] : std::make_shared<DHistogram>(*hcs[n][c])) : nullptr;
        tasks.push_back(std::move(t));
      }
    parallel_for((int64_t)tasks.size(), [&](int64_t ti) {
      Task& t = tasks[ti];
      std::vector<int32_t> ibuf;
      int64_t lo = t.w * rangeLen, hi = std::min(nchunks, lo + rangeLen);
      for (int64_t id = lo; id < hi; id++) computeChunk(t.col, id, t.lh, ibuf);
    });
    for (Task& t : tasks) {
      if (t.w == 0) continue;
      for (int n = 0; n < numLeafs; n++) {
        HistPtr& h1 = hcs[n][t.col];
        if (h1 && t.lh[n]) h1->add(*t.lh[n]);
      }
    }
    // postGlobal :455-466
    for (auto& ary : hcs)
      for (auto& dh : ary)
        if (dh) dh->reducePrecision();
  }

  template <class F>
  void parallel_for(int64_t n, F f) {
    int nt = (int)std::max<int64_t>(1, std::min<int64_t>(nthreads, n));
    if (nt <= 1) {
      for (int64_t i = 0; i < n; i++) f(i);
      return;
    }
    std::vector<std::thread> th;
    std::atomic<int64_t> next{0};
    for (int t = 0; t < nt; t++)
      th.emplace_back([&]() {
        for (int64_t i = next.fetch_add(1); i < n; i = next.fetch_add(1)) f(i);
      });
    for (auto& t : th) t.join();
  }
};

// ----------------------------------------------------------------------------------------------
// Distributions — H/DistributionFactory.java:73-134,243-272,524-546 ; H/LinkFunction.java:55-137
// ----------------------------------------------------------------------------------------------
namespace dist {
inline double exp_(double x) { return std::fmin(1e19, std::exp(x)); }  // LogExpUtil.exp :533
inline double log_(double x) {                                          // LogExpUtil.log :541-544
  x = std::fmax(0, x);
  return x == 0 ? -19 : std::fmax(-19, std::log(x));
}
inline double linkInv(int family, double f) {
  switch (family) {
    case GAUSSIAN: case LAPLACE: case QUANTILE: return f;
    case BERNOULLI: return 1 / (1 + exp_(-f));
    default: return exp_(f);
  }
}
inline double link(int family, double f) {
  switch (family) {
    case GAUSSIAN: case LAPLACE: case QUANTILE: return f;
    case BERNOULLI: return log_(f / (1 - f));
    default: return log_(f);
  }
}
// H/DistributionFactory.java: Gaussian :73-103, Bernoulli :105-134, Multinomial :243-272, Poisson :274-307, Gamma :309-341,
// Tweedie :343-379 (p == _tweediePower)
inline double negHalfGradient(int family, double y, double f, double p = 1.5) {
  switch (family) {
    case GAMMA: return y * exp_(-f) - 1;
    case TWEEDIE: return y * exp_(f * (1 - p)) - exp_(f * (2 - p));
    case LAPLACE: return f > y ? -0.5 : 0.5;                    // Laplace :417-420
    case QUANTILE: return y > f ? 0.5 * p : 0.5 * (p - 1);      // Quantile :432-433 (p == _quantileAlpha)
    default: return y - linkInv(family, f);
  }
}
// Quantile.Histo.findQuantile + computeQuantile / linearInterpolate (h2o-core/src/main/java/hex/quantile/Quantile.java:267-350) for
// unit weights and CombineMethod.INTERPOLATE. The reference refines a 1024-bin histogram until the bins that hold rows r2 = (long)
// (prob (n - 1)) and r2 + 1 are constant; what it then has are the two order statistics, which do not depend on how they were
// found -- here the values are sorted. `v` is reordered.
inline double quantileInterpolated(std::vector<double>& v, double prob) {
  const double nrows = (double)v.size();
  if (v.empty()) return NAN;
  std::sort(v.begin(), v.end());
  const double p2 = prob * (nrows - 1);
  const long long r2 = (long long)p2;
  const double lo = v[(size_t)r2];
  if (r2 + 1 >= (long long)v.size()) return lo;
  const double hi = v[(size_t)r2 + 1];
  if (lo == hi) return lo;                                      // computeQuantile :331
  const double row = (double)r2;
  const double plo = (row + 0) / (nrows - 1), phi = (row + 1) / (nrows - 1);  // linearInterpolate :344-350
  return lo + (hi - lo) * (prob - plo) / (phi - plo);
}
inline double initFNum(int family, double w, double o, double y, double p = 1.5) {
  switch (family) {
    case POISSON: return w * y;
    case GAMMA: return w * y * exp_(-o);
    case TWEEDIE: return w * y * exp_(o * (1 - p));
    default: return w * (y - o);
  }
}
inline double initFDenom(int family, double w, double o, double, double p = 1.5) {
  switch (family) {
    case POISSON: return w * exp_(o);
    case TWEEDIE: return w * exp_(o * (2 - p));
    default: return w;
  }
}
inline double gammaNum(int family, double w, double y, double z, double f, double p = 1.5) {
  switch (family) {
    case POISSON: return w * y;
    case GAMMA: return w * (z + 1);
    case TWEEDIE: return w * y * exp_(f * (1 - p));
    default: return w * z;
  }
}
inline double gammaDenom(int family, double w, double y, double z, double f, double p = 1.5) {
  switch (family) {
    case GAUSSIAN: return w;
    case BERNOULLI: { double ff = y - z; return w * ff * (1 - ff); }
    case POISSON: return w * (y - z);
    case GAMMA: return w;
    case TWEEDIE: return w * exp_(f * (2 - p));
    default: { double absz = std::fabs(z); return w * (absz * (1 - absz)); }
  }
}
}  // namespace dist

// G/GenModel.java:671-681 ; preds[1..K]
inline double log_rescale(double* preds, int len) {
  double maxval = -INFINITY;
  for (int k = 1; k < len; k++) maxval = std::fmax(maxval, preds[k]);
  double dsum = 0;
  for (int k = 1; k < len; k++) dsum += (preds[k] = std::exp(preds[k] - maxval));
  return dsum;
}

// ----------------------------------------------------------------------------------------------
// GBM driver — T/gbm/GBM.java:254-312 (init), :464-591 (buildNextKTrees), :607-680 (growTrees),
// :813-847 (fitBestConstants), :1032-1081 (ComputePredAndRes), :1286-1407 (GammaPass),
// :1410-1470 (AddTreeContributions); T/SharedTree.java:554-710 (buildLayer/ScoreBuildOneTree),
// :1138-1185 (InitialValue); T/Sample.java:22-36.
// ----------------------------------------------------------------------------------------------
struct TreeOut {  // flattened finished tree (what DTree.compress would serialise; T/DTree.java:891-933)
  std::vector<Node> nodes;
  int leaves_start = 0;
  int depth = 0;
};

struct GBMModel {
  Params parms;
  double init_f = 0;
  std::vector<std::vector<TreeOut>> trees;  // [tree][k]
  std::vector<std::vector<double>> F;       // Tree_k columns after training (float-valued), [k][row]
  std::vector<double> oobt;                 // DRF: OUT_BAG_TREES column (sum of weights of the trees that held the row out)
  std::vector<LevelCapture> captures;
  std::vector<float> varimp;
  // timing of the restated CPU path (bench.py cpu_baseline leg only)
  double t_hist_s = 0;        // wall seconds inside ScoreBuildHistogram2 (score + histogram build)
  double t_total_s = 0;       // wall seconds for all trees
  double hist_rowcols = 0;    // sum over levels of rows x active cols histogrammed
};

struct GBM {
  const Frame& fr;
  Params parms;
  int K;         // number of class-trees (numClassTrees :597-599)
  int nclass;
  std::vector<std::vector<double>> tree_;  // Tree_k
  std::vector<std::vector<double>> work_;  // Work_k
  std::vector<std::vector<int32_t>> nids_; // NIDs_k
  PCGRNG rand_;
  int ntrees_built = 0;
  bool capture = false;
  GBMModel model;
  int mtry = 0, mtry_per_tree = 0;

  GBM(const Frame& f, const Params& p) : fr(f), parms(p), rand_(p.seed) {
    nclass = parms.family == MULTINOMIAL ? parms.nclass : (parms.family == BERNOULLI ? 2 : 1);
    K = nclass == 2 ? 1 : nclass;
    const int64_t N = fr.nrows();
    tree_.assign(nclass > 2 ? nclass : 1, std::vector<double>(N, 0.0));
    work_.assign(tree_.size(), std::vector<double>(N, 0.0));
    nids_.assign(tree_.size(), std::vector<int32_t>(N, 0));
    model.parms = parms;
    model.varimp.assign(fr.ncols(), 0.f);
  }

  double getInitialValue() {  // T/SharedTree.java:1138-1185
    double num = 0, denom = 0;
    for (int64_t i = 0; i < fr.nrows(); ++i) {
      if (std::isnan(fr.y[i])) continue;
      double w = fr.w.empty() ? 1 : fr.w[i];
      if (w == 0) continue;
      double y = fr.y[i];
      double o = fr.off.empty() ? 0 : fr.off[i];
      num += dist::initFNum(parms.family, w, o, y, parms.tweedie_power);
      denom += dist::initFDenom(parms.family, w, o, y, parms.tweedie_power);
    }
    if (parms.family == MULTINOMIAL) return -0.5 * dist::link(BERNOULLI, num / denom);
    return dist::link(parms.family, num / denom);
  }

  void init() {  // T/gbm/GBM.java:254-312
    const int ncols = fr.ncols();
    mtry_per_tree = std::max(1, j_d2i(parms.col_sample_rate_per_tree * ncols));
    mtry = std::max(1, j_d2i(parms.col_sample_rate * parms.col_sample_rate_per_tree * ncols));
    double initialPrediction = nclass > 2 ? 0 : getInitialValue();
    if (parms.family == LAPLACE || parms.family == QUANTILE) {  // getInitialValueQuantile :272-276, 355-381: quantile of y - offset
      if (!fr.w.empty()) throw std::runtime_error("oracle: laplace / quantile with observation weights is not restated");
      std::vector<double> v;
      for (int64_t row = 0; row < fr.nrows(); row++)
        if (!std::isnan(fr.y[row])) v.push_back(fr.y[row] - (fr.off.empty() ? 0 : fr.off[row]));
      initialPrediction = dist::quantileInterpolated(v, parms.family == LAPLACE ? 0.5 : parms.tweedie_power);
      if (std::isnan(initialPrediction)) initialPrediction = 0;
    }
    model.init_f = initialPrediction;
    if (initialPrediction != 0.0) std::fill(tree_[0].begin(), tree_[0].end(), initialPrediction);
    // MarkDecidedRows :288-349
    for (int64_t row = 0; row < fr.nrows(); row++) {
      double w = fr.w.empty() ? 1 : fr.w[row];
      if (w == 0 || std::isnan(fr.y[row]))
        for (auto& n : nids_) n[row] = DECIDED_ROW;
    }
  }

  void computePredAndRes() {  // :1032-1081
    const int64_t N = fr.nrows();
    std::vector<double> fs(nclass + 1);
    for (int64_t row = 0; row < N; row++) {
      double weight = fr.w.empty() ? 1 : fr.w[row];
      if (weight == 0) continue;
      if (std::isnan(fr.y[row])) continue;
      double f = tree_[0][row] + (fr.off.empty() ? 0 : fr.off[row]);
      double y = fr.y[row];
      if (parms.family == MULTINOMIAL) {
        // score1static :1486-1513
        fs[1] = f;
        for (int k = 1; k < nclass; k++) fs[k + 1] = tree_[k][row];
        double sum = log_rescale(fs.data(), nclass + 1);
        if (std::isinf(sum)) {
          for (int k = 0; k < nclass; k++) {
            float p = std::isinf(fs[k + 1]) ? 1.0f : 0.0f;
            work_[k][row] = (float)(((int)y == k ? 1.0f : 0.0f) - (double)p);
          }
        } else {
          for (int k = 0; k < nclass; k++) {
            float p = (float)(fs[k + 1] / sum);
            work_[k][row] = (float)((double)((int)y == k ? 1.0f : 0.0f) - (double)p);
          }
        }
      } else {
        work_[0][row] = (float)dist::negHalfGradient(parms.family, y, f, parms.tweedie_power);
      }
    }
  }

  void sample(int k, const DTree& tree) {  // T/Sample.java:22-36
    PCGRNG rand(tree._seed);
    for (int64_t row = 0; row < fr.nrows(); row++) {
      bool skip = std::isnan(fr.y[row]);
      if (!skip) {
        const double rate = parms.sample_rate_per_class.empty() ? parms.sample_rate : parms.sample_rate_per_class[(size_t)(int)fr.y[row]];
        rand.setSeed((int64_t)((uint64_t)tree._seed + (uint64_t)row));
        skip = rand.nextFloat() >= rate;
      }
      if (skip) nids_[k][row] = OUT_OF_BAG;
    }
  }

  // buildLayer (T/SharedTree.java:554-628) + ScoreBuildOneTree.onCompletion (:684-710) for one k
  bool buildLayerK(int k, DTree& tree, int& leafOffset, std::vector<std::vector<HistPtr>>& hcs) {
    SBH2 sbh{fr, parms, tree, leafOffset, hcs, work_[k], nids_[k], parms.nthreads,
             std::max<int64_t>(1, chunk_rows_)};
    auto t0 = std::chrono::steady_clock::now();
    sbh.run();
    model.t_hist_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    model.hist_rowcols += (double)fr.nrows() * fr.ncols();
    if (capture) captureLevel(k, tree, leafOffset, hcs);
    bool did_split = false;
    int tmax = tree.len();
    for (int leaf = leafOffset; leaf < tmax; leaf++) {
      tree.decide(leaf, hcs[leaf - leafOffset]);
      const Node& dn = tree.node(leaf);
      if (!dn._split) tree.doNotSplit(leaf);
      else {
        did_split = true;
        float improvement = (float)(dn._split->pre_split_se() - dn._split->se());
        model.varimp[dn._split->_col] += improvement;
      }
    }
    leafOffset = tmax;
    int new_leafs = tree.len() - tmax;
    hcs.assign(new_leafs, {});
    for (int nl = tmax; nl < tree.len(); nl++) hcs[nl - tmax] = tree.node(nl)._hs;
    if (did_split) tree._depth++;
    return did_split;
  }

  void captureLevel(int k, const DTree& tree, int leaf, const std::vector<std::vector<HistPtr>>& hcs) {
    LevelCapture cap;
    cap.tree = ntrees_built;
    cap.k = k;
    cap.leaf = leaf;
    cap.nleaves = (int)hcs.size();
    const int ncols = fr.ncols();
    cap.entries.resize((size_t)cap.nleaves * ncols);
    for (int n = 0; n < cap.nleaves; n++) {
      const Node& u = tree.node(n + leaf);
      for (int c = 0; c < ncols; c++) {
        auto& e = cap.entries[(size_t)n * ncols + c];
        const HistPtr& h = hcs[n][c];
        bool tracked = u._scoreColsNull || std::find(u._scoreCols.begin(), u._scoreCols.end(), c) != u._scoreCols.end();
        if (!h || !tracked) continue;
        e.present = true;
        e.has_vals = h->_has_vals;
        e.isInt = h->_isInt;
        e.intOpt = h->_intOpt;
        e.nbin = h->_nbin;
        e.min = h->_min;
        e.maxEx = h->_maxEx;
        e.step = h->_step;
        e.min2 = h->_min2;
        e.maxIn = h->_maxIn;
        e.vals = h->_vals;
      }
    }
    cap.nids_after = nids_[k];
    model.captures.push_back(std::move(cap));
  }

  int64_t chunk_rows_ = 1 << 20;

  void buildNextKTrees() {  // :464-591
    std::vector<std::unique_ptr<DTree>> ktrees(tree_.size());
    std::vector<int> leaves(tree_.size(), 0);
    computePredAndRes();
    // growTrees :607-680
    std::vector<std::vector<std::vector<HistPtr>>> hcs(tree_.size());
    int adj_nbins = std::max(parms.nbins_top_level, parms.nbins);
    int64_t rseed = rand_.nextLong();
    for (int k = 0; k < K; k++) {
      ktrees[k] = std::make_unique<DTree>(fr.ncols(), mtry, mtry_per_tree, rseed, &parms);
      hcs[k].resize(1);
      hcs[k][0] = initialHist(fr, adj_nbins, rseed, parms, false);
      ktrees[k]->newUndecided(NO_PARENT, hcs[k][0]);
      hcs[k][0] = ktrees[k]->node(0)._hs;
    }
    if (parms.sample_rate < 1 || !parms.sample_rate_per_class.empty())  // useRowSampling (T/SharedTreeModel.java:86-88)
      for (int k = 0; k < K; k++) sample(k, *ktrees[k]);
    for (int depth = 0; depth < parms.max_depth; depth++) {
      bool did_split = false;
      for (int k = 0; k < K; k++) {
        if (!ktrees[k]) continue;
        if (buildLayerK(k, *ktrees[k], leaves[k], hcs[k])) did_split = true;
      }
      if (!did_split) break;
    }
    for (int k = 0; k < K; k++) {  // :655-679
      DTree& tree = *ktrees[k];
      int leaf = tree.len();
      leaves[k] = leaf;
      for (int nid = 0; nid < leaf; nid++) {
        if (tree.node(nid).kind == DECIDED) {
          if (!tree.node(nid)._split) {
            if (nid == 0) tree.newLeafAt(NO_PARENT, 0);
            continue;
          }
          for (int i = 0; i < 2; i++) {
            int cnid = tree.node(nid)._nids[i];
            if (cnid == UNDECIDED_CHILD_NODE_ID || tree.node(cnid).kind == UNDECIDED ||
                (tree.node(cnid).kind == DECIDED && !tree.node(cnid)._split)) {
              int ln = tree.newLeaf(nid);
              tree.node(nid)._nids[i] = ln;
            }
          }
        }
      }
    }
    // GammaPass :1286-1407
    std::vector<std::vector<double>> num(K), denom(K);
    for (int k = 0; k < K; k++) {
      DTree& tree = *ktrees[k];
      const int leaf = leaves[k];
      denom[k].assign(tree.len() - leaf, 0.0);
      num[k].assign(tree.len() - leaf, 0.0);
      if (tree.node(0).kind == LEAF) continue;
      for (int64_t row = 0; row < fr.nrows(); row++) {
        double w = fr.w.empty() ? 1 : fr.w[row];
        if (w == 0) continue;
        double y = fr.y[row];
        if (std::isnan(y)) continue;
        int nid = nids_[k][row];
        const bool wasOOBRow = isOOBRow(nid);
        if (wasOOBRow) nid = oob2Nid(nid);
        if (nid < 0) continue;
        const Node* dn = &tree.node(nid);
        if (!dn->_split) dn = &tree.node(dn->_pid);
        int leafnid = getChildNodeID(*dn, fr, row);
        nids_[k][row] = leafnid;
        if (wasOOBRow) continue;
        double z = work_[k][row];
        double f = tree_[k][row] + (fr.off.empty() ? 0 : fr.off[row]);
        int idx = leafnid - leaf;
        num[k][idx] += dist::gammaNum(parms.family, w, y, z, f, parms.tweedie_power);
        denom[k][idx] += dist::gammaDenom(parms.family, w, y, z, f, parms.tweedie_power);
      }
    }
    // fitBestConstants :813-847
    double m1class = (nclass > 1 && parms.family == MULTINOMIAL) ? (double)(nclass - 1) / nclass : 1.0;
    double eff_lr = parms.learn_rate * std::pow(parms.learn_rate_annealing, (double)(ntrees_built - 1));  // :725-727
    if (parms.family == LAPLACE || parms.family == QUANTILE) {
      // fitBestConstantsQuantile :522-527, 729-751: StratifiedQuantilesTask over diff = y - (f + o) (ComputeDiff :1132-1147) with the
      // NIDs column as strata -- every row that GammaPass moved onto a leaf takes part, out-of-bag rows included
      DTree& tree = *ktrees[0];
      const double prob = parms.family == LAPLACE ? 0.5 : parms.tweedie_power;
      const int nleaf = tree.len() - leaves[0];
      std::vector<std::vector<double>> diffs((size_t)std::max(nleaf, 0));
      if (tree.node(0).kind != LEAF)
        for (int64_t row = 0; row < fr.nrows(); row++) {
          const int nid = nids_[0][row];
          if (nid < leaves[0] || std::isnan(fr.y[row])) continue;
          diffs[(size_t)(nid - leaves[0])].push_back(fr.y[row] - (tree_[0][row] + (fr.off.empty() ? 0 : fr.off[row])));
        }
      for (int i = 0; i < nleaf; i++) {
        const double q = dist::quantileInterpolated(diffs[(size_t)i], prob);
        if (std::isnan(q)) continue;  // no active rows for this NID
        double val = eff_lr * q;
        if (val > parms.max_abs_leafnode_pred) val = parms.max_abs_leafnode_pred;
        if (val < -parms.max_abs_leafnode_pred) val = -parms.max_abs_leafnode_pred;
        tree.node(leaves[0] + i)._pred = (float)val;
      }
    }
    for (int k = 0; k < K && parms.family != LAPLACE && parms.family != QUANTILE; k++) {
      DTree& tree = *ktrees[k];
      for (int i = 0; i < tree.len() - leaves[k]; i++) {
        Node& leafNode = tree.node(leaves[k] + i);
        double gamma = denom[k][i] == 0 ? 0 : num[k][i] / denom[k][i];
        if (denom[k][i] != 0 && (parms.family == POISSON || parms.family == GAMMA || parms.family == TWEEDIE))
          gamma = dist::link(parms.family, gamma);  // GammaPass.gamma(double) :1315-1323
        double gf = eff_lr * m1class * gamma;
        if (parms.family == MULTINOMIAL) {
          if (gf > 1e4) gf = 1e4f;
          else if (gf < -1e4) gf = -1e4f;
        }
        if (std::isnan(gf)) gf = 0;
        else if (std::isinf(gf)) gf = (gf > 0 ? 1.0 : -1.0) * 1e4f;
        if (gf > parms.max_abs_leafnode_pred) gf = parms.max_abs_leafnode_pred;
        if (gf < -parms.max_abs_leafnode_pred) gf = -parms.max_abs_leafnode_pred;
        leafNode._pred = (float)gf;
      }
    }
    // AddTreeContributions :1410-1470
    for (int k = 0; k < K; k++) {
      DTree& tree = *ktrees[k];
      PCGRNG noise_rng(parms.seed);
      // (0xDECAF + _seed) * (0xFAAAAAAB + k * _ntrees1 + _ntrees2): the second factor is Java INT arithmetic (0xFAAAAAAB is a
      // negative int literal) widened with its sign, the product is long; _ntrees2 = trees built before this iteration
      const int32_t f2 = (int32_t)(0xFAAAAAABu + (uint32_t)(k * parms.ntrees + ntrees_built));
      const int64_t baseseed = (int64_t)(((uint64_t)0xDECAF + (uint64_t)parms.seed) * (uint64_t)(int64_t)f2);
      for (int64_t row = 0; row < fr.nrows(); row++) {
        int nid = nids_[k][row];
        nids_[k][row] = FRESH;
        if (nid < 0) {
          double w = fr.w.empty() ? 1 : fr.w[row];
          if (w == 0 || std::isnan(fr.y[row])) nids_[k][row] = DECIDED_ROW;
          continue;
        }
        double factor = 1;
        if (parms.pred_noise_bandwidth != 0) {
          noise_rng.setSeed((int64_t)((uint64_t)baseseed + (uint64_t)(int64_t)nid));
          factor += noise_rng.nextGaussianAfterSeed() * parms.pred_noise_bandwidth;
        }
        tree_[k][row] = (float)(tree_[k][row] + factor * (double)tree.node(nid)._pred);
      }
    }
    std::vector<TreeOut> outs(K);
    for (int k = 0; k < K; k++) {
      outs[k].nodes = ktrees[k]->_ns;
      for (auto& n : outs[k].nodes) n._hs.clear();
      outs[k].leaves_start = leaves[k];
      outs[k].depth = ktrees[k]->_depth;
    }
    model.trees.push_back(std::move(outs));
    ntrees_built++;
  }

  GBMModel& train() {
    init();
    auto t0 = std::chrono::steady_clock::now();
    for (int t = 0; t < parms.ntrees; t++) buildNextKTrees();
    model.t_total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    model.F = tree_;
    return model;
  }
};

// ----------------------------------------------------------------------------------------------
// DRF driver — T/drf/DRF.java:114-330 on top of the shared level loop (SharedTree.buildLayer + SBH2).
// Differences from GBM that matter for the hot path: the work columns are filled once (class indicator or
// float(response)), histograms use the float-split correction (_checkFloatSplits, :201), every tree samples rows
// (Sample also sends NA-response rows out of bag), per-node mtries column draws, leaf value = the split's child mean
// (DecidedNode.pred, T/DTree.java:765), and only OUT-OF-BAG rows receive the tree's vote (CollectPreds :255-330).
// ----------------------------------------------------------------------------------------------
struct DRF : GBM {
  std::vector<double> oobt;
  std::vector<double> priorClassDist;

  DRF(const Frame& f, const Params& p) : GBM(f, p) { oobt.assign(fr.nrows(), 0.0); }
  bool isClassifier() const { return nclass > 1; }

  void initDRF() {  // initializeModelSpecifics :114-153
    const int ncols = fr.ncols();
    mtry_per_tree = std::max(1, j_d2i(parms.col_sample_rate_per_tree * ncols));
    if (parms.mtries == -2) mtry = ncols;
    else if (parms.mtries == -1) mtry = isClassifier() ? std::max(j_d2i(std::sqrt((double)ncols)), 1) : std::max(ncols / 3, 1);
    else mtry = parms.mtries;
    model.init_f = isClassifier() ? 0 : getInitialValue();
    // class prior (Model.Output._priorClassDist): weighted class frequencies over rows with a response
    if (isClassifier()) {
      priorClassDist.assign(nclass, 0.0);
      double tot = 0;
      for (int64_t i = 0; i < fr.nrows(); i++) {
        if (std::isnan(fr.y[i])) continue;
        double w = fr.w.empty() ? 1 : fr.w[i];
        priorClassDist[(int)fr.y[i]] += w;
        tot += w;
      }
      for (double& v : priorClassDist) v /= tot;
    }
    for (int64_t i = 0; i < fr.nrows(); i++) {  // fill work columns :139-153
      if (std::isnan(fr.y[i])) continue;
      if (isClassifier()) {
        int cls = (int)fr.y[i];
        if (cls < (int)work_.size()) work_[cls][i] = 1.0;
      } else {
        work_[0][i] = (float)fr.y[i];
      }
    }
  }

  void buildNextKTreesDRF() {  // :157-250
    std::vector<std::unique_ptr<DTree>> ktrees(tree_.size());
    std::vector<int> leaves(tree_.size(), 0);
    std::vector<std::vector<std::vector<HistPtr>>> hcs(tree_.size());
    int adj_nbins = std::max(parms.nbins_top_level, parms.nbins);
    int64_t rseed = rand_.nextLong();
    for (int k = 0; k < K; k++) {
      ktrees[k] = std::make_unique<DTree>(fr.ncols(), mtry, mtry_per_tree, rseed, &parms);
      hcs[k].resize(1);
      hcs[k][0] = initialHist(fr, adj_nbins, rseed, parms, true);
      ktrees[k]->newUndecided(NO_PARENT, hcs[k][0]);
      hcs[k][0] = ktrees[k]->node(0)._hs;
    }
    for (int k = 0; k < K; k++) sample(k, *ktrees[k]);  // always (T/drf/DRF.java:206-210)
    for (int depth = 0; depth < parms.max_depth; depth++) {
      bool did_split = false;
      for (int k = 0; k < K; k++)
        if (buildLayerK(k, *ktrees[k], leaves[k], hcs[k])) did_split = true;
      if (!did_split) break;
    }
    for (int k = 0; k < K; k++) {  // :222-248
      DTree& tree = *ktrees[k];
      int leaf = leaves[k] = tree.len();
      for (int nid = 0; nid < leaf; nid++) {
        if (tree.node(nid).kind != DECIDED) continue;
        if (!tree.node(nid)._split) {
          if (nid == 0) {
            tree.newLeafAt(NO_PARENT, 0);
            tree.node(0)._pred = (float)(isClassifier() ? priorClassDist[k] : model.init_f);
          }
          continue;
        }
        for (int i = 0; i < 2; i++) {
          int cnid = tree.node(nid)._nids[i];
          if (cnid == UNDECIDED_CHILD_NODE_ID || tree.node(cnid).kind == UNDECIDED ||
              (tree.node(cnid).kind == DECIDED && !tree.node(cnid)._split)) {
            int ln = tree.newLeaf(nid);
            const Split& sp = *tree.node(nid)._split;
            tree.node(ln)._pred = (float)(i == 0 ? sp._p0 : sp._p1);
            tree.node(nid)._nids[i] = ln;
          }
        }
      }
    }
    // CollectPreds :255-330 (votes / sse bookkeeping omitted: not part of the model)
    for (int64_t row = 0; row < fr.nrows(); row++) {
      double weight = fr.w.empty() ? 1 : fr.w[row];
      const bool wasOOBRow = isOOBRow(nids_[0][row]);
      for (int k = 0; k < K; k++) {
        if (weight != 0) {
          DTree& tree = *ktrees[k];
          int nid = nids_[k][row];
          if (wasOOBRow) {
            nid = oob2Nid(nid);
            if (tree.node(nid).kind == UNDECIDED) nid = tree.node(nid)._pid;
            int leafnid;
            if (tree.node(0).kind == LEAF) leafnid = 0;
            else {
              const Node* dn = &tree.node(nid);
              if (!dn->_split) dn = &tree.node(tree.node(nid)._pid);
              leafnid = getChildNodeID(*dn, fr, row);
            }
            double prediction = tree.node(leafnid)._pred;
            tree_[k][row] = (float)(tree_[k][row] + prediction);
          }
        }
        nids_[k][row] = 0;
      }
      if (wasOOBRow) oobt[row] += weight;
    }
    std::vector<TreeOut> outs(K);
    for (int k = 0; k < K; k++) {
      outs[k].nodes = ktrees[k]->_ns;
      for (auto& n : outs[k].nodes) n._hs.clear();
      outs[k].leaves_start = leaves[k];
      outs[k].depth = ktrees[k]->_depth;
    }
    model.trees.push_back(std::move(outs));
    ntrees_built++;
  }

  GBMModel& trainDRF() {
    initDRF();
    auto t0 = std::chrono::steady_clock::now();
    for (int t = 0; t < parms.ntrees; t++) buildNextKTreesDRF();
    model.t_total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    model.F = tree_;
    model.oobt = oobt;
    return model;
  }
};

// Score one row through a finished tree (T/CompressedTree / G/algos/tree/SharedTreeMojoModel.java:129-240
// semantics on the uncompressed node table — used by tests for testPubDev6356-style checks).
inline float scoreTree(const TreeOut& t, const std::function<double(int)>& x) {
  int nid = 0;
  while (t.nodes[nid].kind != LEAF) {
    const Node& dn = t.nodes[nid];
    const Split& sp = *dn._split;
    double d = x(sp._col);
    int bin = -1;
    bool isNA = std::isnan(d);
    if (!isNA) {
      if (sp._nasplit == NAvsREST) bin = 0;
      else if (sp._equal == 0) bin = d >= dn._splat ? 1 : 0;
      else {
        int b = j_d2i(d);
        if (sp._bs->isInRange(b)) bin = sp._bs->contains(b) ? 1 : 0;
        else isNA = true;
      }
    }
    if (isNA) bin = (sp._nasplit == NALeft || sp._nasplit == Left) ? 0 : 1;
    nid = dn._nids[bin];
  }
  return t.nodes[nid]._pred;
}

}  // namespace h2o_oracle
