// TEST INFRASTRUCTURE ONLY (see oracle/h2o_tree_oracle.hpp). CPU restatement of the reference's streaming AUC:
//   AUC2.AUCBuilder.perRow / reduce / mergeOneBin / find_smallest_impl / dups / compute_delta_error / combine_centers
//                                           h2o-core/src/main/java/hex/AUC2.java:356-581
//   AUC2(AUCBuilder) + compute_auc          h2o-core/src/main/java/hex/AUC2.java:158-190, 262-277
//   AUC_Impl (one builder per chunk, reduced pairwise)   h2o-core/src/main/java/hex/AUC2.java:341-353
// The builder is order dependent by construction (a streaming 400-bin histogram merged greedily), so a result is defined by the
// chunking: `aucOverChunks` builds one AUCBuilder per chunk of consecutive rows, in row order, and folds the builders left to
// right. Parity status: pinned by the reference's fixture-free known answer AUCBuilderTest.testPerRow (a 10-bin builder's
// thresholds after 100 rows: the merge path), by the inputs of AUC2Test.perfectAUC (exact AUCs: with at most NBINS distinct
// predictions the builder never merges and the trapezoid sum is the exact AUC), and by source (tests/test_oracle_compress_auc.py).
// AUCBuilderTest.testPerRow_compat (fast path == insert + mergeOneBin after every row of h2o-core/src/test/resources/hex/
// aucbuilder.csv.gz, 240 618 rows) is re-run against this restatement where the reference tree is present.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace h2o_oracle {

struct AUCBuilder {
  static constexpr int NBINS = 400;  // AUC2.java:36
  int _nBins, _n = 0, _ssx = -1;
  bool _useFastPath = true;  // :372-373, 385: false only in AUCBuilderTest.testPerRow_compat (insert, then mergeOneBin, every time)
  std::vector<double> _ths, _sqe, _tps, _fps;
  explicit AUCBuilder(int nBins = NBINS) : _nBins(nBins), _ths(2 * nBins, 0.0), _sqe(2 * nBins, 0.0), _tps(2 * nBins, 0.0), _fps(2 * nBins, 0.0) {}

  double k(int idx) const { return _tps[idx] + _fps[idx]; }
  static double compute_delta_error(double ths1, double n1, double ths0, double n0) {  // :568-578
    const double delta = (double)((float)ths1 - (float)ths0);
    if (delta == 0) return 0;
    return delta * delta * n0 * n1 / (n0 + n1);
  }
  static double combine_centers(double ths1, double n1, double ths0, double n0) {  // :474-481
    const double center = (ths0 * n0 + ths1 * n1) / (n0 + n1);
    if (std::isnan(center) || std::isinf(center)) return (ths0 + ths1) / 2;
    return center;
  }
  // java.util.Arrays.binarySearch(double[], from, to, key) for a key that is not NaN (-0.0 < 0.0 as in Double.compare)
  static int jcompare(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    int64_t x, y;
    memcpy(&x, &a, 8);
    memcpy(&y, &b, 8);
    return x == y ? 0 : (x < y ? -1 : 1);
  }
  int binary_search(double key) const {
    int low = 0, high = _n - 1;
    while (low <= high) {
      const int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
      const int c = jcompare(_ths[mid], key);
      if (c < 0) low = mid + 1;
      else if (c > 0) high = mid - 1;
      else return mid;
    }
    return -(low + 1);
  }
  int find_smallest_impl() const {  // :523-554
    if (_n == 1) return 0;
    double minSQE = 1.7976931348623157e308;
    int minI = -1;
    for (int i = 0; i < _n - 1; i++) {
      const double derr = compute_delta_error(_ths[i + 1], k(i + 1), _ths[i], k(i));
      if (derr == 0) return i;
      const double sqe = _sqe[i] + _sqe[i + 1] + derr;
      if (sqe < minSQE) { minI = i; minSQE = sqe; }
    }
    if (minI == -1) {
      minI = 0;
      double minDist = _ths[1] - _ths[0];
      for (int i = 1; i < _n - 1; i++) {
        const double dist = _ths[i + 1] - _ths[i];
        if (dist < minDist) { minDist = dist; minI = i; }
      }
    }
    return minI;
  }
  int find_smallest() {
    if (_ssx == -1) _ssx = find_smallest_impl();
    return _ssx;
  }
  bool dups() {  // :556-565
    for (int i = 0; i < _n - 1; i++)
      if (compute_delta_error(_ths[i + 1], k(i + 1), _ths[i], k(i)) == 0) { _ssx = i; return true; }
    return false;
  }
  void mergeOneBin() {  // :483-505
    const int ssx = find_smallest();
    const double k0 = k(ssx), k1 = k(ssx + 1);
    _sqe[ssx] = _sqe[ssx] + _sqe[ssx + 1] + compute_delta_error(_ths[ssx + 1], k1, _ths[ssx], k0);
    _ths[ssx] = combine_centers(_ths[ssx], k0, _ths[ssx + 1], k1);
    _tps[ssx] += _tps[ssx + 1];
    _fps[ssx] += _fps[ssx + 1];
    const int cnt = _n - ssx - 2;
    if (cnt > 0) {
      memmove(&_ths[ssx + 1], &_ths[ssx + 2], sizeof(double) * cnt);
      memmove(&_sqe[ssx + 1], &_sqe[ssx + 2], sizeof(double) * cnt);
      memmove(&_tps[ssx + 1], &_tps[ssx + 2], sizeof(double) * cnt);
      memmove(&_fps[ssx + 1], &_fps[ssx + 2], sizeof(double) * cnt);
    }
    _n--;
    _ssx = -1;
  }
  void perRow(double pred, int act, double w) {  // :384-446
    int idx = binary_search(pred);
    if (idx >= 0) {
      if (act == 0) _fps[idx] += w; else _tps[idx] += w;
      _ssx = -1;
      return;
    }
    idx = -idx - 1;
    if (_useFastPath && _n == _nBins && idx > 0 && idx < _n && _ths[idx - 1] != _ths[idx]) {
      const int ssx = find_smallest();
      const double dssx = _sqe[ssx] + _sqe[ssx + 1] + compute_delta_error(_ths[ssx + 1], k(ssx + 1), _ths[ssx], k(ssx));
      const double d0 = _sqe[idx - 1] + compute_delta_error(pred, w, _ths[idx - 1], k(idx - 1));
      const double d1 = _sqe[idx] + compute_delta_error(_ths[idx], k(idx), pred, w);
      if (d0 < dssx || d1 < dssx) {
        if (d0 <= d1) idx--;
        if (ssx == idx - 1 || ssx == idx) _ssx = -1;
        const double kk = k(idx);
        if (act == 0) _fps[idx] += w; else _tps[idx] += w;
        _sqe[idx] = _sqe[idx] + compute_delta_error(pred, w, _ths[idx], kk);
        _ths[idx] = combine_centers(_ths[idx], kk, pred, w);
        return;
      }
    }
    if (idx == 0 || idx == _n || idx == _ssx) _ssx = -1;
    else if (idx < _ssx) _ssx++;
    const int cnt = _n - idx;
    if (cnt > 0) {
      memmove(&_ths[idx + 1], &_ths[idx], sizeof(double) * cnt);
      memmove(&_sqe[idx + 1], &_sqe[idx], sizeof(double) * cnt);
      memmove(&_tps[idx + 1], &_tps[idx], sizeof(double) * cnt);
      memmove(&_fps[idx + 1], &_fps[idx], sizeof(double) * cnt);
    }
    _ths[idx] = pred;
    _sqe[idx] = 0;
    if (act == 0) { _tps[idx] = 0; _fps[idx] = w; }
    else { _tps[idx] = w; _fps[idx] = 0; }
    _n++;
    if (_n > _nBins) mergeOneBin();
  }
  void reduce(const AUCBuilder& b) {  // :448-472
    int x = _n - 1, y = b._n - 1;
    while (x + y + 1 >= 0) {
      const bool self_is_larger = y < 0 || (x >= 0 && _ths[x] >= b._ths[y]);
      const AUCBuilder& src = self_is_larger ? *this : b;
      const int idx = self_is_larger ? x : y;
      _ths[x + y + 1] = src._ths[idx];
      _sqe[x + y + 1] = src._sqe[idx];
      _tps[x + y + 1] = src._tps[idx];
      _fps[x + y + 1] = src._fps[idx];
      if (self_is_larger) x--; else y--;
    }
    _n += b._n;
    _ssx = -1;
    while (_n > _nBins || dups()) mergeOneBin();
  }
  // AUC2(AUCBuilder) + compute_auc (:158-190, 262-277)
  double auc() const {
    if (_n < 1) return NAN;
    std::vector<double> tps(_n), fps(_n);
    for (int i = 0; i < _n; i++) { tps[i] = _tps[_n - 1 - i]; fps[i] = _fps[_n - 1 - i]; }
    double p = 0, n = 0;
    for (int i = 0; i < _n; i++) { p += tps[i]; tps[i] = p; n += fps[i]; fps[i] = n; }
    if (fps[_n - 1] == 0) return 1.0;
    if (tps[_n - 1] == 0) return 0.0;
    double tp0 = 0, fp0 = 0, area = 0;
    for (int i = 0; i < _n; i++) {
      area += (fps[i] - fp0) * (tps[i] + tp0) / 2.0;
      tp0 = tps[i];
      fp0 = fps[i];
    }
    return area / p / n;
  }
};

// One builder per chunk of `chunk_rows` consecutive rows, folded left to right. Rows with a NaN prediction or actual are skipped
// (AUC2.java:347-349), and so are rows of weight 0 / NaN (MetricBuilderBinomial.perRow returns before _auc.perRow,
// h2o-core/src/main/java/hex/ModelMetricsBinomial.java). Returns the final builder.
inline AUCBuilder aucOverChunks(const double* pred, const double* act, const double* w, int64_t n, int64_t chunk_rows) {
  AUCBuilder total;
  bool first = true;
  for (int64_t r0 = 0; r0 < n; r0 += chunk_rows) {
    AUCBuilder b;
    const int64_t r1 = r0 + chunk_rows < n ? r0 + chunk_rows : n;
    for (int64_t r = r0; r < r1; r++) {
      if (std::isnan(pred[r]) || std::isnan(act[r])) continue;
      if (w && (w[r] == 0 || std::isnan(w[r]))) continue;
      b.perRow(pred[r], (int)act[r], w ? w[r] : 1.0);
    }
    if (first) { total = b; first = false; }
    else total.reduce(b);
  }
  return total;
}

}  // namespace h2o_oracle
