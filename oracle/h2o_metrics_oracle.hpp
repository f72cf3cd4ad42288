// TEST INFRASTRUCTURE ONLY (see oracle/h2o_tree_oracle.hpp). CPU restatement of what hex.tree.Score feeds to the metric builders for
// a GBM's training rows, from the model's Tree_k columns (link space):
//   GbmMojoModel.unifyPreds / GBMModel.score0 tail     h2o-genmodel/src/main/java/hex/genmodel/algos/gbm/GbmMojoModel.java:43-63
//   GenModel.log_rescale / GBM_rescale                 h2o-genmodel/src/main/java/hex/genmodel/GenModel.java:671-688
//   GenModel.GLM_logitInv                              h2o-genmodel/src/main/java/hex/genmodel/GenModel.java:693
//   MetricBuilderBinomial.perRow                       h2o-core/src/main/java/hex/ModelMetricsBinomial.java:175-206
//   MetricBuilderMultinomial.perRow                    h2o-core/src/main/java/hex/ModelMetricsMultinomial.java:273-300
//   MetricBuilderRegression.perRow                     h2o-core/src/main/java/hex/ModelMetricsRegression.java:130-140
//   MathUtils.logloss / MAXLL                          h2o-core/src/main/java/water/util/MathUtils.java:798-802
// The sums are accumulated row by row in row order, as one MetricBuilder would over one chunk; the device sums them in another
// order (tests compare with a relative tolerance of a few ulps times sqrt(n)).
// Parity status: pinned by source; the formulas are the reference's one-liners.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace h2o_oracle {

struct MetricSums {
  double count = 0, wcount = 0, sumsqe = 0, logloss = 0;
};

inline double metric_logloss(double err) {  // MathUtils.logloss (:798-801)
  const double MAXLL = -std::log(1e-15);
  return std::fmin(MAXLL, -std::log(1.0 - err));
}

// family: 0 gaussian, 1 bernoulli, 2 multinomial (h2o_tree_oracle.hpp Family). F: K columns of n rows (K = 1 unless multinomial),
// already holding init_f (the Tree_k work columns do); y: response (class index for classification), NaN == missing; w nullable.
inline MetricSums scoreMetrics(int family, int K, const double* const* F, const double* y, const double* w, int64_t n) {
  MetricSums s;
  std::vector<double> ds((size_t)K + 2);
  for (int64_t r = 0; r < n; r++) {
    const double wv = w ? w[r] : 1.0;
    if (std::isnan(y[r])) continue;
    if (wv == 0 || std::isnan(wv)) continue;
    double err;
    if (family == 0) {  // regression: ds[0] = linkInv(f) = f
      const double p = F[0][r];
      if (std::isnan(p)) continue;
      err = (double)(float)y[r] - p;  // yact is a float[]
      s.sumsqe += wv * err * err;
    } else {
      const int iact = (int)(float)y[r];
      if (family == 1) {  // bernoulli: ds[2] = logitInv(f), ds[1] = 1 - ds[2]
        if (iact != 0 && iact != 1) continue;
        const double p1 = 1.0 / (std::exp(-F[0][r]) + 1.0);
        ds[2] = p1;
        ds[1] = 1.0 - p1;
        if (std::isnan(p1)) continue;
      } else {  // multinomial: log_rescale + GBM_rescale over ds[1..K]
        double mx = -INFINITY;
        bool nan = false;
        for (int k = 0; k < K; k++) {
          ds[(size_t)k + 1] = F[k][r];
          nan = nan || std::isnan(F[k][r]);
          mx = std::fmax(mx, F[k][r]);
        }
        if (nan) continue;
        double dsum = 0;
        for (int k = 1; k <= K; k++) dsum += (ds[(size_t)k] = std::exp(ds[(size_t)k] - mx));
        for (int k = 1; k <= K; k++) ds[(size_t)k] /= dsum;
      }
      const int ncls = family == 1 ? 2 : K;
      err = iact + 1 < ncls + 1 ? 1.0 - ds[(size_t)iact + 1] : 1.0;
      s.sumsqe += wv * err * err;
      s.logloss += wv * metric_logloss(err);
    }
    s.count += 1;
    s.wcount += wv;
  }
  return s;
}

}  // namespace h2o_oracle
