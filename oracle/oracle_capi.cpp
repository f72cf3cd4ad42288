// TEST INFRASTRUCTURE — C ABI over the CPU oracle (oracle/h2o_tree_oracle.hpp) so that tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg can drive it through
// ctypes. Never linked into, or called from, the product library (h2o-3_b200/).
#include "h2o_tree_oracle.hpp"
#include "h2o_chunk_oracle.hpp"
#include "h2o_compress_oracle.hpp"
#include "h2o_auc_oracle.hpp"
#include "h2o_metrics_oracle.hpp"

using namespace h2o_oracle;

extern "C" {

// ---------------------------------------------------------------- DHistogram
void* orc_hist_new(int nbins, int nbins_cats, int isInt, double min, double maxEx, int intOpt, int initNA,
                   double minSplitImprovement, long long seed, int checkFloatSplits) {
  try {
    return new DHistogram(nbins, nbins_cats, (uint8_t)isInt, min, maxEx, intOpt != 0, initNA != 0, minSplitImprovement,
                          seed, checkFloatSplits != 0);
  } catch (...) {
    return nullptr;
  }
}
// histogram types other than UniformAdaptive: histoType (HistoType), and for QuantilesGlobal the global split points of the
// column / for UniformRobust the custom split points of the node (T/DHistogram.java:343-430); applied by orc_hist_init
void orc_hist_set_type(void* h, int histoType, const double* pts, int npts, int canRefine) {
  DHistogram* d = (DHistogram*)h;
  d->_histoType = histoType;
  d->_globalCanRefine = canRefine != 0;
  std::vector<double> v(pts ? pts : nullptr, pts ? pts + npts : nullptr);
  if (histoType == HT_QuantilesGlobal) d->_globalSplitPts = v;
  if (histoType == HT_UniformRobust) d->_customSplitPoints = v;
}
int orc_hist_split_points(void* h, double* out) {  // _splitPts after init (canonical zero included); returns the count
  const DHistogram* d = (const DHistogram*)h;
  if (out) std::copy(d->_splitPts.begin(), d->_splitPts.end(), out);
  return (int)d->_splitPts.size();
}
int orc_hist_absolute(void* h) { return ((DHistogram*)h)->_absoluteSplitPts ? 1 : 0; }
int orc_hist_type(void* h) { return ((DHistogram*)h)->_histoType; }
int orc_limit_to_range(const double* pts, int n, double min, double maxEx, double* out) {
  const std::vector<double> r = limitToRange(std::vector<double>(pts, pts + n), min, maxEx);
  std::copy(r.begin(), r.end(), out);
  return (int)r.size();
}
int orc_make_unique_and_limit(const double* pts, int n, double min, double maxEx, double* out) {
  const std::vector<double> r = makeUniqueAndLimitToRange(std::vector<double>(pts, pts + n), min, maxEx);
  std::copy(r.begin(), r.end(), out);
  return (int)r.size();
}
int orc_pad_uniformly(const double* pts, int n, int newLength, double* out) {
  const std::vector<double> r = padUniformly(std::vector<double>(pts, pts + n), newLength);
  std::copy(r.begin(), r.end(), out);
  return (int)r.size();
}
void orc_hist_free(void* h) { delete (DHistogram*)h; }
void orc_hist_init(void* h) { ((DHistogram*)h)->init(); }
int orc_hist_nbin(void* h) { return ((DHistogram*)h)->_nbin; }
double orc_hist_step(void* h) { return ((DHistogram*)h)->_step; }
double orc_hist_min2(void* h) { return ((DHistogram*)h)->_min2; }
double orc_hist_maxIn(void* h) { return ((DHistogram*)h)->_maxIn; }
int orc_hist_bin(void* h, double x) { return ((DHistogram*)h)->bin(x); }
void orc_hist_bin_many(void* h, const double* x, long long n, int* out) {
  DHistogram* d = (DHistogram*)h;
  for (long long i = 0; i < n; i++) out[i] = d->bin(x[i]);
}
double orc_hist_binAt(void* h, int b) { return ((DHistogram*)h)->binAt(b); }
int orc_hist_nonEmptyBins(void* h) { return ((DHistogram*)h)->nonEmptyBins(); }
void orc_hist_update(void* h, const double* ws, const double* cs, const double* ys, const int* rows, int hi, int lo) {
  ((DHistogram*)h)->updateHisto(ws, cs, ys, rows, hi, lo);
}
void orc_hist_update_int(void* h, const double* ws, const int* cs, const double* ys, const int* rows, int hi, int lo) {
  ((DHistogram*)h)->updateHistoInt(ws, cs, ys, rows, hi, lo);
}
void orc_hist_reduce_precision(void* h) { ((DHistogram*)h)->reducePrecision(); }
int orc_hist_vals(void* h, double* out) {
  DHistogram* d = (DHistogram*)h;
  if (out) std::copy(d->_vals.begin(), d->_vals.end(), out);
  return (int)d->_vals.size();
}
void orc_hist_set_vals(void* h, const double* in, int n) {
  DHistogram* d = (DHistogram*)h;
  if (!d->_has_vals) d->init();
  for (int i = 0; i < n && i < (int)d->_vals.size(); i++) d->_vals[i] = in[i];
}
double orc_find_maxEx(double maxIn, int isInt) { return DHistogram::find_maxEx(maxIn, isInt); }

struct OrcSplit {
  int col, bin, nasplit, equal;
  double se, se0, se1, n0, n1, p0, p1, tree_p0, tree_p1;
  int bs_nbits, bs_bitoff, bs_nbytes;
  unsigned char bs[132];
};
static void fill_split(const Split& s, OrcSplit* out) {
  out->col = s._col;
  out->bin = s._bin;
  out->nasplit = s._nasplit;
  out->equal = s._equal;
  out->se = s._se;
  out->se0 = s._se0;
  out->se1 = s._se1;
  out->n0 = s._n0;
  out->n1 = s._n1;
  out->p0 = s._p0;
  out->p1 = s._p1;
  out->tree_p0 = s._tree_p0;
  out->tree_p1 = s._tree_p1;
  out->bs_nbits = out->bs_bitoff = out->bs_nbytes = 0;
  memset(out->bs, 0, sizeof(out->bs));
  if (s._bs) {
    out->bs_nbits = s._bs->_nbits;
    out->bs_bitoff = s._bs->_bitoff;
    out->bs_nbytes = (int)s._bs->_val.size();
    for (int i = 0; i < out->bs_nbytes && i < 132; i++) out->bs[i] = s._bs->_val[i];
  }
}
int orc_find_split(void* h, int col, double min_rows, OrcSplit* out) {
  auto s = findBestSplitPoint(*(DHistogram*)h, col, min_rows);
  if (!s) return 0;
  fill_split(*s, out);
  return 1;
}
float orc_splat_numeric(void* h, int bin, int nasplit) {
  Split s;
  s._col = 0;
  s._bin = bin;
  s._nasplit = nasplit;
  s._equal = 0;
  return s.splatNumeric(*(DHistogram*)h);
}

// ---------------------------------------------------------------- RNG / sampler
void orc_pcg_floats(long long seed, int n, float* out) {
  PCGRNG r(seed);
  for (int i = 0; i < n; i++) out[i] = r.nextFloat();
}
void orc_pcg_ints(long long seed, int bound, int n, int* out) {
  PCGRNG r(seed);
  for (int i = 0; i < n; i++) out[i] = r.nextInt(bound);
}
long long orc_pcg_next_long(long long seed, int nth) {
  PCGRNG r(seed);
  long long v = 0;
  for (int i = 0; i <= nth; i++) v = r.nextLong();
  return v;
}
// UndecidedNode.scoreCols (T/DTree.java:462-499) for n_nodes successive nodes of one tree with the same present-mask
void orc_tree_score_cols(int ncols, int mtry, int mtry_per_tree, long long seed, double change_per_level, int depth,
                         const unsigned char* present, int n_nodes, int* out, int* out_null) {
  Params p;
  p.col_sample_rate_change_per_level = change_per_level;
  DTree t(ncols, mtry, mtry_per_tree, seed, &p);
  t._depth = depth;
  for (int i = 0; i < n_nodes; i++) {
    std::vector<HistPtr> hs(ncols);
    for (int c = 0; c < ncols; c++)
      if (present[c]) hs[c] = std::make_shared<DHistogram>(2, 2, 0, 0.0, 1.0, false, false, 0.0, 0, false);
    int nid = t.newUndecided(NO_PARENT, hs);
    const Node& n = t.node(nid);
    out_null[i] = n._scoreColsNull ? 1 : 0;
    for (int c = 0; c < ncols; c++) out[(size_t)i * ncols + c] = c < (int)n._scoreCols.size() ? n._scoreCols[c] : -1;
  }
}
// T/Sample.java:22-36 on rows [row0,row0+n): out[i]=1 if the row is sampled OUT.
void orc_sample(long long tree_seed, double rate, long long row0, long long n, unsigned char* out) {
  PCGRNG rand(tree_seed);
  for (long long i = 0; i < n; i++) {
    rand.setSeed((int64_t)((uint64_t)tree_seed + (uint64_t)(row0 + i)));
    out[i] = rand.nextFloat() >= rate ? 1 : 0;
  }
}
// TT/DHistogramTest.java:173-206 input generator (java.util.Random(42))
void orc_java_random_intopt_inputs(long long seed, int n, double* ys, int* dataInt) {
  JavaRandom r(seed);
  for (int i = 0; i < n; i++) {
    ys[i] = r.nextGaussian();
    dataInt[i] = 13 + r.nextInt(900);
  }
}
int orc_equals_within_one_small_ulp_f(float a, float b) { return equalsWithinOneSmallUlp(a, b) ? 1 : 0; }

// ---------------------------------------------------------------- Frame / GBM
void* orc_frame_new(long long nrows, int ncols) {
  Frame* f = new Frame();
  f->cols.resize(ncols);
  f->y.assign(nrows, 0.0);
  return f;
}
void orc_frame_free(void* f) { delete (Frame*)f; }
void orc_frame_set_col(void* f, int c, const double* v, int isCat, int cardinality) {
  Frame* fr = (Frame*)f;
  Column& col = fr->cols[c];
  col.v.assign(v, v + fr->y.size());
  col.isCat = isCat != 0;
  col.cardinality = cardinality;
  col.rollup();
}
void orc_frame_set_y(void* f, const double* y) { Frame* fr = (Frame*)f; std::copy(y, y + fr->y.size(), fr->y.begin()); }
void orc_frame_set_w(void* f, const double* w) { Frame* fr = (Frame*)f; fr->w.assign(w, w + fr->y.size()); }
void orc_frame_set_off(void* f, const double* o) { Frame* fr = (Frame*)f; fr->off.assign(o, o + fr->y.size()); }

struct OrcParams {
  int ntrees, max_depth, nbins, nbins_cats, nbins_top_level, family, nclass, nthreads;
  double min_rows, min_split_improvement, sample_rate, col_sample_rate, col_sample_rate_per_tree,
      col_sample_rate_change_per_level, learn_rate, learn_rate_annealing, max_abs_leafnode_pred;
  long long seed;
  long long chunk_rows;
  double tweedie_power;
  long long histogram_type;  // HistoType
  double pred_noise_bandwidth;
  long long n_rate_per_class;          // 0: sample_rate_per_class == null
  double sample_rate_per_class[16];
};
static Params to_params(const OrcParams* p) {
  Params q;
  q.ntrees = p->ntrees;
  q.max_depth = p->max_depth;
  q.nbins = p->nbins;
  q.nbins_cats = p->nbins_cats;
  q.nbins_top_level = p->nbins_top_level;
  q.family = p->family;
  q.nclass = p->nclass;
  q.nthreads = p->nthreads;
  q.min_rows = p->min_rows;
  q.min_split_improvement = p->min_split_improvement;
  q.sample_rate = p->sample_rate;
  q.col_sample_rate = p->col_sample_rate;
  q.col_sample_rate_per_tree = p->col_sample_rate_per_tree;
  q.col_sample_rate_change_per_level = p->col_sample_rate_change_per_level;
  q.learn_rate = p->learn_rate;
  q.learn_rate_annealing = p->learn_rate_annealing;
  q.max_abs_leafnode_pred = p->max_abs_leafnode_pred;
  q.seed = p->seed;
  if (p->tweedie_power > 0) q.tweedie_power = p->tweedie_power;
  q.histogram_type = (int)p->histogram_type;
  q.pred_noise_bandwidth = p->pred_noise_bandwidth;
  q.sample_rate_per_class.assign(p->sample_rate_per_class, p->sample_rate_per_class + std::min<long long>(16, std::max<long long>(0, p->n_rate_per_class)));
  return q;
}

void* orc_gbm_train(void* f, const OrcParams* p, int capture) {
  try {
    Params q = to_params(p);
    GBM g(*(Frame*)f, q);
    g.capture = capture != 0;
    if (p->chunk_rows > 0) g.chunk_rows_ = p->chunk_rows;
    g.train();
    return new GBMModel(std::move(g.model));
  } catch (const std::exception& e) {
    fprintf(stderr, "orc_gbm_train: %s\n", e.what());
    return nullptr;
  }
}
// DRF (T/drf/DRF.java): family BERNOULLI == binomial (one tree, binomialOpt), MULTINOMIAL == nclass trees, GAUSSIAN == regression
void* orc_drf_train(void* f, const OrcParams* p, int mtries, int capture) {
  try {
    Params q = to_params(p);
    q.drf = true;
    q.mtries = mtries;
    DRF g(*(Frame*)f, q);
    g.capture = capture != 0;
    if (p->chunk_rows > 0) g.chunk_rows_ = p->chunk_rows;
    g.trainDRF();
    return new GBMModel(std::move(g.model));
  } catch (const std::exception& e) {
    fprintf(stderr, "orc_drf_train: %s\n", e.what());
    return nullptr;
  }
}
void orc_model_oobt(void* m, double* out) {
  const std::vector<double>& o = ((GBMModel*)m)->oobt;
  std::copy(o.begin(), o.end(), out);
}
void orc_model_free(void* m) { delete (GBMModel*)m; }
double orc_model_init_f(void* m) { return ((GBMModel*)m)->init_f; }
int orc_model_ntrees(void* m) { return (int)((GBMModel*)m)->trees.size(); }
int orc_model_K(void* m) { return (int)((GBMModel*)m)->F.size(); }
void orc_model_F(void* m, int k, double* out) {
  auto& F = ((GBMModel*)m)->F[k];
  std::copy(F.begin(), F.end(), out);
}
void orc_model_timing(void* m, double* t_hist, double* t_total, double* rowcols) {
  GBMModel* g = (GBMModel*)m;
  *t_hist = g->t_hist_s;
  *t_total = g->t_total_s;
  *rowcols = g->hist_rowcols;
}
void orc_model_varimp(void* m, float* out) {
  auto& v = ((GBMModel*)m)->varimp;
  std::copy(v.begin(), v.end(), out);
}

struct OrcNode {
  int kind, pid, nid, nid0, nid1;
  float splat, pred;
  int has_split;
  OrcSplit split;
};
int orc_model_tree_len(void* m, int t, int k) { return (int)((GBMModel*)m)->trees[t][k].nodes.size(); }
int orc_model_tree_leaves_start(void* m, int t, int k) { return ((GBMModel*)m)->trees[t][k].leaves_start; }
void orc_model_tree_nodes(void* m, int t, int k, OrcNode* out) {
  const TreeOut& tr = ((GBMModel*)m)->trees[t][k];
  for (size_t i = 0; i < tr.nodes.size(); i++) {
    const Node& n = tr.nodes[i];
    OrcNode& o = out[i];
    memset(&o, 0, sizeof(o));
    o.kind = n.kind;
    o.pid = n._pid;
    o.nid = n._nid;
    o.nid0 = n._nids[0];
    o.nid1 = n._nids[1];
    o.splat = n._splat;
    o.pred = n._pred;
    o.has_split = n._split ? 1 : 0;
    if (n._split) fill_split(*n._split, &o.split);
  }
}
float orc_model_score_tree(void* m, int t, int k, const double* x) {
  const TreeOut& tr = ((GBMModel*)m)->trees[t][k];
  return scoreTree(tr, [&](int c) { return x[c]; });
}

// DTree.compress of tree (t, k): returns the number of bytes; copies them when out != nullptr (aux: the _abAux stream)
long long orc_model_tree_compress(void* m, int t, int k, int aux, unsigned char* out) {
  try {
    const CompressedTreeBytes c = compressTree(((GBMModel*)m)->trees[t][k]);
    const std::vector<uint8_t>& v = aux ? c.aux : c.bits;
    if (out) memcpy(out, v.data(), v.size());
    return (long long)v.size();
  } catch (...) {
    return -1;
  }
}
// SharedTreeMojoModel.scoreTree over compressed bytes, for n rows of a row-major [n][ncols] matrix
void orc_score_compressed(const unsigned char* bits, long long len, const double* X, long long n, int ncols, float* out) {
  const std::vector<uint8_t> v(bits, bits + len);
  for (long long r = 0; r < n; r++) out[r] = scoreCompressedTree(v, [&](int c) { return X[r * ncols + c]; });
}
// AUC2: one AUCBuilder per chunk of chunk_rows consecutive rows, folded left to right; returns the AUC and the builder's bins
double orc_auc2(const double* pred, const double* act, const double* w, long long n, long long chunk_rows, int* out_nbins, double* out_ths,
                double* out_tps, double* out_fps) {
  const AUCBuilder b = aucOverChunks(pred, act, w, n, chunk_rows);
  if (out_nbins) *out_nbins = b._n;
  for (int i = 0; i < b._n; i++) {
    if (out_ths) out_ths[i] = b._ths[i];
    if (out_tps) out_tps[i] = b._tps[i];
    if (out_fps) out_fps[i] = b._fps[i];
  }
  return b.auc();
}

// hex.tree.Score's sums over the model's Tree_k columns (F: K pointers to n doubles): out = {count, wcount, sumsqe, logloss}
void orc_score_metrics(int family, int K, const double* const* F, const double* y, const double* w, long long n, double* out) {
  const MetricSums s = scoreMetrics(family, K, F, y, w, n);
  out[0] = s.count;
  out[1] = s.wcount;
  out[2] = s.sumsqe;
  out[3] = s.logloss;
}

// A compressed tree written elsewhere (the reference's MOJO fixtures) decoded into a node table and serialised again.
// Returns 0, or -1 with the reason in `err`; `reserved` (one int per decided node, nullable) receives the auxiliary records' second word.
int orc_recompress_tree(const unsigned char* bits, long long nbits, const unsigned char* aux, long long naux, unsigned char* out_bits,
                        long long* out_nbits, unsigned char* out_aux, long long* out_naux, int* reserved, int* kinds, char* err, int err_cap) {
  try {
    std::vector<uint8_t> b(bits, bits + nbits), a(aux, aux + naux);
    std::vector<int> res;
    const TreeOut t = decompressTree(b, a, &res);
    const CompressedTreeBytes c = compressTree(t);
    if ((long long)c.bits.size() > *out_nbits || (long long)c.aux.size() > *out_naux) throw std::runtime_error("output buffers too small");
    memcpy(out_bits, c.bits.data(), c.bits.size());
    memcpy(out_aux, c.aux.data(), c.aux.size());
    *out_nbits = (long long)c.bits.size();
    *out_naux = (long long)c.aux.size();
    if (reserved)
      for (size_t i = 0; i < res.size(); i++) reserved[i] = res[i];
    if (kinds) {  // decided nodes by split kind: numeric, 32-bit bitset, long bitset, NA vs rest
      kinds[0] = kinds[1] = kinds[2] = kinds[3] = 0;
      for (const Node& n : t.nodes)
        if (n.kind == DECIDED) kinds[n._split->_nasplit == NAvsREST ? 3 : (n._split->_equal <= 1 ? 0 : n._split->_equal - 1)]++;
    }
    return 0;
  } catch (const std::exception& e) {
    if (err && err_cap > 0) snprintf(err, (size_t)err_cap, "%s", e.what());
    return -1;
  }
}

// One AUCBuilder of nbins bins fed row by row (AUCBuilderTest drives the class this way); returns _n, fills the first _n bins.
int orc_auc_builder(int nbins, const double* pred, const int* act, const double* w, long long n, double* out_ths, double* out_tps, double* out_fps,
                    double* out_sqe) {
  AUCBuilder b(nbins);
  for (long long i = 0; i < n; i++) b.perRow(pred[i], act[i], w ? w[i] : 1.0);
  for (int i = 0; i < b._n; i++) {
    if (out_ths) out_ths[i] = b._ths[i];
    if (out_tps) out_tps[i] = b._tps[i];
    if (out_fps) out_fps[i] = b._fps[i];
    if (out_sqe) out_sqe[i] = b._sqe[i];
  }
  return b._n;
}

// AUCBuilderTest.testPerRow_compat: the fast path of perRow against insert + mergeOneBin, compared bin for bin after every row.
// Returns the first row after which the two builders differ, or -1.
long long orc_auc_builder_compat(int nbins, const double* pred, const int* act, long long n) {
  AUCBuilder fast(nbins), slow(nbins);
  slow._useFastPath = false;
  for (long long i = 0; i < n; i++) {
    fast.perRow(pred[i], act[i], 1.0);
    slow.perRow(pred[i], act[i], 1.0);
    if (fast._n != slow._n) return i;
    for (int j = 0; j < nbins; j++)
      if (fast._ths[j] != slow._ths[j] || fast._tps[j] != slow._tps[j] || fast._fps[j] != slow._fps[j] || fast._sqe[j] != slow._sqe[j]) return i;
  }
  return -1;
}

// ---- level captures (histogram-monitor analogue)
int orc_model_ncaptures(void* m) { return (int)((GBMModel*)m)->captures.size(); }
void orc_capture_info(void* m, int i, int* tree, int* k, int* leaf, int* nleaves) {
  const LevelCapture& c = ((GBMModel*)m)->captures[i];
  *tree = c.tree;
  *k = c.k;
  *leaf = c.leaf;
  *nleaves = c.nleaves;
}
struct OrcHistDesc {
  int present, has_vals, isInt, intOpt, nbin;
  double min, maxEx, step, min2, maxIn;
};
void orc_capture_desc(void* m, int i, int n, int c, int ncols, OrcHistDesc* out) {
  const LevelCapture::Entry& e = ((GBMModel*)m)->captures[i].entries[(size_t)n * ncols + c];
  out->present = e.present;
  out->has_vals = e.has_vals;
  out->isInt = e.isInt;
  out->intOpt = e.intOpt;
  out->nbin = e.nbin;
  out->min = e.min;
  out->maxEx = e.maxEx;
  out->step = e.step;
  out->min2 = e.min2;
  out->maxIn = e.maxIn;
}
int orc_capture_vals(void* m, int i, int n, int c, int ncols, double* out) {
  const LevelCapture::Entry& e = ((GBMModel*)m)->captures[i].entries[(size_t)n * ncols + c];
  if (out) std::copy(e.vals.begin(), e.vals.end(), out);
  return (int)e.vals.size();
}
void orc_capture_nids(void* m, int i, int* out) {
  const auto& v = ((GBMModel*)m)->captures[i].nids_after;
  std::copy(v.begin(), v.end(), out);
}

// ---- chunk codecs (h2o_chunk_oracle.hpp)
static int64_t copy_out(const std::vector<uint8_t>& m, uint8_t* out, int64_t cap) {
  if ((int64_t)m.size() > cap) return -(int64_t)m.size();
  if (!m.empty()) std::memcpy(out, m.data(), m.size());
  return (int64_t)m.size();
}
int64_t orc_chunk_rows(int codec, const uint8_t* mem, int64_t len) { return orc_chunk::rows_of(codec, mem, len); }
// Chunk.getDoubles: every row through atd_impl
void orc_chunk_decode(int codec, const uint8_t* mem, int64_t len, int64_t n, double* out) {
  for (int64_t i = 0; i < n; i++) out[i] = orc_chunk::atd(codec, mem, len, i);
}
int64_t orc_chunk_enc_int(int log, int64_t n, const int64_t* v, const uint8_t* na, uint8_t* out, int64_t cap) {
  return copy_out(orc_chunk::enc_int(log, std::vector<int64_t>(v, v + n), std::vector<uint8_t>(na, na + n)), out, cap);
}
int64_t orc_chunk_enc_scaled(int log, int64_t n, const int64_t* man, const int32_t* exp, const uint8_t* na, uint8_t* out, int64_t cap) {
  return copy_out(orc_chunk::enc_scaled(log, std::vector<int64_t>(man, man + n), std::vector<int32_t>(exp, exp + n),
                                        std::vector<uint8_t>(na, na + n)), out, cap);
}
int64_t orc_chunk_enc_c0(int is_long, double dcon, int64_t lcon, int32_t len, uint8_t* out, int64_t cap) {
  return copy_out(orc_chunk::enc_c0(is_long != 0, dcon, lcon, len), out, cap);
}
int64_t orc_chunk_enc_cbs(int64_t n, const uint8_t* v, const uint8_t* na, int bpv, uint8_t* out, int64_t cap) {
  return copy_out(orc_chunk::enc_cbs(std::vector<uint8_t>(v, v + n), std::vector<uint8_t>(na, na + n), bpv), out, cap);
}
int64_t orc_chunk_enc_cud(int64_t n, const double* v, uint8_t* out, int64_t cap) {
  return copy_out(orc_chunk::enc_cud(std::vector<double>(v, v + n)), out, cap);
}
int64_t orc_chunk_enc_cxi(int32_t len, int id_sz, int val_sz, int na_sparse, int64_t nnz, const int32_t* ids, const int64_t* vals,
                          const uint8_t* na, uint8_t* out, int64_t cap) {
  return copy_out(orc_chunk::enc_cxi(len, id_sz, val_sz, na_sparse != 0, std::vector<int32_t>(ids, ids + nnz),
                                     std::vector<int64_t>(vals, vals + nnz), std::vector<uint8_t>(na, na + nnz)), out, cap);
}
int64_t orc_chunk_enc_cxf(int32_t len, int valsz, int na_sparse, int64_t nnz, const int32_t* ids, const double* vals, uint8_t* out,
                          int64_t cap) {
  return copy_out(orc_chunk::enc_cxf(len, valsz, na_sparse != 0, std::vector<int32_t>(ids, ids + nnz), std::vector<double>(vals, vals + nnz)),
                  out, cap);
}
double orc_pow10(int64_t m, int e) { return orc_chunk::pow10(m, e); }

}  // extern "C"
