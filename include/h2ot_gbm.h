/*
 * h2ot_gbm.h — C ABI of the host-side GBM driver shipped in libh2otree_b200.so.
 *
 * In an H2O-3 deployment this layer IS the unchanged Java code (hex.tree.SharedTree / gbm.GBM / DTree /
 * DHistogram): it keeps the trees, builds the per-level DHistogram descriptors and sequences the levels,
 * calling the engine (include/h2ot.h) where ScoreBuildHistogram2 and the GBM row passes used to run.
 * No JDK exists in this build image, so the same sequencing is written in C++ (h2o-3_b200/csrc/host_gbm.cpp)
 * strictly on top of the public C ABI of h2ot.h — it is what the tests, the benchmark and non-JVM users call.
 * Parameter names and meanings are those of GBMV3/SharedTreeV3
 * (h2o-algos/src/main/java/hex/schemas/GBMV3.java:13-75, hex/tree/SharedTreeModel.java:51-98, hex/tree/gbm/GBMModel.java:24-41).
 */
#ifndef H2OT_GBM_H
#define H2OT_GBM_H
#include "h2ot.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct h2ot_gbm h2ot_gbm;

#define H2OT_ALGO_GBM 0
#define H2OT_ALGO_DRF 1 /* hex.tree.drf.DRF: family BERNOULLI == binomial (one tree, binomialOpt), MULTINOMIAL == nclass trees, GAUSSIAN == regression */

typedef struct h2ot_gbm_params {
  int32_t ntrees, max_depth, nbins, nbins_cats, nbins_top_level, family, nclass, algo;
  int32_t mtries; /* DRF: -1 sqrt(ncols) / ncols/3, -2 all columns, else the count (T/drf/DRF.java:117-125) */
  int32_t histogram_type; /* SharedTreeParameters._histogram_type: 0 AUTO / UniformAdaptive, 1 Random (per-node random split points,
                             T/DHistogram.java:347-352, 742-749); QuantilesGlobal / UniformRobust need split points from outside
                             this driver (h2ot_set_split_points serves them at the engine level) */
  double min_rows, min_split_improvement, sample_rate, col_sample_rate, col_sample_rate_per_tree,
      col_sample_rate_change_per_level, learn_rate, learn_rate_annealing, max_abs_leafnode_pred;
  int64_t seed;
  double tweedie_power; /* family H2OT_FAMILY_TWEEDIE (default 1.5 when <= 0) */
} h2ot_gbm_params;

/* Flat view of one DTree node (DecidedNode / LeafNode, T/DTree.java:587-965) */
typedef struct h2ot_tree_node {
  int32_t kind; /* 0 undecided (never in a finished tree), 1 decided, 2 leaf */
  int32_t pid, nid, nid0, nid1;
  float splat, pred;
  int32_t has_split;
  int32_t col, bin, nasplit, equal;
  double se, se0, se1, n0, n1, p0, p1, tree_p0, tree_p1;
  int32_t bs_nbits, bs_bitoff, bs_nbytes;
  uint8_t bs[H2OT_MAX_BITSET_BYTES];
} h2ot_tree_node;

typedef struct h2ot_gbm_timing {
  double total_s;        /* wall seconds spent in h2ot_gbm_train */
  double level_s;        /* wall seconds inside the per-level engine calls (score + histogram + finalize + split search) */
  double hist_rows_cols; /* sum over level passes of global rows x active columns */
  int64_t levels;        /* number of level passes */
} h2ot_gbm_timing;

int h2ot_gbm_create(const h2ot_gbm_params* p, int32_t device, int64_t nrows, int32_t ncols, int64_t row_offset,
                    int64_t nrows_global, int32_t has_weights, int32_t has_offset, h2ot_gbm** out);
int h2ot_gbm_destroy(h2ot_gbm* g);
/* one call per column: full local column, NaN == NA; categorical columns carry integer level codes */
int h2ot_gbm_set_column(h2ot_gbm* g, int32_t col, const double* values, int32_t is_categorical, int32_t cardinality);
/* The same column handed over as H2O chunks (include/h2ot.h: h2ot_load_chunk): one call per Chunk of the Vec, row0 = the chunk's
 * first row (Chunk._start). col may be H2OT_COL_RESPONSE / H2OT_COL_WEIGHTS / H2OT_COL_OFFSET. Call h2ot_gbm_chunks_done(col, ...)
 * once all chunks of a column are in. */
int h2ot_gbm_load_chunk(h2ot_gbm* g, int32_t col, int64_t row0, int64_t n, int32_t codec, const uint8_t* mem, int64_t mem_len);
int h2ot_gbm_chunks_done(h2ot_gbm* g, int32_t col, int32_t is_categorical, int32_t cardinality);
int h2ot_gbm_set_response(h2ot_gbm* g, const double* y);
int h2ot_gbm_set_weights(h2ot_gbm* g, const double* w);
int h2ot_gbm_set_offset(h2ot_gbm* g, const double* o);
int h2ot_gbm_comm_init(h2ot_gbm* g, int32_t world_size, int32_t rank, const uint8_t id[H2OT_COMM_ID_BYTES]);
/* Frame finalisation, rollups, init_f, MarkDecidedRows (GBM.initializeModelSpecifics, T/gbm/GBM.java:254-312) */
int h2ot_gbm_init(h2ot_gbm* g);
/* Builds the next n iterations (GBM.buildNextKTrees, T/gbm/GBM.java:464-591).
 * bernoulli / multinomial GBMs whose levels run on the device (h2ot.h: h2ot_tree_levels_*) are built in a PIPELINE: the call may
 * return while the last iteration is still on the device (its trees built, fitted and added to the predictions there) and its
 * host mirror not yet made; every accessor below waits for it by itself, h2ot_gbm_flush does so explicitly (e.g. to stop a
 * clock). H2OT_PIPELINE=0 in the environment builds the same trees one synchronisation at a time. */
int h2ot_gbm_train(h2ot_gbm* g, int32_t n_iterations);
int h2ot_gbm_flush(h2ot_gbm* g);
int h2ot_gbm_capture_levels(h2ot_gbm* g, int32_t on); /* histogram monitor analogue (T/SharedTree.java:1253-1289) */

double h2ot_gbm_init_f(h2ot_gbm* g);
int32_t h2ot_gbm_ntrees(h2ot_gbm* g);
int32_t h2ot_gbm_tree_len(h2ot_gbm* g, int32_t tree, int32_t k);
int32_t h2ot_gbm_tree_leaves_start(h2ot_gbm* g, int32_t tree, int32_t k);
int h2ot_gbm_tree_nodes(h2ot_gbm* g, int32_t tree, int32_t k, h2ot_tree_node* out);
/* DTree.compress (h2o-algos/src/main/java/hex/tree/DTree.java:845-933, 972-982; IcedBitSet.compress2/3,
 * h2o-core/src/main/java/water/util/IcedBitSet.java:92-111): CompressedTree._bits of tree (tree, k) -- the byte stream
 * SharedTreeMojoModel.scoreTree and the MOJO writer consume -- and, when aux != NULL, the auxiliary stream (DTree._abAux: per
 * decided node, in the same pre-order, {nid, decided nodes in the left subtree, n0, n1, p0, p1, se0, se1, nid0, nid1}, 40 bytes).
 * Little-endian like AutoBuffer on x86. Call once with bits == NULL to get the lengths, then with buffers of at least that size. */
int h2ot_gbm_tree_compress(h2ot_gbm* g, int32_t tree, int32_t k, uint8_t* bits, int64_t bits_cap, int64_t* bits_len, uint8_t* aux,
                           int64_t aux_cap, int64_t* aux_len);
int h2ot_gbm_predictions(h2ot_gbm* g, int32_t k, double* out);
int h2ot_gbm_oobt(h2ot_gbm* g, double* out); /* DRF: OUT_BAG_TREES column */
/* Scores NEW rows with the trees built so far (SharedTreeMojoModel.scoreAllTrees, h2o-genmodel .../SharedTreeMojoModel.java:1043-1076):
 * columns[c] -> nrows doubles (NaN == NA), same column order / categorical coding as the training frame. out[k * nrows + r] =
 * sum over trees of tree_k(row r) in double, in build order, plus init_f for GBM (GbmMojoModel.score0 adds it after the sum); for DRF
 * the plain sum of the trees' votes (DrfMojoModel.score0 divides by the number of trees). Runs on the model's device with a
 * temporary engine (h2ot_score_tree); link-space, no probabilities. */
int h2ot_gbm_predict(h2ot_gbm* g, int64_t nrows, const double* const* columns, double* out);
/* Training metrics of the current GBM model over all ranks (h2ot_score_metrics + sum): out_mse, out_logloss nullable */
int h2ot_gbm_score(h2ot_gbm* g, double* out_mse, double* out_logloss);
/* AUC of the current bernoulli GBM on its training rows, computed like hex.AUC2 (h2o-core/src/main/java/hex/AUC2.java): the device
 * builds one AUCBuilder per chunk of rows (h2ot_auc_builders), this call folds them in row order -- and across ranks in rank
 * order -- with AUCBuilder.reduce and evaluates AUC2.compute_auc. out_ths / out_tps / out_fps (nullable, 400 doubles each) receive
 * the final builder's bins, *out_nbins their number, *out_chunk_rows the rows per chunk the device used. */
int h2ot_gbm_auc(h2ot_gbm* g, double* out_auc, int32_t* out_nbins, double* out_ths, double* out_tps, double* out_fps, int64_t* out_chunk_rows);
/* Two model parameters that are not in h2ot_gbm_params (set after h2ot_gbm_create, before h2ot_gbm_train):
 * GBMParameters._pred_noise_bandwidth (h2o-algos/src/main/java/hex/tree/gbm/GBMModel.java:28; AddTreeContributions multiplies a
 * leaf's value by 1 + nextGaussian * bandwidth, gbm/GBM.java:1450-1463) and SharedTreeParameters._sample_rate_per_class
 * (hex/tree/SharedTreeModel.java:59; Sample.java:14-30: one rate per response class, n == number of classes; n == 0: null). */
int h2ot_gbm_set_pred_noise_bandwidth(h2ot_gbm* g, double bandwidth);
int h2ot_gbm_set_sample_rate_per_class(h2ot_gbm* g, const double* rates, int32_t n);
int h2ot_gbm_varimp(h2ot_gbm* g, float* out);
int h2ot_gbm_timing_get(h2ot_gbm* g, h2ot_gbm_timing* out);
h2ot_engine* h2ot_gbm_engine(h2ot_gbm* g);

/* host-logic probes for the CPU test-suite (no device involved): the DHistogram ctor rule (T/DHistogram.java:190-269; returns 0
 * for StepOutOfRangeException), find_maxEx (:453-459), the PCG generator (water/util/RandomUtils.java:50-119) and the per-node
 * column draws of UndecidedNode.scoreCols (T/DTree.java:462-499) exactly as this driver computes them */
int h2ot_hostcheck_hist(int32_t nbins, int32_t nbins_cats, int32_t is_int, double min, double max_ex, int32_t* out_nbin, double* out_step);
double h2ot_hostcheck_find_max_ex(double max_in, int32_t is_int);
int64_t h2ot_hostcheck_pcg_next_long(int64_t seed, int32_t nth);
void h2ot_hostcheck_pcg_bounded(int64_t seed, int32_t bound, int32_t n, int32_t* out);
void h2ot_hostcheck_score_cols(int32_t ncols, int32_t mtry, int32_t mtry_per_tree, int64_t seed, double change_per_level, int32_t depth,
                               const uint8_t* present, int32_t n_nodes, int32_t* out, int32_t* out_null);

/* captured levels */
int32_t h2ot_gbm_ncaptures(h2ot_gbm* g);
int h2ot_gbm_capture_info(h2ot_gbm* g, int32_t i, int32_t* tree, int32_t* k, int32_t* leaf, int32_t* nleaves);
int h2ot_gbm_capture_desc(h2ot_gbm* g, int32_t i, int32_t n, int32_t c, h2ot_hist_desc* desc, double* min2, double* maxIn,
                          int32_t* has_vals, int64_t* nvals);
int h2ot_gbm_capture_vals(h2ot_gbm* g, int32_t i, int32_t n, int32_t c, double* out);
int h2ot_gbm_capture_nids(h2ot_gbm* g, int32_t i, int32_t* out);

#ifdef __cplusplus
}
#endif
#endif
