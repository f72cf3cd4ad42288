/*
 * h2ot.h — C ABI of libh2otree_b200.so: the B200-native engine behind H2O-3's hex.tree
 * ScoreBuildHistogram2 seam (per-level score-rows + build-histograms + find-splits, plus the GBM
 * per-tree row passes).
 *
 * The reference has no FFI for this path: the seam is the Java class boundary where
 * SharedTree.ScoreBuildOneTree.compute2 constructs ScoreBuildHistogram2 and calls dfork2
 * (h2o-algos/src/main/java/hex/tree/SharedTree.java:671-683). A Java subclass
 * (ScoreBuildHistogramGPU) binds these entry points over JNI / Panama FFM.
 * Each entry point below cites the reference code it replaces (paths relative to the reference root;
 * T/ = h2o-algos/src/main/java/hex/tree/, W/ = h2o-core/src/main/java/water/).
 *
 * Conventions: plain C types only; every function returns 0 (H2OT_OK) or a negative error code and
 * never throws; h2ot_last_error() returns a thread-local UTF-8 message for the last failure on the
 * calling thread. Handles are opaque; calls on one handle must be serialised by the caller;
 * different handles are independent. All calls block until their device work is complete unless
 * stated otherwise. There is NO CPU fallback: with no CUDA device h2ot_create fails.
 */
#ifndef H2OT_H
#define H2OT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define H2OT_OK 0
#define H2OT_ERR_INVALID -1   /* bad argument / call order */
#define H2OT_ERR_CUDA -2      /* CUDA runtime error (message has the CUDA string) */
#define H2OT_ERR_NOMEM -3
#define H2OT_ERR_UNSUPPORTED -4
#define H2OT_ERR_COMM -5      /* NCCL error */

/* hex.genmodel.utils.DistributionFamily subset handled on device (H/DistributionFactory.java:73-134,243-272) */
#define H2OT_FAMILY_GAUSSIAN 0
#define H2OT_FAMILY_BERNOULLI 1
#define H2OT_FAMILY_MULTINOMIAL 2
#define H2OT_FAMILY_POISSON 3  /* H/DistributionFactory.java:274-307 (log link) */
#define H2OT_FAMILY_GAMMA 4    /* :309-341 */
#define H2OT_FAMILY_TWEEDIE 5  /* :343-379, power in h2ot_config.tweedie_power (1 < p < 2) */
#define H2OT_FAMILY_LAPLACE 6  /* :406-421: negHalfGradient = sign(y - f) / 2; leaves are fitted with h2ot_leaf_quantiles(0.5) */
#define H2OT_FAMILY_QUANTILE 7 /* :423-434, alpha in h2ot_config.tweedie_power; leaves are fitted with h2ot_leaf_quantiles(alpha) */

/* T/DHistogram.java:152-168 NASplitDir */
#define H2OT_NA_NONE 0
#define H2OT_NA_VS_REST 1
#define H2OT_NA_LEFT 2
#define H2OT_NA_RIGHT 3
#define H2OT_LEFT 4
#define H2OT_RIGHT 5

/* T/ScoreBuildHistogram.java:60-67 row markers kept in the NIDs_k column */
#define H2OT_DECIDED_ROW (-1)
#define H2OT_OUT_OF_BAG (-2)

/* column storage chosen by the ingest kernel (lossless, order preserving) */
#define H2OT_STORE_F32 0 /* every non-NA value is fp32-exact */
#define H2OT_STORE_F64 1
#define H2OT_STORE_U8 2  /* integer valued, max-min < 255 : code = x - min, 255 = NA */
#define H2OT_STORE_U16 3 /* integer valued, max-min < 65535 : code = x - min, 65535 = NA */

typedef struct h2ot_engine h2ot_engine;

typedef struct h2ot_config {
  int32_t device;        /* CUDA device ordinal */
  int32_t ncols;         /* predictor columns */
  int64_t nrows;         /* rows held by THIS engine (this rank's shard) */
  int64_t row_offset;    /* global index of local row 0 (T/Sample.java:31 seeds by global row) */
  int64_t nrows_global;  /* rows over all ranks (fixes the histogram fixed-point scale); 0 => nrows */
  int32_t nclass_trees;  /* K: class-trees built per iteration (1 for regression/binomial) */
  int32_t nclass;        /* 1 regression, 2 binomial, >2 multinomial */
  int32_t family;        /* H2OT_FAMILY_* */
  int32_t has_weights;
  int32_t has_offset;
  int32_t reserved;
  double tweedie_power;  /* H2OT_FAMILY_TWEEDIE: Model.Parameters._tweedie_power (default 1.5); H2OT_FAMILY_QUANTILE: _quantile_alpha */
} h2ot_config;

/* Per (leaf, column) histogram descriptor == the DHistogram fields the CPU kernel reads
 * (T/DHistogram.java:56-62, ctor :190-269). Built by the host (Java keeps DHistogram/DTree). */
typedef struct h2ot_hist_desc {
  double min;      /* _min */
  double maxEx;    /* _maxEx */
  double step;     /* _step */
  int32_t nbin;    /* _nbin (NA bucket excluded) */
  int16_t order;   /* position of the column in the leaf's _scoreCols (column index if null): bestCol tie order, T/DTree.java:631-654 */
  uint8_t flags;   /* H2OT_DESC_* */
  uint8_t pad;
  int32_t splitpts_off; /* H2OT_DESC_SPLITPTS: index of _splitPts[0] in the pool given to h2ot_set_split_points */
  int32_t splitpts_n;   /* _splitPts.length (== nbin for Random / QuantilesGlobal, nbin + 1 for UniformRobust's custom points) */
} h2ot_hist_desc;
#define H2OT_DESC_ACTIVE 1      /* DHistogram non-null AND column in the leaf's _scoreCols (T/ScoreBuildHistogram2.java:421-426) */
#define H2OT_DESC_ISINT 2       /* _isInt > 0 */
#define H2OT_DESC_ISCAT 4       /* _isInt == 2 */
#define H2OT_DESC_INTOPT 8      /* _intOpt (only changes the untouched _min2/_maxIn sentinels, T/DHistogram.java:677-703) */
#define H2OT_DESC_CHECKFLOAT 16 /* _checkFloatSplits (DRF, T/drf/DRF.java:201) */
#define H2OT_DESC_INITNA 32     /* _initNA */
#define H2OT_DESC_SPLITPTS 64   /* _splitPts != null: bin by Arrays.binarySearch over the split points (histogram_type Random /
                                   QuantilesGlobal / UniformRobust, T/DHistogram.java:283-286) instead of (int)pos */
#define H2OT_DESC_ABSOLUTE 128  /* _absoluteSplitPts: the points are column values, not positions relative to _min in steps */

/* One DecidedNode as score_decide / GammaPass read it (T/DTree.java:587-763). */
typedef struct h2ot_decided_node {
  int32_t nid;
  int32_t pid;        /* -1 for the root */
  int32_t col;        /* _split._col, -1 when _split == null */
  int32_t equal;      /* _split._equal: 0 range split, 2/3 bitset split */
  int32_t nasplit;    /* H2OT_NA_* */
  float splat;        /* _splat (float; the row test is double >= float, T/DTree.java:738) */
  int32_t child[2];   /* _nids */
  int32_t bs_off;     /* byte offset of the IcedBitSet bytes in the bitset pool (equal >= 2) */
  int32_t bs_nbits;   /* IcedBitSet._nbits */
  int32_t bs_bitoff;  /* IcedBitSet._bitoff */
  int32_t pad;
} h2ot_decided_node;

#define H2OT_MAX_BITSET_BYTES 512 /* nbins_cats <= 4096 */

/* DTree.Split (T/DTree.java:182-209) + _splat, as produced by the device split search. */
typedef struct h2ot_split {
  int32_t col;      /* -1: no split found (DecidedNode._split == null) */
  int32_t bin;
  int32_t nasplit;
  int32_t equal;
  float splat;      /* Split.splat(hs) (T/DTree.java:240-291); NaN when abandoned */
  int32_t bs_nbits, bs_bitoff, bs_nbytes;
  double se, se0, se1, n0, n1, p0, p1, tree_p0, tree_p1;
  uint8_t bs[H2OT_MAX_BITSET_BYTES];
} h2ot_split;

typedef struct h2ot_column_info { /* rollups: W/fvec/RollupStats.java:31-48 */
  double min, max;   /* over non-NA values (NaN if all NA) */
  int64_t nacnt;
  int32_t is_int;    /* every non-NA value integral */
  int32_t storage;   /* H2OT_STORE_* */
} h2ot_column_info;

typedef struct h2ot_stats { /* diagnostics: device time per kernel family (ms) and launch counts since create / last reset */
  double ms_score, ms_hist, ms_finalize, ms_split, ms_residuals, ms_gamma, ms_addtree, ms_sample, ms_ingest;
  int64_t launches;          /* kernels launched by this library */
  int64_t h2d_bytes, d2h_bytes;
  int64_t hist_rows_cols;    /* sum over level passes of local rows x active columns */
  int64_t hist_algo_bytes;   /* algorithmic bytes of those passes (SURVEY.md section 8d formula) */
  int64_t hist_launches;     /* histogram-accumulation kernel launches (ms_hist / hist_launches = mean launch time) */
  int64_t sparse_launches;   /* of those: launches of the leaf-major kernel used on levels with per-node column sampling */
  double ms_exchange;        /* cross-rank exchange of the level histograms (peer-memory kernel or the NCCL all-reduces), waits for the slowest rank included */
} h2ot_stats;

/* ---- lifetime: one engine per model build per GPU (created after T/SharedTree.java:385, destroyed in the finally at :407-425) */
int h2ot_version(void);
const char* h2ot_last_error(void);
int h2ot_device_count(int32_t* out);
int h2ot_create(const h2ot_config* cfg, h2ot_engine** out);
int h2ot_destroy(h2ot_engine* h);

/* ---- frame ingest: the host pushes what Chunk.getDoubles hands to DHistogram.extractData
 * (T/DHistogram.java:715-728, W/fvec/Chunk.java:785-794); NaN == NA. Columns may arrive in pieces
 * (one call per Chunk). is_categorical/cardinality mirror Vec.isCategorical()/domain().length. */
int h2ot_load_column(h2ot_engine* h, int32_t col, int64_t row0, int64_t n, const double* values);
int h2ot_load_response(h2ot_engine* h, int64_t row0, int64_t n, const double* values);
int h2ot_load_weights(h2ot_engine* h, int64_t row0, int64_t n, const double* values);
int h2ot_load_offset(h2ot_engine* h, int64_t row0, int64_t n, const double* values);

/* ---- frame ingest straight from H2O's compressed chunks: the host hands over Chunk._mem as is and the
 * decode that Chunk.atd_impl / getDoubles would do on the CPU runs on the device. One codec id per Chunk subclass
 * (W/fvec/<name>.java); layouts are little endian (UnsafeUtils, x86). */
enum {
  H2OT_CODEC_C0D = 0,  /* W/fvec/C0DChunk.java:11-31   [double con][int len]: every row == con (NaN == all NA) */
  H2OT_CODEC_C0L = 1,  /* W/fvec/C0LChunk.java:8-22    [long con][int len] */
  H2OT_CODEC_C1 = 2,   /* W/fvec/C1Chunk.java:8-23     unsigned bytes, 0xFF == NA */
  H2OT_CODEC_C1N = 3,  /* W/fvec/C1NChunk.java:8-12    unsigned bytes, no NA */
  H2OT_CODEC_C1S = 4,  /* W/fvec/C1SChunk.java:31-33 + W/fvec/CSChunk.java:18-62: [long bias][int scale][int szLog | decimal] then unsigned bytes */
  H2OT_CODEC_C2 = 5,   /* W/fvec/C2Chunk.java:9-21     shorts, Short.MIN_VALUE == NA */
  H2OT_CODEC_C2S = 6,  /* W/fvec/C2SChunk.java:32-37 + CSChunk: header then shorts */
  H2OT_CODEC_C4 = 7,   /* W/fvec/C4Chunk.java:9-19     ints, Integer.MIN_VALUE == NA */
  H2OT_CODEC_C4S = 8,  /* W/fvec/C4SChunk.java:24-30 + CSChunk: header then ints */
  H2OT_CODEC_C4F = 9,  /* W/fvec/C4FChunk.java:8-18    floats, NaN == NA */
  H2OT_CODEC_C8 = 10,  /* W/fvec/C8Chunk.java:8-19     longs, Long.MIN_VALUE == NA */
  H2OT_CODEC_C8D = 11, /* W/fvec/C8DChunk.java:8-16    doubles */
  H2OT_CODEC_CBS = 12, /* W/fvec/CBSChunk.java:9-80    [gap][bpv] then 1 or 2 bits per row, 2 == NA */
  H2OT_CODEC_CUD = 13, /* W/fvec/CUDChunk.java:12-47   [int len][int nuniq][nuniq doubles][len bytes: index - 128] */
  H2OT_CODEC_CXI = 14, /* W/fvec/CXIChunk.java:7-120   sparse ints: [int len][id size][value size][NA-sparse] then (id, value) pairs */
  H2OT_CODEC_CXF = 15  /* W/fvec/CXFChunk.java:7-36    sparse float / double values */
};
#define H2OT_COL_RESPONSE (-1)
#define H2OT_COL_WEIGHTS (-2)
#define H2OT_COL_OFFSET (-3)
/* Decodes n rows of one chunk into rows [row0, row0 + n) of predictor `col` (or the response / weights / offset column).
 * mem, mem_len: the chunk's byte array (host memory). n must equal the chunk's own row count (as Chunk.initFromBytes derives
 * it); a malformed chunk (sizes that disagree, an unknown sparse element layout, ids out of range) is H2OT_ERR_INVALID. */
int h2ot_load_chunk(h2ot_engine* h, int32_t col, int64_t row0, int64_t n, int32_t codec, const uint8_t* mem, int64_t mem_len);
/* rollups (min/max/isInt/naCnt) + lossless narrow encoding of every loaded column, resident in HBM */
int h2ot_finalize_frame(h2ot_engine* h);
int h2ot_column_info_get(h2ot_engine* h, int32_t col, h2ot_column_info* out);

/* ---- model state sync with the Java work vecs (Tree_k / Work_k / NIDs_k; T/SharedTree.java:364-385,464-475) */
/* GBM.initializeModelSpecifics (T/gbm/GBM.java:282-311): Tree_k <- init_f[k]; rows with w==0 or NA response <- DECIDED_ROW. */
int h2ot_init_model(h2ot_engine* h, const double* init_f /*[K]*/, int64_t* out_skipped);
/* SharedTree.InitialValue (T/SharedTree.java:1138-1185) on the resident response / weights / offset: num = sum w (y - offset),
 * den = sum w over rows with a response and w != 0 (local rows; sum over ranks before dividing). Optional extras for DRF's set-up:
 * class_weight[nclass] = sum of w per response class, abs_max = max |(float) y|. All outputs nullable. */
int h2ot_initial_stats(h2ot_engine* h, double* out_num, double* out_den, double* out_class_weight, double* out_abs_max);
int h2ot_set_predictions(h2ot_engine* h, int32_t k, const double* f);
int h2ot_get_predictions(h2ot_engine* h, int32_t k, double* f);
int h2ot_get_work(h2ot_engine* h, int32_t k, double* work);
int h2ot_set_work(h2ot_engine* h, int32_t k, const double* work, double abs_bound);
int h2ot_get_nids(h2ot_engine* h, int32_t k, int32_t* nids);
int h2ot_set_nids(h2ot_engine* h, int32_t k, const int32_t* nids);

/* ---- per tree */
/* GBM.ComputePredAndRes (T/gbm/GBM.java:1032-1081): Work_k <- (float) negHalfGradient for all K.
 * out_abs_max[k] receives max |Work_k| over local rows (fixes the fixed-point scale; nullable). */
int h2ot_compute_residuals(h2ot_engine* h, double* out_abs_max /*[K]*/);
/* Bounds of the per-row GammaPass terms |gammaNum / w| and |gammaDenom / w| seen by the last h2ot_compute_residuals (they fix
 * the fixed-point scale of the leaf sums). With several ranks take the max over ranks and set it back before h2ot_gamma_pass. */
int h2ot_get_gamma_bounds(h2ot_engine* h, double out[2]);
int h2ot_set_gamma_bounds(h2ot_engine* h, const double in[2]);
/* Starts class-tree k: clears its node table, fixes the accumulator scale from work_abs_bound
 * (>= max |Work_k| over ALL ranks), and runs Sample (T/Sample.java:22-36) when sample_rate < 1. */
int h2ot_begin_tree(h2ot_engine* h, int32_t k, int64_t tree_seed, double sample_rate, double work_abs_bound);
/* Sample._rate_per_class (T/Sample.java:14, 30; SharedTreeModel.java:59): rates[n] indexed by the response class, used by every
 * later h2ot_begin_tree / h2ot_sample instead of their sample_rate (which then only has to be passed); n == 0 goes back to null. */
int h2ot_set_sample_rate_per_class(h2ot_engine* h, const double* rates, int32_t n);

/* Split points of the NEXT h2ot_build_level / h2ot_level_step of class-tree k: one pool of doubles, referenced by the
 * descriptors' splitpts_off / splitpts_n (DHistogram._splitPts as DHistogram.init leaves them, T/DHistogram.java:343-401).
 * The pool is copied; -0.0 is stored as 0.0 like the reference does (:389-401). */
int h2ot_set_split_points(h2ot_engine* h, int32_t k, const double* points, int64_t n);

/* ---- per level: the drop-in for new ScoreBuildHistogram2(...).dfork2(fr2) (T/SharedTree.java:681-682).
 * decided[0..n_decided): the DecidedNodes created by the previous level's onCompletion (nid order);
 * descs[n_leaves][ncols]: the UndecidedNodes' DHistograms for nids leaf_offset..leaf_offset+n_leaves-1.
 * Runs score_decide (T/ScoreBuildHistogram2.java:139-168,201-211) then the histogram build
 * (T/DHistogram.java:609-704) for every active (leaf, col). Results stay on the device.
 * With several ranks and the peer-memory exchange, the level of every class-tree is accumulated in ONE buffer the peers have mapped:
 * h2ot_build_level(k') before h2ot_finish_level(k) of another class-tree fails with H2OT_ERR_INVALID (h2ot_level_step[_begin] keep
 * that order by themselves). If the call fails after the rows have been moved (a histogram that cannot be placed: H2OT_ERR_UNSUPPORTED
 * / H2OT_ERR_INVALID from the planner), NIDs_k is one level ahead of the caller's tree: every later level call of this class-tree
 * fails until h2ot_begin_tree restarts it. */
int h2ot_build_level(h2ot_engine* h, int32_t k, int32_t leaf_offset, int32_t n_leaves, int32_t n_decided,
                     const h2ot_decided_node* decided, const uint8_t* bitset_pool, int64_t bitset_bytes,
                     const h2ot_hist_desc* descs);
/* Device buffers of the level just built, for the cross-rank exchange that replaces MRTask.reduce /
 * ScoreBuildHistogram.reduce / DHistogram.add (T/ScoreBuildHistogram.java:74-88, T/DHistogram.java:434-447):
 * sums[n_sums] int64 (allreduce SUM; rows per leaf first, then w, wY, wYY per bin) and minmax[n_minmax] uint64 order-preserving keys laid out so that
 * an allreduce MAX merges both bounds (min is stored complemented). Single rank: skip. */
int h2ot_level_buffers(h2ot_engine* h, int32_t k, void** d_sums, int64_t* n_sums, void** d_minmax, int64_t* n_minmax);
/* Built-in exchange over NCCL (h2ot_comm_init must have been called); no-op for world size 1. */
int h2ot_allreduce_level(h2ot_engine* h, int32_t k);
/* postGlobal (T/ScoreBuildHistogram2.java:455-466): fixed-point -> fp64, reducePrecision (T/DHistogram.java:734-740).
 * Optional host copies (all nullable): vals packed per active (leaf,col) in [leaf][col] order, (nbin+1)*3 doubles each;
 * val_offsets[n_leaves*ncols] = start (in doubles) of each block or -1; min2/maxIn/wNA[n_leaves*ncols] (wNA = weight of the
 * NA bucket, all the host needs for DHistogram.hasNABin in Split.nextLevelHistos, T/DTree.java:401); has_vals = _vals != null. */
int h2ot_finish_level(h2ot_engine* h, int32_t k, double* out_vals, int64_t out_vals_capacity, int64_t* out_val_offsets,
                      double* out_min2, double* out_maxIn, double* out_wNA, uint8_t* out_has_vals);
/* DecidedNode ctor's bestCol / FindSplits / findBestSplitPoint / splat (T/DTree.java:616-726, 984-1487, 240-291)
 * on the histograms of the level just finished. out_splits[n_leaves]. */
int h2ot_find_splits(h2ot_engine* h, int32_t k, double min_rows, double min_split_improvement, h2ot_split* out_splits);

/* The whole level in one call and ONE host synchronisation: h2ot_build_level, h2ot_allreduce_level (no-op on one rank),
 * h2ot_finish_level (without the histogram download) and h2ot_find_splits back to back on the engine's stream. What a
 * caller that lets the device search the splits uses (host_gbm.cpp); the separate calls remain for callers that want the
 * histograms themselves (the minimal drop-in of INTEGRATION.md section 2, step 3a). */
int h2ot_level_step(h2ot_engine* h, int32_t k, int32_t leaf_offset, int32_t n_leaves, int32_t n_decided,
                    const h2ot_decided_node* decided, const uint8_t* bitset_pool, int64_t bitset_bytes, const h2ot_hist_desc* descs,
                    double min_rows, double min_split_improvement, double* out_min2, double* out_maxIn, double* out_wNA,
                    uint8_t* out_has_vals, h2ot_split* out_splits);

/* The same level in two halves, for callers that build the K class-trees of a multinomial iteration side by side: _begin
 * enqueues the whole level of class-tree k and returns without waiting (its inputs have been copied); _end waits for THAT level's
 * results only, so the levels of the other class-trees keep the device busy while the host turns these splits into the next
 * level's descriptors. One _begin per class-tree may be outstanding; no other call on the handle between a _begin and its _end
 * except _begin / _end of other class-trees. */
int h2ot_level_step_begin(h2ot_engine* h, int32_t k, int32_t leaf_offset, int32_t n_leaves, int32_t n_decided,
                          const h2ot_decided_node* decided, const uint8_t* bitset_pool, int64_t bitset_bytes,
                          const h2ot_hist_desc* descs, double min_rows, double min_split_improvement);
int h2ot_level_step_end(h2ot_engine* h, int32_t k, double* out_min2, double* out_maxIn, double* out_wNA, uint8_t* out_has_vals,
                        h2ot_split* out_splits);

/* ---- a whole tree with ONE host synchronisation: every level of class-tree k (h2ot_level_step for depth 0 .. max_depth - 1) is
 * enqueued back to back, and what the Java side does between two levels -- the DecidedNode constructor and Split.nextLevelHistos with the
 * DHistogram constructor rule (h2o-algos/src/main/java/hex/tree/DTree.java:312-417, 703-726; DHistogram.java:190-269, 453-459) -- runs
 * on the device. For builds without per-node column sampling (UndecidedNode._scoreCols == null: mtries == ncols), with
 * uniform-adaptive histograms and no float-split correction; root_descs[ncols] are the root's DHistograms.
 * Node ids are STATIC inside these two calls: the nodes of depth d are the slots 2^d - 1 .. 2^(d+1) - 2 of a complete binary tree,
 * child w of slot s being slot 2 s + 1 + w; NIDs_k holds such ids afterwards. _end waits for the tree and returns
 * out_splits[2^max_depth - 1] (the split of every slot of depth < max_depth; col == -1: none) and out_exists[2^(max_depth+1) - 1]
 * (1: the slot holds a node). The caller renumbers the nodes in creation order for its own tree (DTree numbers them as they are
 * made: level by level, in leaf order, left child first) and passes the finished tree to h2ot_set_tree in static ids: decided
 * nodes under their slot, LeafNodes under ids >= 2^(max_depth+1) - 1. Several class-trees may be begun before the first is ended. */
int h2ot_tree_levels_begin(h2ot_engine* h, int32_t k, int32_t max_depth, const h2ot_hist_desc* root_descs, double min_rows,
                           double min_split_improvement, int32_t nbins, int32_t nbins_cats);
int h2ot_tree_levels_end(h2ot_engine* h, int32_t k, h2ot_split* out_splits, uint8_t* out_exists);

/* ---- trees in a pipeline: no host synchronisation between the trees of a model. After h2ot_tree_levels_begin,
 * h2ot_tree_finish_enqueue queues the rest of the tree behind its levels: the leaf insertion (every child pointer that does not lead
 * to a split node becomes a LeafNode, T/gbm/GBM.java:655-679) with the leaves numbered by their static slot (leaf id =
 * 2^(max_depth+1) - 1 + slot of the child), GammaPass (:1286-1407; per-leaf sums all-reduced over the communicator), the leaf values
 * of fitBestConstants for gaussian / bernoulli / multinomial (:813-847: rate x num / den with rate = annealed learn_rate x (K-1)/K,
 * the 1e4 clamp for multinomial, NaN -> 0, max_abs_leafnode_pred) and AddTreeContributions (:1410-1470). Then the next tree's
 * h2ot_compute_residuals_async / h2ot_begin_tree / h2ot_tree_levels_begin may be enqueued before this tree is ended: up to TWO trees
 * per class-tree may be in flight, _end returns the oldest. h2ot_tree_levels_end2 also returns the leaf values,
 * out_leaf_pred[2^(max_depth+1) - 1] indexed by the child's slot (0 where the slot is no leaf).
 * h2ot_compute_residuals_async is h2ot_compute_residuals without waiting for max |residual|: the caller passes a bound it knows
 * to h2ot_begin_tree and h2ot_set_gamma_bounds -- |y - p| <= 1 for bernoulli and multinomial; for gaussian a bound derived from an
 * EARLIER tree's maxima (a leaf's value is a weighted mean of its residuals, so |r'| <= (1 + learn_rate) max |r|). With
 * want_bounds != 0 the maxima (all-reduced over the communicator) are copied to pinned memory behind the kernel and
 * h2ot_residual_bounds_wait returns them, waiting for that copy only: out_abs_max[K], out_gamma_bounds[2] (both nullable). */
int h2ot_tree_finish_enqueue(h2ot_engine* h, int32_t k, double rate, double max_abs_leafnode_pred);
/* Quantile.StratifiedQuantilesTask for GBM's laplace / quantile leaf fitting (h2o-algos/src/main/java/hex/tree/gbm/GBM.java:729-751;
 * h2o-core/src/main/java/hex/quantile/Quantile.java:118-175, 267-350): out[s] = the prob-quantile (interpolated, CombineMethod
 * INTERPOLATE) of y - (Tree_k + offset) over the rows whose NIDs_k is leaf_start + s, s < n_strata (call it after h2ot_gamma_pass
 * has moved the rows onto their leaves); NaN where a leaf has no rows. With use_tree == 0 and n_strata == 1: the quantile of
 * y - offset over all rows, GBM.getInitialValueQuantile (:355-381). Unit weights and a single rank only (H2OT_ERR_UNSUPPORTED). */
int h2ot_leaf_quantiles(h2ot_engine* h, int32_t k, int32_t leaf_start, int32_t n_strata, int32_t use_tree, double prob, double* out);
int h2ot_tree_levels_end2(h2ot_engine* h, int32_t k, h2ot_split* out_splits, uint8_t* out_exists, float* out_leaf_pred);
int h2ot_compute_residuals_async(h2ot_engine* h, int32_t want_bounds);
int h2ot_residual_bounds_wait(h2ot_engine* h, double* out_abs_max, double out_gamma_bounds[2]);

/* ---- end of tree */
/* Uploads the finished tree's DecidedNodes (children now pointing at LeafNodes; T/gbm/GBM.java:655-679). */
int h2ot_set_tree(h2ot_engine* h, int32_t k, int32_t n_nodes, const h2ot_decided_node* nodes, const uint8_t* bitset_pool,
                  int64_t bitset_bytes);
/* GBM.GammaPass (T/gbm/GBM.java:1286-1407): final descent (NIDs_k <- leaf nid) and per-leaf sums.
 * out_num/out_denom[n_leaf_nodes] for leaf nids leaf_start..leaf_start+n_leaf_nodes-1 (local rows only;
 * h2ot_gamma_buffers exposes the device int64 buffers for a cross-rank SUM before reading). */
int h2ot_gamma_pass(h2ot_engine* h, int32_t k, int32_t leaf_start, int32_t n_leaf_nodes, double* out_num, double* out_denom);
/* GBM.AddTreeContributions (T/gbm/GBM.java:1410-1470): Tree_k <- (float)(Tree_k + pred[nid]); NIDs_k <- 0 / DECIDED_ROW. */
int h2ot_add_tree(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred /*[n_nodes], leaf nids used*/);
/* The same with GBM's pred_noise_bandwidth (T/gbm/GBM.java:1450-1463): Tree_k <- (float)(Tree_k + node_factor[nid] * pred[nid]); the
 * caller draws the per-node factors 1 + nextGaussian * bandwidth (they depend on tree, class and node only, not on the row). */
int h2ot_add_tree_scaled(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred, const double* node_factor);

/* ---- scoring a frame with finished trees: SharedTreeMojoModel.scoreTree + scoreTreeRange's
 * `preds[k] += tree value` (h2o-genmodel .../algos/tree/SharedTreeMojoModel.java:129-240, 1043-1076). Every row of the engine's frame
 * walks the tree given to h2ot_set_tree from the root (nid 0) to a leaf (nid >= leaf_start, or a root that never split) with the
 * DecidedNode.getChildNodeID rules, and Tree_k <- Tree_k + (double)node_pred[leaf] WITHOUT the per-tree float rounding of
 * AddTreeContributions. An engine holding a validation / test frame (h2ot_init_model with 0, then one call per tree in build
 * order, then + init_f as GbmMojoModel.score0 does) reproduces the model's link-space predictions on new rows. NIDs_k is untouched. */
int h2ot_score_tree(h2ot_engine* h, int32_t k, int32_t leaf_start, int32_t n_nodes, const float* node_pred);

/* ---- training metrics on the resident frame: hex.tree.Score (T/Score.java:57-141) for GBM. The link-space
 * predictions Tree_k (+ offset) are turned into class probabilities as score1/score2 do (T/gbm/GBM.java:1486-1513,
 * T/SharedTree.java:782-789) and fed to MetricBuilder.perRow (H/ModelMetricsBinomial.java, ModelMetricsMultinomial, ModelMetricsRegression):
 * sums over local rows with a response and weight != 0. MSE = sumsqe / wcount, logloss = logloss / wcount (after summing over ranks).
 * The AUC comes from h2ot_auc_builders below; confusion matrices stay on the host. */
typedef struct h2ot_metrics {
  double count;   /* rows scored */
  double wcount;  /* sum of weights */
  double sumsqe;  /* sum w * err^2 ; err = y - f (gaussian) or 1 - p(actual class) */
  double logloss; /* sum w * min(-log(1e-15), -log(p(actual class))) ; 0 for gaussian */
} h2ot_metrics;
int h2ot_score_metrics(h2ot_engine* h, h2ot_metrics* out);
/* AUC2.AUCBuilder (h2o-core/src/main/java/hex/AUC2.java:356-581) over the resident bernoulli model: the local rows are cut into
 * at most max_chunks chunks of *out_chunk_rows consecutive rows (a multiple of 32; the last one may be short) and every chunk's
 * rows go, in row order, through perRow(p1, actual, weight) of a builder of its own -- what AUC_Impl.map does per Chunk (:343-350).
 * out[c * H2OT_AUC_STRIDE]: number of bins n of chunk c, then ths[400], tps[400], fps[400], sqe[400] (the builder's arrays:
 * ascending bin centres, true / false positive weights, squared error inside the bin; first n valid).
 * The caller folds the builders in chunk order with AUCBuilder.reduce (:448-472) and, across ranks, in rank order. Rows without a
 * response or with weight 0 are skipped (ModelMetricsBinomial.MetricBuilderBinomial.perRow). */
#define H2OT_AUC_BINS 400
#define H2OT_AUC_STRIDE (1 + 4 * H2OT_AUC_BINS)
int h2ot_auc_builders(h2ot_engine* h, int32_t max_chunks, int32_t* out_n_chunks, int64_t* out_chunk_rows, double* out);

/* ---- DRF (T/drf/DRF.java). The level loop is the same; what differs is around it. */
/* DRF.initializeModelSpecifics (:139-153) + the work vecs of SharedTree.java:364-385: Work_k <- class indicator (nclass > 1) or
 * (float) response, Tree_k <- 0, NIDs_k <- 0, OUT_BAG_TREES <- 0. No row is marked DECIDED (that is GBM-only). */
int h2ot_drf_init_model(h2ot_engine* h);
/* Sample (T/Sample.java:22-36) unconditionally: DRF samples every tree, and rows with an NA response go out of bag even at rate 1. */
int h2ot_sample(h2ot_engine* h, int32_t k, int64_t tree_seed, double sample_rate);
/* DRF.CollectPreds (:255-330) for class-tree k after h2ot_set_tree: out-of-bag rows with weight != 0 get
 * Tree_k <- (float)(Tree_k + pred[leaf]); every NIDs_k row <- 0; OUT_BAG_TREES += weight for out-of-bag rows when update_oobt
 * (pass 1 for exactly one k per iteration). root_is_leaf: the tree never split (leaf id 0). */
int h2ot_drf_collect_preds(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred, int32_t root_is_leaf, int32_t update_oobt);
int h2ot_get_oobt(h2ot_engine* h, double* out);

/* ---- multi-GPU plumbing (one process per GPU; rows sharded; replaces the MRTask node tree, W/MRTask.java:695-756,871-930) */
#define H2OT_COMM_ID_BYTES 128
int h2ot_comm_unique_id(uint8_t out_id[H2OT_COMM_ID_BYTES]);
/* A communicator is created the first time an id is seen in the process and is shared by every later engine that passes
 * the same id on the same device (an id can start only one NCCL communicator); it lives until the process exits. */
int h2ot_comm_init(h2ot_engine* h, int32_t world_size, int32_t rank, const uint8_t id[H2OT_COMM_ID_BYTES]);
int h2ot_allreduce_host_f64(h2ot_engine* h, double* inout, int64_t n, int32_t op /*0 sum, 1 max*/);
/* Sizes the buffer the ranks exchange their level histograms through (the MRTask.reduce of ScoreBuildHistogram2,
 * h2o-algos/src/main/java/hex/tree/ScoreBuildHistogram.java:74-88) once, for the widest level the model can have:
 * max_level_bins = bins (NA buckets included) over all (leaf, column) histograms of one level, max_level_pairs = leaves x columns,
 * max_leaves = leaves of one level. COLLECTIVE: every rank calls it with the same numbers, after h2ot_comm_init. Optional -- a level
 * that outgrows the buffer regrows it (also collectively: all ranks reach that level together) -- but mapping the peers' buffers
 * costs milliseconds, which belongs in the set-up and not in the first trees. No-op on one GPU or with H2OT_XREDUCE=0.
 * h2ot_destroy of engines that share a communicator is collective as well (the mapped buffers are released in step). */
int h2ot_reserve_exchange(h2ot_engine* h, int64_t max_level_bins, int64_t max_level_pairs, int64_t max_leaves);
/* max |weight| over the local rows (known after h2ot_finalize_frame). With several ranks the host takes the max over
 * ranks and sets it back, so that every rank uses the same fixed-point scale for the w accumulator. */
int h2ot_get_weight_bound(h2ot_engine* h, double* out);
int h2ot_set_weight_bound(h2ot_engine* h, double bound);

/* ---- diagnostics / parity hooks */
/* Per-row bin index of column col under desc (DHistogram.bin, T/DHistogram.java:272-306), uint16 codes to the host. */
int h2ot_debug_bin_codes(h2ot_engine* h, int32_t col, const h2ot_hist_desc* desc, uint16_t* out_codes);
/* Rows [row0, row0 + n) of a finalized predictor column, decoded back to doubles (NaN == NA): storage is lossless, so this is
 * what was loaded (tests compare it with Chunk.atd of the oracle). */
int h2ot_debug_get_column(h2ot_engine* h, int32_t col, int64_t row0, int64_t n, double* out);
int h2ot_stats_get(h2ot_engine* h, h2ot_stats* out);
int h2ot_stats_reset(h2ot_engine* h);
/* 0: default (shared-memory privatised histograms; leaf-major sparse kernel on levels with per-node column sampling),
 * 1: global (L2) atomics only, 2: column-major shared-memory kernels only — 1 and 2 are used by tests to cross-check. */
int h2ot_set_hist_mode(h2ot_engine* h, int32_t mode);
/* on != 0: h2ot_begin_tree, h2ot_sample, h2ot_set_tree and h2ot_add_tree[_scaled] -- calls that hand nothing back to the host -- return
 * without waiting for the stream; their work, kernel times and any error surface at the next call that does wait (every call that
 * returns data). A model builder that sequences whole trees saves three stream synchronisations per tree. on == 0: waits now. */
int h2ot_set_lazy_sync(h2ot_engine* h, int32_t on);
/* How much of h2ot_stats' per-stage times is measured. A timed section costs two CUDA event records in the stream (~3 us; 45
 * sections per tree are 9 % of a tree on an eighth of the Higgs frame, 2 % on the whole frame). 2 (default): every stage;
 * 1: the histogram launches only (ms_hist, hist_launches, hist_algo_bytes: what a roofline needs); 0: nothing. The model builder
 * (h2ot_gbm) selects 1 unless H2OT_STAGE_TIMERS is set in the environment. */
int h2ot_set_timers(h2ot_engine* h, int32_t mode);

/* ---- histogram of one column for callers that cannot allocate device memory.
 * No counterpart in the reference: it exists for embedders whose process may not call cudaMalloc once it is running (a JVM or
 * service that gave the whole GPU to its own pooling allocator, a real-time path where an allocation is an unbounded stall, a
 * plug-in host that forbids allocation after start-up). The caller owns every buffer it passes — device pointers on the CURRENT
 * device: d_x[n] (fp32 column, NaN == NA), d_work[n] (fp32 response/residual), d_leaf_code[n] (leaf of each row, >= n_leaves ==
 * skip; NULL == all rows in leaf 0), d_out_vals (packed per leaf, (nbin+1)*3 doubles each: w, wY, wYY with the float cast of
 * reducePrecision) — and descs[n_leaves] on the host (min, step, nbin are read). The kernel's accumulators live in a fixed
 * 1.5 MiB device array inside the library (present on every GPU that loads it), so the call allocates nothing.
 * Limits: n_leaves <= H2OT_NOALLOC_MAX_LEAVES and sum(nbin+1) <= H2OT_NOALLOC_MAX_BINS, else H2OT_ERR_UNSUPPORTED.
 * The array is shared by all calls on a GPU and keeps the last call's contents; each call clears what it uses, so calls
 * whose GPU work overlaps would corrupt each other: issue them one after another. The first call on a GPU resolves the
 * array's address. Work is enqueued on cuda_stream (a cudaStream_t, NULL == default stream) and the call returns after that
 * stream has drained: it synchronises the stream, so it must not be issued on a stream that is being captured. Unit weights only.
 * work_abs_bound: at least max |d_work[i]| over the rows that take part (0 == 1.0). It fixes the fixed-point scale of the
 * sums: a smaller value than the true maximum overflows the 64-bit accumulators silently; a NaN in d_work poisons its bin. */
#define H2OT_NOALLOC_MAX_LEAVES 64
#define H2OT_NOALLOC_MAX_BINS 65536
int h2ot_hist_column_noalloc(const float* d_x, const float* d_work, const uint16_t* d_leaf_code, int64_t n, int32_t n_leaves,
                             const h2ot_hist_desc* descs, double work_abs_bound, double* d_out_vals, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* H2OT_H */
