// Host-side GBM driver of libh2otree_b200.so (C ABI: include/h2ot_gbm.h).
//
// In an H2O-3 deployment this role is played by the unchanged Java classes hex.tree.SharedTree, gbm.GBM,
// DTree and DHistogram; only the calls below the ScoreBuildHistogram2 seam go to the engine. No JDK exists
// in this image, so the same sequencing is written here in C++ strictly on top of the engine's public C ABI
// (include/h2ot.h): it keeps the tree (node table), builds the per-(leaf, column) DHistogram descriptors,
// draws the per-node column samples, and sequences levels and per-tree row passes. It never touches rows:
// every O(rows) step is a device kernel behind an h2ot_* call.
//
// Reference code mirrored here (paths relative to the reference root; T/ = h2o-algos/src/main/java/hex/tree/):
//   DHistogram ctor / make / initialHist / useIntOpt / find_maxEx   T/DHistogram.java:190-269, 453-459, 474-554
//   DTree ctor, UndecidedNode.scoreCols, doNotSplit, DecidedNode ctor   T/DTree.java:56-86, 443-511, 703-726
//   Split.nextLevelHistos                                            T/DTree.java:312-417
//   SharedTree.buildLayer / ScoreBuildOneTree.onCompletion          T/SharedTree.java:554-628, 684-710
//   GBM.initializeModelSpecifics / buildNextKTrees / growTrees / fitBestConstants   T/gbm/GBM.java:254-312, 464-680, 813-847
//   SharedTree.InitialValue                                         T/SharedTree.java:1138-1185
//   PCGRNG                                                           h2o-core/src/main/java/water/util/RandomUtils.java:50-119
#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <string>
#include <vector>

#include "../../include/h2ot_gbm.h"

namespace h2ot {
void set_error(const std::string& msg);
}

namespace {

constexpr int kNoParent = -1;
constexpr int kUndecidedChild = -1;  // T/ScoreBuildHistogram.java:65

#define GFAIL(code, msg)      \
  do {                        \
    ::h2ot::set_error(msg);   \
    return (code);            \
  } while (0)
#define GTRY(expr)                \
  do {                            \
    int _r = (expr);              \
    if (_r != H2OT_OK) return _r; \
  } while (0)

// ---- Java numeric casts (JLS 5.1.3)
inline int32_t d2i(double d) {
  if (std::isnan(d)) return 0;
  if (d >= 2147483647.0) return INT32_MAX;
  if (d <= -2147483648.0) return INT32_MIN;
  return (int32_t)d;
}
inline int64_t d2l(double d) {
  if (std::isnan(d)) return 0;
  if (d >= 9223372036854775807.0) return INT64_MAX;
  if (d <= -9223372036854775808.0) return INT64_MIN;
  return (int64_t)d;
}
inline double ulp_d(double d) {  // Math.ulp(double)
  if (std::isnan(d)) return d;
  d = std::fabs(d);
  if (std::isinf(d)) return d;
  if (d == DBL_MAX) return std::ldexp(1.0, 971);
  uint64_t u;  // next representable value above a finite d >= 0: the bit pattern plus one (no libm call on this hot host path)
  memcpy(&u, &d, sizeof u);
  u++;
  double nx;
  memcpy(&nx, &u, sizeof nx);
  return nx - d;
}
inline bool within_one_small_ulp(double a, double b) {  // water/util/MathUtils.java:391-397
  if (std::isnan(a) && std::isnan(b)) return true;
  return std::fabs(a - b) <= std::fmin(ulp_d(a), ulp_d(b));
}

// ---- PCG-XSH-RR 64/32 as java.util.Random subclass (RandomUtils.java:50-119)
struct Pcg {
  uint64_t state = 0, inc = 3;
  explicit Pcg(int64_t seed) { set_seed(seed); }
  void set_seed(int64_t seed) {
    state = 0;
    next_int();
    state += (uint64_t)seed;
    next_int();
  }
  uint32_t next_int() {
    const uint64_t old = state;
    state = old * 6364136223846793005ULL + inc;
    const uint32_t xs = (uint32_t)(((old >> 18) ^ old) >> 27);
    const uint32_t rot = (uint32_t)(old >> 59);
    return (xs >> rot) | (xs << ((-rot) & 31));
  }
  int64_t next_long() {
    const uint64_t hi = next_int();
    const uint64_t lo = next_int();
    return (int64_t)((hi << 32) | lo);
  }
  float next_float() {  // java.util.Random.nextFloat over PCGRNG.next(24) (:86)
    return (float)(int32_t)(next_long() & ((1LL << 24) - 1)) / (float)(1 << 24);
  }
  double next_double() {  // java.util.Random.nextDouble over PCGRNG.next(bits)
    const int64_t a = next_long() & ((1LL << 26) - 1);
    const int64_t b = next_long() & ((1LL << 27) - 1);
    return (double)((a << 27) + b) * 0x1.0p-53;
  }
  double next_gaussian_after_seed() {  // java.util.Random.nextGaussian, first value after setSeed (polar method)
    double v1, v2, s;
    do {
      v1 = 2 * next_double() - 1;
      v2 = 2 * next_double() - 1;
      s = v1 * v1 + v2 * v2;
    } while (s >= 1 || s == 0);
    return v1 * std::sqrt(-2 * std::log(s) / s);
  }
  int32_t next_bounded(int32_t bound) {  // :88-118
    const uint64_t threshold = (uint64_t)((-(int64_t)bound) % (int64_t)bound) & 0xFFFFFFFFULL;
    for (;;) {
      const uint64_t r = next_int();
      if (r >= threshold) return (int32_t)(r % (uint64_t)bound);
    }
  }
};

// ---- DHistogram metadata: everything of a DHistogram except the bins (those live on the device)
struct Hist {
  bool present = false;  // false == Java null
  uint8_t is_int = 0;    // 0 real, 1 int, 2 categorical
  bool int_opt = false, init_na = false, check_float = false;
  int nbin = 0;
  double min = 0, max_ex = 0, step = 0;
  int64_t seed = 0;
  // filled by the level pass
  bool has_vals = false;
  double min2 = DBL_MAX, max_in = -DBL_MAX, w_na = 0;

  bool has_na_bin() const { return has_vals ? w_na > 0 : init_na; }  // T/DHistogram.java:335-340
  double find_max_ex() const { return find_max_ex_of(max_in, is_int); }
  static double find_max_ex_of(double max_in, int is_int) {  // :453-459
    double ulp = ulp_d(max_in);
    if (is_int > 0 && 1 > ulp) ulp = 1;
    const double res = max_in + ulp;
    return std::isinf(res) ? max_in : res;
  }
};

// DHistogram ctor (T/DHistogram.java:190-269, UniformAdaptive). Returns false for StepOutOfRangeException.
bool make_hist(Hist& h, int nbins, int nbins_cats, uint8_t is_int, double min, double max_ex, bool int_opt, bool init_na,
               int64_t seed, bool check_float) {
  h = Hist{};
  h.is_int = is_int;
  h.min = min;
  h.max_ex = max_ex;
  h.init_na = init_na;
  h.int_opt = int_opt;
  h.seed = seed;
  h.check_float = check_float;
  int xbins = is_int == 2 ? nbins_cats : nbins;
  if (is_int > 0 && max_ex - min <= xbins) {
    xbins = (uint16_t)(d2l(max_ex) - d2l(min));  // (char) cast
    h.step = 1.0;
  } else {
    h.step = xbins / (max_ex - min);
    if (h.step <= 0 || std::isinf(h.step) || std::isnan(h.step)) return false;
  }
  h.nbin = (uint16_t)xbins;
  h.present = true;
  return true;
}

struct ColMeta {  // Vec rollups + type (RollupStats.java:31-48)
  double min = NAN, max = NAN;
  int64_t nacnt = 0;
  bool is_int = true, is_cat = false;
  int cardinality = 0;
  bool loaded = false;
};

struct Node {
  int kind = 0;  // 0 undecided, 1 decided, 2 leaf
  int pid = kNoParent, nid = 0;
  int sid = -1;  // slot of the node in the engine's static numbering when the tree's levels ran back to back on the device
  std::vector<Hist> hs;  // undecided only ([ncols]); empty == no histograms
  bool score_cols_null = true;
  std::vector<int> score_cols;
  bool has_split = false;
  h2ot_split split{};
  float splat = NAN;
  int nids[2] = {-1, -1};
  float pred = 0;
};

struct Tree {
  int ncols = 0, depth = 0, leaves = 0;
  int64_t seed = 0;
  int mtrys = 0, mtrys_per_tree = 0;
  double change_per_level = 1.0;
  Pcg rand;
  std::vector<int> cols;
  std::vector<Node> ns;
  std::vector<int> dirty;  // decided nodes whose device copy is stale
  std::vector<int> draw_scratch;

  Tree(int ncols_, int mtrys_, int mtrys_per_tree_, int64_t seed_, double change)
      : ncols(ncols_), seed(seed_), mtrys(mtrys_), mtrys_per_tree(mtrys_per_tree_), change_per_level(change), rand(seed_) {
    // T/DTree.java:60-86
    std::vector<int> active(ncols);
    for (int i = 0; i < ncols; i++) active[i] = i;
    int len = ncols;
    if (mtrys_per_tree < ncols) {
      Pcg col_rng((int64_t)((uint64_t)seed * 0xDA7AULL));
      for (int i = 0; i < mtrys_per_tree; i++) {
        if (len == 0) break;
        const int idx2 = col_rng.next_bounded(len);
        const int col = active[idx2];
        active[idx2] = active[--len];
        active[len] = col;
      }
      active.assign(active.begin() + len, active.end());
    }
    cols = active;
  }
  int len() const { return (int)ns.size(); }
  int actual_mtries() const {  // :56-58
    return std::min(std::max(1, d2i((double)mtrys * std::pow(change_per_level, depth))), ncols);
  }
  // UndecidedNode ctor + scoreCols (:443-499)
  int new_undecided(int pid, std::vector<Hist> hs) {
    if (ns.capacity() == ns.size()) ns.reserve(std::max<size_t>(1024, ns.size() * 4));  // a Node carries its split record (bitset included)
    ns.emplace_back();
    Node& n = ns.back();
    n.kind = 0;
    n.pid = pid;
    n.nid = len() - 1;
    n.hs = std::move(hs);
    const int hslen = (int)n.hs.size();
    if (actual_mtries() == hslen && mtrys_per_tree == hslen) {
      n.score_cols_null = true;
    } else {
      std::vector<int>& c = draw_scratch;
      c.resize(cols.size());
      int l = 0;
      for (size_t i = 0; i < cols.size(); i++) {
        if (!n.hs[cols[i]].present) continue;
        c[l++] = cols[i];
      }
      const int choices = l;
      const int mtries = actual_mtries();
      if (choices > 0)
        for (int i = 0; i < mtries; i++) {
          if (l == 0) break;
          const int idx2 = rand.next_bounded(l);
          const int col = c[idx2];
          c[idx2] = c[--l];
          c[l] = col;
        }
      n.score_cols_null = false;
      n.score_cols.assign(c.begin() + l, c.begin() + choices);
    }
    return n.nid;
  }
  // an UndecidedNode whose histograms nobody will read (children made by the last level of the tree: they turn into leaves)
  int new_undecided_stub(int pid) {
    if (ns.capacity() == ns.size()) ns.reserve(std::max<size_t>(1024, ns.size() * 4));
    ns.emplace_back();
    Node& n = ns.back();
    n.kind = 0;
    n.pid = pid;
    n.nid = len() - 1;
    return n.nid;
  }
  void do_not_split(int nid) {  // :504-511
    const int pid = ns[nid].pid;
    if (pid == kNoParent) return;
    for (int i = 0; i < 2; i++)
      if (ns[pid].nids[i] == nid) {
        ns[pid].nids[i] = kUndecidedChild;
        dirty.push_back(pid);
        return;
      }
  }
  int new_leaf(int pid) {
    if (ns.capacity() == ns.size()) ns.reserve(std::max<size_t>(1024, ns.size() * 4));
    ns.emplace_back();
    Node& n = ns.back();
    n.kind = 2;
    n.pid = pid;
    n.nid = len() - 1;
    leaves++;
    return n.nid;
  }
};

struct Capture {
  int tree, k, leaf, nleaves;
  std::vector<h2ot_hist_desc> descs;
  std::vector<double> vals, min2, max_in;
  std::vector<int64_t> offs;
  std::vector<uint8_t> has;
  std::vector<int32_t> nids;
};

struct FinishedTree {
  std::vector<Node> nodes;
  int leaves_start = 0;
};

}  // namespace

struct h2ot_gbm {
  h2ot_gbm_params p{};
  h2ot_engine* eng = nullptr;
  int64_t nrows = 0, nrows_global = 0, row_offset = 0;
  int ncols = 0, K = 1, nclass = 1;
  bool has_w = false, has_off = false, inited = false;
  int device = 0;
  int world = 1, rank = 0;
  std::vector<ColMeta> cols;
  bool have_y = false, have_w = false, have_off = false;
  double init_f = 0;
  std::vector<double> prior_class_dist;  // DRF classification (Model.Output._priorClassDist)
  double work_bound = 1.0;               // DRF: max |Work| (fixed for the whole build)
  int mtry = 0, mtry_per_tree = 0;
  Pcg rand{0};
  int ntrees_built = 0;
  std::vector<std::vector<FinishedTree>> trees;  // [tree][k]
  std::vector<float> varimp;
  bool capture = false;
  double pred_noise_bandwidth = 0;            // GBMParameters._pred_noise_bandwidth (T/gbm/GBMModel.java:28)
  std::vector<double> sample_rate_per_class;  // SharedTreeParameters._sample_rate_per_class (empty: null)
  std::vector<Capture> captures;
  h2ot_gbm_timing timing{};
  // scratch
  // per class-tree scratch of the level in flight (the K levels of a multinomial iteration are enqueued side by side)
  struct LevelScratch {
    std::vector<h2ot_hist_desc> descs;
    std::vector<double> split_pts;  // histogram_type Random: the level's split points
    std::vector<double> min2, max_in, w_na;
    std::vector<uint8_t> has;
    std::vector<h2ot_split> splits;
    std::vector<char> col_active;
    double t0 = 0, t_descs_done = 0;
  };
  std::vector<LevelScratch> scr;
  // iterations whose trees are on the device (built, fitted, added to the predictions) but not yet mirrored on the host
  struct PendingIter {
    std::vector<std::unique_ptr<Tree>> kt;
    std::vector<int> ncol_active;
  };
  std::deque<PendingIter> pending;
  // gaussian: the fixed-point bound of a tree's residuals follows from the PREVIOUS tree's measured maximum (see residual_bound)
  double y_abs_max = 0;      // max |response| over all rows
  double prev_absmax = -1;   // max |Work| measured at the previous iteration (-1: none yet)
  double prev_lr = 0;        // that iteration's effective learn rate
};

namespace {

// T/DHistogram.java:526-554
bool use_int_opt(const ColMeta& v, const h2ot_gbm_params& p) {
  if (p.histogram_type == 1) return false;  // Random: no integer fast path (T/DHistogram.java:537-542)
  if (v.is_cat) return v.cardinality < p.nbins_cats;
  if (v.is_int) {
    const double len = v.max - v.min;
    return len < p.nbins && d2i(v.max) - d2i(v.min) == d2i(len);
  }
  return false;
}

// T/DHistogram.java:474-502
std::vector<Hist> initial_hist(const h2ot_gbm* g, int nbins, int64_t seed, bool check_float) {
  std::vector<Hist> hs(g->ncols);
  for (int c = 0; c < g->ncols; c++) {
    const ColMeta& v = g->cols[c];
    const double min_in = v.is_cat ? 0 : std::fmax(v.min, -DBL_MAX);
    const double max_in = v.is_cat ? v.cardinality - 1 : std::fmin(v.max, DBL_MAX);
    const double max_ex = v.is_cat ? v.cardinality : Hist::find_max_ex_of(max_in, v.is_int ? 1 : 0);
    const bool all_na = v.nacnt == g->nrows_global;
    const bool is_const = v.min == v.max && v.nacnt == 0;
    if (all_na || is_const) continue;
    const uint8_t type = (uint8_t)(v.is_cat ? 2 : (v.is_int ? 1 : 0));
    Hist h;
    if (make_hist(h, nbins, g->p.nbins_cats, type, min_in, max_ex, use_int_opt(v, g->p), v.nacnt > 0, seed, check_float)) hs[c] = h;
  }
  return hs;
}

// Split.nextLevelHistos (T/DTree.java:312-417; no constraints). Empty result == Java null.
// any_only: stop at the first column that gets a histogram and return a one-element vector (the caller only asks whether the child
// exists -- the last level of a tree, whose children become leaves whatever their histograms would have been).
std::vector<Hist> next_level_histos(const h2ot_gbm* g, const h2ot_split& s, const std::vector<Hist>& cur, int way, double splat,
                                    bool any_only = false) {
  const double n = way == 0 ? s.n0 : s.n1;
  if (n < g->p.min_rows) return {};
  const double se = way == 0 ? s.se0 : s.se1;
  if (se <= 1e-30) return {};
  int cnt = 0;
  std::vector<Hist> nh(any_only ? 1 : cur.size());
  for (size_t j = 0; j < cur.size(); j++) {
    const Hist& h = cur[j];
    if (!h.present) continue;
    const int adj_nbins = std::max(h.nbin >> 1, g->p.nbins);
    double min, max_ex;
    if (!h.has_vals || s.equal > 1) {
      min = h.min;
      max_ex = h.max_ex;
    } else {
      min = h.min2;
      if (h.max_in == min) continue;
      max_ex = h.find_max_ex();
    }
    if (s.nasplit == H2OT_NA_VS_REST && way == 1) continue;
    if (s.col == (int)j) {
      if (s.equal == 0) {
        double split = splat;
        if (h.is_int > 0) split = (float)std::ceil(split);
        if (s.nasplit != H2OT_NA_VS_REST) {
          if (way == 0) max_ex = split;
          else min = split;
        }
      } else if (s.equal == 1) {
        if (way == 1) continue;
      }
    }
    if (min > max_ex) continue;
    if (within_one_small_ulp(min, max_ex)) continue;
    {  // Double.isInfinite(adj_nbins / (maxEx - min)): only a denormal-sized range can overflow the quotient -- no division otherwise
      const double range = max_ex - min;
      if (range < 1e-280 && std::isinf(adj_nbins / range)) continue;
    }
    if (h.is_int > 0 && !(min + 1 < max_ex)) continue;
    const bool has_nas = ((s.nasplit == H2OT_NA_LEFT && way == 0) || (s.nasplit == H2OT_NA_RIGHT && way == 1)) && h.has_na_bin();
    Hist c;
    // DHistogram.make (:504-512): seed = h._seed*0xDECAF+(way+1)
    if (!make_hist(c, adj_nbins, g->p.nbins_cats, h.is_int, min, max_ex, h.int_opt, has_nas,
                   (int64_t)((uint64_t)h.seed * 0xDECAFULL + (uint64_t)(way + 1)), h.check_float))
      continue;
    if (any_only) {
      nh[0] = c;
      return nh;
    }
    nh[j] = c;
    cnt++;
  }
  if (cnt == 0) return {};
  return nh;
}

void to_device_node(const Node& n, h2ot_decided_node& d, std::vector<uint8_t>& pool) {
  memset(&d, 0, sizeof(d));
  d.nid = n.nid;
  d.pid = n.pid;
  d.col = n.has_split ? n.split.col : -1;
  d.equal = n.has_split ? n.split.equal : 0;
  d.nasplit = n.has_split ? n.split.nasplit : H2OT_NA_NONE;
  d.splat = n.splat;
  d.child[0] = n.nids[0];
  d.child[1] = n.nids[1];
  if (n.has_split && n.split.equal >= 2) {
    d.bs_off = (int32_t)pool.size();
    d.bs_nbits = n.split.bs_nbits;
    d.bs_bitoff = n.split.bs_bitoff;
    pool.insert(pool.end(), n.split.bs, n.split.bs + n.split.bs_nbytes);
  }
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Builds without per-node column sampling and with uniform-adaptive histograms can run all levels of a tree back to back on the
// device (h2ot_tree_levels_begin / _end): _scoreCols == null for every node (T/DTree.java:455-461), so no host-side draw sits
// between two levels. H2OT_DEVICE_LOOP=0 keeps the level-by-level host loop (the two must grow identical trees).
// laplace / quantile: leaves are fitted with the residuals' median / alpha-quantile (fitBestConstantsQuantile, T/gbm/GBM.java:522-527, 729-751)
bool quantile_family(const h2ot_gbm* g) {
  return g->p.algo == H2OT_ALGO_GBM && (g->p.family == H2OT_FAMILY_LAPLACE || g->p.family == H2OT_FAMILY_QUANTILE);
}
double quantile_prob(const h2ot_gbm* g) {
  if (g->p.family == H2OT_FAMILY_LAPLACE) return 0.5;
  return g->p.tweedie_power > 0.0 && g->p.tweedie_power < 1.0 ? g->p.tweedie_power : 0.5;  // (as the engine reads the slot)
}

bool use_device_loop(const h2ot_gbm* g) {
  const char* e = getenv("H2OT_DEVICE_LOOP");  // read per tree: tests switch it between two builds of one process
  const bool off = e != nullptr && atoi(e) == 0;
  return !off && !g->capture && g->p.algo == H2OT_ALGO_GBM && g->p.histogram_type == 0 && g->mtry == g->ncols &&
         g->mtry_per_tree == g->ncols && g->p.col_sample_rate_change_per_level == 1.0 && g->p.max_depth >= 1 && g->p.max_depth <= 8;
}

void hist_to_desc(const Hist& h, int order, h2ot_hist_desc& d) {
  d = h2ot_hist_desc{};
  if (!h.present) return;
  d.min = h.min;
  d.maxEx = h.max_ex;
  d.step = h.step;
  d.nbin = h.nbin;
  d.order = (int16_t)order;
  d.flags = H2OT_DESC_ACTIVE | (h.is_int > 0 ? H2OT_DESC_ISINT : 0) | (h.is_int == 2 ? H2OT_DESC_ISCAT : 0) |
            (h.int_opt ? H2OT_DESC_INTOPT : 0) | (h.check_float ? H2OT_DESC_CHECKFLOAT : 0) | (h.init_na ? H2OT_DESC_INITNA : 0);
}

// DecidedNode record in the engine's static numbering (device-loop trees): decided nodes under their slot, LeafNodes under
// leaf_base + (leaf nid - first leaf nid)
void to_device_node_static(const std::vector<Node>& ns, int nid, int first_leaf, int leaf_base, h2ot_decided_node& d, std::vector<uint8_t>& pool) {
  const Node& n = ns[(size_t)nid];
  memset(&d, 0, sizeof(d));
  d.nid = n.sid;
  d.pid = n.pid == kNoParent ? -1 : ns[(size_t)n.pid].sid;
  d.col = n.has_split ? n.split.col : -1;
  d.equal = n.has_split ? n.split.equal : 0;
  d.nasplit = n.has_split ? n.split.nasplit : H2OT_NA_NONE;
  d.splat = n.splat;
  for (int i = 0; i < 2; i++) {
    const int c = n.nids[i];
    d.child[i] = c < 0 ? -1 : (ns[(size_t)c].kind == 2 ? leaf_base + (c - first_leaf) : ns[(size_t)c].sid);
  }
  if (n.has_split && n.split.equal >= 2) {
    d.bs_off = (int32_t)pool.size();
    d.bs_nbits = n.split.bs_nbits;
    d.bs_bitoff = n.split.bs_bitoff;
    pool.insert(pool.end(), n.split.bs, n.split.bs + n.split.bs_nbytes);
  }
}

// All levels of the K class-trees of one iteration on the device, then the host's mirror of each tree: nodes in H2O's creation
// order (level by level, leaf order, left child first -- what build_layer_k's loop over the leaves produces), splits, varimp.
int enqueue_levels_on_device(h2ot_gbm* g, std::vector<std::unique_ptr<Tree>>& kt, std::vector<int>& ncol_active) {
  const int K = g->K, C = g->ncols, D = g->p.max_depth;
  std::vector<h2ot_hist_desc> descs((size_t)C);
  ncol_active.assign((size_t)K, 0);
  for (int k = 0; k < K; k++) {
    const Node& root = kt[k]->ns[0];
    for (int c = 0; c < C; c++) {
      hist_to_desc(root.hs.empty() ? Hist{} : root.hs[(size_t)c], c, descs[(size_t)c]);
      ncol_active[(size_t)k] += (descs[(size_t)c].flags & H2OT_DESC_ACTIVE) ? 1 : 0;
    }
    GTRY(h2ot_tree_levels_begin(g->eng, k, D, descs.data(), g->p.min_rows, g->p.min_split_improvement, g->p.nbins, g->p.nbins_cats));
  }
  return H2OT_OK;
}

// Waits for the K trees (oldest in flight per class) and mirrors them. slot_pred (optional): [K][S] leaf values by static child slot.
int mirror_levels_on_device(h2ot_gbm* g, std::vector<std::unique_ptr<Tree>>& kt, const std::vector<int>& ncol_active,
                            std::vector<float>* slot_pred) {
  const int K = g->K, D = g->p.max_depth;
  const size_t S = ((size_t)1 << (D + 1)) - 1, nS = ((size_t)1 << D) - 1;
  std::vector<h2ot_split> splits(nS);
  std::vector<uint8_t> exists(S);
  if (slot_pred) slot_pred->assign((size_t)K * S, 0.f);
  // variable importance is a float sum whose order the reference fixes: per level, class-tree by class-tree, leaf by leaf
  // (ScoreBuildOneTree.onCompletion of the K trees of one buildLayer, T/SharedTree.java:563-607, 697-700)
  std::vector<std::vector<std::pair<int, float>>> vi((size_t)D * K);
  for (int k = 0; k < K; k++) {
    GTRY(h2ot_tree_levels_end2(g->eng, k, splits.data(), exists.data(), slot_pred ? slot_pred->data() + (size_t)k * S : nullptr));
    Tree& t = *kt[k];
    t.ns[0].sid = 0;
    std::vector<int> cur{0}, next;
    for (int depth = 0; depth < D && !cur.empty(); depth++) {
      g->timing.levels++;
      g->timing.hist_rows_cols += (double)g->nrows_global * (double)ncol_active[(size_t)k];
      bool did = false;
      next.clear();
      for (int nid : cur) {
        const int sid = t.ns[(size_t)nid].sid;
        const h2ot_split& sp = splits[(size_t)sid];
        {
          Node& n = t.ns[(size_t)nid];
          n.kind = 1;
          n.hs.clear();
          n.score_cols.clear();
          if (sp.col < 0) {
            n.has_split = false;
            n.splat = NAN;
            n.nids[0] = n.nids[1] = kUndecidedChild;
            t.do_not_split(nid);
            continue;
          }
          n.has_split = true;
          n.split = sp;
          n.splat = sp.splat;
        }
        did = true;
        vi[(size_t)depth * K + k].emplace_back(sp.col, (float)(sp.se - (sp.se0 + sp.se1)));
        for (int way = 0; way < 2; way++) {
          const int csid = 2 * sid + 1 + way;
          int child = kUndecidedChild;
          if (exists[(size_t)csid]) {
            child = t.new_undecided_stub(nid);  // (may move t.ns)
            t.ns[(size_t)child].sid = csid;
            next.push_back(child);
          }
          t.ns[(size_t)nid].nids[way] = child;
        }
      }
      if (did) t.depth++;
      cur.swap(next);
    }
    t.dirty.clear();
  }
  for (const auto& lvl : vi)
    for (const auto& e : lvl) g->varimp[e.first] += e.second;
  return H2OT_OK;
}

int build_levels_on_device(h2ot_gbm* g, std::vector<std::unique_ptr<Tree>>& kt) {
  const double t0 = now_s();
  std::vector<int> ncol_active;
  GTRY(enqueue_levels_on_device(g, kt, ncol_active));
  GTRY(mirror_levels_on_device(g, kt, ncol_active, nullptr));
  g->timing.level_s += now_s() - t0;
  return H2OT_OK;
}

// leaf insertion (T/gbm/GBM.java:655-679); returns the id of the first LeafNode. slot_pred: the leaves' values by static child
// slot when the device fitted them (pipelined trees), else nullptr
int insert_leaves(Tree& t, const float* slot_pred) {
  const int leaf = t.len();
  for (int nid = 0; nid < leaf; nid++) {
    if (t.ns[nid].kind != 1) continue;
    if (!t.ns[nid].has_split) {
      if (nid == 0) {  // the root never split: it becomes a LeafNode in place
        Node n;
        n.kind = 2;
        n.pid = kNoParent;
        n.nid = 0;
        t.ns[0] = n;
        t.leaves++;
      }
      continue;
    }
    for (int i = 0; i < 2; i++) {
      const int cn = t.ns[nid].nids[i];
      if (cn == kUndecidedChild || t.ns[cn].kind == 0 || (t.ns[cn].kind == 1 && !t.ns[cn].has_split)) {
        const int ln = t.new_leaf(nid);  // (may move t.ns)
        t.ns[nid].nids[i] = ln;
        if (slot_pred) t.ns[ln].pred = slot_pred[2 * t.ns[nid].sid + 1 + i];
      }
    }
  }
  return leaf;
}

// One ScoreBuildHistogram2 pass + ScoreBuildOneTree.onCompletion for class-tree k (T/SharedTree.java:661-710).
// last_level: the tree stops after this level (depth == max_depth - 1), so the children only need to exist with the right node
// ids; their histograms and column draws (tree._rand is not used again) are skipped -- the widest level of a deep tree otherwise
// spends more host time describing children nobody splits than the GPU spends on the level.
// phase 0: the whole level; 1: descriptors + enqueue only (h2ot_level_step_begin); 2: wait for the results + onCompletion.
int build_layer_k(h2ot_gbm* g, int k, Tree& tree, int& leaf_offset, bool& did_split, bool last_level = false, int phase = 0) {
  if ((int)g->scr.size() < g->K) g->scr.resize((size_t)g->K);
  h2ot_gbm::LevelScratch& sc = g->scr[(size_t)k];
  did_split = false;
  const int C = g->ncols;
  const int tmax = tree.len();
  const int L = tmax - leaf_offset;
  if (L <= 0) return H2OT_OK;
  const size_t LC = (size_t)L * C;
  std::vector<char>& col_active = sc.col_active;
  if (phase != 2) {
  sc.t0 = now_s();
  // descriptors of the undecided leaves
  sc.descs.assign((size_t)L * C, h2ot_hist_desc{});
  sc.split_pts.clear();
  int64_t active_pairs = 0;
  col_active.assign(C, 0);
  for (int l = 0; l < L; l++) {
    const Node& u = tree.ns[leaf_offset + l];
    std::vector<int16_t> order(C, -1);
    if (u.score_cols_null)
      for (int c = 0; c < C; c++) order[c] = (int16_t)c;
    else
      for (size_t i = 0; i < u.score_cols.size(); i++) order[u.score_cols[i]] = (int16_t)i;
    for (int c = 0; c < C; c++) {
      h2ot_hist_desc& d = sc.descs[(size_t)l * C + c];
      if (u.hs.empty() || !u.hs[c].present || order[c] < 0) continue;
      const Hist& h = u.hs[c];
      d.min = h.min;
      d.maxEx = h.max_ex;
      d.step = h.step;
      d.nbin = h.nbin;
      d.order = order[c];
      d.flags = H2OT_DESC_ACTIVE | (h.is_int > 0 ? H2OT_DESC_ISINT : 0) | (h.is_int == 2 ? H2OT_DESC_ISCAT : 0) |
                (h.int_opt ? H2OT_DESC_INTOPT : 0) | (h.check_float ? H2OT_DESC_CHECKFLOAT : 0) | (h.init_na ? H2OT_DESC_INITNA : 0);
      if (g->p.histogram_type == 1) {  // Random: DHistogram.init draws the node's split points (T/DHistogram.java:347-352, 742-749)
        const double sd = ((h.step + 0.324) * h.min + 8.3425) + 89.342 * h.max_ex;
        int64_t bits;
        memcpy(&bits, &sd, sizeof bits);
        Pcg rng((int64_t)((uint64_t)bits + 0xDECAFULL * (uint64_t)h.nbin + 0xC0FFEEULL * (uint64_t)h.is_int + (uint64_t)h.seed));
        const size_t off = sc.split_pts.size();
        sc.split_pts.resize(off + (size_t)h.nbin);
        double* pts = sc.split_pts.data() + off;
        pts[0] = 0;
        for (int i = 1; i < h.nbin; i++) pts[i] = (double)(rng.next_float() * (float)h.nbin);
        std::sort(pts, pts + h.nbin);
        d.flags |= H2OT_DESC_SPLITPTS;
        d.splitpts_off = (int32_t)off;
        d.splitpts_n = h.nbin;
      }
      active_pairs++;
      col_active[c] = 1;
    }
  }
  if (g->p.histogram_type == 1) GTRY(h2ot_set_split_points(g->eng, k, sc.split_pts.data(), (int64_t)sc.split_pts.size()));
  sc.t_descs_done = now_s();
  // decided nodes the device does not have yet
  std::vector<h2ot_decided_node> dn;
  std::vector<uint8_t> pool;
  std::sort(tree.dirty.begin(), tree.dirty.end());
  tree.dirty.erase(std::unique(tree.dirty.begin(), tree.dirty.end()), tree.dirty.end());
  for (int nid : tree.dirty) {
    dn.emplace_back();
    to_device_node(tree.ns[nid], dn.back(), pool);
  }
  tree.dirty.clear();
  sc.min2.resize(LC);
  sc.max_in.resize(LC);
  sc.w_na.resize(LC);
  sc.has.resize(LC);
  sc.splits.resize(L);
  if (g->capture) {
    GTRY(h2ot_build_level(g->eng, k, leaf_offset, L, (int)dn.size(), dn.data(), pool.data(), (int64_t)pool.size(), sc.descs.data()));
    if (g->world > 1) GTRY(h2ot_allreduce_level(g->eng, k));
    Capture cap;
    cap.tree = g->ntrees_built;
    cap.k = k;
    cap.leaf = leaf_offset;
    cap.nleaves = L;
    cap.descs = sc.descs;
    int64_t nvals = 0;
    for (const h2ot_hist_desc& d : sc.descs)
      if (d.flags & H2OT_DESC_ACTIVE) nvals += (int64_t)(d.nbin + 1) * 3;
    cap.vals.resize((size_t)std::max<int64_t>(1, nvals));
    cap.offs.resize(LC);
    GTRY(h2ot_finish_level(g->eng, k, cap.vals.data(), (int64_t)cap.vals.size(), cap.offs.data(), sc.min2.data(), sc.max_in.data(),
                           sc.w_na.data(), sc.has.data()));
    cap.min2 = sc.min2;
    cap.max_in = sc.max_in;
    cap.has = sc.has;
    cap.nids.resize((size_t)g->nrows);
    GTRY(h2ot_get_nids(g->eng, k, cap.nids.data()));
    g->captures.push_back(std::move(cap));
    GTRY(h2ot_find_splits(g->eng, k, g->p.min_rows, g->p.min_split_improvement, sc.splits.data()));
  } else {
    GTRY(h2ot_level_step_begin(g->eng, k, leaf_offset, L, (int)dn.size(), dn.data(), pool.data(), (int64_t)pool.size(), sc.descs.data(),
                               g->p.min_rows, g->p.min_split_improvement));
    if (phase == 1) return H2OT_OK;
  }
  }  // phase != 2
  if (!g->capture)
    GTRY(h2ot_level_step_end(g->eng, k, sc.min2.data(), sc.max_in.data(), sc.w_na.data(), sc.has.data(), sc.splits.data()));
  const double t0 = sc.t0, t_descs_done = sc.t_descs_done;
  const double t_call_end = now_s();
  g->timing.level_s += t_call_end - t0;
  g->timing.levels++;
  int64_t ncol_active = 0;
  for (int c = 0; c < C; c++) ncol_active += col_active[c];
  g->timing.hist_rows_cols += (double)g->nrows_global * (double)ncol_active;

  // onCompletion: DecidedNode ctor per leaf (T/DTree.java:703-726). First the children's histogram descriptors
  // (nextLevelHistos: 2 x ncols DHistogram.make per split), then node ids and the per-node column draws (tree._rand) in leaf
  // order. (Preparing the descriptors on several host threads was measured and dropped: the hand-over of the descriptor
  // vectors between cores cost what the threads saved, 1.42 vs 1.16 ms at 614 leaves.)
  struct Kids { std::vector<Hist> hs, nh[2]; };
  std::vector<Kids> kids((size_t)L);
  for (int l = 0; l < L; l++) {
    const int nid = leaf_offset + l;
    Kids& kd = kids[(size_t)l];
    kd.hs = std::move(tree.ns[nid].hs);
    std::vector<Hist>& hs = kd.hs;
    for (int c = 0; c < C && !hs.empty(); c++) {
      Hist& h = hs[c];
      const size_t i = (size_t)l * C + c;
      if (!h.present || !(sc.descs[i].flags & H2OT_DESC_ACTIVE)) continue;
      h.has_vals = sc.has[i] != 0;
      h.min2 = sc.min2[i];
      h.max_in = sc.max_in[i];
      h.w_na = sc.w_na[i];
    }
    const h2ot_split& s = sc.splits[l];
    if (s.col < 0 || hs.empty()) continue;
    for (int way = 0; way < 2; way++) kd.nh[way] = next_level_histos(g, s, hs, way, (double)s.splat, last_level);
  }
  const double t_kids_done = now_s();
  for (int l = 0; l < L; l++) {
    const int nid = leaf_offset + l;
    Kids& kd = kids[(size_t)l];
    Node& dnode = tree.ns[nid];
    dnode.kind = 1;
    dnode.hs.clear();
    dnode.score_cols.clear();
    const h2ot_split& s = sc.splits[l];
    tree.dirty.push_back(nid);
    if (s.col < 0 || kd.hs.empty()) {
      dnode.has_split = false;
      dnode.splat = NAN;
      dnode.nids[0] = dnode.nids[1] = kUndecidedChild;
      tree.do_not_split(nid);
      continue;
    }
    dnode.has_split = true;
    dnode.split = s;
    dnode.splat = s.splat;
    did_split = true;
    g->varimp[s.col] += (float)(s.se - (s.se0 + s.se1));  // T/SharedTree.java:697-700
    for (int way = 0; way < 2; way++) {
      const int child = kd.nh[way].empty() ? kUndecidedChild
                                           : (last_level ? tree.new_undecided_stub(nid) : tree.new_undecided(nid, std::move(kd.nh[way])));
      tree.ns[nid].nids[way] = child;
    }
  }
  leaf_offset = tmax;
  if (did_split) tree.depth++;
  static const bool trace = getenv("H2OT_TRACE") != nullptr;
  if (trace)
    fprintf(stderr, "[h2ot-host] L=%d: descs+call %.3f ms (call starts %.3f ms after entry), decide %.3f ms (children's descriptors %.3f)\n", L,
            1e3 * (t_call_end - t0), 1e3 * (t_descs_done - t0), 1e3 * (now_s() - t_call_end), 1e3 * (t_kids_done - t_call_end));
  return H2OT_OK;
}

int allreduce(h2ot_gbm* g, double* v, int64_t n, int op) {
  if (g->world <= 1) return H2OT_OK;
  return h2ot_allreduce_host_f64(g->eng, v, n, op);
}

// AddTreeContributions' per-node factor 1 + nextGaussian * pred_noise_bandwidth, the generator re-seeded per (tree, class, node)
// (T/gbm/GBM.java:1450, 1458-1463; the second factor of the seed is Java int arithmetic widened with its sign)
double noise_factor(const h2ot_gbm* g, int k, int nid) {
  const int32_t f2 = (int32_t)(0xFAAAAAABu + (uint32_t)(k * g->p.ntrees + g->ntrees_built));
  const int64_t baseseed = (int64_t)(((uint64_t)0xDECAF + (uint64_t)g->p.seed) * (uint64_t)(int64_t)f2);
  Pcg r(g->p.seed);
  r.set_seed((int64_t)((uint64_t)baseseed + (uint64_t)(int64_t)nid));
  return 1.0 + r.next_gaussian_after_seed() * g->pred_noise_bandwidth;
}

// ---- trees in a pipeline: the device runs an iteration start to finish (residuals, levels, leaf insertion, GammaPass, leaf values,
// AddTreeContributions) with nothing waiting on the host, and the host mirrors iteration t while the device builds t + 1.
// What lets the host step aside: static node ids (the device loop), residual bounds known before the residuals are (|y - p| <= 1;
// gaussian: residual_bound below, from the previous tree's maximum), leaf values that need no transcendental (poisson / gamma /
// tweedie take a log), no per-node random numbers (pred_noise_bandwidth seeds by H2O node id). H2OT_PIPELINE=0 switches it off;
// the trees are the same either way.
bool use_pipeline(const h2ot_gbm* g) {
  const char* e = getenv("H2OT_PIPELINE");
  const bool off = e != nullptr && atoi(e) == 0;
  const int f = g->p.family;
  return !off && use_device_loop(g) && g->pred_noise_bandwidth == 0 &&
         (f == H2OT_FAMILY_BERNOULLI || f == H2OT_FAMILY_MULTINOMIAL || (f == H2OT_FAMILY_GAUSSIAN && !g->has_off));
}

// Gaussian residual bound for the fixed-point scale of iteration t, known before iteration t's residuals have been looked at:
// a leaf's value is a weighted mean of its rows' residuals, |gamma| <= max |r|, so after adding learn_rate x gamma every residual
// obeys |r'| <= (1 + learn_rate) max |r|; F is rounded to float once per tree, which moves a residual by at most 2^-24 |F| and
// |F| <= max |y| + max |r|. The first iteration uses its own measured maximum. Both the pipelined and the tree-at-a-time build
// take their scale from here, so they accumulate the same integers.
double residual_bound(const h2ot_gbm* g, double measured_now) {
  if (g->prev_absmax < 0) return measured_now;
  return (1.0 + g->prev_lr) * g->prev_absmax + std::ldexp(g->y_abs_max + g->prev_absmax, -23);
}

int mirror_oldest_iteration(h2ot_gbm* g) {
  if (g->pending.empty()) return H2OT_OK;
  const double t0 = now_s();
  h2ot_gbm::PendingIter it = std::move(g->pending.front());
  g->pending.pop_front();
  const int K = g->K;
  const size_t S = ((size_t)1 << (g->p.max_depth + 1)) - 1;
  std::vector<float> slot_pred;
  GTRY(mirror_levels_on_device(g, it.kt, it.ncol_active, &slot_pred));
  std::vector<FinishedTree> outs(K);
  for (int k = 0; k < K; k++) {
    Tree& t = *it.kt[k];
    outs[k].leaves_start = insert_leaves(t, slot_pred.data() + (size_t)k * S);
    outs[k].nodes = std::move(t.ns);
    for (Node& n : outs[k].nodes) n.hs.clear();
  }
  g->trees.push_back(std::move(outs));
  g->timing.level_s += now_s() - t0;
  return H2OT_OK;
}

int flush_pending(h2ot_gbm* g) {
  while (!g->pending.empty()) GTRY(mirror_oldest_iteration(g));
  return H2OT_OK;
}

int build_next_k_trees_pipelined(h2ot_gbm* g) {
  const int K = g->K;
  const double t0 = now_s();
  const bool gaussian = g->p.family == H2OT_FAMILY_GAUSSIAN;
  const double eff_lr_now = g->p.learn_rate * std::pow(g->p.learn_rate_annealing, (double)(g->ntrees_built - 1));
  GTRY(h2ot_compute_residuals_async(g->eng, gaussian ? 1 : 0));
  double bound = 1.0, a_now = -1;
  if (gaussian) {
    // the maximum of THIS iteration's residuals is fetched after the tree has been enqueued (it serves the next iteration);
    // only the very first iteration has to wait for its own
    if (g->prev_absmax < 0) GTRY(h2ot_residual_bounds_wait(g->eng, &a_now, nullptr));
    bound = residual_bound(g, a_now);
    if (!(bound > 0)) bound = 1.0;
    const double gb[2] = {bound, 1.0};  // |w z| / w <= bound, w / w = 1
    GTRY(h2ot_set_gamma_bounds(g->eng, gb));
  } else {
    const double gb[2] = {1.0, 0.25};  // |w z| / w <= 1, p (1 - p) <= 1/4
    GTRY(h2ot_set_gamma_bounds(g->eng, gb));
  }
  h2ot_gbm::PendingIter it;
  it.kt.resize((size_t)K);
  const int adj_nbins = std::max(g->p.nbins_top_level, g->p.nbins);
  const int64_t rseed = g->rand.next_long();
  for (int k = 0; k < K; k++) {
    it.kt[k] = std::make_unique<Tree>(g->ncols, g->mtry, g->mtry_per_tree, rseed, g->p.col_sample_rate_change_per_level);
    it.kt[k]->new_undecided(kNoParent, initial_hist(g, adj_nbins, rseed, false));
    GTRY(h2ot_begin_tree(g->eng, k, rseed, g->p.sample_rate, bound));
  }
  GTRY(enqueue_levels_on_device(g, it.kt, it.ncol_active));
  const double m1class = (g->nclass > 1 && g->p.family == H2OT_FAMILY_MULTINOMIAL) ? (double)(g->nclass - 1) / g->nclass : 1.0;
  const double eff_lr = g->p.learn_rate * std::pow(g->p.learn_rate_annealing, (double)(g->ntrees_built - 1));
  for (int k = 0; k < K; k++) GTRY(h2ot_tree_finish_enqueue(g->eng, k, eff_lr * m1class, g->p.max_abs_leafnode_pred));
  g->pending.push_back(std::move(it));
  g->ntrees_built++;
  if (gaussian) {  // (returns once the previous tree and this iteration's residual kernel are done: where the mirror below waits anyway)
    if (a_now < 0) GTRY(h2ot_residual_bounds_wait(g->eng, &a_now, nullptr));
    g->prev_absmax = a_now;
    g->prev_lr = eff_lr_now;
  }
  const double t1 = now_s();
  g->timing.level_s += t1 - t0;
  while (g->pending.size() > 1) GTRY(mirror_oldest_iteration(g));
  static const bool trace = getenv("H2OT_TRACE_PIPE") != nullptr;
  if (trace) fprintf(stderr, "[h2ot pipe] tree %d: enqueue %.3f ms, wait + mirror of the previous %.3f ms\n", g->ntrees_built, 1e3 * (t1 - t0), 1e3 * (now_s() - t1));
  return H2OT_OK;
}

// GBM.buildNextKTrees (T/gbm/GBM.java:464-591)
int build_next_k_trees(h2ot_gbm* g) {
  if (use_pipeline(g)) return build_next_k_trees_pipelined(g);
  GTRY(flush_pending(g));
  const int K = g->K;
  std::vector<double> absmax(K + 2, 0.0);
  GTRY(h2ot_compute_residuals(g->eng, absmax.data()));
  if (g->p.family == H2OT_FAMILY_BERNOULLI || g->p.family == H2OT_FAMILY_MULTINOMIAL) {
    // bounds known in advance (|y - p| <= 1, p (1 - p) <= 1/4): the same fixed-point scale as the pipelined build, hence the
    // same bits, and no all-reduce of the observed maxima
    for (int k = 0; k < K; k++) absmax[k] = 1.0;
    absmax[K] = 1.0;
    absmax[K + 1] = 0.25;
  } else {
    GTRY(h2ot_get_gamma_bounds(g->eng, absmax.data() + K));
    GTRY(allreduce(g, absmax.data(), K + 2, 1));
    if (g->p.family == H2OT_FAMILY_GAUSSIAN && !g->has_off && g->p.algo == H2OT_ALGO_GBM) {
      // the scale the pipelined build would use for this tree (residual_bound): same integers, same bits
      const double a_now = absmax[0];
      const double b = residual_bound(g, a_now);
      absmax[0] = b > 0 ? b : 1.0;
      absmax[K] = absmax[0];
      absmax[K + 1] = 1.0;
      g->prev_absmax = a_now;
      g->prev_lr = g->p.learn_rate * std::pow(g->p.learn_rate_annealing, (double)(g->ntrees_built - 1));
    }
  }
  GTRY(h2ot_set_gamma_bounds(g->eng, absmax.data() + K));
  std::vector<std::unique_ptr<Tree>> kt(K);
  std::vector<int> leaves(K, 0);
  const int adj_nbins = std::max(g->p.nbins_top_level, g->p.nbins);
  const int64_t rseed = g->rand.next_long();
  for (int k = 0; k < K; k++) {
    kt[k] = std::make_unique<Tree>(g->ncols, g->mtry, g->mtry_per_tree, rseed, g->p.col_sample_rate_change_per_level);
    kt[k]->new_undecided(kNoParent, initial_hist(g, adj_nbins, rseed, false));
    GTRY(h2ot_begin_tree(g->eng, k, rseed, g->p.sample_rate, absmax[k]));
  }
  const bool on_device = use_device_loop(g);
  if (on_device) GTRY(build_levels_on_device(g, kt));
  for (int depth = 0; depth < g->p.max_depth && !on_device; depth++) {
    bool any = false;
    const bool last = depth == g->p.max_depth - 1;
    if (K > 1 && !g->capture) {
      // the K class-trees' levels are enqueued back to back; while the host turns tree k's splits into its next level the
      // device works on the levels of the trees behind it
      bool did = false;
      for (int k = 0; k < K; k++) GTRY(build_layer_k(g, k, *kt[k], leaves[k], did, last, 1));
      for (int k = 0; k < K; k++) {
        did = false;
        GTRY(build_layer_k(g, k, *kt[k], leaves[k], did, last, 2));
        any = any || did;
      }
    } else {
      for (int k = 0; k < K; k++) {
        bool did = false;
        GTRY(build_layer_k(g, k, *kt[k], leaves[k], did, last));
        any = any || did;
      }
    }
    if (!any) break;
  }
  // leaf insertion (:655-679)
  for (int k = 0; k < K; k++) leaves[k] = insert_leaves(*kt[k], nullptr);
  // GammaPass + fitBestConstants (:1286-1407, 813-847)
  const double m1class = (g->nclass > 1 && g->p.family == H2OT_FAMILY_MULTINOMIAL) ? (double)(g->nclass - 1) / g->nclass : 1.0;
  const double eff_lr = g->p.learn_rate * std::pow(g->p.learn_rate_annealing, (double)(g->ntrees_built - 1));
  for (int k = 0; k < K; k++) {
    Tree& t = *kt[k];
    const int nleaf = t.len() - leaves[k];
    if (t.ns[0].kind == 2 || nleaf <= 0) continue;
    std::vector<h2ot_decided_node> dn;
    std::vector<uint8_t> pool;
    const int leaf_base = on_device ? (1 << (g->p.max_depth + 1)) - 1 : leaves[k];  // the engine's id of the first LeafNode
    for (int nid = 0; nid < leaves[k]; nid++)
      if (t.ns[nid].kind == 1) {
        dn.emplace_back();
        if (on_device) to_device_node_static(t.ns, nid, leaves[k], leaf_base, dn.back(), pool);
        else to_device_node(t.ns[nid], dn.back(), pool);
      }
    GTRY(h2ot_set_tree(g->eng, k, (int)dn.size(), dn.data(), pool.data(), (int64_t)pool.size()));
    std::vector<double> num(nleaf), den(nleaf);
    GTRY(h2ot_gamma_pass(g->eng, k, leaf_base, nleaf, num.data(), den.data()));
    if (quantile_family(g)) {  // fitBestConstantsQuantile: the rows now sit on their leaves (NIDs_k), which are the strata
      std::vector<double> q((size_t)nleaf);
      GTRY(h2ot_leaf_quantiles(g->eng, k, leaf_base, nleaf, 1, quantile_prob(g), q.data()));
      for (int i = 0; i < nleaf; i++) {
        if (std::isnan(q[(size_t)i])) continue;  // no active rows for this NID (:745)
        double val = eff_lr * q[(size_t)i];
        if (val > g->p.max_abs_leafnode_pred) val = g->p.max_abs_leafnode_pred;
        if (val < -g->p.max_abs_leafnode_pred) val = -g->p.max_abs_leafnode_pred;
        t.ns[leaves[k] + i].pred = (float)val;
      }
      continue;
    }
    for (int i = 0; i < nleaf; i++) {
      double gamma = den[i] == 0 ? 0 : num[i] / den[i];
      if (den[i] != 0 && g->p.family >= H2OT_FAMILY_POISSON && g->p.family <= H2OT_FAMILY_TWEEDIE) {  // GammaPass.gamma(double): link for poisson / gamma / tweedie
        const double x = std::fmax(0.0, gamma);
        gamma = x == 0 ? -19.0 : std::fmax(-19.0, std::log(x));  // LogExpUtil.log
      }
      double gf = eff_lr * m1class * gamma;
      if (g->p.family == H2OT_FAMILY_MULTINOMIAL) {
        if (gf > 1e4) gf = 1e4f;
        else if (gf < -1e4) gf = -1e4f;
      }
      if (std::isnan(gf)) gf = 0;
      else if (std::isinf(gf)) gf = (gf > 0 ? 1.0 : -1.0) * 1e4f;
      if (gf > g->p.max_abs_leafnode_pred) gf = g->p.max_abs_leafnode_pred;
      if (gf < -g->p.max_abs_leafnode_pred) gf = -g->p.max_abs_leafnode_pred;
      t.ns[leaves[k] + i].pred = (float)gf;
    }
  }
  // AddTreeContributions (:1410-1470)
  std::vector<FinishedTree> outs(K);
  for (int k = 0; k < K; k++) {
    Tree& t = *kt[k];
    if (on_device) {  // rows hold static ids: predictions under the same ids (LeafNodes from leaf_base on)
      const int leaf_base = (1 << (g->p.max_depth + 1)) - 1;
      const int nleaf = t.len() - leaves[k];
      std::vector<float> pred((size_t)leaf_base + (size_t)std::max(nleaf, 0), 0.f);
      for (int i = 0; i < leaves[k]; i++)
        if (t.ns[i].sid >= 0) pred[(size_t)t.ns[i].sid] = t.ns[i].pred;
      for (int i = 0; i < nleaf; i++) pred[(size_t)leaf_base + i] = t.ns[leaves[k] + i].pred;
      if (g->pred_noise_bandwidth != 0) {
        std::vector<double> fac(pred.size(), 1.0);
        for (int i = 0; i < nleaf; i++) fac[(size_t)leaf_base + i] = noise_factor(g, k, leaves[k] + i);
        if (nleaf <= 0) fac[0] = noise_factor(g, k, 0);
        GTRY(h2ot_add_tree_scaled(g->eng, k, (int)pred.size(), pred.data(), fac.data()));
      } else GTRY(h2ot_add_tree(g->eng, k, (int)pred.size(), pred.data()));
    } else {
      std::vector<float> pred(t.len());
      for (int i = 0; i < t.len(); i++) pred[i] = t.ns[i].pred;
      if (g->pred_noise_bandwidth != 0) {
        std::vector<double> fac(pred.size());
        for (int i = 0; i < t.len(); i++) fac[i] = noise_factor(g, k, i);
        GTRY(h2ot_add_tree_scaled(g->eng, k, t.len(), pred.data(), fac.data()));
      } else GTRY(h2ot_add_tree(g->eng, k, t.len(), pred.data()));
    }
    outs[k].nodes = std::move(t.ns);
    for (Node& n : outs[k].nodes) n.hs.clear();
    outs[k].leaves_start = leaves[k];
  }
  g->trees.push_back(std::move(outs));
  g->ntrees_built++;
  return H2OT_OK;
}

// DRF.buildNextKTrees / growTrees / CollectPreds (T/drf/DRF.java:157-330)
int build_next_k_trees_drf(h2ot_gbm* g) {
  const int K = g->K;
  std::vector<std::unique_ptr<Tree>> kt(K);
  std::vector<int> leaves(K, 0);
  const int adj_nbins = std::max(g->p.nbins_top_level, g->p.nbins);
  const int64_t rseed = g->rand.next_long();
  for (int k = 0; k < K; k++) {
    kt[k] = std::make_unique<Tree>(g->ncols, g->mtry, g->mtry_per_tree, rseed, g->p.col_sample_rate_change_per_level);
    kt[k]->new_undecided(kNoParent, initial_hist(g, adj_nbins, rseed, true));  // _checkFloatSplits (:201)
    GTRY(h2ot_begin_tree(g->eng, k, rseed, 1.0, g->work_bound));
    GTRY(h2ot_sample(g->eng, k, rseed, g->p.sample_rate));  // always: NA responses go out of bag even at rate 1
  }
  for (int depth = 0; depth < g->p.max_depth; depth++) {
    bool any = false;
    const bool last = depth == g->p.max_depth - 1;
    if (K > 1 && !g->capture) {
      // the K class-trees' levels are enqueued back to back; while the host turns tree k's splits into its next level the
      // device works on the levels of the trees behind it
      bool did = false;
      for (int k = 0; k < K; k++) GTRY(build_layer_k(g, k, *kt[k], leaves[k], did, last, 1));
      for (int k = 0; k < K; k++) {
        did = false;
        GTRY(build_layer_k(g, k, *kt[k], leaves[k], did, last, 2));
        any = any || did;
      }
    } else {
      for (int k = 0; k < K; k++) {
        bool did = false;
        GTRY(build_layer_k(g, k, *kt[k], leaves[k], did, last));
        any = any || did;
      }
    }
    if (!any) break;
  }
  const bool classifier = g->nclass > 1;
  std::vector<FinishedTree> outs(K);
  for (int k = 0; k < K; k++) {  // leaf insertion with the split's child means (:222-248)
    Tree& t = *kt[k];
    const int leaf = leaves[k] = t.len();
    for (int nid = 0; nid < leaf; nid++) {
      if (t.ns[nid].kind != 1) continue;
      if (!t.ns[nid].has_split) {
        if (nid == 0) {
          Node n;
          n.kind = 2;
          n.pid = kNoParent;
          n.nid = 0;
          n.pred = (float)(classifier ? g->prior_class_dist[k] : g->init_f);
          t.ns[0] = n;
          t.leaves++;
        }
        continue;
      }
      for (int i = 0; i < 2; i++) {
        const int cn = t.ns[nid].nids[i];
        if (cn == kUndecidedChild || t.ns[cn].kind == 0 || (t.ns[cn].kind == 1 && !t.ns[cn].has_split)) {
          const int ln = t.new_leaf(nid);
          t.ns[ln].pred = (float)(i == 0 ? t.ns[nid].split.p0 : t.ns[nid].split.p1);
          t.ns[nid].nids[i] = ln;
        }
      }
    }
    const bool root_is_leaf = t.ns[0].kind == 2;
    if (!root_is_leaf) {
      std::vector<h2ot_decided_node> dn;
      std::vector<uint8_t> pool;
      for (int nid = 0; nid < leaf; nid++)
        if (t.ns[nid].kind != 2) {  // decided nodes, and undecided ones (chopped for depth) as "no split" records
          dn.emplace_back();
          to_device_node(t.ns[nid], dn.back(), pool);
        }
      GTRY(h2ot_set_tree(g->eng, k, (int)dn.size(), dn.data(), pool.data(), (int64_t)pool.size()));
    }
    std::vector<float> pred(t.len());
    for (int i = 0; i < t.len(); i++) pred[i] = t.ns[i].pred;
    GTRY(h2ot_drf_collect_preds(g->eng, k, t.len(), pred.data(), root_is_leaf ? 1 : 0, k == 0 ? 1 : 0));
    outs[k].nodes = std::move(t.ns);
    for (Node& n : outs[k].nodes) n.hs.clear();
    outs[k].leaves_start = leaves[k];
  }
  g->trees.push_back(std::move(outs));
  g->ntrees_built++;
  return H2OT_OK;
}

}  // namespace

extern "C" {

int h2ot_gbm_create(const h2ot_gbm_params* p, int32_t device, int64_t nrows, int32_t ncols, int64_t row_offset, int64_t nrows_global,
                    int32_t has_weights, int32_t has_offset, h2ot_gbm** out) {
  if (!p || !out) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (p->family < H2OT_FAMILY_GAUSSIAN || p->family > H2OT_FAMILY_QUANTILE) GFAIL(H2OT_ERR_UNSUPPORTED, "unsupported distribution family");
  if (p->algo == H2OT_ALGO_DRF && p->family > H2OT_FAMILY_MULTINOMIAL) GFAIL(H2OT_ERR_UNSUPPORTED, "DRF takes gaussian / bernoulli / multinomial");
  if (p->max_depth < 1 || p->nbins < 2 || p->nbins_cats < 2 || p->ntrees < 0) GFAIL(H2OT_ERR_INVALID, "bad tree parameters");
  if (p->nbins_cats > H2OT_MAX_BITSET_BYTES * 8) GFAIL(H2OT_ERR_UNSUPPORTED, "nbins_cats above 4096");
  if (p->algo != H2OT_ALGO_GBM && p->algo != H2OT_ALGO_DRF) GFAIL(H2OT_ERR_UNSUPPORTED, "unknown algo");
  h2ot_gbm* g = new h2ot_gbm();
  g->p = *p;
  g->nrows = nrows;
  g->nrows_global = nrows_global > 0 ? nrows_global : nrows;
  g->row_offset = row_offset;
  g->ncols = ncols;
  g->nclass = p->family == H2OT_FAMILY_MULTINOMIAL ? p->nclass : (p->family == H2OT_FAMILY_BERNOULLI ? 2 : 1);
  g->K = g->nclass == 2 ? 1 : g->nclass;
  if (g->K < 1) {
    delete g;
    GFAIL(H2OT_ERR_INVALID, "multinomial needs nclass >= 2");
  }
  g->has_w = has_weights != 0;
  g->has_off = has_offset != 0;
  g->device = device;
  g->cols.resize(ncols);
  g->varimp.assign(ncols, 0.f);
  g->rand = Pcg(p->seed);
  h2ot_config cfg{};
  cfg.device = device;
  cfg.ncols = ncols;
  cfg.nrows = nrows;
  cfg.row_offset = row_offset;
  cfg.nrows_global = g->nrows_global;
  cfg.nclass_trees = g->K;
  cfg.nclass = g->nclass;
  cfg.family = p->family;
  cfg.has_weights = has_weights;
  cfg.has_offset = has_offset;
  cfg.tweedie_power = p->tweedie_power;
  const int rc = h2ot_create(&cfg, &g->eng);
  if (rc != H2OT_OK) {
    delete g;
    return rc;
  }
  *out = g;
  return H2OT_OK;
}

int h2ot_gbm_destroy(h2ot_gbm* g) {
  if (!g) return H2OT_OK;
  flush_pending(g);  // (drains the trees in flight; an error here has nobody left to hear it)
  h2ot_destroy(g->eng);
  delete g;
  return H2OT_OK;
}

int h2ot_gbm_set_column(h2ot_gbm* g, int32_t col, const double* values, int32_t is_categorical, int32_t cardinality) {
  if (!g || !values) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (col < 0 || col >= g->ncols) GFAIL(H2OT_ERR_INVALID, "column index out of range");
  GTRY(h2ot_load_column(g->eng, col, 0, g->nrows, values));
  g->cols[col].is_cat = is_categorical != 0;
  g->cols[col].cardinality = cardinality;
  g->cols[col].loaded = true;
  return H2OT_OK;
}
int h2ot_gbm_load_chunk(h2ot_gbm* g, int32_t col, int64_t row0, int64_t n, int32_t codec, const uint8_t* mem, int64_t mem_len) {
  if (!g || !mem) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (col >= g->ncols) GFAIL(H2OT_ERR_INVALID, "column index out of range");
  return h2ot_load_chunk(g->eng, col, row0, n, codec, mem, mem_len);
}
int h2ot_gbm_chunks_done(h2ot_gbm* g, int32_t col, int32_t is_categorical, int32_t cardinality) {
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (col >= g->ncols) GFAIL(H2OT_ERR_INVALID, "column index out of range");
  if (col >= 0) {
    g->cols[col].is_cat = is_categorical != 0;
    g->cols[col].cardinality = cardinality;
    g->cols[col].loaded = true;
  } else if (col == H2OT_COL_RESPONSE) {
    g->have_y = true;
  } else if (col == H2OT_COL_WEIGHTS) {
    if (!g->has_w) GFAIL(H2OT_ERR_INVALID, "created without weights");
    g->have_w = true;
  } else if (col == H2OT_COL_OFFSET) {
    if (!g->has_off) GFAIL(H2OT_ERR_INVALID, "created without offset");
    g->have_off = true;
  } else {
    GFAIL(H2OT_ERR_INVALID, "column index out of range");
  }
  return H2OT_OK;
}
int h2ot_gbm_set_response(h2ot_gbm* g, const double* y) {
  if (!g || !y) GFAIL(H2OT_ERR_INVALID, "null argument");
  g->have_y = true;
  return h2ot_load_response(g->eng, 0, g->nrows, y);
}
int h2ot_gbm_set_weights(h2ot_gbm* g, const double* w) {
  if (!g || !w) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (!g->has_w) GFAIL(H2OT_ERR_INVALID, "created without weights");
  g->have_w = true;
  return h2ot_load_weights(g->eng, 0, g->nrows, w);
}
int h2ot_gbm_set_offset(h2ot_gbm* g, const double* o) {
  if (!g || !o) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (!g->has_off) GFAIL(H2OT_ERR_INVALID, "created without offset");
  g->have_off = true;
  return h2ot_load_offset(g->eng, 0, g->nrows, o);
}

int h2ot_gbm_comm_init(h2ot_gbm* g, int32_t world_size, int32_t rank, const uint8_t id[H2OT_COMM_ID_BYTES]) {
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  g->world = world_size;
  g->rank = rank;
  return h2ot_comm_init(g->eng, world_size, rank, id);
}

int h2ot_gbm_init(h2ot_gbm* g) {
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (g->inited) return H2OT_OK;
  if (!g->have_y) GFAIL(H2OT_ERR_INVALID, "response not set");
  if (g->has_w && !g->have_w) GFAIL(H2OT_ERR_INVALID, "weights not set");
  if (g->has_off && !g->have_off) GFAIL(H2OT_ERR_INVALID, "offset not set");
  GTRY(h2ot_finalize_frame(g->eng));
  const int C = g->ncols;
  // rollups, merged over ranks: [min as -min | max | nacnt | not_int]
  std::vector<double> mx(2 * (size_t)C), sm(2 * (size_t)C);
  for (int c = 0; c < C; c++) {
    h2ot_column_info ci;
    GTRY(h2ot_column_info_get(g->eng, c, &ci));
    mx[c] = std::isnan(ci.min) ? -INFINITY : -ci.min;
    mx[C + c] = std::isnan(ci.max) ? -INFINITY : ci.max;
    sm[c] = (double)ci.nacnt;
    sm[C + c] = ci.is_int ? 0.0 : 1.0;
  }
  GTRY(allreduce(g, mx.data(), 2 * (int64_t)C, 1));
  GTRY(allreduce(g, sm.data(), 2 * (int64_t)C, 0));
  for (int c = 0; c < C; c++) {
    ColMeta& m = g->cols[c];
    m.min = std::isinf(mx[c]) && mx[c] < 0 ? NAN : -mx[c];
    m.max = std::isinf(mx[C + c]) && mx[C + c] < 0 ? NAN : mx[C + c];
    m.nacnt = (int64_t)sm[c];
    m.is_int = sm[C + c] == 0.0;
  }
  // weights bound (fixes the fixed-point scale of the w accumulator; must agree on every rank)
  if (g->has_w) {
    double wb = 0;
    GTRY(h2ot_get_weight_bound(g->eng, &wb));
    GTRY(allreduce(g, &wb, 1, 1));
    GTRY(h2ot_set_weight_bound(g->eng, wb));
  }
  // Size the cross-rank exchange of the level histograms once, for the widest level this model can have: 2^d leaves at depth
  // d, each column with max(nbins_top_level >> d, nbins) bins (nbins_cats for categoricals, fewer for narrow integer columns)
  // plus the NA bucket. Levels beyond the cap (very deep forests) regrow the buffer when they get there.
  if (g->world > 1) {
    int64_t best_bins = 0, best_leaves = 1;
    const int64_t cap_words = (int64_t)32 << 20;
    for (int d = 0; d < g->p.max_depth; d++) {
      const int64_t leaves = (int64_t)1 << std::min(d, 40);
      int64_t per_leaf = 0;
      for (int c = 0; c < C; c++) {
        const ColMeta& m = g->cols[c];
        int64_t nb = m.is_cat ? g->p.nbins_cats : std::max(g->p.nbins_top_level >> std::min(d, 30), g->p.nbins);
        if ((m.is_int || m.is_cat) && !std::isnan(m.min)) nb = std::min<int64_t>(nb, (int64_t)std::max(1.0, (m.is_cat ? m.cardinality : m.max - m.min + 1)));
        per_leaf += nb + 1;
      }
      const int64_t words = leaves * per_leaf * 3 + leaves * (2 * C + 1);
      if (words > cap_words) break;
      if (leaves * per_leaf > best_bins) best_bins = leaves * per_leaf;
      best_leaves = std::max(best_leaves, leaves);
    }
    GTRY(h2ot_reserve_exchange(g->eng, best_bins, best_leaves * C, best_leaves));
  }
  // GBM.init / initializeModelSpecifics (T/gbm/GBM.java:254-312) ; DRF: T/drf/DRF.java:114-153
  const bool drf = g->p.algo == H2OT_ALGO_DRF;
  g->mtry_per_tree = std::max(1, d2i(g->p.col_sample_rate_per_tree * C));
  if (!drf) g->mtry = std::max(1, d2i(g->p.col_sample_rate * g->p.col_sample_rate_per_tree * C));
  else if (g->p.mtries == -2) g->mtry = C;
  else if (g->p.mtries == -1) g->mtry = g->nclass > 1 ? std::max(d2i(std::sqrt((double)C)), 1) : std::max(C / 3, 1);
  else g->mtry = g->p.mtries;
  if (g->mtry < 1 || g->mtry > C) GFAIL(H2OT_ERR_INVALID, "computed mtry outside 1..ncols");
  // InitialValue (T/SharedTree.java:1138-1185) and, for DRF, the class prior and the bound of the work column: one device
  // reduction over the resident response / weights / offset, merged over ranks
  double nd[2] = {0, 0};
  {
    std::vector<double> cls((size_t)std::max(1, g->nclass) + 1, 0.0);
    double ymax = 0;
    GTRY(h2ot_initial_stats(g->eng, &nd[0], &nd[1], (drf && g->nclass > 1) ? cls.data() : nullptr, &ymax));
    GTRY(allreduce(g, nd, 2, 0));
    GTRY(allreduce(g, &ymax, 1, 1));
    g->y_abs_max = ymax;
    if (drf) {
      GTRY(allreduce(g, cls.data(), (int64_t)g->nclass, 0));
      g->prior_class_dist.assign(std::max(1, g->nclass), 0.0);
      double tot = 0;
      for (int c = 0; c < g->nclass; c++) tot += cls[c];
      if (g->nclass > 1)
        for (int c = 0; c < g->nclass; c++) g->prior_class_dist[c] = cls[c] / tot;
      g->work_bound = g->nclass > 1 ? 1.0 : (ymax > 0 ? ymax : 1.0);
    }
  }
  const double mean = nd[0] / nd[1];
  if (drf) g->init_f = g->nclass > 1 ? 0 : mean;  // T/drf/DRF.java:131
  else if (g->nclass > 2) g->init_f = 0;
  else if (g->p.family == H2OT_FAMILY_BERNOULLI) g->init_f = std::log(mean / (1 - mean));  // LogitFunction.link
  else if (g->p.family >= H2OT_FAMILY_POISSON && g->p.family <= H2OT_FAMILY_TWEEDIE) {  // LogFunction.link == LogExpUtil.log
    const double x = std::fmax(0.0, mean);
    g->init_f = x == 0 ? -19.0 : std::fmax(-19.0, std::log(x));
  } else if (quantile_family(g)) {
    // GBM.getInitialValueQuantile (T/gbm/GBM.java:272-276, 355-381): the median / alpha-quantile of y - offset
    GTRY(h2ot_leaf_quantiles(g->eng, 0, 0, 1, 0, quantile_prob(g), &g->init_f));
    if (std::isnan(g->init_f)) g->init_f = 0;
  } else g->init_f = mean;
  if (drf) {
    GTRY(h2ot_drf_init_model(g->eng));
  } else {
    std::vector<double> f0(g->K, 0.0);
    f0[0] = g->init_f;
    int64_t skipped = 0;
    GTRY(h2ot_init_model(g->eng, f0.data(), &skipped));
  }
  // this driver sequences whole trees: calls that return nothing need not wait for the stream
  GTRY(h2ot_set_lazy_sync(g->eng, 1));
  GTRY(h2ot_set_timers(g->eng, getenv("H2OT_STAGE_TIMERS") ? 2 : 1));
  g->inited = true;
  return H2OT_OK;
}

int h2ot_gbm_train(h2ot_gbm* g, int32_t n_iterations) {
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (!g->inited) GTRY(h2ot_gbm_init(g));
  const double t0 = now_s();
  for (int i = 0; i < n_iterations; i++) GTRY(g->p.algo == H2OT_ALGO_DRF ? build_next_k_trees_drf(g) : build_next_k_trees(g));
  g->timing.total_s += now_s() - t0;
  return H2OT_OK;
}

int h2ot_gbm_capture_levels(h2ot_gbm* g, int32_t on) {
  if (g) GTRY(flush_pending(g));
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  g->capture = on != 0;
  return H2OT_OK;
}

double h2ot_gbm_init_f(h2ot_gbm* g) { return g ? g->init_f : NAN; }
int32_t h2ot_gbm_ntrees(h2ot_gbm* g) {
  if (g) flush_pending(g);
  return g ? (int32_t)g->trees.size() : 0;
}
// Waits for the trees still in flight on the device and mirrors them (every accessor does so by itself).
int h2ot_gbm_flush(h2ot_gbm* g) {
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  return flush_pending(g);
}
int32_t h2ot_gbm_tree_len(h2ot_gbm* g, int32_t tree, int32_t k) {
  flush_pending(g);
  return (int32_t)g->trees[tree][k].nodes.size();
}
int32_t h2ot_gbm_tree_leaves_start(h2ot_gbm* g, int32_t tree, int32_t k) {
  flush_pending(g);
  return g->trees[tree][k].leaves_start;
}

int h2ot_gbm_tree_nodes(h2ot_gbm* g, int32_t tree, int32_t k, h2ot_tree_node* out) {
  if (g) GTRY(flush_pending(g));
  if (!g || !out || tree < 0 || tree >= (int)g->trees.size() || k < 0 || k >= g->K) GFAIL(H2OT_ERR_INVALID, "bad tree index");
  const std::vector<Node>& ns = g->trees[tree][k].nodes;
  for (size_t i = 0; i < ns.size(); i++) {
    const Node& n = ns[i];
    h2ot_tree_node& o = out[i];
    memset(&o, 0, sizeof(o));
    o.kind = n.kind;
    o.pid = n.pid;
    o.nid = n.nid;
    o.nid0 = n.nids[0];
    o.nid1 = n.nids[1];
    o.splat = n.splat;
    o.pred = n.pred;
    o.has_split = n.has_split ? 1 : 0;
    if (n.has_split) {
      const h2ot_split& s = n.split;
      o.col = s.col; o.bin = s.bin; o.nasplit = s.nasplit; o.equal = s.equal;
      o.se = s.se; o.se0 = s.se0; o.se1 = s.se1; o.n0 = s.n0; o.n1 = s.n1; o.p0 = s.p0; o.p1 = s.p1;
      o.tree_p0 = s.tree_p0; o.tree_p1 = s.tree_p1;
      o.bs_nbits = s.bs_nbits; o.bs_bitoff = s.bs_bitoff; o.bs_nbytes = s.bs_nbytes;
      memcpy(o.bs, s.bs, (size_t)std::max(0, std::min<int>(s.bs_nbytes, H2OT_MAX_BITSET_BYTES)));  // bytes past bs_nbytes are unspecified
    } else {
      o.col = -1;
    }
  }
  return H2OT_OK;
}

// DTree.compress (T/DTree.java:972-982): the byte stream SharedTreeMojoModel.scoreTree walks, plus the auxiliary stream
// (_abAux, :905-916). DecidedNode.size / compress (:845-933), LeafNode.compress (:918-921), IcedBitSet.compress2 / compress3
// (W/util/IcedBitSet.java:92-111). Children are always created after their parent, so one backwards sweep over the node table
// gives every subtree's size, node count and type byte; an explicit stack then writes the nodes in pre-order.
// Two-call protocol: with bits_cap / aux_cap too small nothing is written and the needed lengths come back.
int h2ot_gbm_tree_compress(h2ot_gbm* g, int32_t tree, int32_t k, uint8_t* bits, int64_t bits_cap, int64_t* bits_len, uint8_t* aux,
                           int64_t aux_cap, int64_t* aux_len) {
  if (g) GTRY(flush_pending(g));
  if (!g || tree < 0 || tree >= (int)g->trees.size() || k < 0 || k >= g->K) GFAIL(H2OT_ERR_INVALID, "bad tree index");
  const std::vector<Node>& ns = g->trees[tree][k].nodes;
  const int n = (int)ns.size();
  if (n == 0) GFAIL(H2OT_ERR_INVALID, "empty tree");
  std::vector<int64_t> size((size_t)n, 0);
  std::vector<int32_t> inner((size_t)n, 0);  // decided nodes of the subtree (leaves count 0)
  std::vector<uint8_t> type((size_t)n, 0);
  int64_t n_decided = 0;
  for (int i = n - 1; i >= 0; i--) {
    const Node& nd = ns[(size_t)i];
    if (nd.kind == 2) {
      size[(size_t)i] = 4;
      continue;
    }
    if (nd.kind != 1 || !nd.has_split) continue;  // unreachable leftovers (a child that never split was replaced by a leaf)
    const h2ot_split& sp = nd.split;
    const int l = nd.nids[0], r = nd.nids[1];
    if (l <= i || r <= i || l >= n || r >= n || size[(size_t)l] == 0 || size[(size_t)r] == 0) GFAIL(H2OT_ERR_INVALID, "tree is not finished");
    uint8_t ty = sp.equal == 0 ? 0 : (sp.equal == 1 ? 4 : (sp.equal == 2 ? 8 : 12));
    int64_t sz = (sp.equal == 3 ? 9 + sp.bs_nbytes : 7) + 1;  // type, column, split value or group, NA direction
    if (sp.nasplit == H2OT_NA_VS_REST) sz -= 4;
    if (ns[(size_t)l].kind == 2) ty |= 48;
    else {
      const int64_t lsz = size[(size_t)l];
      const int slen = lsz < 256 ? 0 : (lsz < 65535 ? 1 : (lsz < (1 << 24) ? 2 : 3));
      ty |= (uint8_t)slen;
      sz += slen + 1;
    }
    if (ns[(size_t)r].kind == 2) ty |= (uint8_t)(48 << 2);
    size[(size_t)i] = sz + size[(size_t)l] + size[(size_t)r];
    inner[(size_t)i] = 1 + inner[(size_t)l] + inner[(size_t)r];
    type[(size_t)i] = ty;
    n_decided++;
  }
  const bool stump = ns[0].kind == 2;
  const int64_t need_bits = size[0] + (stump ? 3 : 0), need_aux = (int64_t)inner[0] * 40;
  (void)n_decided;
  if (bits_len) *bits_len = need_bits;
  if (aux_len) *aux_len = need_aux;
  if (!bits || bits_cap < need_bits || (aux && aux_cap < need_aux)) return H2OT_OK;
  uint8_t* b = bits;
  uint8_t* a = aux;
  auto put = [](uint8_t*& p, uint32_t v, int nbytes) {
    for (int i = 0; i < nbytes; i++) *p++ = (uint8_t)(v >> (8 * i));  // AutoBuffer writes in native (little-endian) order
  };
  auto putf = [&](uint8_t*& p, float f) {
    uint32_t v;
    memcpy(&v, &f, 4);
    put(p, v, 4);
  };
  if (stump) {
    put(b, 0, 1);
    put(b, 65535, 2);
  }
  std::vector<int> stack;
  stack.push_back(0);
  while (!stack.empty()) {
    const int i = stack.back();
    stack.pop_back();
    const Node& nd = ns[(size_t)i];
    if (nd.kind == 2) {
      putf(b, nd.pred);
      continue;
    }
    const h2ot_split& sp = nd.split;
    put(b, type[(size_t)i], 1);
    put(b, (uint32_t)sp.col, 2);
    put(b, (uint32_t)sp.nasplit, 1);
    if (sp.nasplit != H2OT_NA_VS_REST) {
      if (sp.equal == 0 || sp.equal == 1) putf(b, nd.splat);
      else if (sp.equal == 2) {
        for (int j = 0; j < 4; j++) put(b, sp.bs[j], 1);
      } else {
        put(b, (uint32_t)sp.bs_bitoff, 2);
        put(b, (uint32_t)sp.bs_nbits, 4);
        for (int j = 0; j < sp.bs_nbytes; j++) put(b, sp.bs[j], 1);
      }
    }
    if (a) {
      put(a, (uint32_t)nd.nid, 4);
      put(a, (uint32_t)inner[(size_t)nd.nids[0]], 4);
      putf(a, (float)sp.n0);
      putf(a, (float)sp.n1);
      putf(a, (float)sp.p0);
      putf(a, (float)sp.p1);
      putf(a, (float)sp.se0);
      putf(a, (float)sp.se1);
      put(a, (uint32_t)nd.nids[0], 4);
      put(a, (uint32_t)nd.nids[1], 4);
    }
    if ((type[(size_t)i] & 48) == 0) {
      const int64_t lsz = size[(size_t)nd.nids[0]];
      put(b, (uint32_t)lsz, lsz < 256 ? 1 : (lsz < 65535 ? 2 : (lsz < (1 << 24) ? 3 : 4)));
    }
    stack.push_back(nd.nids[1]);  // the left subtree is written first
    stack.push_back(nd.nids[0]);
  }
  if (b - bits != need_bits || (aux && a - aux != need_aux)) GFAIL(H2OT_ERR_INVALID, "compressed tree size differs from DecidedNode.size()");
  return H2OT_OK;
}

int h2ot_gbm_predictions(h2ot_gbm* g, int32_t k, double* out) {
  if (g) GTRY(flush_pending(g));
  if (!g || !out) GFAIL(H2OT_ERR_INVALID, "null argument");
  return h2ot_get_predictions(g->eng, k, out);
}
int h2ot_gbm_predict(h2ot_gbm* g, int64_t nrows, const double* const* columns, double* out) {
  if (g) GTRY(flush_pending(g));
  if (!g || !columns || !out || nrows <= 0) GFAIL(H2OT_ERR_INVALID, "bad argument");
  if (!g->inited) GFAIL(H2OT_ERR_INVALID, "model not initialised");
  h2ot_config cfg{};
  cfg.device = g->device;
  cfg.ncols = g->ncols;
  cfg.nrows = nrows;
  cfg.nrows_global = nrows;
  cfg.nclass_trees = g->K;
  cfg.nclass = g->nclass;
  cfg.family = g->p.family;
  cfg.tweedie_power = g->p.tweedie_power;
  h2ot_engine* e = nullptr;
  GTRY(h2ot_create(&cfg, &e));
  auto run = [&]() -> int {
    for (int c = 0; c < g->ncols; c++) {
      if (!columns[c]) GFAIL(H2OT_ERR_INVALID, "null column");
      GTRY(h2ot_load_column(e, c, 0, nrows, columns[c]));
    }
    std::vector<double> zeros((size_t)nrows, 0.0);
    GTRY(h2ot_load_response(e, 0, nrows, zeros.data()));
    GTRY(h2ot_finalize_frame(e));
    std::vector<double> f0((size_t)g->K, 0.0);
    GTRY(h2ot_init_model(e, f0.data(), nullptr));
    for (const std::vector<FinishedTree>& tk : g->trees)
      for (int k = 0; k < g->K; k++) {
        const FinishedTree& ft = tk[(size_t)k];
        std::vector<h2ot_decided_node> dn;
        std::vector<uint8_t> pool;
        std::vector<float> pred(ft.nodes.size());
        for (size_t i = 0; i < ft.nodes.size(); i++) {
          pred[i] = ft.nodes[i].pred;
          if ((int)i < ft.leaves_start && ft.nodes[i].kind == 1) {
            dn.emplace_back();
            to_device_node(ft.nodes[i], dn.back(), pool);
          }
        }
        GTRY(h2ot_set_tree(e, k, (int)dn.size(), dn.data(), pool.data(), (int64_t)pool.size()));
        GTRY(h2ot_score_tree(e, k, ft.nodes[0].kind == 2 ? 0 : ft.leaves_start, (int)pred.size(), pred.data()));
      }
    for (int k = 0; k < g->K; k++) {
      double* o = out + (size_t)k * (size_t)nrows;
      GTRY(h2ot_get_predictions(e, k, o));
      if (g->p.algo == H2OT_ALGO_GBM && g->init_f != 0)
        for (int64_t r = 0; r < nrows; r++) o[r] += g->init_f;
    }
    return H2OT_OK;
  };
  const int rc = run();
  h2ot_destroy(e);
  return rc;
}

int h2ot_gbm_oobt(h2ot_gbm* g, double* out) {
  if (g) GTRY(flush_pending(g));
  if (!g || !out) GFAIL(H2OT_ERR_INVALID, "null argument");
  return h2ot_get_oobt(g->eng, out);
}
int h2ot_gbm_score(h2ot_gbm* g, double* out_mse, double* out_logloss) {
  if (g) GTRY(flush_pending(g));
  if (!g) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (g->p.algo != H2OT_ALGO_GBM) GFAIL(H2OT_ERR_UNSUPPORTED, "device scoring covers GBM only");
  h2ot_metrics m{};
  GTRY(h2ot_score_metrics(g->eng, &m));
  double v[3] = {m.wcount, m.sumsqe, m.logloss};
  GTRY(allreduce(g, v, 3, 0));
  if (out_mse) *out_mse = v[1] / v[0];
  if (out_logloss) *out_logloss = v[2] / v[0];
  return H2OT_OK;
}
// ---- AUC2 on the host side of the seam (h2o-core/src/main/java/hex/AUC2.java): the device fills one AUCBuilder per chunk of rows
// (h2ot_auc_builders == AUC_Impl.map); here they are folded like MRTask.reduce does -- AUCBuilder.reduce (:448-472), mergeOneBin
// (:483-505), find_smallest_impl (:523-554), dups (:556-565) -- and the AUC2 constructor + compute_auc (:158-190, 262-277) turn
// the final builder into the number the model reports.
namespace {
struct AucFold {
  int n = 0, ssx = -1;
  double ths[2 * H2OT_AUC_BINS], sqe[2 * H2OT_AUC_BINS], tps[2 * H2OT_AUC_BINS], fps[2 * H2OT_AUC_BINS];
  double k(int i) const { return tps[i] + fps[i]; }
  static double delta_error(double ths1, double n1, double ths0, double n0) {
    const double delta = (double)((float)ths1 - (float)ths0);
    if (delta == 0) return 0;
    return delta * delta * n0 * n1 / (n0 + n1);
  }
  static double combine(double ths1, double n1, double ths0, double n0) {
    const double c = (ths0 * n0 + ths1 * n1) / (n0 + n1);
    return (std::isnan(c) || std::isinf(c)) ? (ths0 + ths1) / 2 : c;
  }
  int smallest() {
    if (ssx != -1) return ssx;
    if (n == 1) return ssx = 0;
    double best = DBL_MAX;
    int at = -1;
    for (int i = 0; i + 1 < n; i++) {
      const double derr = delta_error(ths[i + 1], k(i + 1), ths[i], k(i));
      if (derr == 0) return ssx = i;
      const double s = sqe[i] + sqe[i + 1] + derr;
      if (s < best) { at = i; best = s; }
    }
    if (at == -1) {
      at = 0;
      double md = ths[1] - ths[0];
      for (int i = 1; i + 1 < n; i++)
        if (ths[i + 1] - ths[i] < md) { md = ths[i + 1] - ths[i]; at = i; }
    }
    return ssx = at;
  }
  bool dups() {
    for (int i = 0; i + 1 < n; i++)
      if (delta_error(ths[i + 1], k(i + 1), ths[i], k(i)) == 0) { ssx = i; return true; }
    return false;
  }
  void merge_one() {
    const int s = smallest();
    const double k0 = k(s), k1 = k(s + 1);
    sqe[s] = sqe[s] + sqe[s + 1] + delta_error(ths[s + 1], k1, ths[s], k0);
    ths[s] = combine(ths[s], k0, ths[s + 1], k1);
    tps[s] += tps[s + 1];
    fps[s] += fps[s + 1];
    for (int i = s + 1; i + 1 < n; i++) { ths[i] = ths[i + 1]; sqe[i] = sqe[i + 1]; tps[i] = tps[i + 1]; fps[i] = fps[i + 1]; }
    n--;
    ssx = -1;
  }
  // this <- reduce(this, other)
  void reduce(int on, const double* oths, const double* osqe, const double* otps, const double* ofps) {
    int x = n - 1, y = on - 1;
    while (x + y + 1 >= 0) {
      const bool self = y < 0 || (x >= 0 && ths[x] >= oths[y]);
      const int d = x + y + 1;
      if (self) { ths[d] = ths[x]; sqe[d] = sqe[x]; tps[d] = tps[x]; fps[d] = fps[x]; x--; }
      else { ths[d] = oths[y]; sqe[d] = osqe[y]; tps[d] = otps[y]; fps[d] = ofps[y]; y--; }
    }
    n += on;
    ssx = -1;
    while (n > H2OT_AUC_BINS || dups()) merge_one();
  }
  double auc() const {
    if (n < 1) return NAN;
    double p = 0, q = 0;
    std::vector<double> tp(n), fp(n);
    for (int i = 0; i < n; i++) {  // reversed: thresholds from 1 down to 0, rolled up
      p += tps[n - 1 - i];
      q += fps[n - 1 - i];
      tp[i] = p;
      fp[i] = q;
    }
    if (fp[n - 1] == 0) return 1.0;
    if (tp[n - 1] == 0) return 0.0;
    double tp0 = 0, fp0 = 0, area = 0;
    for (int i = 0; i < n; i++) {
      area += (fp[i] - fp0) * (tp[i] + tp0) / 2.0;
      tp0 = tp[i];
      fp0 = fp[i];
    }
    return area / p / q;
  }
};
}  // namespace

int h2ot_gbm_auc(h2ot_gbm* g, double* out_auc, int32_t* out_nbins, double* out_ths, double* out_tps, double* out_fps, int64_t* out_chunk_rows) {
  if (g) GTRY(flush_pending(g));
  if (!g || !out_auc) GFAIL(H2OT_ERR_INVALID, "null argument");
  if (g->p.algo != H2OT_ALGO_GBM || g->p.family != H2OT_FAMILY_BERNOULLI) GFAIL(H2OT_ERR_UNSUPPORTED, "AUC2 covers bernoulli GBM");
  const int max_chunks = 4096;
  std::vector<double> buf((size_t)max_chunks * H2OT_AUC_STRIDE);
  int32_t nch = 0;
  int64_t chunk_rows = 0;
  GTRY(h2ot_auc_builders(g->eng, max_chunks, &nch, &chunk_rows, buf.data()));
  if (out_chunk_rows) *out_chunk_rows = chunk_rows;
  std::unique_ptr<AucFold> f(new AucFold());
  for (int c = 0; c < nch; c++) {
    const double* o = buf.data() + (size_t)c * H2OT_AUC_STRIDE;
    const int on = (int)o[0];
    const double* oths = o + 1;
    const double* otps = o + 1 + H2OT_AUC_BINS;
    const double* ofps = o + 1 + 2 * H2OT_AUC_BINS;
    const double* osqe = o + 1 + 3 * H2OT_AUC_BINS;
    if (c == 0) {
      f->n = on;
      for (int i = 0; i < on; i++) { f->ths[i] = oths[i]; f->sqe[i] = osqe[i]; f->tps[i] = otps[i]; f->fps[i] = ofps[i]; }
    } else f->reduce(on, oths, osqe, otps, ofps);
  }
  if (g->world > 1) {  // every rank's folded builder travels to every rank (slots of a summed array), then folds in rank order
    const size_t stride = 1 + 4 * (size_t)H2OT_AUC_BINS;
    std::vector<double> all((size_t)g->world * stride, 0.0);
    double* mine = all.data() + (size_t)g->rank * stride;
    mine[0] = f->n;
    for (int i = 0; i < f->n; i++) {
      mine[1 + i] = f->ths[i];
      mine[1 + H2OT_AUC_BINS + i] = f->tps[i];
      mine[1 + 2 * H2OT_AUC_BINS + i] = f->fps[i];
      mine[1 + 3 * H2OT_AUC_BINS + i] = f->sqe[i];
    }
    GTRY(allreduce(g, all.data(), (int64_t)all.size(), 0));
    f.reset(new AucFold());
    for (int r = 0; r < g->world; r++) {
      const double* o = all.data() + (size_t)r * stride;
      const int on = (int)o[0];
      if (r == 0) {
        f->n = on;
        for (int i = 0; i < on; i++) { f->ths[i] = o[1 + i]; f->tps[i] = o[1 + H2OT_AUC_BINS + i]; f->fps[i] = o[1 + 2 * H2OT_AUC_BINS + i]; f->sqe[i] = o[1 + 3 * H2OT_AUC_BINS + i]; }
      } else f->reduce(on, o + 1, o + 1 + 3 * H2OT_AUC_BINS, o + 1 + H2OT_AUC_BINS, o + 1 + 2 * H2OT_AUC_BINS);
    }
  }
  *out_auc = f->auc();
  if (out_nbins) *out_nbins = f->n;
  for (int i = 0; i < f->n; i++) {
    if (out_ths) out_ths[i] = f->ths[i];
    if (out_tps) out_tps[i] = f->tps[i];
    if (out_fps) out_fps[i] = f->fps[i];
  }
  return H2OT_OK;
}
int h2ot_gbm_set_pred_noise_bandwidth(h2ot_gbm* g, double bandwidth) {
  if (g) GTRY(flush_pending(g));
  if (!g || !(bandwidth >= 0)) GFAIL(H2OT_ERR_INVALID, "pred_noise_bandwidth must be >= 0");  // T/gbm/GBM.java:189-190
  if (g->p.algo != H2OT_ALGO_GBM) GFAIL(H2OT_ERR_UNSUPPORTED, "pred_noise_bandwidth is a GBM parameter");
  g->pred_noise_bandwidth = bandwidth;
  return H2OT_OK;
}
int h2ot_gbm_set_sample_rate_per_class(h2ot_gbm* g, const double* rates, int32_t n) {
  if (g) GTRY(flush_pending(g));
  if (!g || n < 0 || (n > 0 && !rates)) GFAIL(H2OT_ERR_INVALID, "bad rates");
  if (n > 0 && g->nclass < 2) GFAIL(H2OT_ERR_INVALID, "sample_rate_per_class needs a categorical response");
  if (n > 0 && n != g->nclass) GFAIL(H2OT_ERR_INVALID, "sample_rate_per_class needs one rate per response class");
  for (int i = 0; i < n; i++)
    if (!(rates[i] >= 0 && rates[i] <= 1)) GFAIL(H2OT_ERR_INVALID, "sample rates must be within [0, 1]");
  g->sample_rate_per_class.assign(rates, rates + n);
  return h2ot_set_sample_rate_per_class(g->eng, rates, n);
}
int h2ot_gbm_varimp(h2ot_gbm* g, float* out) {
  if (g) GTRY(flush_pending(g));
  if (!g || !out) GFAIL(H2OT_ERR_INVALID, "null argument");
  std::copy(g->varimp.begin(), g->varimp.end(), out);
  return H2OT_OK;
}
int h2ot_gbm_timing_get(h2ot_gbm* g, h2ot_gbm_timing* out) {
  if (g) GTRY(flush_pending(g));
  if (!g || !out) GFAIL(H2OT_ERR_INVALID, "null argument");
  *out = g->timing;
  return H2OT_OK;
}
h2ot_engine* h2ot_gbm_engine(h2ot_gbm* g) {
  if (g) flush_pending(g);  // (the caller is about to use the engine directly)
  return g ? g->eng : nullptr;
}

// ---- host-logic probes (no device needed): let the CPU test-suite compare this file's DHistogram / DTree / PCG bookkeeping
// with an independent CPU restatement of the reference
int h2ot_hostcheck_hist(int32_t nbins, int32_t nbins_cats, int32_t is_int, double min, double max_ex, int32_t* out_nbin, double* out_step) {
  Hist h;
  if (!make_hist(h, nbins, nbins_cats, (uint8_t)is_int, min, max_ex, false, false, 0, false)) return 0;
  if (out_nbin) *out_nbin = h.nbin;
  if (out_step) *out_step = h.step;
  return 1;
}
double h2ot_hostcheck_find_max_ex(double max_in, int32_t is_int) { return Hist::find_max_ex_of(max_in, is_int); }
int64_t h2ot_hostcheck_pcg_next_long(int64_t seed, int32_t nth) {
  Pcg r(seed);
  int64_t v = 0;
  for (int i = 0; i <= nth; i++) v = r.next_long();
  return v;
}
void h2ot_hostcheck_pcg_bounded(int64_t seed, int32_t bound, int32_t n, int32_t* out) {
  Pcg r(seed);
  for (int i = 0; i < n; i++) out[i] = r.next_bounded(bound);
}
// n_nodes successive UndecidedNode constructions on one tree (all with the same present-mask): out[i*ncols ..] = the node's
// _scoreCols padded with -1; out_null[i] = 1 when _scoreCols == null (all columns)
void h2ot_hostcheck_score_cols(int32_t ncols, int32_t mtry, int32_t mtry_per_tree, int64_t seed, double change_per_level, int32_t depth,
                               const uint8_t* present, int32_t n_nodes, int32_t* out, int32_t* out_null) {
  Tree t(ncols, mtry, mtry_per_tree, seed, change_per_level);
  t.depth = depth;
  for (int i = 0; i < n_nodes; i++) {
    std::vector<Hist> hs(ncols);
    for (int c = 0; c < ncols; c++) hs[c].present = present[c] != 0;
    const int nid = t.new_undecided(kNoParent, std::move(hs));
    const Node& n = t.ns[nid];
    out_null[i] = n.score_cols_null ? 1 : 0;
    for (int c = 0; c < ncols; c++) out[(size_t)i * ncols + c] = c < (int)n.score_cols.size() ? n.score_cols[c] : -1;
  }
}

int32_t h2ot_gbm_ncaptures(h2ot_gbm* g) { return g ? (int32_t)g->captures.size() : 0; }
int h2ot_gbm_capture_info(h2ot_gbm* g, int32_t i, int32_t* tree, int32_t* k, int32_t* leaf, int32_t* nleaves) {
  if (!g || i < 0 || i >= (int)g->captures.size()) GFAIL(H2OT_ERR_INVALID, "bad capture index");
  const Capture& c = g->captures[i];
  if (tree) *tree = c.tree;
  if (k) *k = c.k;
  if (leaf) *leaf = c.leaf;
  if (nleaves) *nleaves = c.nleaves;
  return H2OT_OK;
}
int h2ot_gbm_capture_desc(h2ot_gbm* g, int32_t i, int32_t n, int32_t c, h2ot_hist_desc* desc, double* min2, double* maxIn,
                          int32_t* has_vals, int64_t* nvals) {
  if (!g || i < 0 || i >= (int)g->captures.size()) GFAIL(H2OT_ERR_INVALID, "bad capture index");
  const Capture& cap = g->captures[i];
  if (n < 0 || n >= cap.nleaves || c < 0 || c >= g->ncols) GFAIL(H2OT_ERR_INVALID, "bad (leaf, col)");
  const size_t p = (size_t)n * g->ncols + c;
  if (desc) *desc = cap.descs[p];
  if (min2) *min2 = cap.min2[p];
  if (maxIn) *maxIn = cap.max_in[p];
  if (has_vals) *has_vals = cap.has[p];
  if (nvals) *nvals = (cap.descs[p].flags & H2OT_DESC_ACTIVE) ? (int64_t)(cap.descs[p].nbin + 1) * 3 : 0;
  return H2OT_OK;
}
int h2ot_gbm_capture_vals(h2ot_gbm* g, int32_t i, int32_t n, int32_t c, double* out) {
  if (!g || i < 0 || i >= (int)g->captures.size() || !out) GFAIL(H2OT_ERR_INVALID, "bad capture index");
  const Capture& cap = g->captures[i];
  if (n < 0 || n >= cap.nleaves || c < 0 || c >= g->ncols) GFAIL(H2OT_ERR_INVALID, "bad (leaf, col)");
  const size_t p = (size_t)n * g->ncols + c;
  if (!(cap.descs[p].flags & H2OT_DESC_ACTIVE) || cap.offs[p] < 0) GFAIL(H2OT_ERR_INVALID, "histogram not present");
  const int64_t nv = (int64_t)(cap.descs[p].nbin + 1) * 3;
  std::copy(cap.vals.begin() + cap.offs[p], cap.vals.begin() + cap.offs[p] + nv, out);
  return H2OT_OK;
}
int h2ot_gbm_capture_nids(h2ot_gbm* g, int32_t i, int32_t* out) {
  if (!g || i < 0 || i >= (int)g->captures.size() || !out) GFAIL(H2OT_ERR_INVALID, "bad capture index");
  std::copy(g->captures[i].nids.begin(), g->captures[i].nids.end(), out);
  return H2OT_OK;
}

}  // extern "C"
