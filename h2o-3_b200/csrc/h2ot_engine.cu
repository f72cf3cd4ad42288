// libh2otree_b200.so — engine state + the C ABI declared in include/h2ot.h.
// One engine == one model build on one B200 (one process per GPU). No CPU fallback anywhere.
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <queue>
#include <vector>

#include "common.cuh"
#include "k_level.cuh"
#include "k_sparse.cuh"
#include "k_ingest.cuh"
#include "k_rows.cuh"
#include "k_split.cuh"
#include "k_xreduce.cuh"
#include "k_auc.cuh"
#include "k_next.cuh"
#include "k_quant.cuh"

#include <cub/device/device_radix_sort.cuh>

namespace h2ot {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
template <class T>
struct DBuf {  // grow-only device buffer
  T* p = nullptr;
  size_t cap = 0;
  int ensure(size_t n) {
    if (n <= cap) return H2OT_OK;
    if (p) CUDA_TRY(cudaFree(p));
    p = nullptr;
    cap = 0;
    size_t want = std::max<size_t>(n, 16);
    CUDA_TRY(cudaMalloc(&p, want * sizeof(T)));
    cap = want;
    return H2OT_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

struct Timer {  // CUDA-event timing of one kernel family inside an API call
  cudaEvent_t a = nullptr, b = nullptr;
  double* acc = nullptr;
};

struct Column {
  void* data = nullptr;
  double* staging = nullptr;
  int storage = H2OT_STORE_F32;
  double base = 0;
  h2ot_column_info info{};
  bool finalized = false;
};

struct TreeState {  // per class-tree k
  double* F = nullptr;
  float* work = nullptr;
  int32_t* nids = nullptr;
  uint16_t* code = nullptr;
  DBuf<h2ot_decided_node> nodes;       // device node table indexed by nid
  std::vector<h2ot_decided_node> h_nodes;  // host mirror
  DBuf<uint8_t> bitsets;
  std::vector<uint8_t> h_bitsets;
  int n_nodes = 0;
  double work_bound = 1.0;
  int ey = 0, eyy = 0;
  // level state
  int L = 0, leaf_offset = 0;
  bool built = false, finished = false;
  std::vector<DescDev> h_descs;
  std::vector<GroupDev> h_groups;
  std::vector<double> h_conc;
  std::vector<int32_t> h_group_cols;
  int bucket_begin[4] = {0, 0, 0, 0}, bucket_end[4] = {0, 0, 0, 0};  // smem groups per storage type
  int bucket_ctas[4] = {0, 0, 0, 0};                                  // CTAs of each storage type's launch
  int n_smem_groups = 0;
  unsigned char* stage = nullptr;  // pinned host staging of this class-tree's level results (h2ot_level_step[_begin/_end])
  size_t stage_cap = 0;
  cudaEvent_t done_ev = nullptr;   // recorded after the results' download was enqueued
  bool step_pending = false;
  std::vector<double> h_sp;  // split points of the next level (h2ot_set_split_points), -0.0 canonicalised
  DBuf<double> d_sp;
  bool sp_dirty = false, any_splitpts = false;
  // rows expected in each leaf of the level being planned: the parents' split counts (n0 / n1) of the previous level_step
  std::vector<double> prev_n0, prev_n1, est_rows;
  int prev_leaf_offset = -1;
  int n_sparse_groups = 0, sparse_ctas = 0;  // leaf-range groups of hist_sparse_kernel (levels with per-node column sampling)
  double sparse_bytes_per_row = 0, sparse_cols_per_row = 0;
  size_t smem_bytes = 0;
  int64_t total_bins = 0, total_vals = 0;
  size_t payload_words = 0;  // 64-bit words of the level's payload (sums + leaf rows + min/max keys)
  int active_cols = 0;
  DBuf<unsigned char> d_lvl;  // per-level upload: [DescDev LC][GroupDev G][int32 group columns]
  unsigned char* lvl_host = nullptr;
  size_t lvl_host_cap = 0;
  DescDev* p_descs = nullptr;
  GroupDev* p_groups = nullptr;
  int32_t* p_group_cols = nullptr;
  DBuf<long long> d_sums;  // [3*total_bins | L leaf_rows]
  long long* sums = nullptr;          // this level's payload: d_sums, or the peer-mapped exchange buffer
  unsigned long long* mm = nullptr;
  bool in_xbuf = false;
  bool best_pending = false;  // h2ot_find_splits left DecidedNode.bestCol to next_level_kernel
  bool fin_pending = false;   // h2ot_finish_level left the finalize kernel to the split search (fin_args)
  FinalizeArgs fin_args{};
  bool level_failed = false;  // a level call failed after score_decide had moved the rows: the tree has to be restarted
  // ---- levels enqueued back to back with the step between them on the device (h2ot_tree_levels_begin / _end, k_next.cuh)
  struct DevLoop {
    bool active = false;        // h2ot_build_level is being driven by the device loop
    int depth = 0, max_depth = 0;
    int nbins = 0, nbins_cats = 0;
    double min_rows = 0;
    DBuf<unsigned char> lvl_alt, pack_alt;   // the previous level's descriptors / results (ping-pong with d_lvl / d_pack)
    DBuf<double> maxex[2];                   // _maxEx of the levels' histograms [slot][col] (ping-pong)
    // what the host reads back after a tree, in ONE device block laid out like a half of out_stage:
    // [tree_splits: 2^D - 1][exists: 2^(D+1) - 1 bytes][err][root concentration: C doubles][leaf values: 2^(D+1) - 1 floats]
    DBuf<unsigned char> out_dev;
    struct { uint8_t* p = nullptr; } exists;            // [static id]
    struct { h2ot_split* p = nullptr; } tree_splits;    // [static id < 2^D - 1]
    struct { int32_t* p = nullptr; } err;
    double* p_conc_out = nullptr;
    float* p_pred_out = nullptr;
    DBuf<int32_t> nb_caps;                   // [depth][col] largest nbin the level's layout holds
    DBuf<double> root_maxex;                 // the root histograms' _maxEx [col]
    std::vector<int32_t> h_caps;             // host copies of the two: uploaded again only when they change
    std::vector<double> h_root_maxex;
    std::vector<unsigned char*> lvl_stage;   // pinned staging of each level's upload (the levels are enqueued ahead of the GPU)
    std::vector<size_t> lvl_stage_cap;
    std::vector<char> lvl_stage_own;         // 1: a block of its own (not a slice of the arena)
    unsigned char* arena = nullptr;          // pinned arena the levels' staging slices come from
    size_t arena_cap = 0;
    unsigned char* out_stage = nullptr;      // pinned: splits, exists, err, root concentration
    size_t out_stage_cap = 0;
    // the previous level as the step kernel reads it
    const DescDev* prev_descs = nullptr;
    const double *prev_min2 = nullptr, *prev_maxIn = nullptr, *prev_wna = nullptr, *prev_maxex = nullptr;
    const uint8_t* prev_has = nullptr;
    const h2ot_split* prev_best = nullptr;
    const double* conc_src = nullptr;  // the root level's per-column concentration (finalize), copied out by the first step kernel
    // Two trees of a class may be in flight (the host mirrors tree t while the device builds tree t + 1): every pinned staging
    // area has two halves, a tree uses half `slot`
    int next_slot = 0, n_pend = 0;
    int pend_slots[2] = {0, 0};                  // oldest first
    cudaEvent_t ev[2] = {nullptr, nullptr};      // recorded after a tree's last download was enqueued
    size_t timer_mark[2] = {0, 0};               // engine timers recorded before ev[slot]
    bool has_pred[2] = {false, false};           // h2ot_tree_finish_enqueue ran for the tree in this half
    size_t arena_half = 0, out_half = 0;
    std::vector<size_t> lvl_off;                 // offset of each level's slice inside an arena half
    size_t o_exists = 0, o_err = 0, o_conc = 0, o_pred = 0;  // offsets inside an out_stage half
  } dev;
  DBuf<double> d_vals;
  // small per-level results, packed in one device buffer so that h2ot_level_step brings them back with one copy:
  // [min2 | maxIn | wNA | conc : LC doubles each][best split per leaf][has_vals : LC bytes]
  DBuf<unsigned char> d_pack;
  double *p_min2 = nullptr, *p_maxIn = nullptr, *p_wna = nullptr, *p_conc = nullptr;
  h2ot_split* p_best = nullptr;
  uint8_t* p_has = nullptr;
  size_t pack_bytes = 0;
  DBuf<h2ot_split> d_cand;
  DBuf<long long> d_gamma;
  DBuf<float> d_pred;
  DBuf<double> d_factor;  // per node: 1 + noise of GBM's pred_noise_bandwidth (h2ot_add_tree_scaled)
  int max_split_nbin = 0;
  bool any_cat = false;
  bool any_checkfloat = false;  // some descriptor of the level asks for the float-split correction (DRF)
};

// minimal NCCL surface, resolved with dlopen so that the library has no link-time NCCL dependency
struct IdByValue { char internal[H2OT_COMM_ID_BYTES]; };
struct Nccl {
  void* lib = nullptr;
  void* comm = nullptr;
  int world = 1, rank = 0;
  bool shared = false;  // the communicator belongs to the process-wide table (never destroyed by an engine)
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, IdByValue, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

}  // namespace h2ot

using namespace h2ot;

struct h2ot_engine {
  h2ot_config cfg{};
  int64_t n = 0, n_pad = 0;
  int C = 0, K = 0;
  int sm_count = 148;
  int max_smem_optin = 0;
  int split_smem_max = 0;
  cudaStream_t stream = nullptr;
  std::vector<Column> cols;
  DBuf<ColRef> d_cols;
  double* resp = nullptr;
  double* w = nullptr;
  double* off = nullptr;
  double* resp_staging = nullptr;
  double* oobt = nullptr;  // DRF: OUT_BAG_TREES column
  double* staging_all = nullptr;
  DBuf<uint8_t> d_chunk;  // device copy of the Chunk._mem being decoded (h2ot_load_chunk), grow-only
  DBuf<int32_t> d_chunk_err;
  unsigned char* col_block = nullptr;  // all encoded columns (each 256-byte aligned)
  double w_max = 1.0;
  double gamma_bound[2] = {1.0, 1.0};  // max |gammaNum / w|, max |gammaDenom / w| of the current residuals
  bool frame_final = false;
  std::vector<TreeState> ts;
  // device pointer tables
  DBuf<double*> d_Fp;
  DBuf<float*> d_workp;
  DBuf<int32_t*> d_nidsp;
  DBuf<unsigned int> d_absmax;
  DBuf<unsigned long long> d_scalar;
  DBuf<RollupDev> d_rollup;
  h2ot_stats stats{};
  std::vector<Timer> timers;
  std::vector<cudaEvent_t> event_pool;
  bool defer_sync = false;
  bool timer_open = false;
  int timer_mode = 2;  // h2ot_set_timers
  DBuf<uint64_t> q_keys[2];       // h2ot_leaf_quantiles: sort buffers (value keys, strata) and CUB's scratch
  DBuf<uint32_t> q_strata[2];
  DBuf<unsigned char> q_tmp;
  DBuf<unsigned long long> q_counts;
  DBuf<long long> q_idx;
  unsigned int* bounds_host = nullptr;  // pinned: residual maxima of h2ot_compute_residuals_async
  cudaEvent_t bounds_ev = nullptr;
  bool bounds_pending = false;
  bool in_tree_loop = false;  // inside h2ot_tree_levels_begin: per-level housekeeping copies are done once per tree
  bool lazy_sync = false;          // h2ot_set_lazy_sync: calls that return nothing to the host do not wait for the stream
  std::vector<double> rate_per_class;  // Sample's _rate_per_class (empty: null)
  DBuf<double> d_rate_per_class;
  std::vector<double> col_conc;    // per column: chance that two rows of the root share a bin (0 until a root level ran)
  int steps_pending = 0;  // h2ot_level_step_begin calls not yet ended
  int hist_mode = 0;
  bool no_agg = false;             // H2OT_NO_AGG=1: never aggregate few-valued columns inside the warp (A/B timing)
  int smem_budget_kb = 0;  // 0: automatic
  Nccl nccl;
  // peer-memory exchange (k_xreduce.cuh): [in: xr_cap words][out: xr_cap words][flags: kXrMaxWorld words]
  bool xr_enabled = false, xr_failed = false, xr_coop = true;
  bool xr_ll = true;  // the push exchange with flagged lines (xr_ll_kernel); H2OT_XREDUCE=3 / 2 select the older pull versions
  unsigned long long xr_ll_epoch = 0;
  unsigned long long* xr_buf = nullptr;
  size_t xr_cap = 0;
  XrPeers xr_peers{};
  void* xr_opened[kXrMaxWorld] = {nullptr};
  unsigned long long xr_epoch = 0;
  std::string xr_key;  // communicator id + device: the process-wide parking slot of this engine's exchange buffer
  int xr_owner = -1;  // class-tree whose level lives in xr_buf between its build and its finish (one at a time)
  int* xr_err = nullptr;
  int* xr_err_host = nullptr;  // pinned mirror of xr_err, refreshed after every exchange
  // persistent scratch of the small host-side all-reduces (h2ot_allreduce_host_f64) and of the rank barrier
  DBuf<double> ar_dev;
  double* ar_host = nullptr;  // pinned
  size_t ar_host_cap = 0;
};

namespace {

int check(h2ot_engine* h) {
  if (!h) H2OT_FAIL(H2OT_ERR_INVALID, "null engine handle");
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) H2OT_FAIL(H2OT_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
  return H2OT_OK;
}
#define CHECK_ENGINE(h)         \
  do {                          \
    int _r = check(h);          \
    if (_r != H2OT_OK) return _r; \
  } while (0)
#define CHECK_K(h, k) \
  if ((k) < 0 || (k) >= (h)->K) H2OT_FAIL(H2OT_ERR_INVALID, "class-tree index out of range")
#define TRY(expr)               \
  do {                          \
    int _r = (expr);            \
    if (_r != H2OT_OK) return _r; \
  } while (0)

void xr_release(h2ot_engine* h);
bool xr_park(h2ot_engine* h);
int xr_rank_barrier(h2ot_engine* h);
int xr_ensure(h2ot_engine* h, size_t words, bool exact = false);
int xr_allreduce(h2ot_engine* h, long long n_sums, long long n);

int t_begin(h2ot_engine* h, double* acc) {
  // Every timed section costs two event records in the stream (~3 us: 0.14 ms of a 1.5 ms tree on an eighth of the Higgs frame):
  // mode 2 times every stage, mode 1 the histogram launches only (what the roofline needs), mode 0 nothing.
  h->timer_open = h->timer_mode >= 2 || (h->timer_mode == 1 && acc == &h->stats.ms_hist);
  if (!h->timer_open) return H2OT_OK;
  Timer t;
  for (cudaEvent_t* e : {&t.a, &t.b}) {
    if (!h->event_pool.empty()) {
      *e = h->event_pool.back();
      h->event_pool.pop_back();
    } else {
      CUDA_TRY(cudaEventCreate(e));
    }
  }
  t.acc = acc;
  CUDA_TRY(cudaEventRecord(t.a, h->stream));
  h->timers.push_back(t);
  return H2OT_OK;
}
int t_end(h2ot_engine* h) {
  if (!h->timer_open) return H2OT_OK;
  h->timer_open = false;
  CUDA_TRY(cudaEventRecord(h->timers.back().b, h->stream));
  return H2OT_OK;
}
// synchronise the stream and fold the per-family event times into the stats
int finish_call(h2ot_engine* h) {
  if (h->defer_sync) return H2OT_OK;  // inside h2ot_level_step: one synchronisation at the end of the whole level
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  for (Timer& t : h->timers) {
    float ms = 0;
    if (t.a && t.b && cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) *t.acc += ms;
    else cudaGetLastError();  // a call that failed half way leaves an unrecorded event behind: not an error of this call
    if (t.a) h->event_pool.push_back(t.a);
    if (t.b) h->event_pool.push_back(t.b);
  }
  h->timers.clear();
  for (TreeState& ts : h->ts) ts.dev.timer_mark[0] = ts.dev.timer_mark[1] = 0;
  CUDA_TRY(cudaGetLastError());
  if (h->xr_err_host && *h->xr_err_host) H2OT_FAIL(H2OT_ERR_COMM, "peer-memory exchange timed out waiting for another rank");
  return H2OT_OK;
}
// end of a call that hands nothing back to the host: with lazy sync on, the stream keeps running and the kernel timers and any
// error are picked up by the next call that does synchronise
int finish_call_lazy(h2ot_engine* h) {
  if (h->lazy_sync) {
    CUDA_TRY(cudaGetLastError());
    return H2OT_OK;
  }
  return finish_call(h);
}
int h2d(h2ot_engine* h, void* dst, const void* src, size_t bytes) {
  if (!bytes) return H2OT_OK;
  CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, h->stream));
  h->stats.h2d_bytes += (int64_t)bytes;
  return H2OT_OK;
}
int d2h(h2ot_engine* h, void* dst, const void* src, size_t bytes) {
  if (!bytes) return H2OT_OK;
  CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, h->stream));
  h->stats.d2h_bytes += (int64_t)bytes;
  return H2OT_OK;
}
int grid_for(h2ot_engine* h, int64_t n, int block, int per_sm = 8) {
  int64_t need = (n + block - 1) / block;
  int64_t cap = (int64_t)h->sm_count * per_sm;
  return (int)std::max<int64_t>(1, std::min(need, cap));
}
// 2^e scale such that N * bound * 2^e <= 2^62 and, when cap_bits > 0, bound * 2^e <= 2^cap_bits
int exponent_for(double bound, int bitsN, int cap_bits = 0) {
  if (!(bound > 0) || std::isinf(bound)) bound = 1.0;
  int e2;
  if (std::frexp(bound, &e2) == 0.5) e2--;  // bound <= 2^e2 (an exact power of two is its own bound: |y - p| <= 1 keeps e2 = 0)
  int e = 62 - bitsN - e2;
  if (cap_bits > 0) e = std::min(e, cap_bits - e2);
  return std::max(-900, std::min(900, e));
}

}  // namespace

extern "C" {

int h2ot_version(void) { return 100; }
const char* h2ot_last_error(void) { return g_last_error.c_str(); }

int h2ot_device_count(int32_t* out) {
  int n = 0;
  CUDA_TRY(cudaGetDeviceCount(&n));
  *out = n;
  return H2OT_OK;
}

int h2ot_destroy(h2ot_engine* h);
static int create_into(const h2ot_config* cfg, h2ot_engine* h) {
  h->cfg = *cfg;
  if (h->cfg.nrows_global <= 0) h->cfg.nrows_global = cfg->nrows;
  if (h->cfg.family == H2OT_FAMILY_QUANTILE) {  // the slot holds _quantile_alpha (default 0.5, H/Model.java Parameters)
    if (!(h->cfg.tweedie_power > 0.0 && h->cfg.tweedie_power < 1.0)) h->cfg.tweedie_power = 0.5;
  } else if (!(h->cfg.tweedie_power > 1.0 && h->cfg.tweedie_power < 2.0)) h->cfg.tweedie_power = 1.5;
  h->n = cfg->nrows;
  h->n_pad = (cfg->nrows + kRowPad - 1) / kRowPad * kRowPad;
  h->C = cfg->ncols;
  h->K = cfg->nclass_trees;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, cfg->device));
  h->sm_count = prop.multiProcessorCount;
  h->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  if (const char* s = getenv("H2OT_SMEM_BUDGET_KB")) h->smem_budget_kb = atoi(s);
  if (const char* s = getenv("H2OT_NO_AGG")) h->no_agg = atoi(s) != 0;
  CUDA_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  h->cols.resize(h->C);
  h->col_conc.assign(h->C, 0.0);
  h->ts.resize(h->K);
  const size_t np = (size_t)h->n_pad;
  CUDA_TRY(cudaMalloc(&h->resp, np * sizeof(double)));
  CUDA_TRY(cudaMemsetAsync(h->resp, 0xFF, np * sizeof(double), h->stream));  // all-ones == NaN == NA
  if (cfg->has_weights) {
    CUDA_TRY(cudaMalloc(&h->w, np * sizeof(double)));
    CUDA_TRY(cudaMemsetAsync(h->w, 0, np * sizeof(double), h->stream));
  }
  if (cfg->has_offset) {
    CUDA_TRY(cudaMalloc(&h->off, np * sizeof(double)));
    CUDA_TRY(cudaMemsetAsync(h->off, 0, np * sizeof(double), h->stream));
  }
  std::vector<double*> Fp(h->K);
  std::vector<float*> wp(h->K);
  std::vector<int32_t*> np_(h->K);
  for (int k = 0; k < h->K; k++) {
    TreeState& t = h->ts[k];
    CUDA_TRY(cudaMalloc(&t.F, np * sizeof(double)));
    CUDA_TRY(cudaMalloc(&t.work, np * sizeof(float)));
    CUDA_TRY(cudaMalloc(&t.nids, np * sizeof(int32_t)));
    CUDA_TRY(cudaMalloc(&t.code, np * sizeof(uint16_t)));
    CUDA_TRY(cudaMemsetAsync(t.F, 0, np * sizeof(double), h->stream));
    CUDA_TRY(cudaMemsetAsync(t.work, 0, np * sizeof(float), h->stream));
    CUDA_TRY(cudaMemsetAsync(t.nids, 0, np * sizeof(int32_t), h->stream));
    CUDA_TRY(cudaMemsetAsync(t.code, 0xFF, np * sizeof(uint16_t), h->stream));
    Fp[k] = t.F;
    wp[k] = t.work;
    np_[k] = t.nids;
  }
  TRY(h->d_Fp.ensure(h->K));
  TRY(h->d_workp.ensure(h->K));
  TRY(h->d_nidsp.ensure(h->K));
  TRY(h->d_absmax.ensure(h->K + 2));
  TRY(h->d_scalar.ensure(8));
  TRY(h->d_rollup.ensure(1));
  TRY(h->d_cols.ensure(h->C));
  CUDA_TRY(cudaMemcpyAsync(h->d_Fp.p, Fp.data(), h->K * sizeof(double*), cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaMemcpyAsync(h->d_workp.p, wp.data(), h->K * sizeof(float*), cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaMemcpyAsync(h->d_nidsp.p, np_.data(), h->K * sizeof(int32_t*), cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  // opt in to the large dynamic shared memory carve-out once
#define H2OT_OPTIN(ST)                                                                                                                      \
  CUDA_TRY(cudaFuncSetAttribute(hist_smem_kernel<5, false, ST, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin)); \
  CUDA_TRY(cudaFuncSetAttribute(hist_smem_kernel<7, true, ST, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin));  \
  CUDA_TRY(cudaFuncSetAttribute(hist_smem_kernel<5, false, ST, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin));  \
  CUDA_TRY(cudaFuncSetAttribute(hist_smem_kernel<7, true, ST, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin));   \
  CUDA_TRY(cudaFuncSetAttribute(hist_smem_kernel<5, false, ST, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin));  \
  CUDA_TRY(cudaFuncSetAttribute(hist_smem_kernel<7, true, ST, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin))
  H2OT_OPTIN(H2OT_STORE_F32);
  H2OT_OPTIN(H2OT_STORE_F64);
  H2OT_OPTIN(H2OT_STORE_U8);
  H2OT_OPTIN(H2OT_STORE_U16);
#undef H2OT_OPTIN
#define H2OT_OPTIN_SPARSE(W, HW, CF, SP) \
  CUDA_TRY(cudaFuncSetAttribute(hist_sparse_kernel<W, HW, CF, SP>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem_optin))
  H2OT_OPTIN_SPARSE(5, false, false, false);
  H2OT_OPTIN_SPARSE(5, false, true, false);
  H2OT_OPTIN_SPARSE(7, true, false, false);
  H2OT_OPTIN_SPARSE(7, true, true, false);
  H2OT_OPTIN_SPARSE(5, false, false, true);
  H2OT_OPTIN_SPARSE(5, false, true, true);
  H2OT_OPTIN_SPARSE(7, true, false, true);
  H2OT_OPTIN_SPARSE(7, true, true, true);
#undef H2OT_OPTIN_SPARSE
  {
    cudaFuncAttributes fa;
    CUDA_TRY(cudaFuncGetAttributes(&fa, find_split_kernel));
    h->split_smem_max = h->max_smem_optin - (int)fa.sharedSizeBytes;  // the opt-in limit covers static + dynamic
    CUDA_TRY(cudaFuncSetAttribute(find_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, h->split_smem_max));
  }
  return H2OT_OK;
}

int h2ot_create(const h2ot_config* cfg, h2ot_engine** out) {
  if (!cfg || !out) H2OT_FAIL(H2OT_ERR_INVALID, "null argument");
  if (cfg->nrows <= 0 || cfg->ncols <= 0 || cfg->nclass_trees <= 0) H2OT_FAIL(H2OT_ERR_INVALID, "nrows/ncols/nclass_trees must be positive");
  if (cfg->nrows > ((int64_t)1 << 32) - 2 * kRowPad)  // row ids travel as 32-bit values inside the level kernels; a shard is far smaller anyway
    H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "more than 2^32 rows on one GPU: shard the rows over more engines");
  if (cfg->nclass_trees > 64) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "at most 64 class-trees");
  if (cfg->family < H2OT_FAMILY_GAUSSIAN || cfg->family > H2OT_FAMILY_QUANTILE) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "unknown distribution family");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    H2OT_FAIL(H2OT_ERR_CUDA, std::string("no CUDA device available (there is no CPU fallback): ") + cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= ndev) H2OT_FAIL(H2OT_ERR_INVALID, "device ordinal out of range");
  CUDA_TRY(cudaSetDevice(cfg->device));
  h2ot_engine* h = new h2ot_engine();
  const int rc = create_into(cfg, h);
  if (rc != H2OT_OK) {  // give back whatever was allocated before the failure (out of memory on a large shard, most likely)
    const std::string why = g_last_error;
    h2ot_destroy(h);
    set_error(why);
    return rc;
  }
  *out = h;
  return H2OT_OK;
}

int h2ot_destroy(h2ot_engine* h) {
  if (!h) return H2OT_OK;
  cudaSetDevice(h->cfg.device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  if (!xr_park(h)) xr_release(h);  // (release is collective when peers have mapped the buffer: needs the communicator)
  if (h->nccl.comm && h->nccl.CommDestroy && !h->nccl.shared) h->nccl.CommDestroy(h->nccl.comm);
  h->ar_dev.release();
  if (h->ar_host) cudaFreeHost(h->ar_host);
  for (Column& c : h->cols) {
    c.data = nullptr;
  }
  if (h->col_block) cudaFree(h->col_block);
  if (h->staging_all) cudaFree(h->staging_all);
  h->d_chunk.release();
  h->d_chunk_err.release();
  if (h->resp) cudaFree(h->resp);
  if (h->w) cudaFree(h->w);
  if (h->off) cudaFree(h->off);
  if (h->oobt) cudaFree(h->oobt);
  if (h->xr_err) cudaFree(h->xr_err);
  if (h->xr_err_host) cudaFreeHost(h->xr_err_host);
  h->q_keys[0].release(); h->q_keys[1].release(); h->q_strata[0].release(); h->q_strata[1].release(); h->q_tmp.release();
  h->q_counts.release(); h->q_idx.release();
  if (h->bounds_host) cudaFreeHost(h->bounds_host);
  if (h->bounds_ev) cudaEventDestroy(h->bounds_ev);
  for (TreeState& t : h->ts) {
    if (t.F) cudaFree(t.F);
    if (t.work) cudaFree(t.work);
    if (t.nids) cudaFree(t.nids);
    if (t.code) cudaFree(t.code);
    t.nodes.release(); t.bitsets.release(); t.d_lvl.release();
    if (t.lvl_host) cudaFreeHost(t.lvl_host);
    if (t.stage) cudaFreeHost(t.stage);
    if (t.done_ev) cudaEventDestroy(t.done_ev);
    t.d_sums.release(); t.d_vals.release(); t.d_pack.release();
    t.d_cand.release(); t.d_gamma.release(); t.d_pred.release(); t.d_factor.release(); t.d_sp.release();
    t.dev.lvl_alt.release(); t.dev.pack_alt.release(); t.dev.maxex[0].release(); t.dev.maxex[1].release(); t.dev.out_dev.release();
    t.dev.nb_caps.release(); t.dev.root_maxex.release();
    for (size_t i = 0; i < t.dev.lvl_stage.size(); i++) if (t.dev.lvl_stage[i] && t.dev.lvl_stage_own[i]) cudaFreeHost(t.dev.lvl_stage[i]);
    if (t.dev.arena) cudaFreeHost(t.dev.arena);
    if (t.dev.out_stage) cudaFreeHost(t.dev.out_stage);
    for (cudaEvent_t e : t.dev.ev) if (e) cudaEventDestroy(e);
  }
  h->d_Fp.release(); h->d_workp.release(); h->d_nidsp.release(); h->d_absmax.release(); h->d_scalar.release();
  h->d_rollup.release(); h->d_cols.release(); h->d_rate_per_class.release();
  for (cudaEvent_t e : h->event_pool) cudaEventDestroy(e);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return H2OT_OK;
}

// ------------------------------------------------------------------------------------------------ ingest
static int load_rows(h2ot_engine* h, double* dst, int64_t row0, int64_t n, const double* values) {
  if (!values || row0 < 0 || n < 0 || row0 + n > h->n) H2OT_FAIL(H2OT_ERR_INVALID, "row range out of bounds");
  TRY(h2d(h, dst + row0, values, (size_t)n * sizeof(double)));
  return finish_call(h);
}

int h2ot_load_column(h2ot_engine* h, int32_t col, int64_t row0, int64_t n, const double* values) {
  CHECK_ENGINE(h);
  if (col < 0 || col >= h->C) H2OT_FAIL(H2OT_ERR_INVALID, "column index out of range");
  if (h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame already finalized");
  Column& c = h->cols[col];
  if (!h->staging_all) {  // one staging area for the whole frame (freed by h2ot_finalize_frame); unloaded rows are NA
    const size_t bytes = (size_t)h->C * (size_t)h->n * sizeof(double);
    CUDA_TRY(cudaMalloc(&h->staging_all, bytes));
    CUDA_TRY(cudaMemsetAsync(h->staging_all, 0xFF, bytes, h->stream));
  }
  c.staging = h->staging_all + (size_t)col * (size_t)h->n;
  return load_rows(h, c.staging, row0, n, values);
}
int h2ot_load_response(h2ot_engine* h, int64_t row0, int64_t n, const double* values) {
  CHECK_ENGINE(h);
  return load_rows(h, h->resp, row0, n, values);
}
int h2ot_load_weights(h2ot_engine* h, int64_t row0, int64_t n, const double* values) {
  CHECK_ENGINE(h);
  if (!h->w) H2OT_FAIL(H2OT_ERR_INVALID, "engine was created without a weights column");
  return load_rows(h, h->w, row0, n, values);
}
int h2ot_load_offset(h2ot_engine* h, int64_t row0, int64_t n, const double* values) {
  CHECK_ENGINE(h);
  if (!h->off) H2OT_FAIL(H2OT_ERR_INVALID, "engine was created without an offset column");
  return load_rows(h, h->off, row0, n, values);
}

namespace {
inline int32_t rd_i32(const uint8_t* p) { int32_t v; memcpy(&v, p, 4); return v; }
inline int64_t rd_i64(const uint8_t* p) { int64_t v; memcpy(&v, p, 8); return v; }
inline double rd_f64(const uint8_t* p) { double v; memcpy(&v, p, 8); return v; }
// PrettyPrint.pow10(int) for exp >= 0 (W/util/PrettyPrint.java:80-129): a table up to 1e10, Math.pow(10, exp) beyond -- exact
// powers of ten are representable up to 1e22, past that the C library's pow stands in for the JVM's
inline double pow10_nonneg(int e) {
  static const double t[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15,
                             1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  return e <= 22 ? t[e] : std::pow(10.0, (double)e);
}
}  // namespace

int h2ot_load_chunk(h2ot_engine* h, int32_t col, int64_t row0, int64_t n, int32_t codec, const uint8_t* mem, int64_t mem_len) {
  CHECK_ENGINE(h);
  if (!mem || mem_len < 0 || n < 0 || row0 < 0 || row0 + n > h->n) H2OT_FAIL(H2OT_ERR_INVALID, "bad chunk / row range out of bounds");
  double* dst = nullptr;
  if (col >= 0) {
    if (col >= h->C) H2OT_FAIL(H2OT_ERR_INVALID, "column index out of range");
    if (h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame already finalized");
  } else if (col == H2OT_COL_RESPONSE) {
    dst = h->resp;
  } else if (col == H2OT_COL_WEIGHTS) {
    if (!h->w) H2OT_FAIL(H2OT_ERR_INVALID, "engine was created without a weights column");
    dst = h->w;
  } else if (col == H2OT_COL_OFFSET) {
    if (!h->off) H2OT_FAIL(H2OT_ERR_INVALID, "engine was created without an offset column");
    dst = h->off;
  } else {
    H2OT_FAIL(H2OT_ERR_INVALID, "column index out of range");
  }
  // ---- header, as Chunk.initFromBytes reads it; the chunk's own row count must be n
  ChunkArgs a{};
  a.codec = codec;
  a.n = n;
  int64_t rows = -1;
  auto cs_header = [&](int log) -> bool {  // CSChunk.initFromBytes :57-64
    if (mem_len < kCSOff) return false;
    a.bias = rd_i64(mem);
    const int x = rd_i32(mem + 8), sz = rd_i32(mem + 12);
    a.decimal = sz < 0 ? 1 : 0;
    if ((sz < 0 ? -sz - 1 : sz) != log) return false;
    const int e = a.decimal ? -x : x;  // _scale = PrettyPrint.pow10(1, e): e < 0 ? 1 / pow10(-e) : 1 * pow10(e)
    if (e < -330 || e > 330) return false;
    a.scale = e < 0 ? 1.0 / pow10_nonneg(-e) : pow10_nonneg(e);
    rows = (mem_len - kCSOff) >> log;
    return ((mem_len - kCSOff) & ((1 << log) - 1)) == 0;
  };
  bool ok = true;
  switch (codec) {
    case H2OT_CODEC_C0D: ok = mem_len == 12; if (ok) { a.con = rd_f64(mem); rows = rd_i32(mem + 8); } break;
    case H2OT_CODEC_C0L: ok = mem_len == 12; if (ok) { a.con = (double)rd_i64(mem); rows = rd_i32(mem + 8); } break;
    case H2OT_CODEC_C1:
    case H2OT_CODEC_C1N: rows = mem_len; break;
    case H2OT_CODEC_C1S: ok = cs_header(0); break;
    case H2OT_CODEC_C2: ok = (mem_len & 1) == 0; rows = mem_len >> 1; break;
    case H2OT_CODEC_C2S: ok = cs_header(1); break;
    case H2OT_CODEC_C4:
    case H2OT_CODEC_C4F: ok = (mem_len & 3) == 0; rows = mem_len >> 2; break;
    case H2OT_CODEC_C4S: ok = cs_header(2); break;
    case H2OT_CODEC_C8:
    case H2OT_CODEC_C8D: ok = (mem_len & 7) == 0; rows = mem_len >> 3; break;
    case H2OT_CODEC_CBS:  // CBSChunk.initFromBytes :155-160
      ok = mem_len >= 2 && (mem[1] == 1 || mem[1] == 2) && mem[0] < 8;
      if (ok) { a.bpv = mem[1]; rows = ((mem_len - 2) * 8 - mem[0]) / a.bpv; }
      break;
    case H2OT_CODEC_CUD:  // CUDChunk.initFromBytes :91-96, computeByteSize :14-18
      ok = mem_len >= 8;
      if (ok) {
        rows = rd_i32(mem);
        a.nuniq = rd_i32(mem + 4);
        ok = a.nuniq >= 0 && a.nuniq <= 256 && rows >= 0 && mem_len >= 8 + ((int64_t)a.nuniq << 3) + rows;
        for (int64_t i = 0; ok && i < rows; i++) ok = (int)(int8_t)mem[8 + ((int64_t)a.nuniq << 3) + i] + 128 < a.nuniq;
      }
      break;
    case H2OT_CODEC_CXI:
    case H2OT_CODEC_CXF:  // CXIChunk.initFromBytes :58-66
      ok = mem_len >= 8;
      if (ok) {
        rows = rd_i32(mem);
        a.id_sz = mem[4];
        a.val_sz = mem[5];
        a.na_sparse = mem[6] == 1 ? 1 : 0;
        const int id = a.id_sz, vs = a.val_sz;
        if (codec == H2OT_CODEC_CXI) ok = (id == 2 && (vs == 0 || vs == 2)) || (id == 4 && (vs == 0 || vs == 4 || vs == 8));
        else ok = id == 4 && (vs == 4 || vs == 8);
        if (ok) {
          ok = (mem_len - 8) % (id + vs) == 0;
          a.n_elems = (mem_len - 8) / (id + vs);
        }
      }
      break;
    default: H2OT_FAIL(H2OT_ERR_INVALID, "unknown chunk codec");
  }
  if (!ok) H2OT_FAIL(H2OT_ERR_INVALID, "malformed chunk (header and byte length disagree, or an unknown element layout)");
  if (rows != n) H2OT_FAIL(H2OT_ERR_INVALID, "chunk holds " + std::to_string(rows) + " rows, caller said " + std::to_string(n));
  if (col >= 0) {
    Column& c = h->cols[col];
    if (!h->staging_all) {
      const size_t bytes = (size_t)h->C * (size_t)h->n * sizeof(double);
      CUDA_TRY(cudaMalloc(&h->staging_all, bytes));
      CUDA_TRY(cudaMemsetAsync(h->staging_all, 0xFF, bytes, h->stream));
    }
    c.staging = h->staging_all + (size_t)col * (size_t)h->n;
    dst = c.staging;
  }
  if (n == 0) return H2OT_OK;
  TRY(h->d_chunk.ensure((size_t)std::max<int64_t>(mem_len, 16)));
  TRY(h->d_chunk_err.ensure(1));
  TRY(h2d(h, h->d_chunk.p, mem, (size_t)mem_len));
  CUDA_TRY(cudaMemsetAsync(h->d_chunk_err.p, 0, sizeof(int32_t), h->stream));
  a.mem = h->d_chunk.p;
  a.dst = dst + row0;
  a.err = h->d_chunk_err.p;
  chunk_decode_kernel<<<grid_for(h, n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  if ((codec == H2OT_CODEC_CXI || codec == H2OT_CODEC_CXF) && a.n_elems > 0) {
    chunk_scatter_kernel<<<grid_for(h, a.n_elems, 256), 256, 0, h->stream>>>(a);
    h->stats.launches++;
    int32_t err = 0;
    CUDA_TRY(cudaMemcpyAsync(&err, h->d_chunk_err.p, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    TRY(finish_call(h));
    if (err) H2OT_FAIL(H2OT_ERR_INVALID, "sparse chunk: row id out of range");
    return H2OT_OK;
  }
  return finish_call(h);
}

int h2ot_debug_get_column(h2ot_engine* h, int32_t col, int64_t row0, int64_t n, double* out) {
  CHECK_ENGINE(h);
  if (col < 0 || col >= h->C || !out || row0 < 0 || n < 0 || row0 + n > h->n) H2OT_FAIL(H2OT_ERR_INVALID, "bad column / row range");
  if (!h->cols[col].finalized) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  if (n == 0) return H2OT_OK;
  DBuf<double> tmp;
  TRY(tmp.ensure((size_t)n));
  ColRef cr{h->cols[col].data, h->cols[col].base, h->cols[col].storage, 0};
  column_values_kernel<<<grid_for(h, n, 256), 256, 0, h->stream>>>(cr, row0, n, tmp.p);
  h->stats.launches++;
  int rc = d2h(h, out, tmp.p, (size_t)n * sizeof(double));
  if (rc == H2OT_OK) rc = finish_call(h);
  else cudaStreamSynchronize(h->stream);
  tmp.release();
  return rc;
}

int h2ot_finalize_frame(h2ot_engine* h) {
  CHECK_ENGINE(h);
  if (h->frame_final) return H2OT_OK;
  TRY(t_begin(h, &h->stats.ms_ingest));
  const int C = h->C;
  for (int ci = 0; ci < C; ci++)
    if (!h->cols[ci].staging) H2OT_FAIL(H2OT_ERR_INVALID, "column " + std::to_string(ci) + " was never loaded");
  // rollups of every column (+ the weights), one round trip
  TRY(h->d_rollup.ensure((size_t)C + 1));
  CUDA_TRY(cudaMemsetAsync(h->d_rollup.p, 0, ((size_t)C + 1) * sizeof(RollupDev), h->stream));
  for (int ci = 0; ci < C; ci++) {
    rollup_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(h->cols[ci].staging, h->n, h->d_rollup.p + ci);
    h->stats.launches++;
  }
  if (h->w) {
    rollup_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(h->w, h->n, h->d_rollup.p + C);
    h->stats.launches++;
  }
  std::vector<RollupDev> ru((size_t)C + 1);
  CUDA_TRY(cudaMemcpyAsync(ru.data(), h->d_rollup.p, ru.size() * sizeof(RollupDev), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  std::vector<ColRef> refs(C);
  std::vector<size_t> col_off(C, 0);
  size_t block_bytes = 0;
  for (int ci = 0; ci < C; ci++) {  // storage decision first, so that all encoded columns share ONE allocation
    const RollupDev& r = ru[ci];
    const bool any = r.max_key != 0;
    const bool is_int = !r.not_int;
    int storage = r.not_f32 ? H2OT_STORE_F64 : H2OT_STORE_F32;
    if (any && is_int) {
      const double span = dkey_inv(r.max_key) - dkey_inv(~r.min_key);
      if (span < 255) storage = H2OT_STORE_U8;
      else if (span < 65535) storage = H2OT_STORE_U16;
    }
    h->cols[ci].storage = storage;
    const size_t esz = storage == H2OT_STORE_F32 ? 4 : storage == H2OT_STORE_F64 ? 8 : storage == H2OT_STORE_U8 ? 1 : 2;
    col_off[ci] = block_bytes;
    block_bytes += ((size_t)h->n_pad * esz + 255) / 256 * 256;
  }
  CUDA_TRY(cudaMalloc(&h->col_block, std::max<size_t>(block_bytes, 256)));
  for (int ci = 0; ci < C; ci++) {
    Column& c = h->cols[ci];
    const RollupDev& r = ru[ci];
    c.info.nacnt = (int64_t)r.nacnt;
    const bool any = r.max_key != 0;
    c.info.min = any ? dkey_inv(~r.min_key) : NAN;
    c.info.max = any ? dkey_inv(r.max_key) : NAN;
    c.info.is_int = r.not_int ? 0 : 1;
    int storage = r.not_f32 ? H2OT_STORE_F64 : H2OT_STORE_F32;
    double base = 0;
    if (any && c.info.is_int) {
      const double span = c.info.max - c.info.min;
      if (span < 255) { storage = H2OT_STORE_U8; base = c.info.min; }
      else if (span < 65535) { storage = H2OT_STORE_U16; base = c.info.min; }
    }
    c.storage = storage;
    c.base = base;
    c.info.storage = storage;
    c.data = h->col_block + col_off[ci];
    encode_kernel<<<grid_for(h, h->n_pad, 256), 256, 0, h->stream>>>(c.staging, h->n, h->n_pad, storage, base, c.data);
    h->stats.launches++;
    c.finalized = true;
    refs[ci].data = c.data;
    refs[ci].base = base;
    refs[ci].storage = storage;
    refs[ci].pad = 0;
  }
  CUDA_TRY(cudaMemcpyAsync(h->d_cols.p, refs.data(), refs.size() * sizeof(ColRef), cudaMemcpyHostToDevice, h->stream));
  if (h->w) {
    const RollupDev& r = ru[C];
    h->w_max = r.max_key ? std::max(dkey_inv(r.max_key), -dkey_inv(~r.min_key)) : 1.0;
    if (!(h->w_max > 0)) h->w_max = 1.0;
  }
  TRY(t_end(h));
  h->frame_final = true;
  TRY(finish_call(h));
  for (Column& c : h->cols) c.staging = nullptr;
  CUDA_TRY(cudaFree(h->staging_all));
  h->staging_all = nullptr;
  return H2OT_OK;
}

int h2ot_column_info_get(h2ot_engine* h, int32_t col, h2ot_column_info* out) {
  CHECK_ENGINE(h);
  if (col < 0 || col >= h->C || !out) H2OT_FAIL(H2OT_ERR_INVALID, "bad column / null out");
  if (!h->cols[col].finalized) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  *out = h->cols[col].info;
  return H2OT_OK;
}

// ------------------------------------------------------------------------------------------------ model state
namespace {
__global__ void fill_f64_kernel(double* p, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void f32_to_f64_kernel(const float* s, double* d, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) d[i] = (double)s[i];
}
__global__ void f64_to_f32_kernel(const double* s, float* d, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) d[i] = (float)s[i];
}
}  // namespace

int h2ot_init_model(h2ot_engine* h, const double* init_f, int64_t* out_skipped) {
  CHECK_ENGINE(h);
  if (!init_f) H2OT_FAIL(H2OT_ERR_INVALID, "null init_f");
  for (int k = 0; k < h->K; k++) {
    fill_f64_kernel<<<grid_for(h, h->n_pad, 256), 256, 0, h->stream>>>(h->ts[k].F, h->n_pad, init_f[k]);
    h->stats.launches++;
  }
  CUDA_TRY(cudaMemsetAsync(h->d_scalar.p, 0, sizeof(unsigned long long), h->stream));
  InitArgs a{h->n, h->n_pad, h->K, h->resp, h->w, h->d_nidsp.p, h->d_scalar.p};
  init_rows_kernel<<<grid_for(h, h->n_pad, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  unsigned long long skipped = 0;
  TRY(d2h(h, &skipped, h->d_scalar.p, sizeof(skipped)));
  TRY(finish_call(h));
  if (out_skipped) *out_skipped = (int64_t)skipped;
  return H2OT_OK;
}

int h2ot_initial_stats(h2ot_engine* h, double* out_num, double* out_den, double* out_class_weight, double* out_abs_max) {
  CHECK_ENGINE(h);
  const int grid = grid_for(h, h->n, 256);
  const int nc = (out_class_weight && h->cfg.nclass > 1) ? h->cfg.nclass : 0;
  if (nc > kInitStatsClasses) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "class weights are computed for at most 1024 classes");
  DBuf<double> tmp;
  TRY(tmp.ensure((size_t)grid * 2 + nc + 1));
  CUDA_TRY(cudaMemsetAsync(tmp.p, 0, ((size_t)grid * 2 + nc + 1) * sizeof(double), h->stream));
  InitStatsArgs a{h->n, h->cfg.nclass, h->cfg.family, h->cfg.tweedie_power, h->resp, h->w, h->off, tmp.p, nc ? tmp.p + (size_t)grid * 2 : nullptr,
                  reinterpret_cast<unsigned int*>(tmp.p + (size_t)grid * 2 + nc)};
  init_stats_kernel<<<grid, 256, 0, h->stream>>>(a);
  h->stats.launches++;
  std::vector<double> hbuf((size_t)grid * 2 + nc + 1);
  TRY(d2h(h, hbuf.data(), tmp.p, hbuf.size() * sizeof(double)));
  const int rc = finish_call(h);
  tmp.release();
  if (rc != H2OT_OK) return rc;
  double num = 0, den = 0;
  for (int b = 0; b < grid; b++) {  // fixed order
    num += hbuf[2 * (size_t)b];
    den += hbuf[2 * (size_t)b + 1];
  }
  if (out_num) *out_num = num;
  if (out_den) *out_den = den;
  if (out_class_weight)
    for (int c = 0; c < nc; c++) out_class_weight[c] = hbuf[(size_t)grid * 2 + c];
  if (out_abs_max) {
    unsigned int bits;
    memcpy(&bits, &hbuf[(size_t)grid * 2 + nc], sizeof(bits));
    float f;
    memcpy(&f, &bits, sizeof(f));
    *out_abs_max = (double)f;
  }
  return H2OT_OK;
}

int h2ot_set_predictions(h2ot_engine* h, int32_t k, const double* f) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TRY(h2d(h, h->ts[k].F, f, (size_t)h->n * sizeof(double)));
  return finish_call(h);
}
int h2ot_get_predictions(h2ot_engine* h, int32_t k, double* f) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TRY(d2h(h, f, h->ts[k].F, (size_t)h->n * sizeof(double)));
  return finish_call(h);
}
int h2ot_get_work(h2ot_engine* h, int32_t k, double* work) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  DBuf<double> tmp;
  TRY(tmp.ensure(h->n));
  f32_to_f64_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(h->ts[k].work, tmp.p, h->n);
  h->stats.launches++;
  TRY(d2h(h, work, tmp.p, (size_t)h->n * sizeof(double)));
  int r = finish_call(h);
  tmp.release();
  return r;
}
int h2ot_set_work(h2ot_engine* h, int32_t k, const double* work, double abs_bound) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  DBuf<double> tmp;
  TRY(tmp.ensure(h->n));
  TRY(h2d(h, tmp.p, work, (size_t)h->n * sizeof(double)));
  f64_to_f32_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(tmp.p, h->ts[k].work, h->n);
  h->stats.launches++;
  h->ts[k].work_bound = abs_bound;
  int r = finish_call(h);
  tmp.release();
  return r;
}
int h2ot_get_nids(h2ot_engine* h, int32_t k, int32_t* nids) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TRY(d2h(h, nids, h->ts[k].nids, (size_t)h->n * sizeof(int32_t)));
  return finish_call(h);
}
int h2ot_set_nids(h2ot_engine* h, int32_t k, const int32_t* nids) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TRY(h2d(h, h->ts[k].nids, nids, (size_t)h->n * sizeof(int32_t)));
  return finish_call(h);
}

// ------------------------------------------------------------------------------------------------ per tree
int h2ot_compute_residuals(h2ot_engine* h, double* out_abs_max) {
  CHECK_ENGINE(h);
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  CUDA_TRY(cudaMemsetAsync(h->d_absmax.p, 0, (h->K + 2) * sizeof(unsigned int), h->stream));
  ResidArgs a{h->n, h->K, h->cfg.nclass, h->cfg.family, h->resp, h->w, h->off, h->d_Fp.p, h->d_workp.p, h->d_absmax.p, h->cfg.tweedie_power};
  TRY(t_begin(h, &h->stats.ms_residuals));
  residuals_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  TRY(t_end(h));
  std::vector<unsigned int> bits(h->K + 2);
  TRY(d2h(h, bits.data(), h->d_absmax.p, (h->K + 2) * sizeof(unsigned int)));
  TRY(finish_call(h));
  for (int q = 0; q < 2; q++) {
    float f;
    memcpy(&f, &bits[h->K + q], 4);
    h->gamma_bound[q] = f > 0 ? (double)f : 1.0;
  }
  for (int k = 0; k < h->K; k++) {
    float f;
    memcpy(&f, &bits[k], 4);
    if (out_abs_max) out_abs_max[k] = (double)f;
    h->ts[k].work_bound = (double)f;
  }
  return H2OT_OK;
}

int h2ot_get_gamma_bounds(h2ot_engine* h, double out[2]) {
  if (!h || !out) H2OT_FAIL(H2OT_ERR_INVALID, "null argument");
  out[0] = h->gamma_bound[0];
  out[1] = h->gamma_bound[1];
  return H2OT_OK;
}
int h2ot_set_gamma_bounds(h2ot_engine* h, const double in[2]) {
  if (!h || !in || !(in[0] > 0) || !(in[1] > 0)) H2OT_FAIL(H2OT_ERR_INVALID, "gamma bounds must be positive");
  h->gamma_bound[0] = in[0];
  h->gamma_bound[1] = in[1];
  return H2OT_OK;
}

int h2ot_begin_tree(h2ot_engine* h, int32_t k, int64_t tree_seed, double sample_rate, double work_abs_bound) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  t.n_nodes = 0;
  t.built = t.finished = false;
  t.level_failed = false;
  if (h->xr_owner == k) h->xr_owner = -1;
  t.L = 0;
  t.prev_leaf_offset = -1;
  if (work_abs_bound > 0) t.work_bound = work_abs_bound;
  const int bitsN = ceil_log2_u64((uint64_t)h->cfg.nrows_global + 1);
  const double b = t.work_bound > 0 ? t.work_bound : 1.0;
  t.ey = exponent_for(h->w_max * b, bitsN);
  t.eyy = exponent_for(h->w_max * b * b, bitsN);
  if (sample_rate < 1.0 || !h->rate_per_class.empty()) {  // SharedTreeParameters.useRowSampling (T/SharedTreeModel.java:86-88)
    TRY(t_begin(h, &h->stats.ms_sample));
    sample_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(h->n, h->cfg.row_offset, (long long)tree_seed, (float)sample_rate, sample_rate,
                                                                 h->rate_per_class.empty() ? nullptr : h->d_rate_per_class.p,
                                                                 (int)h->rate_per_class.size(), h->resp, t.nids);
    h->stats.launches++;
    TRY(t_end(h));
  }
  return finish_call_lazy(h);
}


// ------------------------------------------------------------------------------------------------ per level
}  // extern "C" (helpers below are internal)

namespace {

// Merge caller-supplied DecidedNodes (+ their bitset bytes) into the per-tree node table (host mirror + device).
int upload_nodes(h2ot_engine* h, TreeState& t, int n_decided, const h2ot_decided_node* decided, const uint8_t* pool,
                 int64_t pool_bytes, bool replace_all) {
  if (replace_all) {
    t.h_nodes.clear();
    t.h_bitsets.clear();
    t.n_nodes = 0;
  }
  if (n_decided <= 0) return H2OT_OK;
  if (!decided) H2OT_FAIL(H2OT_ERR_INVALID, "null decided-node array");
  const size_t bs_base = t.h_bitsets.size();
  if (pool_bytes > 0) {
    if (!pool) H2OT_FAIL(H2OT_ERR_INVALID, "null bitset pool");
    t.h_bitsets.insert(t.h_bitsets.end(), pool, pool + pool_bytes);
  }
  int min_nid = INT32_MAX, max_nid = -1;
  for (int i = 0; i < n_decided; i++) {
    const h2ot_decided_node& d = decided[i];
    if (d.nid < 0) H2OT_FAIL(H2OT_ERR_INVALID, "negative node id");
    if (d.col >= h->C) H2OT_FAIL(H2OT_ERR_INVALID, "split column out of range");
    if (d.col >= 0 && d.equal >= 2) {
      const int nbytes = ((d.bs_nbits - 1) >> 3) + 1;
      if (d.bs_off < 0 || (int64_t)d.bs_off + nbytes > pool_bytes) H2OT_FAIL(H2OT_ERR_INVALID, "bitset outside the pool");
    }
    min_nid = std::min(min_nid, d.nid);
    max_nid = std::max(max_nid, d.nid);
  }
  if ((size_t)max_nid + 1 > t.h_nodes.size()) {
    h2ot_decided_node blank;
    memset(&blank, 0xFF, sizeof(blank));  // col = pid = -1
    t.h_nodes.resize((size_t)max_nid + 1, blank);
  }
  for (int i = 0; i < n_decided; i++) {
    h2ot_decided_node d = decided[i];
    if (d.col >= 0 && d.equal >= 2) d.bs_off += (int32_t)bs_base;
    t.h_nodes[d.nid] = d;
  }
  t.n_nodes = (int)t.h_nodes.size();
  if (t.h_nodes.size() > t.nodes.cap) {
    TRY(t.nodes.ensure(std::max<size_t>(t.h_nodes.size() * 2, 256)));
    min_nid = 0;
    max_nid = (int)t.h_nodes.size() - 1;
  }
  TRY(h2d(h, t.nodes.p + min_nid, t.h_nodes.data() + min_nid, (size_t)(max_nid - min_nid + 1) * sizeof(h2ot_decided_node)));
  if (!t.h_bitsets.empty()) {
    if (t.h_bitsets.size() > t.bitsets.cap) {
      TRY(t.bitsets.ensure(std::max<size_t>(t.h_bitsets.size() * 2, 4096)));
      TRY(h2d(h, t.bitsets.p, t.h_bitsets.data(), t.h_bitsets.size()));
    } else if (pool_bytes > 0) {
      TRY(h2d(h, t.bitsets.p + bs_base, t.h_bitsets.data() + bs_base, (size_t)pool_bytes));
    }
  } else {
    TRY(t.bitsets.ensure(16));
  }
  return H2OT_OK;
}


int nccl_rc(h2ot_engine* h, int rc, const char* what) {
  if (rc == 0) return H2OT_OK;
  std::string m = std::string(what) + " failed: ";
  m += h->nccl.GetErrorString ? h->nccl.GetErrorString(rc) : "nccl error";
  H2OT_FAIL(H2OT_ERR_COMM, m);
}

// Every rank has finished all work it enqueued before this call (a one-word all-reduce on the stream, then a synchronise:
// the collective completes on this rank only once every rank's stream has reached it).
int xr_rank_barrier(h2ot_engine* h) {
  if (h->nccl.world <= 1 || !h->nccl.comm || !h->nccl.AllReduce) return H2OT_OK;
  TRY(h->ar_dev.ensure(64));
  TRY(nccl_rc(h, h->nccl.AllReduce(h->ar_dev.p, h->ar_dev.p, 1, /*ncclFloat64*/ 8, /*ncclSum*/ 0, h->nccl.comm, h->stream), "ncclAllReduce(barrier)"));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return H2OT_OK;
}

// Mapping a peer's buffer is the slow part of a sharded model's start (cudaIpcOpenMemHandle: ~80 ms for 7 peers), and a process
// that builds one model after another over the same communicator needs it once: an engine that is destroyed PARKS its exchange
// buffer (allocation, peer mappings, epochs: the flags inside the buffer live on) under its communicator + device, and the next
// engine on that communicator adopts it instead of allocating. Every rank parks and adopts at the same calls (the ranks run the
// same sequence of models), so the decision is collective without a message. Parked buffers are freed when a larger one is needed
// (collectively, xr_release) or with the process.
struct XrParked {
  unsigned long long* buf = nullptr;
  size_t cap = 0;
  XrPeers peers{};
  void* opened[kXrMaxWorld] = {nullptr};
  unsigned long long epoch = 0, ll_epoch = 0;
  int* err = nullptr;
  int* err_host = nullptr;
};
static std::map<std::string, XrParked> g_xr_parked;
static std::mutex g_xr_mutex;

bool xr_park(h2ot_engine* h) {
  if (!h->xr_buf || h->nccl.world <= 1 || h->xr_key.empty() || !h->nccl.shared) return false;
  if (h->xr_err_host && *h->xr_err_host) return false;  // an exchange timed out: the epochs of the ranks may disagree
  std::lock_guard<std::mutex> lock(g_xr_mutex);
  if (g_xr_parked.count(h->xr_key)) return false;  // (two engines alive on one communicator: the second one frees its own)
  XrParked p;
  p.buf = h->xr_buf;
  p.cap = h->xr_cap;
  p.peers = h->xr_peers;
  for (int i = 0; i < kXrMaxWorld; i++) p.opened[i] = h->xr_opened[i];
  p.epoch = h->xr_epoch;
  p.ll_epoch = h->xr_ll_epoch;
  p.err = h->xr_err;
  p.err_host = h->xr_err_host;
  g_xr_parked[h->xr_key] = p;
  h->xr_buf = nullptr;
  h->xr_cap = 0;
  for (int i = 0; i < kXrMaxWorld; i++) h->xr_opened[i] = nullptr;
  h->xr_err = nullptr;
  h->xr_err_host = nullptr;
  return true;
}
bool xr_adopt(h2ot_engine* h) {
  if (h->xr_buf || h->xr_key.empty()) return false;
  std::lock_guard<std::mutex> lock(g_xr_mutex);
  auto it = g_xr_parked.find(h->xr_key);
  if (it == g_xr_parked.end()) return false;
  const XrParked& p = it->second;
  h->xr_buf = p.buf;
  h->xr_cap = p.cap;
  h->xr_peers = p.peers;
  for (int i = 0; i < kXrMaxWorld; i++) h->xr_opened[i] = p.opened[i];
  h->xr_epoch = p.epoch;
  h->xr_ll_epoch = p.ll_epoch;
  if (h->xr_err) cudaFree(h->xr_err);
  if (h->xr_err_host) cudaFreeHost(h->xr_err_host);
  h->xr_err = p.err;
  h->xr_err_host = p.err_host;
  g_xr_parked.erase(it);
  return true;
}

// Unmaps the peers' exchange buffers and frees this rank's. A buffer that other processes have mapped must not be freed
// while a peer may still read it or still holds the mapping, so the release is collective: barrier (nobody reads any more),
// close the imported handles, barrier (nobody holds a mapping of ours any more), free.
void xr_release(h2ot_engine* h) {
  const bool shared = h->xr_buf && h->nccl.world > 1;
  if (shared) {
    if (h->stream) cudaStreamSynchronize(h->stream);
    xr_rank_barrier(h);
  }
  for (int p = 0; p < kXrMaxWorld; p++)
    if (h->xr_opened[p]) {
      cudaIpcCloseMemHandle(h->xr_opened[p]);
      h->xr_opened[p] = nullptr;
    }
  if (shared) xr_rank_barrier(h);
  if (h->xr_buf) cudaFree(h->xr_buf);
  h->xr_buf = nullptr;
  h->xr_cap = 0;
}

// Makes sure the exchange buffer holds `words` 64-bit words per half. Growing is collective (every rank plans the same
// level, so every rank grows at the same call): allocate, trade CUDA IPC handles over the NCCL communicator, map the peers.
// H2OT_ERR_UNSUPPORTED: peer mapping is not possible here; the caller falls back to the NCCL collectives for good.
int xr_ensure(h2ot_engine* h, size_t words, bool exact) {
  if (words <= h->xr_cap) return H2OT_OK;
  if (xr_adopt(h) && words <= h->xr_cap) return H2OT_OK;  // (too small: released below like a buffer of this engine's own)
  const auto t_start = std::chrono::steady_clock::now();
  const int W = h->nccl.world, R = h->nccl.rank;
  if (W > kXrMaxWorld || !h->nccl.AllGather) {
    h->xr_failed = true;
    return H2OT_ERR_UNSUPPORTED;
  }
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  xr_release(h);  // collective: every rank grows at the same call, and nobody frees a buffer a peer still reads or maps
  // [in: cap][out: cap][flags: 64][reduce-scatter lines: 2 parities x cap lines x 2 words][all-gather lines: the same]
  // (a caller that reserves the model's largest level asks for exactly that; growth on demand doubles)
  const size_t cap = std::max<size_t>(((exact ? words + words / 8 : words * 2) + 63) & ~(size_t)63, 1 << 16);
  const size_t total = (10 * cap + 64) * sizeof(unsigned long long);
  CUDA_TRY(cudaMalloc(&h->xr_buf, total));
  CUDA_TRY(cudaMemset(h->xr_buf, 0, total));
  if (!h->xr_err) {
    CUDA_TRY(cudaMalloc(&h->xr_err, sizeof(int)));
    CUDA_TRY(cudaMemset(h->xr_err, 0, sizeof(int)));
    CUDA_TRY(cudaMallocHost(&h->xr_err_host, sizeof(int)));
    *h->xr_err_host = 0;
  }
  h->xr_epoch = 0;
  h->xr_ll_epoch = 0;
  cudaIpcMemHandle_t mine;
  CUDA_TRY(cudaIpcGetMemHandle(&mine, h->xr_buf));
  DBuf<unsigned char> d_handles;
  TRY(d_handles.ensure((size_t)W * sizeof(cudaIpcMemHandle_t)));
  CUDA_TRY(cudaMemcpy(d_handles.p + (size_t)R * sizeof(mine), &mine, sizeof(mine), cudaMemcpyHostToDevice));
  TRY(nccl_rc(h, h->nccl.AllGather(d_handles.p + (size_t)R * sizeof(mine), d_handles.p, sizeof(mine), /*ncclUint8*/ 1, h->nccl.comm, h->stream),
              "ncclAllGather(ipc handles)"));
  std::vector<cudaIpcMemHandle_t> all(W);
  CUDA_TRY(cudaMemcpyAsync(all.data(), d_handles.p, (size_t)W * sizeof(mine), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  d_handles.release();
  bool ok = true;
  for (int p = 0; p < W; p++) {
    if (p == R) {
      h->xr_peers.buf[p] = h->xr_buf;
      continue;
    }
    void* ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      ok = false;
      break;
    }
    h->xr_opened[p] = ptr;
    h->xr_peers.buf[p] = static_cast<unsigned long long*>(ptr);
  }
  // all ranks must agree on the outcome
  DBuf<int> d_ok;
  TRY(d_ok.ensure(1));
  const int mine_ok = ok ? 1 : 0;
  CUDA_TRY(cudaMemcpy(d_ok.p, &mine_ok, sizeof(int), cudaMemcpyHostToDevice));
  TRY(nccl_rc(h, h->nccl.AllReduce(d_ok.p, d_ok.p, 1, /*ncclInt32*/ 2, /*ncclMin*/ 3, h->nccl.comm, h->stream), "ncclAllReduce(ipc ok)"));
  int all_ok = 0;
  CUDA_TRY(cudaMemcpyAsync(&all_ok, d_ok.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  d_ok.release();
  if (!all_ok) {
    xr_release(h);
    h->xr_failed = true;
    return H2OT_ERR_UNSUPPORTED;
  }
  h->xr_cap = cap;
  if (getenv("H2OT_TRACE_XR"))
    fprintf(stderr, "[h2ot xr] rank %d: exchange buffer %.1f MB mapped from %d peers in %.1f ms\n", R, total / 1e6, W - 1,
            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count());
  return H2OT_OK;
}

int xr_allreduce(h2ot_engine* h, long long n_sums, long long n) {
  const int W = h->nccl.world, R = h->nccl.rank;
  const size_t flag_off = 2 * h->xr_cap;
  const unsigned long long timeout = 20ULL * 1000 * 1000 * 1000;  // ~10 s of SM clocks
  const long long slice = (n + W - 1) / W;
  const int g_rs = (int)std::max<long long>(1, std::min<long long>((slice + 255) / 256, (long long)h->sm_count * 4));
  const int g_ag = (int)std::max<long long>(1, std::min<long long>((n + 255) / 256, (long long)h->sm_count * 4));
  if (h->xr_ll && W >= 2 && n >= W && (size_t)n + (size_t)W <= h->xr_cap) {  // pushed lines with the epoch inside, no barriers
    const uint32_t ep = (uint32_t)(++h->xr_ll_epoch);
    static const int per = getenv("H2OT_XR_PER") ? atoi(getenv("H2OT_XR_PER")) : 512;
    const int grid = (int)std::max<long long>(1, std::min<long long>((n + per - 1) / per, (long long)h->sm_count * 2));
    xr_ll_kernel<<<grid, 256, 0, h->stream>>>(h->xr_peers, W, R, 2 * h->xr_cap + 64, 6 * h->xr_cap + 64, 2 * h->xr_cap, n_sums, n, ep, timeout,
                                              h->xr_err);
    h->stats.launches += 1;
  } else if (h->xr_coop) {  // one cooperative launch; the grid must be co-resident: at most 2 CTAs of 256 threads per SM here
    size_t cap = h->xr_cap, fo = flag_off;
    unsigned long long ep = h->xr_epoch + 1, to = timeout;
    h->xr_epoch += 2;
    int w = W, r = R;
    int grid = (int)std::max<long long>(1, std::min<long long>((slice + 255) / 256, (long long)h->sm_count * 2));
    void* args[] = {&h->xr_peers, &w, &r, &cap, &fo, &n_sums, &n, &ep, &to, &h->xr_err};
    CUDA_TRY(cudaLaunchCooperativeKernel((const void*)xr_fused_kernel, dim3(grid), dim3(256), args, 0, h->stream));
    h->stats.launches += 1;
  } else {
    xr_barrier_kernel<<<1, 32, 0, h->stream>>>(h->xr_peers, W, R, flag_off, ++h->xr_epoch, timeout, h->xr_err);
    xr_reduce_scatter_kernel<<<g_rs, 256, 0, h->stream>>>(h->xr_peers, W, R, h->xr_cap, n_sums, n);
    xr_barrier_kernel<<<1, 32, 0, h->stream>>>(h->xr_peers, W, R, flag_off, ++h->xr_epoch, timeout, h->xr_err);
    xr_all_gather_kernel<<<g_ag, 256, 0, h->stream>>>(h->xr_peers, W, R, h->xr_cap, n);
    h->stats.launches += 4;
  }
  // (levels enqueued back to back fetch the error word once per tree, h2ot_tree_levels_begin)
  if (!h->in_tree_loop) CUDA_TRY(cudaMemcpyAsync(h->xr_err_host, h->xr_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaGetLastError());
  return H2OT_OK;
}

// split-point fields of one descriptor: placement in the pool and _zeroSplitPntPos = Arrays.binarySearch(_splitPts, 0.0)
// (T/DHistogram.java:389-401; -0.0 is already stored as 0.0)
int desc_split_points(const TreeState& t, const h2ot_hist_desc& s, DescDev& d) {
  if (s.splitpts_off < 0 || s.splitpts_n < 1 || (size_t)s.splitpts_off + (size_t)s.splitpts_n > t.h_sp.size())
    H2OT_FAIL(H2OT_ERR_INVALID, "descriptor's split points lie outside the pool given to h2ot_set_split_points");
  d.sp_off = s.splitpts_off;
  d.sp_n = s.splitpts_n;
  const double* a = t.h_sp.data() + s.splitpts_off;
  int low = 0, high = s.splitpts_n - 1;
  while (low <= high) {
    const int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
    if (a[mid] < 0.0) low = mid + 1;
    else if (a[mid] > 0.0) high = mid - 1;
    else {
      d.zero_pos = mid;
      return H2OT_OK;
    }
  }
  d.zero_pos = -(low + 1);
  return H2OT_OK;
}

// Places every active (leaf, col) histogram: global bin offsets, packed-vals offsets, shared-memory groups.
int plan_level(h2ot_engine* h, TreeState& t, int L, const h2ot_hist_desc* descs) {
  const int C = h->C;
  const int WPB = h->w ? 7 : 5;
  t.h_descs.assign((size_t)L * C, DescDev{});
  t.total_bins = 0;
  t.total_vals = 0;
  t.max_split_nbin = 0;
  t.any_cat = false;
  t.any_checkfloat = false;
  t.any_splitpts = false;
  std::vector<int64_t> col_words(C, 0);
  for (int l = 0; l < L; l++)
    for (int c = 0; c < C; c++) {
      const h2ot_hist_desc& s = descs[(size_t)l * C + c];
      DescDev& d = t.h_descs[(size_t)l * C + c];
      d.min = s.min;
      d.step = s.step;
      d.flags = s.flags;
      d.smem_off = -1;
      d.pad = 0;
      if (!(s.flags & H2OT_DESC_ACTIVE)) {
        d.nbin = -1;
        d.gbin = 0;
        d.voff = 0;
        continue;
      }
      if (s.nbin < 1 || s.nbin > 65534) H2OT_FAIL(H2OT_ERR_INVALID, "nbin out of range (1..65534)");
      if (!(s.step > 0) || std::isinf(s.step)) H2OT_FAIL(H2OT_ERR_INVALID, "histogram step must be positive and finite");
      d.nbin = s.nbin;
      if (s.flags & H2OT_DESC_SPLITPTS) {
        TRY(desc_split_points(t, s, d));
        t.any_splitpts = true;
      }
      d.gbin = t.total_bins;
      d.voff = t.total_vals;
      t.total_bins += s.nbin + 1;
      t.total_vals += (int64_t)(s.nbin + 1) * 3;
      col_words[c] += (int64_t)(s.nbin + 1) * WPB;
      t.max_split_nbin = std::max(t.max_split_nbin, (int)s.nbin);
      if ((s.flags & H2OT_DESC_ISCAT) && s.step == 1.0) t.any_cat = true;
      if (s.flags & H2OT_DESC_CHECKFLOAT) t.any_checkfloat = true;
    }
  // ---- shared-memory groups: columns of one storage type per group (the kernel is specialised on it)
  auto fixed_bytes_l = [&](int ncg, int storage, int nleaves) {
    return (size_t)ncg * nleaves * (kTStride * sizeof(uint32_t) + (storage == H2OT_STORE_F64 ? sizeof(MinMaxS) : 0)) +
           (size_t)ncg * (sizeof(ColRef) + 3 * sizeof(int32_t)) + 16 + 32 * 7 * sizeof(uint32_t);  // + the per-lane sink bins
  };
  auto fixed_bytes = [&](int ncg, int storage) { return fixed_bytes_l(ncg, storage, L); };
  const size_t full_budget = (size_t)h->max_smem_optin;
  // the kernel runs one 1024-thread CTA per SM (64 registers/thread), so a CTA may use the whole opt-in carve-out
  size_t budget = full_budget;
  if (h->smem_budget_kb > 0) budget = std::min<size_t>((size_t)h->smem_budget_kb * 1024, full_budget);
  // ---- sparse level (per-node mtries / col_sample_rate): most (leaf, column) pairs have no histogram. Leaf-range groups for
  // hist_sparse_kernel: a CTA owns consecutive leaves whose histograms + records fit shared memory and gathers, per row, only
  // that row's leaf's columns (the column-major kernel would test rows x all columns, and re-scan per leaf range).
  t.n_sparse_groups = 0;
  t.sparse_ctas = 0;
  if ((h->hist_mode == 0 && L >= 2) || (t.any_splitpts && h->hist_mode != 1)) {
    int64_t pairs = 0, pair_bytes = 0;
    int cols_act = 0;
    bool any64 = false;
    for (int c = 0; c < C; c++) cols_act += col_words[c] ? 1 : 0;
    std::vector<int> m_leaf(L, 0);
    std::vector<size_t> w_leaf(L, 0);
    for (int l = 0; l < L; l++)
      for (int c = 0; c < C; c++) {
        const DescDev& d = t.h_descs[(size_t)l * C + c];
        if (d.nbin < 0) continue;
        pairs++;
        m_leaf[l]++;
        w_leaf[l] += (size_t)(d.nbin + 1) * WPB;
        const int st = h->cols[c].storage;
        pair_bytes += st == H2OT_STORE_F32 ? 4 : st == H2OT_STORE_F64 ? 8 : st == H2OT_STORE_U8 ? 1 : 2;
        any64 |= st == H2OT_STORE_F64;
      }
    // (levels whose descriptors carry split points always take this kernel: it is the one that bins by binary search)
    if (pairs > 0 && (t.any_splitpts || pairs * 5 <= (int64_t)L * cols_act * 4)) {  // at most 80 % of the (leaf, column) pairs are active
      const size_t col_bytes = C <= kSparseColsSmem ? (size_t)C * sizeof(ColRef) : 0;
      auto fixed_sparse = [&](int leaves, int M) {
        return (size_t)leaves * M * (kSStride * sizeof(uint32_t) + (any64 ? sizeof(MinMaxS) : 0)) + col_bytes + kSparseQueueWords * sizeof(uint32_t) + 64;
      };
      const size_t sbudget = (size_t)h->max_smem_optin;
      std::vector<GroupDev> sg;
      bool ok = true;
      size_t smem_max = 0;
      for (int l0 = 0; l0 < L && ok;) {
        size_t words = 0;
        int M = 1, l1 = l0;
        while (l1 < L) {
          const int Mn = std::max(M, m_leaf[l1]);
          if ((words + w_leaf[l1]) * 4 + fixed_sparse(l1 + 1 - l0, Mn) > sbudget || words + w_leaf[l1] >= 0xFFFFu) break;
          words += w_leaf[l1];
          M = Mn;
          l1++;
        }
        if (l1 == l0) { ok = false; break; }  // one leaf alone does not fit: leave the level to the column-major planner
        sg.push_back(GroupDev{0, M, (int32_t)words, 2, any64 ? 1 : 0, 1, 0, l0, l1 - l0, 0});
        smem_max = std::max(smem_max, words * 4 + fixed_sparse(l1 - l0, M));
        l0 = l1;
      }
      if (ok) {
        const int G = (int)sg.size();
        const int n_ctas = std::max(h->sm_count, G);
        int cta = 0;
        // CTA shares: every group scans all leaf codes (1 unit per row) and works on its own rows (~6 units per row and
        // column); its rows are estimated from the parents' split counts, evenly when those are unknown
        std::vector<double> gcost((size_t)G, 0.0);
        {
          double tot = 0;
          for (double v : t.est_rows) tot += v;
          double cost_sum = 0;
          for (int gi = 0; gi < G; gi++) {
            double frac = (double)sg[gi].l_count / L;
            if (tot > 0) {
              double r = 0;
              for (int l = sg[gi].l_begin; l < sg[gi].l_begin + sg[gi].l_count; l++) r += t.est_rows[(size_t)l];
              frac = r / tot;
            }
            gcost[(size_t)gi] = 1.0 + 6.0 * sg[gi].ncols * frac;
            cost_sum += gcost[(size_t)gi];
          }
          int used = 0;
          for (int gi = 0; gi < G; gi++) {
            sg[gi].n_slices = std::max(1, (int)std::floor(n_ctas * gcost[(size_t)gi] / cost_sum));
            used += sg[gi].n_slices;
          }
          for (; used < n_ctas; used++) {
            int best = 0;
            for (int gi = 1; gi < G; gi++)
              if (gcost[(size_t)gi] / sg[gi].n_slices > gcost[(size_t)best] / sg[best].n_slices) best = gi;
            sg[best].n_slices++;
          }
        }
        for (int gi = 0; gi < G; gi++) {
          sg[gi].cta_begin = cta;
          cta += sg[gi].n_slices;
          int off = 0;
          for (int l = sg[gi].l_begin; l < sg[gi].l_begin + sg[gi].l_count; l++)
            for (int c = 0; c < C; c++) {
              DescDev& d = t.h_descs[(size_t)l * C + c];
              if (d.nbin < 0) continue;
              d.smem_off = off;
              off += (d.nbin + 1) * WPB;
            }
        }
        t.h_groups = sg;
        t.h_group_cols.clear();
        t.active_cols = cols_act;
        t.n_sparse_groups = G;
        t.sparse_ctas = cta;
        t.n_smem_groups = G;
        t.smem_bytes = smem_max;
        t.sparse_bytes_per_row = (double)pair_bytes / L;
        t.sparse_cols_per_row = (double)pairs / L;
        for (int st = 0; st < 4; st++) t.bucket_begin[st] = t.bucket_end[st] = t.bucket_ctas[st] = 0;
        return H2OT_OK;
      }
    }
  }
  // Columns whose rows pile up in a few bins (c / L large) serialise their atomics on those addresses. With unit weights the
  // worst of them (eight or more of a warp's 32 rows expected in one bin: binary and few-valued columns on the first levels) are
  // aggregated inside the warp before shared memory is touched -- on frames of 2^17 rows or more, whose fixed-point addends
  // stay below 2^44 (the aggregation sums them in 24-bit limbs); the others get 2 to 32 replicas of their histograms (lanes
  // pick one by lane id; the replicas are summed when the CTA merges) where that fits.
  std::vector<int> col_rep(C, 1), col_agg(C, 0);
  const bool agg_ok = !h->w && !h->no_agg && !t.any_checkfloat && ceil_log2_u64((uint64_t)h->cfg.nrows_global + 1) >= 18;
  for (int c = 0; c < C; c++) {
    if (!col_words[c] || h->hist_mode == 1 || t.any_splitpts) continue;
    static const double agg_min = getenv("H2OT_AGG_MIN") ? atof(getenv("H2OT_AGG_MIN")) : 8.0;  // expected rows of a warp on one bin
    if (agg_ok && h->col_conc[c] * 32.0 / (double)L >= agg_min) {
      col_agg[c] = 1;
      continue;
    }
    const double excess = 8.0 * h->col_conc[c] / (double)L;
    int rep = 1;
    while (rep < 32 && excess / rep >= 0.4) rep <<= 1;
    // as many copies as one CTA can hold for the column alone (a few-valued column then gets a group of its own: measured on the
    // Higgs shape a 3-valued column costs 3.5x a continuous one without copies, 2.4x with 4)
    while (rep > 1 && (size_t)col_words[c] * rep * 4 + fixed_bytes(1, h->cols[c].storage) > budget - 4096) rep >>= 1;
    col_rep[c] = rep;
    col_words[c] *= rep;
  }
  t.h_groups.clear();
  t.h_group_cols.clear();
  t.active_cols = 0;
  std::vector<int> glob_cols;
  std::vector<int> bucket[4], split_cols[4];
  for (int c = 0; c < C; c++) {
    if (!col_words[c]) continue;
    t.active_cols++;
    const int st = h->cols[c].storage;
    if (h->hist_mode == 1 || t.any_splitpts) glob_cols.push_back(c);  // (split points: the sparse plan above did not fit)
    else if ((size_t)col_words[c] * 4 + fixed_bytes(1, st) > budget) {
      // too big for one CTA. Dense column (most leaves histogram it): cut into leaf ranges, each range's CTAs scan all rows
      // (measured 2.5 / 5.2 ms at 64 / 128 leaves x 257 bins vs 15 / 11 ms with L2 atomics). Sparse column (per-node mtries):
      // few rows take part at all, where the L2-atomics kernel that simply skips them is faster (DRF depth 12: 11.9 vs 14.5 ms).
      int act = 0;
      for (int l = 0; l < L; l++) act += t.h_descs[(size_t)l * C + c].nbin >= 0 ? 1 : 0;
      if (2 * act < L) glob_cols.push_back(c);
      else split_cols[st].push_back(c);
    } else bucket[st].push_back(c);
  }
  t.smem_bytes = 0;
  for (int st = 0; st < 4; st++) {
    t.bucket_begin[st] = (int)t.h_groups.size();
    const std::vector<int>& cols = bucket[st];
    if (cols.empty() && split_cols[st].empty()) {
      t.bucket_end[st] = t.bucket_begin[st];
      t.bucket_ctas[st] = 0;
      continue;
    }
    std::vector<std::vector<int>> groups;
    std::vector<double> gcost;
    std::vector<std::pair<int, int>> granges;  // leaf range [begin, count) of each group
    auto cost_of = [&](int c) { return col_agg[c] ? 1.0 : 1.0 + 8.0 * h->col_conc[c] / ((double)L * col_rep[c]); };
    if (!cols.empty()) {
    // Cost model: a column's atomics serialise when rows of a warp share a bin; with c = chance that two rows of the
    // root share a bin (measured by the last root level) and L leaves, cost ~ 1 + 8 c / L (fits the measured 3.6 vs 14
    // wavefronts per atomic of continuous vs 3-valued columns). Groups: fewest that fit, filled longest-first onto the
    // cheapest group (LPT); CTAs: shared out in proportion to the group costs so that all CTAs of the launch end together.
    std::vector<int> order(cols);
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return cost_of(x) > cost_of(y); });
    size_t total_words = 0;
    for (int c : cols) total_words += (size_t)col_words[c];
    // first guess of the group count from the bytes a column really takes (bins + its records); a min-heap on the group cost
    // makes a placement O(log groups) (wide frames: 2000 columns spread over ~200 groups), and a failed attempt grows the
    // count geometrically
    const size_t per_col_fixed = fixed_bytes(2, st) - fixed_bytes(1, st);
    size_t ng0 = (total_words * 4 + cols.size() * per_col_fixed + (budget - fixed_bytes(0, st)) - 1) / (budget - fixed_bytes(0, st));
    for (size_t ng = std::max<size_t>(1, ng0);; ng += std::max<size_t>(1, ng / 16)) {
      groups.assign(ng, {});
      gcost.assign(ng, 0.0);
      std::vector<size_t> gwords(ng, 0);
      typedef std::pair<double, int> Ent;
      std::priority_queue<Ent, std::vector<Ent>, std::greater<Ent>> heap;
      for (size_t g = 0; g < ng; g++) heap.push(Ent(0.0, (int)g));
      bool ok = true;
      auto fits = [&](size_t g, int c) { return (gwords[g] + (size_t)col_words[c]) * 4 + fixed_bytes((int)groups[g].size() + 1, st) <= budget; };
      for (int c : order) {
        int best = -1;
        if (!heap.empty() && fits((size_t)heap.top().second, c)) {
          best = heap.top().second;
          heap.pop();
        } else {  // the cheapest group is full for this column: look at all of them
          for (size_t g = 0; g < ng; g++) {
            if (!fits(g, c)) continue;
            if (best < 0 || gcost[g] < gcost[best]) best = (int)g;
          }
          if (best >= 0) {  // take it out of the heap (rebuild: rare)
            std::vector<Ent> keep;
            while (!heap.empty()) {
              if (heap.top().second != best) keep.push_back(heap.top());
              heap.pop();
            }
            for (const Ent& e : keep) heap.push(e);
          }
        }
        if (best < 0) { ok = false; break; }
        groups[best].push_back(c);
        gwords[best] += (size_t)col_words[c];
        gcost[best] += cost_of(c);
        heap.push(Ent(gcost[best], best));
      }
      if (ok) break;
      if (ng > cols.size()) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "cannot place the level's histograms in shared memory");
    }
    granges.assign(groups.size(), std::make_pair(0, L));
    }  // !cols.empty()
    // a column whose leaves x bins exceed shared memory: one group per leaf range that fits (rows of other ranges are skipped)
    for (int c : split_cols[st]) {
      size_t widest = 0;
      for (int l = 0; l < L; l++) widest = std::max(widest, (size_t)std::max(t.h_descs[(size_t)l * C + c].nbin + 1, 0) * WPB);
      if (widest * 4 + fixed_bytes_l(1, st, 1) > full_budget) {  // a single (leaf, column) histogram is too wide: L2 atomics
        glob_cols.push_back(c);
        continue;
      }
      int l0 = 0;
      while (l0 < L) {
        size_t words = 0;
        int l1 = l0;
        while (l1 < L) {
          const DescDev& d = t.h_descs[(size_t)l1 * C + c];
          const size_t nw = words + (d.nbin >= 0 ? (size_t)(d.nbin + 1) * WPB : 0);
          if (l1 > l0 && nw * 4 + fixed_bytes_l(1, st, l1 + 1 - l0) > budget) break;
          words = nw;
          l1++;
        }
        if (words * 4 + fixed_bytes_l(1, st, l1 - l0) > full_budget)
          H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "one (leaf, column) histogram exceeds shared memory");
        if (words > 0) {
          groups.push_back({c});
          gcost.push_back(cost_of(c) * (double)(l1 - l0) / (double)L);
          granges.push_back(std::make_pair(l0, l1 - l0));
        }
        l0 = l1;
      }
    }
    double cost_sum = 0;
    for (double v : gcost) cost_sum += v;
    int cta = 0;
    const int n_ctas = std::max<int>(h->sm_count, (int)groups.size());
    std::vector<int> slices(groups.size(), 1);
    {
      int used = 0;
      for (size_t g = 0; g < groups.size(); g++) {
        slices[g] = std::max(1, (int)std::floor(n_ctas * gcost[g] / cost_sum));
        used += slices[g];
      }
      // hand the remaining CTAs to the groups with the highest cost per CTA
      for (; used < n_ctas; used++) {
        size_t best = 0;
        for (size_t g = 1; g < groups.size(); g++)
          if (gcost[g] / slices[g] > gcost[best] / slices[best]) best = g;
        slices[best]++;
      }
    }
    for (size_t gi = 0; gi < groups.size(); gi++) {
      std::vector<int>& g = groups[gi];
      std::sort(g.begin(), g.end());
      GroupDev gd{(int32_t)t.h_group_cols.size(), (int32_t)g.size(), 0, 0, st, slices[gi], cta, granges[gi].first, granges[gi].second, 0};
      cta += slices[gi];
      int off = 0;
      for (int c : g) {
        int lg = 0;
        while ((1 << lg) < col_rep[c]) lg++;
        t.h_group_cols.push_back(c | (lg << 24) | (col_agg[c] << 29));
        for (int l = granges[gi].first; l < granges[gi].first + granges[gi].second; l++) {
          DescDev& d = t.h_descs[(size_t)l * C + c];
          if (d.nbin < 0) continue;
          d.smem_off = off;
          off += (d.nbin + 1) * WPB * col_rep[c];
        }
      }
      gd.smem_words = off;
      t.smem_bytes = std::max(t.smem_bytes, fixed_bytes_l((int)g.size(), st, granges[gi].second) + (size_t)off * 4);
      t.h_groups.push_back(gd);
    }
    t.bucket_ctas[st] = cta;
    t.bucket_end[st] = (int)t.h_groups.size();
  }
  t.n_smem_groups = (int)t.h_groups.size();
  for (int c : glob_cols) {
    t.h_groups.push_back(GroupDev{(int32_t)t.h_group_cols.size(), 1, 0, 1, h->cols[c].storage, 1, 0, 0, L, 0});
    t.h_group_cols.push_back(c);
  }
  return H2OT_OK;
}

}  // namespace

extern "C" {

int h2ot_set_split_points(h2ot_engine* h, int32_t k, const double* points, int64_t n) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (n < 0 || (n > 0 && !points) || n > (int64_t)1 << 30) H2OT_FAIL(H2OT_ERR_INVALID, "bad split point pool");
  TreeState& t = h->ts[k];
  t.h_sp.assign(points, points + n);
  for (double& v : t.h_sp)
    if (v == 0.0) v = 0.0;  // -0.0 -> 0.0 (T/DHistogram.java:389-401)
  t.sp_dirty = true;
  return H2OT_OK;
}

int h2ot_build_level(h2ot_engine* h, int32_t k, int32_t leaf_offset, int32_t n_leaves, int32_t n_decided,
                     const h2ot_decided_node* decided, const uint8_t* bitset_pool, int64_t bitset_bytes,
                     const h2ot_hist_desc* descs) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  if (n_leaves <= 0 || n_leaves > 65534) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "n_leaves must be in 1..65534");
  if (!descs) H2OT_FAIL(H2OT_ERR_INVALID, "null descs");
  TreeState& t = h->ts[k];
  if (t.level_failed)
    H2OT_FAIL(H2OT_ERR_INVALID, "an earlier level of this tree failed after the rows had been moved to its leaves: restart the tree with h2ot_begin_tree");
  t.built = t.finished = false;
  TRY(upload_nodes(h, t, n_decided, decided, bitset_pool, bitset_bytes, false));
  if (leaf_offset > 0 && t.n_nodes == 0) H2OT_FAIL(H2OT_ERR_INVALID, "no decided nodes uploaded for a non-root level");
  TRY(t.nodes.ensure(1));
  TRY(t.bitsets.ensure(16));
  t.L = n_leaves;
  t.leaf_offset = leaf_offset;
  const size_t LC = (size_t)n_leaves * h->C;
  // Payload of the level, laid out [rows per leaf : L][w, wY, wYY per bin : 3 x bins][min/max keys : 2 x LC]. Its size needs
  // only the bin counts, so it is cleared and score_decide is launched before the (host-side) planning of the histogram pass.
  {
    int64_t bins = 0;
    for (size_t i = 0; i < LC; i++) {
      if (!(descs[i].flags & H2OT_DESC_ACTIVE)) continue;
      // validated here, before any kernel touches the row state
      if (descs[i].nbin < 1 || descs[i].nbin > 65534) H2OT_FAIL(H2OT_ERR_INVALID, "nbin out of range (1..65534)");
      if (!(descs[i].step > 0) || std::isinf(descs[i].step)) H2OT_FAIL(H2OT_ERR_INVALID, "histogram step must be positive and finite");
      bins += descs[i].nbin + 1;
    }
    const size_t ns = (size_t)bins * 3 + n_leaves, nm = LC * 2;
    t.in_xbuf = false;
    if (h->xr_enabled && !h->xr_failed && h->nccl.world > 1) {
      // every class-tree's level is accumulated in the ONE exchange buffer the peers have mapped: a second class-tree may build
      // only after the first one's h2ot_finish_level has copied its level out (the level_step calls do that by themselves)
      if (h->xr_owner >= 0 && h->xr_owner != k)
        H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_build_level: another class-tree's level is still in the exchange buffer; call h2ot_finish_level for it first");
      const int rc = xr_ensure(h, ns + nm);
      if (rc == H2OT_OK) {
        t.sums = reinterpret_cast<long long*>(h->xr_buf);
        t.mm = h->xr_buf + ns;
        t.in_xbuf = true;
        h->xr_owner = k;
      } else if (rc != H2OT_ERR_UNSUPPORTED) {
        return rc;
      }
    }
    if (!t.in_xbuf) {
      TRY(t.d_sums.ensure(ns + nm));
      t.sums = t.d_sums.p;
      t.mm = reinterpret_cast<unsigned long long*>(t.d_sums.p + ns);
    }
    t.payload_words = ns + nm;
  }
  // K2a: clear the payload and move the rows to this level's leaves. Launched here, ahead of the (host-side) planning it overlaps
  // with -- unless the device loop drives the level: then the step kernel that creates this level's nodes has to run first, and
  // it needs the planned layout (below).
  auto launch_score = [&]() -> int {
    CUDA_TRY(cudaMemsetAsync(t.sums, 0, t.payload_words * sizeof(long long), h->stream));
    ScoreArgs sa{h->n, h->n_pad, leaf_offset, n_leaves, t.nids, t.code, t.nodes.p, t.bitsets.p, h->d_cols.p,
                 reinterpret_cast<unsigned long long*>(t.sums)};
    TRY(t_begin(h, &h->stats.ms_score));
    score_decide_kernel<<<grid_for(h, h->n_pad / 4, 256, 16), 256, 0, h->stream>>>(sa);
    h->stats.launches++;
    TRY(t_end(h));
    return H2OT_OK;
  };
  const bool dev_step = t.dev.active && t.dev.depth > 0;
  if (!dev_step) TRY(launch_score());
  t.est_rows.assign((size_t)n_leaves, 0.0);
  if (t.prev_leaf_offset >= 0 && decided)
    for (int i = 0; i < n_decided; i++) {
      const h2ot_decided_node& dn = decided[i];
      const int pl = dn.nid - t.prev_leaf_offset;
      if (dn.col < 0 || pl < 0 || pl >= (int)t.prev_n0.size()) continue;
      for (int w = 0; w < 2; w++) {
        const int ch = dn.child[w] - leaf_offset;
        if (ch >= 0 && ch < n_leaves) t.est_rows[(size_t)ch] = w == 0 ? t.prev_n0[(size_t)pl] : t.prev_n1[(size_t)pl];
      }
    }
  t.prev_leaf_offset = -1;  // consumed; set again by the level_step that ends this level
  // (score_decide is already in flight -- it overlaps this host-side planning -- so a plan that cannot be made leaves NIDs_k one
  // level ahead of the caller's tree: the tree is marked and has to be restarted with h2ot_begin_tree + h2ot_set_nids / a new build)
  {
    const int prc = plan_level(h, t, n_leaves, descs);
    if (prc != H2OT_OK) {
      t.level_failed = true;
      if (h->xr_owner == k) h->xr_owner = -1;
      return prc;
    }
  }
  // this level's descriptors, groups and column lists travel in ONE copy from pinned staging
  {
    const size_t b_descs = LC * sizeof(DescDev), b_groups = std::max<size_t>(1, t.h_groups.size()) * sizeof(GroupDev);
    const size_t b_cols = std::max<size_t>(1, t.h_group_cols.size()) * sizeof(int32_t);
    const size_t total = b_descs + b_groups + b_cols;
    TRY(t.d_lvl.ensure(total));
    unsigned char* stage = nullptr;
    if (t.dev.active) {
      // the device loop enqueues its levels ahead of the GPU: every level keeps its own slice of one pinned arena until the tree
      // is done (h2ot_tree_levels_begin sized it for all levels; a level that needs more than its slice gets a block of its own)
      const size_t d = (size_t)t.dev.depth;
      if (t.dev.lvl_stage.size() <= d) {
        t.dev.lvl_stage.resize(d + 1, nullptr);
        t.dev.lvl_stage_cap.resize(d + 1, 0);
        t.dev.lvl_stage_own.resize(d + 1, 0);
      }
      // (a block of its own is shared by the two trees that may be in flight: drain the stream before touching it)
      if (total > t.dev.lvl_stage_cap[d] || t.dev.lvl_stage_own[d]) CUDA_TRY(cudaStreamSynchronize(h->stream));
      if (total > t.dev.lvl_stage_cap[d]) {
        if (t.dev.lvl_stage[d] && t.dev.lvl_stage_own[d]) cudaFreeHost(t.dev.lvl_stage[d]);
        t.dev.lvl_stage[d] = nullptr;
        t.dev.lvl_stage_cap[d] = 0;
        CUDA_TRY(cudaMallocHost(&t.dev.lvl_stage[d], total * 2));
        t.dev.lvl_stage_cap[d] = total * 2;
        t.dev.lvl_stage_own[d] = 1;
      }
      stage = t.dev.lvl_stage[d];
    } else {
      if (total > t.lvl_host_cap) {
        if (t.lvl_host) cudaFreeHost(t.lvl_host);
        t.lvl_host = nullptr;
        t.lvl_host_cap = 0;
        CUDA_TRY(cudaMallocHost(&t.lvl_host, total * 2));
        t.lvl_host_cap = total * 2;
      }
      stage = t.lvl_host;
    }
    for (size_t i = 0; i < LC; i++) t.h_descs[i].pad = (uint32_t)(uint16_t)descs[i].order;  // bestCol tie order
    memcpy(stage, t.h_descs.data(), b_descs);
    if (!t.h_groups.empty()) memcpy(stage + b_descs, t.h_groups.data(), t.h_groups.size() * sizeof(GroupDev));
    if (!t.h_group_cols.empty()) memcpy(stage + b_descs + b_groups, t.h_group_cols.data(), t.h_group_cols.size() * sizeof(int32_t));
    t.p_descs = reinterpret_cast<DescDev*>(t.d_lvl.p);
    t.p_groups = reinterpret_cast<GroupDev*>(t.d_lvl.p + b_descs);
    t.p_group_cols = reinterpret_cast<int32_t*>(t.d_lvl.p + b_descs + b_groups);
    TRY(h2d(h, t.d_lvl.p, stage, total));
  }
  if (dev_step) {
    // the step from the previous level: its splits become this level's nodes and descriptors (the layout fields of the
    // descriptors just uploaded stay, everything else is written by the kernel), then the rows can move
    TreeState::DevLoop& dv = t.dev;
    NextArgs na{};
    na.C = h->C;
    na.Lp = n_leaves / 2;
    na.parent_base = na.Lp - 1;
    na.last_level = 0;
    na.nbins = dv.nbins;
    na.nbins_cats = dv.nbins_cats;
    na.min_rows = dv.min_rows;
    na.pdescs = dv.prev_descs;
    na.pmaxex = dv.prev_maxex;
    na.best = dv.prev_best;
    na.min2 = dv.prev_min2;
    na.maxIn = dv.prev_maxIn;
    na.wna = dv.prev_wna;
    na.has_vals = dv.prev_has;
    na.ndescs = t.p_descs;
    na.nmaxex = dv.maxex[dv.depth & 1].p;
    na.nb_cap = dv.nb_caps.p + (size_t)dv.depth * h->C;
    na.nodes = t.nodes.p;
    na.bitsets = t.bitsets.p;
    na.exists = dv.exists.p;
    na.tree_splits = dv.tree_splits.p;
    na.err = dv.err.p;
    if (t.best_pending) {
      na.cand = t.d_cand.p;
      na.best_out = const_cast<h2ot_split*>(dv.prev_best);
      t.best_pending = false;
    }
    if (na.Lp == 1) {
      na.conc_src = dv.conc_src;
      na.conc_dst = dv.p_conc_out;
    }
    next_level_kernel<<<na.Lp, 128, 0, h->stream>>>(na);
    h->stats.launches++;
    TRY(launch_score());
  }
  // K2b
  HistArgs ha{};
  ha.n_pad = h->n_pad;
  ha.L = n_leaves;
  ha.C = h->C;
  ha.group_cols = t.p_group_cols;
  const int64_t n_tiles = h->n_pad / kTileRows;
  ha.code = t.code;
  ha.work = t.work;
  ha.w = h->w;
  ha.cols = h->d_cols.p;
  ha.descs = t.p_descs;
  ha.groups = t.p_groups;
  ha.acc = t.sums + n_leaves;
  ha.minmax = t.mm;
  ha.sy = std::ldexp(1.0, t.ey);
  ha.syy = std::ldexp(1.0, t.eyy);
  const int bitsN = ceil_log2_u64((uint64_t)h->cfg.nrows_global + 1);
  const int ew = h->w ? exponent_for(h->w_max, bitsN) : 0;
  ha.sw = std::ldexp(1.0, ew);
  ha.sp = nullptr;
  if (t.any_splitpts) {
    if (t.sp_dirty) {
      TRY(t.d_sp.ensure(std::max<size_t>(t.h_sp.size(), 1)));
      TRY(h2d(h, t.d_sp.p, t.h_sp.data(), t.h_sp.size() * sizeof(double)));
      t.sp_dirty = false;
    }
    ha.sp = t.d_sp.p;
  }
  TRY(t_begin(h, &h->stats.ms_hist));
  if (t.n_smem_groups > 0 && t.smem_bytes > (size_t)h->max_smem_optin) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "shared-memory plan exceeds the device limit");
  if (t.n_sparse_groups > 0) {
    ha.group_begin = 0;
    ha.n_groups = t.n_sparse_groups;
    ha.n_slices = 0;
    const int grid = t.sparse_ctas;
    const size_t sm = t.smem_bytes;
#define H2OT_LAUNCH_SPARSE(SP)                                                                                    \
    if (h->w) {                                                                                                    \
      if (t.any_checkfloat) hist_sparse_kernel<7, true, true, SP><<<grid, kHistThreads, sm, h->stream>>>(ha);      \
      else hist_sparse_kernel<7, true, false, SP><<<grid, kHistThreads, sm, h->stream>>>(ha);                      \
    } else {                                                                                                       \
      if (t.any_checkfloat) hist_sparse_kernel<5, false, true, SP><<<grid, kHistThreads, sm, h->stream>>>(ha);     \
      else hist_sparse_kernel<5, false, false, SP><<<grid, kHistThreads, sm, h->stream>>>(ha);                     \
    }
    if (t.any_splitpts) { H2OT_LAUNCH_SPARSE(true) } else { H2OT_LAUNCH_SPARSE(false) }
#undef H2OT_LAUNCH_SPARSE
    h->stats.launches++;
    h->stats.hist_launches++;
    h->stats.sparse_launches++;
  }
  for (int st = 0; st < 4; st++) {
    const int ng = t.bucket_end[st] - t.bucket_begin[st];
    if (ng <= 0) continue;
    ha.group_begin = t.bucket_begin[st];
    ha.n_groups = ng;
    ha.n_slices = 0;  // per group: GroupDev::n_slices
    const int grid = t.bucket_ctas[st];
    (void)n_tiles;
    const size_t sm = t.smem_bytes;
#define H2OT_LAUNCH_HIST(ST)                                                                                      \
  if (h->w) {                                                                                                     \
    if (t.any_checkfloat) hist_smem_kernel<7, true, ST, true, false><<<grid, kHistThreads, sm, h->stream>>>(ha);     \
    else if (n_leaves == 1) hist_smem_kernel<7, true, ST, false, true><<<grid, kHistThreads, sm, h->stream>>>(ha);   \
    else hist_smem_kernel<7, true, ST, false, false><<<grid, kHistThreads, sm, h->stream>>>(ha);                     \
  } else {                                                                                                        \
    if (t.any_checkfloat) hist_smem_kernel<5, false, ST, true, false><<<grid, kHistThreads, sm, h->stream>>>(ha);    \
    else if (n_leaves == 1) hist_smem_kernel<5, false, ST, false, true><<<grid, kHistThreads, sm, h->stream>>>(ha);  \
    else hist_smem_kernel<5, false, ST, false, false><<<grid, kHistThreads, sm, h->stream>>>(ha);                    \
  }
    switch (st) {
      case H2OT_STORE_F32: H2OT_LAUNCH_HIST(H2OT_STORE_F32); break;
      case H2OT_STORE_F64: H2OT_LAUNCH_HIST(H2OT_STORE_F64); break;
      case H2OT_STORE_U8: H2OT_LAUNCH_HIST(H2OT_STORE_U8); break;
      default: H2OT_LAUNCH_HIST(H2OT_STORE_U16); break;
    }
#undef H2OT_LAUNCH_HIST
    h->stats.launches++;
    h->stats.hist_launches++;
  }
  for (size_t g = t.n_smem_groups; g < t.h_groups.size(); g++) {
    const int c = t.h_group_cols[t.h_groups[g].cols_off] & 0xFFFFFF;
    const int grid = grid_for(h, h->n_pad / 4, 256, 8);
    if (h->w) hist_global_kernel<true><<<grid, 256, 0, h->stream>>>(ha, c, c + 1);
    else hist_global_kernel<false><<<grid, 256, 0, h->stream>>>(ha, c, c + 1);
    h->stats.launches++;
    h->stats.hist_launches++;
  }
  TRY(t_end(h));
  CUDA_TRY(cudaGetLastError());
  // accounting: N*(sum B_c + nid r [+ nid w] + y [+ w]) + bins*24
  int64_t bytes_per_row = 4 + 4 + (leaf_offset > 0 ? 4 : 0) + (h->w ? 4 : 0);
  for (int c = 0; c < h->C; c++) {
    bool act = false;
    for (int l = 0; l < n_leaves && !act; l++) act = t.h_descs[(size_t)l * h->C + c].nbin >= 0;
    if (!act) continue;
    const int st = h->cols[c].storage;
    bytes_per_row += st == H2OT_STORE_F32 ? 4 : st == H2OT_STORE_F64 ? 8 : st == H2OT_STORE_U8 ? 1 : 2;
  }
  if (t.n_sparse_groups > 0) {  // a row needs only its own leaf's columns: average over the leaves
    bytes_per_row = 4 + 4 + (leaf_offset > 0 ? 4 : 0) + (h->w ? 4 : 0) + (int64_t)std::llround(t.sparse_bytes_per_row);
    h->stats.hist_rows_cols += (int64_t)std::llround((double)h->n * t.sparse_cols_per_row);
  } else {
    h->stats.hist_rows_cols += h->n * (int64_t)t.active_cols;
  }
  h->stats.hist_algo_bytes += h->n * bytes_per_row + t.total_bins * 24;
  t.built = true;
  return finish_call(h);
}

int h2ot_level_buffers(h2ot_engine* h, int32_t k, void** d_sums, int64_t* n_sums, void** d_minmax, int64_t* n_minmax) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  if (!t.built) H2OT_FAIL(H2OT_ERR_INVALID, "no level built");
  if (d_sums) *d_sums = t.sums;
  if (n_sums) *n_sums = t.total_bins * 3 + t.L;
  if (d_minmax) *d_minmax = t.mm;
  if (n_minmax) *n_minmax = (int64_t)t.L * h->C * 2;
  return H2OT_OK;
}

static int nccl_check(h2ot_engine* h, int rc, const char* what) {
  if (rc == 0) return H2OT_OK;
  std::string m = std::string(what) + " failed: ";
  m += h->nccl.GetErrorString ? h->nccl.GetErrorString(rc) : "nccl error";
  H2OT_FAIL(H2OT_ERR_COMM, m);
}

int h2ot_allreduce_level(h2ot_engine* h, int32_t k) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  if (!t.built) H2OT_FAIL(H2OT_ERR_INVALID, "no level built");
  if (h->nccl.world <= 1 || !h->nccl.comm) return H2OT_OK;
  const size_t ns = (size_t)t.total_bins * 3 + t.L, nm = (size_t)t.L * h->C * 2;
  TRY(t_begin(h, &h->stats.ms_exchange));
  if (t.in_xbuf) {
    TRY(xr_allreduce(h, (long long)ns, (long long)(ns + nm)));
    TRY(t_end(h));
    return finish_call(h);
  }
  TRY(nccl_check(h, h->nccl.AllReduce(t.sums, t.sums, ns, /*ncclInt64*/ 4, /*ncclSum*/ 0, h->nccl.comm, h->stream), "ncclAllReduce(sums)"));
  TRY(nccl_check(h, h->nccl.AllReduce(t.mm, t.mm, nm, /*ncclUint64*/ 5, /*ncclMax*/ 2, h->nccl.comm, h->stream), "ncclAllReduce(minmax)"));
  TRY(t_end(h));
  return finish_call(h);
}

int h2ot_finish_level(h2ot_engine* h, int32_t k, double* out_vals, int64_t out_vals_capacity, int64_t* out_val_offsets,
                      double* out_min2, double* out_maxIn, double* out_wNA, uint8_t* out_has_vals) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  if (!t.built) H2OT_FAIL(H2OT_ERR_INVALID, "no level built");
  if (t.in_xbuf && h->xr_owner == k) h->xr_owner = -1;  // the finalize kernel enqueued below reads the payload in stream order
  const size_t LC = (size_t)t.L * h->C;
  TRY(t.d_vals.ensure(std::max<int64_t>(1, t.total_vals)));
  {
    const size_t o_best = 4 * LC * sizeof(double), o_has = o_best + (size_t)t.L * sizeof(h2ot_split);
    t.pack_bytes = o_has + LC;
    TRY(t.d_pack.ensure(t.pack_bytes));
    t.p_min2 = reinterpret_cast<double*>(t.d_pack.p);
    t.p_maxIn = t.p_min2 + LC;
    t.p_wna = t.p_maxIn + LC;
    t.p_conc = t.p_wna + LC;
    t.p_best = reinterpret_cast<h2ot_split*>(t.d_pack.p + o_best);
    t.p_has = t.d_pack.p + o_has;
  }
  FinalizeArgs fa{};
  fa.L = t.L;
  fa.C = h->C;
  fa.descs = t.p_descs;
  fa.acc = t.sums + t.L;
  fa.minmax = t.mm;
  fa.leaf_rows = reinterpret_cast<const unsigned long long*>(t.sums);
  fa.vals = t.d_vals.p;
  fa.min2 = t.p_min2;
  fa.maxIn = t.p_maxIn;
  fa.has_vals = t.p_has;
  fa.wna = t.p_wna;
  fa.conc = t.p_conc;
  const int bitsN = ceil_log2_u64((uint64_t)h->cfg.nrows_global + 1);
  const int ew = h->w ? exponent_for(h->w_max, bitsN) : 0;
  fa.inv_sy = std::ldexp(1.0, -t.ey);
  fa.inv_syy = std::ldexp(1.0, -t.eyy);
  fa.inv_sw = std::ldexp(1.0, -ew);
  // levels enqueued back to back (h2ot_tree_levels_begin): the split search that follows does each pair's finalize itself
  static const bool no_fuse_fin = getenv("H2OT_NO_FUSE_FIN") != nullptr;
  const bool fuse = !no_fuse_fin && h->in_tree_loop && !out_vals && !out_min2 && !out_maxIn && !out_wNA && !out_has_vals;
  t.fin_pending = fuse;
  if (fuse) {
    t.fin_args = fa;
  } else {
    TRY(t_begin(h, &h->stats.ms_finalize));
    finalize_kernel<<<(unsigned)LC, 128, 0, h->stream>>>(fa);
    h->stats.launches++;
    TRY(t_end(h));
  }
  if (out_vals) {
    if (out_vals_capacity < t.total_vals) H2OT_FAIL(H2OT_ERR_INVALID, "out_vals too small: need " + std::to_string(t.total_vals) + " doubles");
    TRY(d2h(h, out_vals, t.d_vals.p, (size_t)t.total_vals * sizeof(double)));
  }
  if (out_min2) TRY(d2h(h, out_min2, t.p_min2, LC * sizeof(double)));
  if (out_maxIn) TRY(d2h(h, out_maxIn, t.p_maxIn, LC * sizeof(double)));
  if (out_wNA) TRY(d2h(h, out_wNA, t.p_wna, LC * sizeof(double)));
  if (out_has_vals) TRY(d2h(h, out_has_vals, t.p_has, LC));
  if (t.L == 1 && !h->defer_sync) {
    t.h_conc.resize(LC);
    TRY(d2h(h, t.h_conc.data(), t.p_conc, LC * sizeof(double)));
  }
  if (out_val_offsets)
    for (size_t i = 0; i < LC; i++) out_val_offsets[i] = t.h_descs[i].nbin >= 0 ? t.h_descs[i].voff : -1;
  t.finished = true;
  TRY(finish_call(h));
  if (t.L == 1 && !h->defer_sync)
    for (int c = 0; c < h->C; c++)
      if (t.h_descs[c].nbin >= 0) h->col_conc[c] = t.h_conc[c];
  return H2OT_OK;
}

int h2ot_find_splits(h2ot_engine* h, int32_t k, double min_rows, double min_split_improvement, h2ot_split* out_splits) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  if (!t.finished) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_finish_level must run before h2ot_find_splits");
  const size_t LC = (size_t)t.L * h->C;
  TRY(t.d_cand.ensure(LC));
  const size_t n1 = (size_t)t.max_split_nbin + 1;
  size_t smem = 8 * 9 * n1;
  if (t.any_cat) smem += 8 * (3 * n1 + n1) + 4 * n1;
  if (smem > (size_t)h->split_smem_max) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "histogram too wide for the split-search kernel");
  if (t.any_cat && ((t.max_split_nbin - 1) >> 3) + 1 > H2OT_MAX_BITSET_BYTES) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "categorical bitset above 4096 bits");
  SplitArgs sa{t.L, h->C, t.p_descs, t.d_vals.p, t.p_min2, t.p_maxIn, t.p_has, min_rows, min_split_improvement, t.d_cand.p,
               t.any_splitpts ? t.d_sp.p : nullptr, 0, FinalizeArgs{}};
  if (t.fin_pending) {
    sa.fused = 1;
    sa.fin = t.fin_args;
    t.fin_pending = false;
  }
  TRY(t_begin(h, &h->stats.ms_split));
  find_split_kernel<<<(unsigned)LC, kSplitThreads, smem, h->stream>>>(sa);
  h->stats.launches++;
  // levels enqueued back to back: next_level_kernel picks each leaf's column itself (nobody else reads p_best before it)
  static const bool no_fuse_best = getenv("H2OT_NO_FUSE_BEST") != nullptr;
  t.best_pending = !no_fuse_best && h->in_tree_loop && !out_splits;
  if (!t.best_pending) {
    best_col_kernel<<<t.L, 32, 0, h->stream>>>(t.L, h->C, t.p_descs, t.d_cand.p, t.p_best);
    h->stats.launches++;
  }
  TRY(t_end(h));
  if (out_splits) TRY(d2h(h, out_splits, t.p_best, (size_t)t.L * sizeof(h2ot_split)));
  return finish_call(h);
}

int h2ot_level_step_begin(h2ot_engine* h, int32_t k, int32_t leaf_offset, int32_t n_leaves, int32_t n_decided,
                          const h2ot_decided_node* decided, const uint8_t* bitset_pool, int64_t bitset_bytes,
                          const h2ot_hist_desc* descs, double min_rows, double min_split_improvement) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (h->ts[k].step_pending) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_level_step_begin: the previous step of this class-tree was not ended");
  static const bool trace = getenv("H2OT_TRACE") != nullptr;
  auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double tr0 = trace ? now() : 0;
  h->defer_sync = true;
  int rc = h2ot_build_level(h, k, leaf_offset, n_leaves, n_decided, decided, bitset_pool, bitset_bytes, descs);
  const double tr1 = trace ? now() : 0;
  if (rc == H2OT_OK) rc = h2ot_allreduce_level(h, k);
  if (rc == H2OT_OK) rc = h2ot_finish_level(h, k, nullptr, 0, nullptr, nullptr, nullptr, nullptr, nullptr);
  if (rc == H2OT_OK) rc = h2ot_find_splits(h, k, min_rows, min_split_improvement, nullptr);
  TreeState& t = h->ts[k];
  if (rc == H2OT_OK && t.pack_bytes > t.stage_cap) {  // pinned staging for the packed results (grow-only)
    if (t.stage) {  // (a download of an earlier step into it has been waited for by that step's _end)
      cudaFreeHost(t.stage);
    }
    t.stage = nullptr;
    t.stage_cap = 0;
    const size_t want = std::max<size_t>(t.pack_bytes * 2, 1 << 20);
    if (cudaMallocHost(&t.stage, want) != cudaSuccess) {
      set_error("cudaMallocHost failed for the level result staging buffer");
      rc = H2OT_ERR_NOMEM;
    } else {
      t.stage_cap = want;
    }
  }
  if (rc == H2OT_OK) rc = d2h(h, t.stage, t.d_pack.p, t.pack_bytes);
  if (rc == H2OT_OK && !t.done_ev && cudaEventCreateWithFlags(&t.done_ev, cudaEventDisableTiming) != cudaSuccess) {
    set_error("cudaEventCreate failed");
    rc = H2OT_ERR_CUDA;
  }
  if (rc == H2OT_OK && cudaEventRecord(t.done_ev, h->stream) != cudaSuccess) {
    set_error("cudaEventRecord failed");
    rc = H2OT_ERR_CUDA;
  }
  h->defer_sync = false;
  if (trace) fprintf(stderr, "[h2ot] level L=%d k=%d: enqueue build %.3f ms, enqueue rest %.3f ms\n", n_leaves, k, tr1 - tr0, now() - tr1);
  if (rc != H2OT_OK) {
    if (h->steps_pending == 0) finish_call(h);  // drain the stream after a failed stage
    return rc;
  }
  t.step_pending = true;
  h->steps_pending++;
  return H2OT_OK;
}

int h2ot_level_step_end(h2ot_engine* h, int32_t k, double* out_min2, double* out_maxIn, double* out_wNA, uint8_t* out_has_vals,
                        h2ot_split* out_splits) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  if (!t.step_pending) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_level_step_end without h2ot_level_step_begin");
  t.step_pending = false;
  h->steps_pending--;
  int rs = H2OT_OK;
  if (h->steps_pending == 0) {
    rs = finish_call(h);  // the last pending step: drain the stream, fold the kernel timers, check the exchange
  } else if (cudaEventSynchronize(t.done_ev) != cudaSuccess) {  // other class-trees' levels keep running behind this one
    set_error("cudaEventSynchronize failed waiting for a level");
    rs = H2OT_ERR_CUDA;
  }
  if (rs == H2OT_OK && h->xr_err_host && *h->xr_err_host) {
    set_error("peer-memory exchange timed out waiting for another rank");
    rs = H2OT_ERR_COMM;
  }
  if (rs != H2OT_OK) return rs;
  const size_t LC = (size_t)t.L * h->C;
  const unsigned char* sp = t.stage;
  if (out_min2) memcpy(out_min2, sp, LC * sizeof(double));
  if (out_maxIn) memcpy(out_maxIn, sp + LC * sizeof(double), LC * sizeof(double));
  if (out_wNA) memcpy(out_wNA, sp + 2 * LC * sizeof(double), LC * sizeof(double));
  if (out_splits) memcpy(out_splits, sp + 4 * LC * sizeof(double), (size_t)t.L * sizeof(h2ot_split));
  if (out_has_vals) memcpy(out_has_vals, sp + 4 * LC * sizeof(double) + (size_t)t.L * sizeof(h2ot_split), LC);
  {  // the children's row counts steer the next level's CTA shares (sparse levels)
    t.prev_n0.resize((size_t)t.L);
    t.prev_n1.resize((size_t)t.L);
    t.prev_leaf_offset = t.leaf_offset;
    for (int l = 0; l < t.L; l++) {
      h2ot_split sv;
      memcpy(&sv, sp + 4 * LC * sizeof(double) + (size_t)l * sizeof(h2ot_split), sizeof(sv));
      t.prev_n0[(size_t)l] = sv.col >= 0 ? sv.n0 : 0.0;
      t.prev_n1[(size_t)l] = sv.col >= 0 ? sv.n1 : 0.0;
    }
  }
  if (t.L == 1) {  // root level: refresh the planner's per-column bin-concentration estimate
    const double* conc = reinterpret_cast<const double*>(sp + 3 * LC * sizeof(double));
    for (int c = 0; c < h->C; c++)
      if (t.h_descs[c].nbin >= 0) h->col_conc[c] = conc[c];
  }
  return H2OT_OK;
}

int h2ot_level_step(h2ot_engine* h, int32_t k, int32_t leaf_offset, int32_t n_leaves, int32_t n_decided,
                    const h2ot_decided_node* decided, const uint8_t* bitset_pool, int64_t bitset_bytes, const h2ot_hist_desc* descs,
                    double min_rows, double min_split_improvement, double* out_min2, double* out_maxIn, double* out_wNA,
                    uint8_t* out_has_vals, h2ot_split* out_splits) {
  TRY(h2ot_level_step_begin(h, k, leaf_offset, n_leaves, n_decided, decided, bitset_pool, bitset_bytes, descs, min_rows,
                            min_split_improvement));
  return h2ot_level_step_end(h, k, out_min2, out_maxIn, out_wNA, out_has_vals, out_splits);
}

// ------------------------------------------------------------------------------------------------ whole tree, levels back to back
// All levels of class-tree k enqueued without a host round trip in between: the step from one level to the next
// (DecidedNode constructor + Split.nextLevelHistos, k_next.cuh) runs on the device, node ids are the slots of a complete
// binary tree, and every level is planned for all of its 2^d slots with the widest histograms they can have.
int h2ot_tree_levels_begin(h2ot_engine* h, int32_t k, int32_t max_depth, const h2ot_hist_desc* root_descs, double min_rows,
                           double min_split_improvement, int32_t nbins, int32_t nbins_cats) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (!root_descs) H2OT_FAIL(H2OT_ERR_INVALID, "null descs");
  if (max_depth < 1 || max_depth > 10) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_tree_levels: max_depth must be in 1..10");
  TreeState& t = h->ts[k];
  TreeState::DevLoop& dv = t.dev;
  if (dv.n_pend >= 2 || t.step_pending)
    H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_tree_levels_begin: two trees of this class-tree are already in flight (or a level was not ended)");
  const int C = h->C, D = max_depth;
  const int slot = dv.next_slot;
  const size_t S = ((size_t)1 << (D + 1)) - 1;  // static ids: levels 0 .. D (the last one only as children that exist or not)
  for (int c = 0; c < C; c++)
    if ((root_descs[c].flags & H2OT_DESC_ACTIVE) && (root_descs[c].flags & (H2OT_DESC_SPLITPTS | H2OT_DESC_CHECKFLOAT)))
      H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_tree_levels: uniform-adaptive histograms without the float-split correction only");
  // ---- buffers of the whole tree, sized once
  TRY(t.nodes.ensure(S));
  TRY(t.bitsets.ensure(S * H2OT_MAX_BITSET_BYTES));
  TRY(dv.nb_caps.ensure((size_t)D * C));
  TRY(dv.root_maxex.ensure((size_t)C));
  TRY(dv.maxex[0].ensure(((size_t)1 << (D - 1)) * C));
  TRY(dv.maxex[1].ensure(((size_t)1 << (D - 1)) * C));
  const size_t n_splits = ((size_t)1 << D) - 1;  // nodes of levels 0 .. D-1
  dv.o_exists = n_splits * sizeof(h2ot_split);
  dv.o_err = dv.o_exists + ((S + 7) & ~(size_t)7);
  dv.o_conc = dv.o_err + 8;
  dv.o_pred = dv.o_conc + (size_t)C * sizeof(double);
  const size_t out_bytes = (dv.o_pred + S * sizeof(float) + 63) & ~(size_t)63;
  if (out_bytes * 2 > dv.out_stage_cap || out_bytes != dv.out_half) {
    if (dv.n_pend > 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_tree_levels_begin: the tree shape changed while a tree is in flight");
    if (dv.out_stage) cudaFreeHost(dv.out_stage);
    dv.out_stage = nullptr;
    dv.out_stage_cap = 0;
    CUDA_TRY(cudaMallocHost(&dv.out_stage, out_bytes * 2));
    dv.out_stage_cap = out_bytes * 2;
    dv.out_half = out_bytes;
  }
  unsigned char* const out_stage = dv.out_stage + (size_t)slot * dv.out_half;
  TRY(dv.out_dev.ensure(dv.out_half));
  dv.tree_splits.p = reinterpret_cast<h2ot_split*>(dv.out_dev.p);
  dv.exists.p = dv.out_dev.p + dv.o_exists;
  dv.err.p = reinterpret_cast<int32_t*>(dv.out_dev.p + dv.o_err);
  dv.p_conc_out = reinterpret_cast<double*>(dv.out_dev.p + dv.o_conc);
  dv.p_pred_out = reinterpret_cast<float*>(dv.out_dev.p + dv.o_pred);
  // widest histogram a column can have at each depth (DHistogram ctor rule: max(parent nbin >> 1, nbins); categoricals keep theirs)
  std::vector<int32_t> caps((size_t)D * C, 0);
  for (int c = 0; c < C; c++) {
    int nb = (root_descs[c].flags & H2OT_DESC_ACTIVE) ? root_descs[c].nbin : 0;
    for (int d = 0; d < D; d++) {
      caps[(size_t)d * C + c] = nb;
      if (nb > 0 && !(root_descs[c].flags & H2OT_DESC_ISCAT)) nb = std::max(nb >> 1, nbins);
      if (nb > 0 && (root_descs[c].flags & H2OT_DESC_ISCAT)) nb = std::min(std::max(nb, 1), nbins_cats);
    }
  }
  {  // one pinned arena for the uploads of all levels: [descriptors of 2^d slots x C] + groups and column lists (bounded by the same)
    std::vector<size_t> need((size_t)D);
    size_t sum = 0;
    for (int d = 0; d < D; d++) {
      need[(size_t)d] = (((size_t)1 << d) * C * sizeof(DescDev) * 2 + ((size_t)C + 64) * 64 + 4095) & ~(size_t)4095;
      sum += need[(size_t)d];
    }
    if (sum * 2 > dv.arena_cap || sum != dv.arena_half || dv.lvl_off.size() != (size_t)D) {
      if (dv.n_pend > 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_tree_levels_begin: the tree shape changed while a tree is in flight");
      CUDA_TRY(cudaStreamSynchronize(h->stream));
      if (dv.arena) cudaFreeHost(dv.arena);
      dv.arena = nullptr;
      dv.arena_cap = 0;
      CUDA_TRY(cudaMallocHost(&dv.arena, sum * 2));
      dv.arena_cap = sum * 2;
      dv.arena_half = sum;
      for (size_t i = 0; i < dv.lvl_stage.size(); i++)
        if (dv.lvl_stage[i] && dv.lvl_stage_own[i]) cudaFreeHost(dv.lvl_stage[i]);
      dv.lvl_stage.assign((size_t)D, nullptr);
      dv.lvl_stage_cap.assign((size_t)D, 0);
      dv.lvl_stage_own.assign((size_t)D, 0);
      dv.lvl_off.assign((size_t)D, 0);
      size_t off = 0;
      for (int d = 0; d < D; d++) {
        dv.lvl_off[(size_t)d] = off;
        dv.lvl_stage_cap[(size_t)d] = need[(size_t)d];
        off += need[(size_t)d];
      }
    }
    for (int d = 0; d < D; d++)  // this tree's half (a level that outgrew its slice keeps its own block)
      if (!dv.lvl_stage_own[(size_t)d]) dv.lvl_stage[(size_t)d] = dv.arena + (size_t)slot * dv.arena_half + dv.lvl_off[(size_t)d];
  }
  // node table: col = pid = -1 everywhere; only the root exists; no error; no concentration measured yet
  tree_init_kernel<<<(unsigned)((S * sizeof(h2ot_decided_node) / 4 + 255) / 256), 256, 0, h->stream>>>(
      reinterpret_cast<uint32_t*>(t.nodes.p), (int)(S * sizeof(h2ot_decided_node) / 4), dv.exists.p, (int)S, dv.err.p, dv.p_conc_out, C);
  h->stats.launches++;
  // the two small tables change only with the model's root histograms: uploaded when they differ from what the device holds
  // (pageable sources: staged by the runtime before the call returns)
  if (caps != dv.h_caps) {
    CUDA_TRY(cudaMemcpyAsync(dv.nb_caps.p, caps.data(), caps.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    dv.h_caps = caps;
  }
  {
    std::vector<double> mx(C);
    for (int c = 0; c < C; c++) mx[c] = root_descs[c].maxEx;
    if (mx.size() != dv.h_root_maxex.size() || memcmp(mx.data(), dv.h_root_maxex.data(), mx.size() * sizeof(double)) != 0) {
      CUDA_TRY(cudaMemcpyAsync(dv.root_maxex.p, mx.data(), (size_t)C * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      dv.h_root_maxex = mx;
    }
  }
  t.n_nodes = (int)S;
  t.h_nodes.clear();
  t.h_bitsets.clear();
  dv.max_depth = D;
  dv.nbins = nbins;
  dv.nbins_cats = nbins_cats;
  dv.min_rows = min_rows;
  TRY(t.d_cand.ensure(((size_t)1 << (D - 1)) * C));  // (never reallocated while a kernel of an earlier level still reads it)
  // (no early return between here and the reset of these flags below: every failure goes through rc)
  dv.active = true;
  h->defer_sync = true;
  h->in_tree_loop = true;
  t.best_pending = t.fin_pending = false;
  int rc = H2OT_OK;
  std::vector<h2ot_hist_desc> tmpl;
  for (int d = 0; d < D && rc == H2OT_OK; d++) {
    const int L = 1 << d;
    dv.depth = d;
    const h2ot_hist_desc* descs = root_descs;
    if (d > 0) {
      std::swap(t.d_lvl, dv.lvl_alt);    // the previous level's descriptors and results stay readable for the step kernel
      std::swap(t.d_pack, dv.pack_alt);
      tmpl.assign((size_t)L * C, h2ot_hist_desc{});
      for (int l = 0; l < L; l++)
        for (int c = 0; c < C; c++) {
          const int nb = caps[(size_t)d * C + c];
          if (nb <= 0) continue;
          h2ot_hist_desc& e = tmpl[(size_t)l * C + c];
          e = root_descs[c];
          e.nbin = nb;
          e.min = 0;
          e.maxEx = 1;
          e.step = 1.0;  // (categorical columns: the planner sizes the split search for a sorted-group scan when step == 1)
          e.order = (int16_t)c;
          e.flags = (uint8_t)(root_descs[c].flags & ~H2OT_DESC_INITNA);
        }
      descs = tmpl.data();
    }
    rc = h2ot_build_level(h, k, L - 1, L, 0, nullptr, nullptr, 0, descs);
    if (rc == H2OT_OK) rc = h2ot_allreduce_level(h, k);
    if (rc == H2OT_OK) rc = h2ot_finish_level(h, k, nullptr, 0, nullptr, nullptr, nullptr, nullptr, nullptr);
    if (rc == H2OT_OK) rc = h2ot_find_splits(h, k, min_rows, min_split_improvement, nullptr);
    if (rc != H2OT_OK) break;
    if (d == 0) dv.conc_src = t.p_conc;  // (the step kernel behind the root level copies it into the tree's output block)
    dv.prev_descs = t.p_descs;
    dv.prev_min2 = t.p_min2;
    dv.prev_maxIn = t.p_maxIn;
    dv.prev_wna = t.p_wna;
    dv.prev_has = t.p_has;
    dv.prev_best = t.p_best;
    dv.prev_maxex = d == 0 ? dv.root_maxex.p : dv.maxex[d & 1].p;
  }
  if (rc == H2OT_OK) {  // the last level's nodes and whether their children exist
    NextArgs na{};
    na.C = C;
    na.Lp = 1 << (D - 1);
    na.parent_base = na.Lp - 1;
    na.last_level = 1;
    na.nbins = nbins;
    na.nbins_cats = nbins_cats;
    na.min_rows = min_rows;
    na.pdescs = dv.prev_descs;
    na.pmaxex = dv.prev_maxex;
    na.best = dv.prev_best;
    na.min2 = dv.prev_min2;
    na.maxIn = dv.prev_maxIn;
    na.wna = dv.prev_wna;
    na.has_vals = dv.prev_has;
    na.nb_cap = dv.nb_caps.p;
    na.nodes = t.nodes.p;
    na.bitsets = t.bitsets.p;
    na.exists = dv.exists.p;
    na.tree_splits = dv.tree_splits.p;
    na.err = dv.err.p;
    if (t.best_pending) {
      na.cand = t.d_cand.p;
      na.best_out = const_cast<h2ot_split*>(dv.prev_best);
      t.best_pending = false;
    }
    if (na.Lp == 1) {
      na.conc_src = dv.conc_src;
      na.conc_dst = dv.p_conc_out;
    }
    next_level_kernel<<<na.Lp, 128, 0, h->stream>>>(na);
    h->stats.launches++;
    if (cudaGetLastError() != cudaSuccess) {
      set_error("next_level_kernel launch failed");
      rc = H2OT_ERR_CUDA;
    }
  }
  if (rc == H2OT_OK) rc = d2h(h, out_stage, dv.out_dev.p, dv.o_pred);  // splits, exists, err, root concentration in one copy
  if (rc == H2OT_OK && h->xr_err && h->xr_err_host && h->nccl.world > 1 &&
      cudaMemcpyAsync(h->xr_err_host, h->xr_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) {
    set_error("cudaMemcpyAsync failed");
    rc = H2OT_ERR_CUDA;
  }
  if (rc == H2OT_OK && !dv.ev[slot] && cudaEventCreateWithFlags(&dv.ev[slot], cudaEventDisableTiming) != cudaSuccess) {
    set_error("cudaEventCreate failed");
    rc = H2OT_ERR_CUDA;
  }
  if (rc == H2OT_OK && cudaEventRecord(dv.ev[slot], h->stream) != cudaSuccess) {
    set_error("cudaEventRecord failed");
    rc = H2OT_ERR_CUDA;
  }
  dv.active = false;
  h->defer_sync = false;
  h->in_tree_loop = false;
  if (rc != H2OT_OK) {
    if (h->steps_pending == 0) finish_call(h);
    return rc;
  }
  dv.next_slot = slot ^ 1;
  dv.pend_slots[dv.n_pend++] = slot;
  dv.timer_mark[slot] = h->timers.size();
  dv.has_pred[slot] = false;
  h->steps_pending++;
  return H2OT_OK;
}

// fold the first n engine timers (their events are known to have completed) into the stats
static void fold_timers(h2ot_engine* h, size_t n) {
  n = std::min(n, h->timers.size());
  for (size_t i = 0; i < n; i++) {
    Timer& t = h->timers[i];
    float ms = 0;
    if (t.a && t.b && cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) *t.acc += ms;
    else cudaGetLastError();
    if (t.a) h->event_pool.push_back(t.a);
    if (t.b) h->event_pool.push_back(t.b);
  }
  h->timers.erase(h->timers.begin(), h->timers.begin() + (ptrdiff_t)n);
  for (TreeState& ts : h->ts)
    for (size_t& m : ts.dev.timer_mark) m = m > n ? m - n : 0;
}

int h2ot_tree_levels_end2(h2ot_engine* h, int32_t k, h2ot_split* out_splits, uint8_t* out_exists, float* out_leaf_pred) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  TreeState::DevLoop& dv = t.dev;
  if (dv.n_pend <= 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_tree_levels_end without h2ot_tree_levels_begin");
  const int slot = dv.pend_slots[0];
  dv.pend_slots[0] = dv.pend_slots[1];
  dv.n_pend--;
  h->steps_pending--;
  int rs = H2OT_OK;
  if (h->steps_pending == 0) rs = finish_call(h);
  else if (cudaEventSynchronize(dv.ev[slot]) != cudaSuccess) {  // younger trees keep running behind this one
    set_error("cudaEventSynchronize failed waiting for a tree");
    rs = H2OT_ERR_CUDA;
  } else fold_timers(h, dv.timer_mark[slot]);
  if (rs == H2OT_OK && h->xr_err_host && *h->xr_err_host) {
    set_error("peer-memory exchange timed out waiting for another rank");
    rs = H2OT_ERR_COMM;
  }
  if (rs != H2OT_OK) return rs;
  const unsigned char* out_stage = dv.out_stage + (size_t)slot * dv.out_half;
  int32_t err = 0;
  memcpy(&err, out_stage + dv.o_err, sizeof(err));
  if (err) {
    t.level_failed = true;
    H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_tree_levels: a histogram outgrew the level's planned layout");
  }
  const int D = dv.max_depth;
  const size_t S = ((size_t)1 << (D + 1)) - 1, n_splits = ((size_t)1 << D) - 1;
  if (out_splits) memcpy(out_splits, out_stage, n_splits * sizeof(h2ot_split));
  if (out_exists) memcpy(out_exists, out_stage + dv.o_exists, S);
  if (out_leaf_pred) {
    if (!dv.has_pred[slot]) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_tree_levels_end2: leaf predictions asked for, but h2ot_tree_finish_enqueue did not run");
    memcpy(out_leaf_pred, out_stage + dv.o_pred, S * sizeof(float));
  }
  const double* conc = reinterpret_cast<const double*>(out_stage + dv.o_conc);
  for (int c = 0; c < h->C; c++)
    if (conc[c] > 0) h->col_conc[c] = conc[c];
  return H2OT_OK;
}
int h2ot_tree_levels_end(h2ot_engine* h, int32_t k, h2ot_split* out_splits, uint8_t* out_exists) {
  return h2ot_tree_levels_end2(h, k, out_splits, out_exists, nullptr);
}

// ------------------------------------------------------------------------------------------------ end of tree
// Quantile of the residuals per leaf (k_quant.cuh). use_tree != 0: strata are the rows' leaf ids leaf_start .. leaf_start + n_strata - 1
// (NIDs_k after h2ot_gamma_pass; out-of-bag rows take part like the others, as in the reference, which stratifies by the NIDs
// column alone) and the value is y - (Tree_k + offset); use_tree == 0: one stratum of all rows, value y - offset (the initial
// prediction). out[s] = NaN for a stratum without rows. Unit weights and a single rank only.
int h2ot_leaf_quantiles(h2ot_engine* h, int32_t k, int32_t leaf_start, int32_t n_strata, int32_t use_tree, double prob, double* out) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (!out || n_strata < 1 || !(prob >= 0 && prob <= 1)) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_leaf_quantiles: bad arguments");
  if (!use_tree && n_strata != 1) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_leaf_quantiles: without a tree there is one stratum");
  if (h->w) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_leaf_quantiles: observation weights are not supported (weighted quantiles)");
  if (h->nccl.world > 1) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_leaf_quantiles: single rank only (the order statistics span the shards)");
  TreeState& t = h->ts[k];
  const size_t np = (size_t)h->n_pad;
  for (int i = 0; i < 2; i++) {
    TRY(h->q_keys[i].ensure(np));
    TRY(h->q_strata[i].ensure(np));
  }
  TRY(h->q_counts.ensure((size_t)n_strata + 1));
  TRY(h->q_idx.ensure((size_t)n_strata * 2));
  CUDA_TRY(cudaMemsetAsync(h->q_counts.p, 0, ((size_t)n_strata + 1) * sizeof(unsigned long long), h->stream));
  QuantKeysArgs qa{};
  qa.n = h->n;
  qa.n_pad = h->n_pad;
  qa.leaf_start = leaf_start;
  qa.n_strata = n_strata;
  qa.use_tree = use_tree ? 1 : 0;
  qa.resp = h->resp;
  qa.w = h->w;
  qa.off = h->off;
  qa.F = use_tree ? t.F : nullptr;
  qa.nids = use_tree ? t.nids : nullptr;
  qa.keys = h->q_keys[0].p;
  qa.strata = h->q_strata[0].p;
  qa.counts = h->q_counts.p;
  quant_keys_kernel<<<grid_for(h, h->n_pad, 256), 256, 0, h->stream>>>(qa);
  h->stats.launches++;
  // value order first, then a stable sort by stratum: rows end up grouped by stratum, ascending inside
  int sbits = 1;
  while ((1 << sbits) < n_strata + 1) sbits++;
  size_t tmp1 = 0, tmp2 = 0;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp1, h->q_keys[0].p, h->q_keys[1].p, h->q_strata[0].p, h->q_strata[1].p, (int64_t)np, 0, 64, h->stream));
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp2, h->q_strata[1].p, h->q_strata[0].p, h->q_keys[1].p, h->q_keys[0].p, (int64_t)np, 0, sbits, h->stream));
  TRY(h->q_tmp.ensure(std::max(tmp1, tmp2) + 256));
  size_t tb = h->q_tmp.cap;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(h->q_tmp.p, tb, h->q_keys[0].p, h->q_keys[1].p, h->q_strata[0].p, h->q_strata[1].p, (int64_t)np, 0, 64, h->stream));
  tb = h->q_tmp.cap;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(h->q_tmp.p, tb, h->q_strata[1].p, h->q_strata[0].p, h->q_keys[1].p, h->q_keys[0].p, (int64_t)np, 0, sbits, h->stream));
  h->stats.launches += 2;
  std::vector<unsigned long long> cnt((size_t)n_strata + 1);
  TRY(d2h(h, cnt.data(), h->q_counts.p, cnt.size() * sizeof(unsigned long long)));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  // Histo.findQuantile: fractional row p2 = prob (n - 1), lo = element r2 = (long) p2, hi = element r2 + 1
  std::vector<long long> idx((size_t)n_strata * 2, -1);
  std::vector<long long> r2s((size_t)n_strata, 0);
  long long off = 0;
  for (int s = 0; s < n_strata; s++) {
    const long long n = (long long)cnt[(size_t)s];
    if (n > 0) {
      const double p2 = prob * ((double)n - 1.0);
      const long long r2 = (long long)p2;
      r2s[(size_t)s] = r2;
      idx[(size_t)s * 2] = off + r2;
      idx[(size_t)s * 2 + 1] = r2 + 1 < n ? off + r2 + 1 : -1;
    }
    off += n;
  }
  TRY(h2d(h, h->q_idx.p, idx.data(), idx.size() * sizeof(long long)));
  uint64_t* d_out = h->q_keys[1].p;  // (free again after the second sort)
  quant_gather_kernel<<<(n_strata * 2 + 127) / 128, 128, 0, h->stream>>>(h->q_keys[0].p, h->q_idx.p, n_strata * 2, d_out);
  h->stats.launches++;
  std::vector<uint64_t> vals((size_t)n_strata * 2);
  TRY(d2h(h, vals.data(), d_out, vals.size() * sizeof(uint64_t)));
  TRY(finish_call(h));
  for (int s = 0; s < n_strata; s++) {
    const double n = (double)cnt[(size_t)s];
    if (!(n > 0)) {
      out[s] = NAN;
      continue;
    }
    const double lo = dkey_inv(vals[(size_t)s * 2]);
    if (idx[(size_t)s * 2 + 1] < 0) {
      out[s] = lo;
      continue;
    }
    const double hi = dkey_inv(vals[(size_t)s * 2 + 1]);
    if (lo == hi) {  // Quantile.computeQuantile (:330-331)
      out[s] = lo;
      continue;
    }
    const double row = (double)r2s[(size_t)s];
    const double plo = (row + 0) / (n - 1), phi = (row + 1) / (n - 1);  // linearInterpolate (:344-350)
    out[s] = lo + (hi - lo) * (prob - plo) / (phi - plo);
  }
  return H2OT_OK;
}

// The rest of a tree without the host (after h2ot_tree_levels_begin): GammaPass with the leaves numbered by their static slot,
// the leaf values, AddTreeContributions, and the download of the leaf values next to the tree's splits. Nothing synchronises.
int h2ot_tree_finish_enqueue(h2ot_engine* h, int32_t k, double rate, double max_abs_leafnode_pred) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TreeState& t = h->ts[k];
  TreeState::DevLoop& dv = t.dev;
  if (dv.n_pend <= 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_tree_finish_enqueue without h2ot_tree_levels_begin");
  const int fam = h->cfg.family;
  if (fam != H2OT_FAMILY_GAUSSIAN && fam != H2OT_FAMILY_BERNOULLI && fam != H2OT_FAMILY_MULTINOMIAL)
    H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_tree_finish_enqueue: gaussian, bernoulli and multinomial only (the others take a log on the host)");
  const int slot = dv.pend_slots[dv.n_pend - 1];
  const int D = dv.max_depth;
  const int S = (1 << (D + 1)) - 1, nS = (1 << D) - 1;
  TRY(t.d_gamma.ensure((size_t)S * 4));
  TRY(t.d_pred.ensure((size_t)S * 2));
  CUDA_TRY(cudaMemsetAsync(t.d_gamma.p, 0, (size_t)S * 4 * sizeof(long long), h->stream));
  leafify_kernel<<<(nS + 127) / 128, 128, 0, h->stream>>>(t.nodes.p, dv.exists.p, nS, S);
  h->stats.launches++;
  GammaArgs ga{};
  ga.n = h->n;
  ga.n_pad = h->n_pad;
  ga.family = fam;
  ga.leaf_start = S;
  ga.n_leaf_nodes = S;
  {
    const int64_t per = (h->n + (int64_t)h->sm_count * 8 - 1) / ((int64_t)h->sm_count * 8);
    ga.rows_per_cta = (int32_t)std::min<int64_t>(65536, std::max<int64_t>(2048, (per + 1023) / 1024 * 1024));
  }
  ga.resp = h->resp;
  ga.w = h->w;
  ga.work = t.work;
  ga.nids = t.nids;
  ga.nodes = t.nodes.p;
  ga.bitsets = t.bitsets.p;
  ga.cols = h->d_cols.p;
  ga.acc = t.d_gamma.p;
  const int e_num = exponent_for(h->w_max * h->gamma_bound[0], 17);
  const int e_den = exponent_for(h->w_max * h->gamma_bound[1], 17);
  ga.F = t.F;
  ga.off = h->off;
  ga.tweedie_power = h->cfg.tweedie_power;
  ga.s_num = std::ldexp(1.0, e_num);
  ga.s_den = std::ldexp(1.0, e_den);
  const int grid = (int)((h->n + ga.rows_per_cta - 1) / ga.rows_per_cta);
  TRY(t_begin(h, &h->stats.ms_gamma));
  gamma_kernel<<<grid, 256, 0, h->stream>>>(ga);
  h->stats.launches++;
  TRY(t_end(h));
  if (h->nccl.world > 1 && h->nccl.comm)
    TRY(nccl_check(h, h->nccl.AllReduce(t.d_gamma.p, t.d_gamma.p, (size_t)S * 4, 4, 0, h->nccl.comm, h->stream), "ncclAllReduce(gamma)"));
  LeafValueArgs lv{};
  lv.S = S;
  lv.acc = t.d_gamma.p;
  lv.e_num = e_num;
  lv.e_den = e_den;
  lv.rate = rate;
  lv.multinomial = fam == H2OT_FAMILY_MULTINOMIAL ? 1 : 0;
  lv.max_abs = max_abs_leafnode_pred;
  lv.pred = t.d_pred.p;
  lv.pred_out = dv.p_pred_out;
  leaf_value_kernel<<<(2 * S + 127) / 128, 128, 0, h->stream>>>(lv);
  h->stats.launches++;
  AddTreeArgs a{h->n, 2 * S, h->resp, h->w, t.nids, t.F, t.d_pred.p, nullptr};
  TRY(t_begin(h, &h->stats.ms_addtree));
  add_tree_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  TRY(t_end(h));
  TRY(d2h(h, dv.out_stage + (size_t)slot * dv.out_half + dv.o_pred, dv.p_pred_out, (size_t)S * sizeof(float)));
  CUDA_TRY(cudaEventRecord(dv.ev[slot], h->stream));
  dv.timer_mark[slot] = h->timers.size();
  dv.has_pred[slot] = true;
  CUDA_TRY(cudaGetLastError());
  return H2OT_OK;
}

// ComputePredAndRes without waiting for max |residual|: for callers that know a bound (bernoulli and multinomial residuals lie in
// [-1, 1]) or can derive one from an EARLIER tree's maxima (gaussian: |r'| <= (1 + learn_rate) max |r|). want_bounds != 0: the
// maxima are all-reduced over the communicator on the stream and copied to pinned memory behind the kernel;
// h2ot_residual_bounds_wait blocks until that copy -- not the rest of the stream -- is done. Nothing synchronises here.
int h2ot_compute_residuals_async(h2ot_engine* h, int32_t want_bounds) {
  CHECK_ENGINE(h);
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  CUDA_TRY(cudaMemsetAsync(h->d_absmax.p, 0, (h->K + 2) * sizeof(unsigned int), h->stream));
  ResidArgs a{h->n, h->K, h->cfg.nclass, h->cfg.family, h->resp, h->w, h->off, h->d_Fp.p, h->d_workp.p, h->d_absmax.p, h->cfg.tweedie_power};
  TRY(t_begin(h, &h->stats.ms_residuals));
  residuals_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  TRY(t_end(h));
  if (want_bounds) {
    if (!h->bounds_host) CUDA_TRY(cudaMallocHost(&h->bounds_host, 64 * sizeof(unsigned int)));
    if (h->K + 2 > 64) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "h2ot_compute_residuals_async: more than 62 class-trees");
    if (!h->bounds_ev) CUDA_TRY(cudaEventCreateWithFlags(&h->bounds_ev, cudaEventDisableTiming));
    // (positive floats order like their bit patterns: an unsigned max over the ranks is the max of the values)
    if (h->nccl.world > 1 && h->nccl.comm)
      TRY(nccl_check(h, h->nccl.AllReduce(h->d_absmax.p, h->d_absmax.p, (size_t)h->K + 2, /*ncclUint32*/ 3, /*ncclMax*/ 2, h->nccl.comm, h->stream),
                     "ncclAllReduce(residual bounds)"));
    TRY(d2h(h, h->bounds_host, h->d_absmax.p, (h->K + 2) * sizeof(unsigned int)));
    CUDA_TRY(cudaEventRecord(h->bounds_ev, h->stream));
    h->bounds_pending = true;
  }
  CUDA_TRY(cudaGetLastError());
  return H2OT_OK;
}
int h2ot_residual_bounds_wait(h2ot_engine* h, double* out_abs_max, double out_gamma_bounds[2]) {
  CHECK_ENGINE(h);
  if (!h->bounds_pending) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_residual_bounds_wait without h2ot_compute_residuals_async(want_bounds)");
  CUDA_TRY(cudaEventSynchronize(h->bounds_ev));
  h->bounds_pending = false;
  for (int i = 0; i < h->K + 2; i++) {
    float f;
    memcpy(&f, &h->bounds_host[i], 4);
    if (i < h->K) {
      if (out_abs_max) out_abs_max[i] = (double)f;
    } else if (out_gamma_bounds) out_gamma_bounds[i - h->K] = f > 0 ? (double)f : 1.0;
  }
  return H2OT_OK;
}

int h2ot_set_tree(h2ot_engine* h, int32_t k, int32_t n_nodes, const h2ot_decided_node* nodes, const uint8_t* bitset_pool,
                  int64_t bitset_bytes) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TRY(upload_nodes(h, h->ts[k], n_nodes, nodes, bitset_pool, bitset_bytes, true));
  return finish_call_lazy(h);  // (the uploads come from the engine's own host mirror of the table)
}

int h2ot_gamma_pass(h2ot_engine* h, int32_t k, int32_t leaf_start, int32_t n_leaf_nodes, double* out_num, double* out_denom) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (n_leaf_nodes <= 0) H2OT_FAIL(H2OT_ERR_INVALID, "n_leaf_nodes must be positive");
  TreeState& t = h->ts[k];
  if (t.n_nodes == 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_set_tree must run before h2ot_gamma_pass");
  TRY(t.d_gamma.ensure((size_t)n_leaf_nodes * 4));
  CUDA_TRY(cudaMemsetAsync(t.d_gamma.p, 0, (size_t)n_leaf_nodes * 4 * sizeof(long long), h->stream));
  GammaArgs ga{};
  ga.n = h->n;
  ga.n_pad = h->n_pad;
  ga.family = h->cfg.family;
  ga.leaf_start = leaf_start;
  ga.n_leaf_nodes = n_leaf_nodes;
  {  // enough CTAs to fill the GPU; at most 2^16 rows per CTA partial (the fixed-point split below assumes it)
    const int64_t per = (h->n + (int64_t)h->sm_count * 8 - 1) / ((int64_t)h->sm_count * 8);
    ga.rows_per_cta = (int32_t)std::min<int64_t>(65536, std::max<int64_t>(2048, (per + 1023) / 1024 * 1024));
  }
  ga.resp = h->resp;
  ga.w = h->w;
  ga.work = t.work;
  ga.nids = t.nids;
  ga.nodes = t.nodes.p;
  ga.bitsets = t.bitsets.p;
  ga.cols = h->d_cols.p;
  ga.acc = t.d_gamma.p;
  const int e_num = exponent_for(h->w_max * h->gamma_bound[0], 17);  // <= 2^16 rows per CTA partial, one spare bit
  const int e_den = exponent_for(h->w_max * h->gamma_bound[1], 17);
  ga.F = t.F;
  ga.off = h->off;
  ga.tweedie_power = h->cfg.tweedie_power;
  ga.s_num = std::ldexp(1.0, e_num);
  ga.s_den = std::ldexp(1.0, e_den);
  const int grid = (int)((h->n + ga.rows_per_cta - 1) / ga.rows_per_cta);
  TRY(t_begin(h, &h->stats.ms_gamma));
  gamma_kernel<<<grid, 256, 0, h->stream>>>(ga);
  h->stats.launches++;
  TRY(t_end(h));
  if (h->nccl.world > 1 && h->nccl.comm)
    TRY(nccl_check(h, h->nccl.AllReduce(t.d_gamma.p, t.d_gamma.p, (size_t)n_leaf_nodes * 4, 4, 0, h->nccl.comm, h->stream), "ncclAllReduce(gamma)"));
  std::vector<long long> acc((size_t)n_leaf_nodes * 4);
  TRY(d2h(h, acc.data(), t.d_gamma.p, acc.size() * sizeof(long long)));
  TRY(finish_call(h));
  for (int i = 0; i < n_leaf_nodes; i++) {
    if (out_num) out_num[i] = std::ldexp((double)acc[i * 4 + 0], kGammaSplitBits - e_num) + std::ldexp((double)acc[i * 4 + 1], -e_num);
    if (out_denom) out_denom[i] = std::ldexp((double)acc[i * 4 + 2], kGammaSplitBits - e_den) + std::ldexp((double)acc[i * 4 + 3], -e_den);
  }
  return H2OT_OK;
}

static int add_tree_impl(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred, const double* node_factor) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (n_nodes <= 0 || !node_pred) H2OT_FAIL(H2OT_ERR_INVALID, "bad node_pred");
  TreeState& t = h->ts[k];
  TRY(t.d_pred.ensure(n_nodes));
  TRY(h2d(h, t.d_pred.p, node_pred, (size_t)n_nodes * sizeof(float)));
  if (node_factor) {
    TRY(t.d_factor.ensure(n_nodes));
    TRY(h2d(h, t.d_factor.p, node_factor, (size_t)n_nodes * sizeof(double)));
  }
  AddTreeArgs a{h->n, n_nodes, h->resp, h->w, t.nids, t.F, t.d_pred.p, node_factor ? t.d_factor.p : nullptr};
  TRY(t_begin(h, &h->stats.ms_addtree));
  add_tree_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  TRY(t_end(h));
  return finish_call_lazy(h);
}
int h2ot_add_tree(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred) { return add_tree_impl(h, k, n_nodes, node_pred, nullptr); }
int h2ot_add_tree_scaled(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred, const double* node_factor) {
  if (!node_factor) H2OT_FAIL(H2OT_ERR_INVALID, "null node_factor");
  return add_tree_impl(h, k, n_nodes, node_pred, node_factor);
}

int h2ot_score_tree(h2ot_engine* h, int32_t k, int32_t leaf_start, int32_t n_nodes, const float* node_pred) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (n_nodes <= 0 || !node_pred || leaf_start < 0) H2OT_FAIL(H2OT_ERR_INVALID, "bad node_pred / leaf_start");
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  TreeState& t = h->ts[k];
  // a root-only tree has no decided node at all: every row takes node_pred[0]
  if (t.n_nodes == 0 && leaf_start > 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_set_tree must run before h2ot_score_tree");
  TRY(t.nodes.ensure(1));
  TRY(t.bitsets.ensure(16));
  TRY(t.d_pred.ensure(n_nodes));
  TRY(h2d(h, t.d_pred.p, node_pred, (size_t)n_nodes * sizeof(float)));
  ScoreTreeArgs a{h->n, n_nodes, std::min(leaf_start, t.n_nodes), t.nodes.p, t.bitsets.p, h->d_cols.p, t.F, t.d_pred.p};
  TRY(t_begin(h, &h->stats.ms_addtree));
  score_tree_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  TRY(t_end(h));
  return finish_call(h);
}

int h2ot_score_metrics(h2ot_engine* h, h2ot_metrics* out) {
  CHECK_ENGINE(h);
  if (!out) H2OT_FAIL(H2OT_ERR_INVALID, "null out");
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  const int grid = grid_for(h, h->n, 256);
  DBuf<double> tmp;
  TRY(tmp.ensure((size_t)grid * 4));
  ScoreArgs2 a{h->n, h->K, h->cfg.nclass, h->cfg.family, 0, h->resp, h->w, h->off, h->d_Fp.p, tmp.p};
  score_metrics_kernel<<<grid, 256, 0, h->stream>>>(a);
  h->stats.launches++;
  std::vector<double> hb((size_t)grid * 4);
  TRY(d2h(h, hb.data(), tmp.p, hb.size() * sizeof(double)));
  const int rc = finish_call(h);
  tmp.release();
  if (rc != H2OT_OK) return rc;
  double acc[4] = {0, 0, 0, 0};
  for (int b = 0; b < grid; b++)
    for (int q = 0; q < 4; q++) acc[q] += hb[(size_t)b * 4 + q];
  out->count = acc[0];
  out->wcount = acc[1];
  out->sumsqe = acc[2];
  out->logloss = acc[3];
  return H2OT_OK;
}

int h2ot_auc_builders(h2ot_engine* h, int32_t max_chunks, int32_t* out_n_chunks, int64_t* out_chunk_rows, double* out) {
  CHECK_ENGINE(h);
  if (!out || !out_n_chunks || !out_chunk_rows || max_chunks < 1) H2OT_FAIL(H2OT_ERR_INVALID, "bad arguments");
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  if (h->cfg.family != H2OT_FAMILY_BERNOULLI) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "AUC2 is the binomial metric: bernoulli models only");
  // two warps (chunks) per SM unless the caller wants fewer; chunks of whole 32-row groups
  int64_t n_chunks = std::min<int64_t>(max_chunks, (int64_t)h->sm_count * 2);
  int64_t chunk_rows = (h->n + n_chunks - 1) / n_chunks;
  chunk_rows = std::max<int64_t>(32, (chunk_rows + 31) / 32 * 32);
  n_chunks = (h->n + chunk_rows - 1) / chunk_rows;
  const size_t stride = 1 + 4 * (size_t)kAucBins;
  DBuf<double> tmp;
  TRY(tmp.ensure((size_t)n_chunks * stride));
  const size_t smem = (size_t)kAucWarpsPerCta * 4 * (kAucCap + 1) * sizeof(double);
  CUDA_TRY(cudaFuncSetAttribute(auc_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  AucArgs a{h->n, chunk_rows, (int32_t)n_chunks, h->cfg.family, h->cfg.nclass, h->resp, h->w, h->off, h->ts[0].F, tmp.p};
  // spread the chunks over the SMs: one warp per CTA first, more warps per CTA only when there are more chunks than SMs
  const int grid = (int)((n_chunks + kAucWarpsPerCta - 1) / kAucWarpsPerCta);
  auc_build_kernel<<<grid, 32 * kAucWarpsPerCta, smem, h->stream>>>(a);
  h->stats.launches++;
  CUDA_TRY(cudaGetLastError());
  TRY(d2h(h, out, tmp.p, (size_t)n_chunks * stride * sizeof(double)));
  const int rc = finish_call(h);
  tmp.release();
  *out_n_chunks = (int32_t)n_chunks;
  *out_chunk_rows = chunk_rows;
  return rc;
}

// ------------------------------------------------------------------------------------------------ DRF
int h2ot_drf_init_model(h2ot_engine* h) {
  CHECK_ENGINE(h);
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  if (!h->oobt) CUDA_TRY(cudaMalloc(&h->oobt, (size_t)h->n_pad * sizeof(double)));
  DrfInitArgs a{h->n, h->n_pad, h->K, h->cfg.nclass, h->resp, h->d_Fp.p, h->d_workp.p, h->d_nidsp.p, h->oobt};
  drf_init_kernel<<<grid_for(h, h->n_pad, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  return finish_call(h);
}

int h2ot_sample(h2ot_engine* h, int32_t k, int64_t tree_seed, double sample_rate) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  TRY(t_begin(h, &h->stats.ms_sample));
  sample_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(h->n, h->cfg.row_offset, (long long)tree_seed, (float)sample_rate, sample_rate,
                                                               h->rate_per_class.empty() ? nullptr : h->d_rate_per_class.p,
                                                               (int)h->rate_per_class.size(), h->resp, h->ts[k].nids);
  h->stats.launches++;
  TRY(t_end(h));
  return finish_call_lazy(h);
}

int h2ot_drf_collect_preds(h2ot_engine* h, int32_t k, int32_t n_nodes, const float* node_pred, int32_t root_is_leaf, int32_t update_oobt) {
  CHECK_ENGINE(h);
  CHECK_K(h, k);
  if (n_nodes <= 0 || !node_pred) H2OT_FAIL(H2OT_ERR_INVALID, "bad node_pred");
  if (!h->oobt) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_drf_init_model must run first");
  TreeState& t = h->ts[k];
  if (!root_is_leaf && t.n_nodes == 0) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_set_tree must run before h2ot_drf_collect_preds");
  TRY(t.nodes.ensure(1));
  TRY(t.bitsets.ensure(16));
  TRY(t.d_pred.ensure(n_nodes));
  TRY(h2d(h, t.d_pred.p, node_pred, (size_t)n_nodes * sizeof(float)));
  CollectArgs a{h->n, n_nodes, root_is_leaf, update_oobt, 0, h->w, t.nids, t.F, h->oobt, t.d_pred.p, t.nodes.p, t.bitsets.p, h->d_cols.p};
  TRY(t_begin(h, &h->stats.ms_addtree));
  drf_collect_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(a);
  h->stats.launches++;
  TRY(t_end(h));
  return finish_call(h);
}

int h2ot_get_oobt(h2ot_engine* h, double* out) {
  CHECK_ENGINE(h);
  if (!h->oobt || !out) H2OT_FAIL(H2OT_ERR_INVALID, "no OUT_BAG_TREES column (not a DRF build) or null out");
  TRY(d2h(h, out, h->oobt, (size_t)h->n * sizeof(double)));
  return finish_call(h);
}

// ------------------------------------------------------------------------------------------------ multi-GPU plumbing
static int nccl_load(Nccl& n) {
  if (n.lib) return H2OT_OK;
  const char* names[] = {getenv("H2OT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    if (!nm || !*nm) continue;
    n.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (n.lib) break;
  }
  if (!n.lib) H2OT_FAIL(H2OT_ERR_COMM, "cannot dlopen libnccl.so.2 (set H2OT_NCCL_LIB to its path)");
  n.GetUniqueId = (int (*)(void*))dlsym(n.lib, "ncclGetUniqueId");
  n.CommInitRank = (int (*)(void**, int, IdByValue, int))dlsym(n.lib, "ncclCommInitRank");
  n.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(n.lib, "ncclAllReduce");
  n.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(n.lib, "ncclAllGather");
  n.CommDestroy = (int (*)(void*))dlsym(n.lib, "ncclCommDestroy");
  n.GetErrorString = (const char* (*)(int))dlsym(n.lib, "ncclGetErrorString");
  if (!n.GetUniqueId || !n.CommInitRank || !n.AllReduce) H2OT_FAIL(H2OT_ERR_COMM, "libnccl is missing required symbols");
  return H2OT_OK;
}

int h2ot_comm_unique_id(uint8_t out_id[H2OT_COMM_ID_BYTES]) {
  static Nccl loader;
  TRY(nccl_load(loader));
  IdByValue id;
  int rc = loader.GetUniqueId(&id);
  if (rc != 0) H2OT_FAIL(H2OT_ERR_COMM, "ncclGetUniqueId failed");
  memcpy(out_id, id.internal, H2OT_COMM_ID_BYTES);
  return H2OT_OK;
}

// A NCCL unique id can start exactly one communicator, so communicators are kept per process and keyed by their id:
// every engine created with the same id (successive model builds of one job) shares the communicator. They live until
// the process exits.
static std::mutex g_comm_mutex;
static std::map<std::string, void*> g_comms;

int h2ot_comm_init(h2ot_engine* h, int32_t world_size, int32_t rank, const uint8_t id[H2OT_COMM_ID_BYTES]) {
  CHECK_ENGINE(h);
  if (world_size < 1 || rank < 0 || rank >= world_size) H2OT_FAIL(H2OT_ERR_INVALID, "bad world size / rank");
  h->nccl.world = world_size;
  h->nccl.rank = rank;
  if (world_size == 1) return H2OT_OK;
  if (!id) H2OT_FAIL(H2OT_ERR_INVALID, "null communicator id");
  TRY(nccl_load(h->nccl));
  std::lock_guard<std::mutex> lock(g_comm_mutex);
  const std::string key(reinterpret_cast<const char*>(id), H2OT_COMM_ID_BYTES);
  auto it = g_comms.find(key + ":" + std::to_string(h->cfg.device));
  h->xr_key = key + ":" + std::to_string(h->cfg.device);
  {
    const char* e = getenv("H2OT_XREDUCE");
    h->xr_enabled = !(e && e[0] == '0');
    h->xr_coop = !(e && e[0] == '2');  // 2: the four-kernel sequence instead of the cooperative launch
    h->xr_ll = !(e && (e[0] == '2' || e[0] == '3'));  // 3: the cooperative pull exchange (barriers) instead of the pushed lines
  }
  if (it != g_comms.end()) {
    h->nccl.comm = it->second;
    h->nccl.shared = true;
    return H2OT_OK;
  }
  IdByValue v;
  memcpy(v.internal, id, H2OT_COMM_ID_BYTES);
  TRY(nccl_check(h, h->nccl.CommInitRank(&h->nccl.comm, world_size, v, rank), "ncclCommInitRank"));
  h->nccl.shared = true;
  g_comms[key + ":" + std::to_string(h->cfg.device)] = h->nccl.comm;
  return H2OT_OK;
}

int h2ot_allreduce_host_f64(h2ot_engine* h, double* inout, int64_t n, int32_t op) {
  CHECK_ENGINE(h);
  if (h->nccl.world <= 1 || !h->nccl.comm) return H2OT_OK;
  if (n <= 0 || !inout) H2OT_FAIL(H2OT_ERR_INVALID, "bad arguments");
  // persistent device + pinned scratch: this runs twice per tree, an allocation per call would cost more than the collective
  TRY(h->ar_dev.ensure(std::max<size_t>((size_t)n, 64)));
  if ((size_t)n > h->ar_host_cap) {
    if (h->ar_host) cudaFreeHost(h->ar_host);
    h->ar_host = nullptr;
    h->ar_host_cap = 0;
    const size_t cap = std::max<size_t>((size_t)n * 2, 256);
    CUDA_TRY(cudaMallocHost(&h->ar_host, cap * sizeof(double)));
    h->ar_host_cap = cap;
  }
  memcpy(h->ar_host, inout, (size_t)n * sizeof(double));
  TRY(h2d(h, h->ar_dev.p, h->ar_host, (size_t)n * sizeof(double)));
  TRY(nccl_check(h, h->nccl.AllReduce(h->ar_dev.p, h->ar_dev.p, (size_t)n, /*ncclFloat64*/ 8, op == 1 ? 2 : 0, h->nccl.comm, h->stream), "ncclAllReduce(f64)"));
  TRY(d2h(h, h->ar_host, h->ar_dev.p, (size_t)n * sizeof(double)));
  const int r = finish_call(h);
  if (r == H2OT_OK) memcpy(inout, h->ar_host, (size_t)n * sizeof(double));
  return r;
}

int h2ot_reserve_exchange(h2ot_engine* h, int64_t max_level_bins, int64_t max_level_pairs, int64_t max_leaves) {
  CHECK_ENGINE(h);
  if (max_level_bins < 0 || max_level_pairs < 0 || max_leaves < 0) H2OT_FAIL(H2OT_ERR_INVALID, "negative sizes");
  if (!(h->xr_enabled && !h->xr_failed && h->nccl.world > 1)) return H2OT_OK;
  const int rc = xr_ensure(h, (size_t)max_level_bins * 3 + (size_t)max_leaves + (size_t)max_level_pairs * 2, true);
  return rc == H2OT_ERR_UNSUPPORTED ? H2OT_OK : rc;  // no peer mapping here: the NCCL collectives serve the build
}

// ------------------------------------------------------------------------------------------------ diagnostics
int h2ot_debug_bin_codes(h2ot_engine* h, int32_t col, const h2ot_hist_desc* desc, uint16_t* out_codes) {
  CHECK_ENGINE(h);
  if (col < 0 || col >= h->C || !desc || !out_codes) H2OT_FAIL(H2OT_ERR_INVALID, "bad arguments");
  if (!h->frame_final) H2OT_FAIL(H2OT_ERR_INVALID, "frame not finalized");
  DBuf<uint16_t> tmp;
  TRY(tmp.ensure(h->n));
  ColRef cr{h->cols[col].data, h->cols[col].base, h->cols[col].storage, 0};
  DescDev dd{};
  dd.min = desc->min;
  dd.step = desc->step;
  dd.nbin = desc->nbin;
  dd.flags = desc->flags;
  const double* sp = nullptr;
  if (desc->flags & H2OT_DESC_SPLITPTS) {  // against class-tree 0's pool (h2ot_set_split_points)
    TreeState& t = h->ts[0];
    TRY(desc_split_points(t, *desc, dd));
    if (t.sp_dirty) {
      TRY(t.d_sp.ensure(std::max<size_t>(t.h_sp.size(), 1)));
      TRY(h2d(h, t.d_sp.p, t.h_sp.data(), t.h_sp.size() * sizeof(double)));
      t.sp_dirty = false;
    }
    sp = t.d_sp.p;
  }
  bin_codes_kernel<<<grid_for(h, h->n, 256), 256, 0, h->stream>>>(h->n, cr, dd, sp, tmp.p);
  h->stats.launches++;
  TRY(d2h(h, out_codes, tmp.p, (size_t)h->n * sizeof(uint16_t)));
  int r = finish_call(h);
  tmp.release();
  return r;
}

int h2ot_stats_get(h2ot_engine* h, h2ot_stats* out) {
  if (!h || !out) H2OT_FAIL(H2OT_ERR_INVALID, "null argument");
  *out = h->stats;
  return H2OT_OK;
}
int h2ot_stats_reset(h2ot_engine* h) {
  if (!h) H2OT_FAIL(H2OT_ERR_INVALID, "null argument");
  h->stats = h2ot_stats{};
  return H2OT_OK;
}
int h2ot_get_weight_bound(h2ot_engine* h, double* out) {
  if (!h || !out) H2OT_FAIL(H2OT_ERR_INVALID, "null argument");
  *out = h->w_max;
  return H2OT_OK;
}
int h2ot_set_weight_bound(h2ot_engine* h, double bound) {
  if (!h || !(bound > 0) || std::isinf(bound)) H2OT_FAIL(H2OT_ERR_INVALID, "weight bound must be positive and finite");
  h->w_max = bound;
  return H2OT_OK;
}
int h2ot_set_timers(h2ot_engine* h, int32_t mode) {
  if (!h || mode < 0 || mode > 2) H2OT_FAIL(H2OT_ERR_INVALID, "h2ot_set_timers: mode must be 0, 1 or 2");
  h->timer_mode = mode;
  return H2OT_OK;
}
int h2ot_set_lazy_sync(h2ot_engine* h, int32_t on) {
  CHECK_ENGINE(h);
  h->lazy_sync = on != 0;
  if (!on) return finish_call(h);
  return H2OT_OK;
}

int h2ot_set_sample_rate_per_class(h2ot_engine* h, const double* rates, int32_t n) {
  CHECK_ENGINE(h);
  if (n < 0 || (n > 0 && !rates)) H2OT_FAIL(H2OT_ERR_INVALID, "bad rates");
  if (n > 0 && n < h->cfg.nclass) H2OT_FAIL(H2OT_ERR_INVALID, "sample_rate_per_class needs one rate per response class");
  h->rate_per_class.assign(rates, rates + n);
  if (n > 0) {
    TRY(h->d_rate_per_class.ensure(n));
    TRY(h2d(h, h->d_rate_per_class.p, h->rate_per_class.data(), (size_t)n * sizeof(double)));
    return finish_call(h);
  }
  return H2OT_OK;
}

int h2ot_set_hist_mode(h2ot_engine* h, int32_t mode) {
  if (!h || mode < 0 || mode > 2) H2OT_FAIL(H2OT_ERR_INVALID, "bad mode");
  h->hist_mode = mode;
  return H2OT_OK;
}

}  // extern "C"
