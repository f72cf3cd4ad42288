// h2ot_hist_column_noalloc — a histogram entry point for callers that must not allocate device memory.
//
// Who needs it: code that runs inside somebody else's memory budget or on a latency-critical path — a JVM or service
// that handed the whole GPU to a pooling allocator at start-up (a later cudaMalloc would fail or fragment the pool), a
// real-time scoring loop where an allocation is an unbounded stall (cudaMalloc/cudaFree synchronise the device), or a
// plug-in loaded into a process whose policy forbids allocation after initialisation. Such callers already own device
// buffers for their inputs and outputs; what they cannot provide is the kernel's private accumulator storage. Here that
// storage is a fixed file-scope device array, so the call allocates nothing: no cudaMalloc, no cudaFree, no library
// routine that allocates.
//
// It is the global-accumulator histogram kernel (hist_global_kernel in k_level.cuh) cut down to one fp32 column with unit
// weights: rows -> (leaf, bin) -> int64 fixed-point sums of 1, y, y*y in the scratch array -> fp64 output with the
// reference's float cast of wY / wYY (DHistogram.updateHisto + reducePrecision, T/DHistogram.java:609-658, 734-740).
// The reference has no counterpart of this entry point.
//
// Limits and sharing: the scratch holds H2OT_NOALLOC_MAX_BINS (65536) bins x 3 int64 = 1.5 MiB, and every GPU on which the
// library is loaded carries that array until the process exits. All calls on one GPU use the same array: it still holds the
// previous call's sums, so each call first zeroes the part it is about to use, and two calls whose GPU work overlaps (different
// streams or host threads) would add into each other's bins — run them one after the other.
// The first call on each GPU looks up the array's device address (cudaGetSymbolAddress) and caches it; make that first
// call outside any stream capture.
#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace h2ot {
namespace {

constexpr int kMaxLeaves = H2OT_NOALLOC_MAX_LEAVES;
constexpr long long kMaxBins = H2OT_NOALLOC_MAX_BINS;

// Accumulator scratch. In host code this name is only a handle for the runtime's symbol table, not a device address:
// host code must go through cudaGetSymbolAddress (see scratch_for_current_device) before handing it to a kernel.
static __device__ __align__(256) long long g_noalloc_scratch[kMaxBins * 3];

struct LeafDesc {
  double min, step;
  int32_t nbin;
  int32_t bin0;  // first bin of this leaf in the scratch / output (in bins)
};
struct NoAllocArgs {
  const float* x;
  const float* work;
  const uint16_t* leaf;  // nullptr: every row belongs to leaf 0
  long long n;
  int32_t n_leaves;
  long long* acc;
  double* out;
  double sy, syy, inv_sy, inv_syy;
  LeafDesc d[kMaxLeaves];
};

__global__ void __launch_bounds__(256) noalloc_clear_kernel(long long* acc, long long n_words) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (long long)gridDim.x * blockDim.x) acc[i] = 0;
}

__global__ void __launch_bounds__(256) noalloc_hist_kernel(NoAllocArgs a) {
  for (long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += (long long)gridDim.x * blockDim.x) {
    const int l = a.leaf ? (int)a.leaf[row] : 0;
    if (l >= a.n_leaves) continue;  // kSkipCode and anything out of range
    const LeafDesc d = a.d[l];
    const double x = (double)a.x[row];
    const double y = (double)a.work[row];
    const int b = isnan(x) ? d.nbin : bin_of(x, d.min, d.step, d.nbin, false);
    const long long qy = __double2ll_rn(__dmul_rn(y, a.sy));
    const long long qyy = __double2ll_rn(__dmul_rn(__dmul_rn(y, y), a.syy));
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(a.acc + (long long)(d.bin0 + b) * 3);
    atomicAdd(dst, 1ULL);
    if (qy) atomicAdd(dst + 1, (unsigned long long)qy);
    if (qyy) atomicAdd(dst + 2, (unsigned long long)qyy);
  }
}

__global__ void __launch_bounds__(256) noalloc_finalize_kernel(const long long* acc, double* out, long long n_bins, double inv_sy,
                                                               double inv_syy) {
  for (long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; b < n_bins; b += (long long)gridDim.x * blockDim.x) {
    out[b * 3 + 0] = (double)acc[b * 3 + 0];
    out[b * 3 + 1] = (double)(float)__dmul_rn((double)acc[b * 3 + 1], inv_sy);
    out[b * 3 + 2] = (double)(float)__dmul_rn((double)acc[b * 3 + 2], inv_syy);
  }
}

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
long long* g_scratch_addr[kMaxDevices] = {nullptr};

// Once per device, with that device current: resolve the scratch array's device address.
int scratch_for_current_device(long long** out) {
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "device ordinal above 63");
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) H2OT_FAIL(H2OT_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    g_scratch_addr[dev] = static_cast<long long*>(p);
  }
  *out = g_scratch_addr[dev];
  return H2OT_OK;
}

int exponent_for_bound(double bound, int bitsN) {  // same rule as the engine: N * bound * 2^e <= 2^62
  if (!(bound > 0) || std::isinf(bound)) bound = 1.0;
  int e2;
  std::frexp(bound, &e2);
  const int e = 62 - bitsN - e2;
  return e < -900 ? -900 : (e > 900 ? 900 : e);
}

}  // namespace
}  // namespace h2ot

using namespace h2ot;

extern "C" int h2ot_hist_column_noalloc(const float* d_x, const float* d_work, const uint16_t* d_leaf_code, int64_t n, int32_t n_leaves,
                                        const h2ot_hist_desc* descs, double work_abs_bound, double* d_out_vals, void* cuda_stream) {
  if (!d_x || !d_work || !descs || !d_out_vals) H2OT_FAIL(H2OT_ERR_INVALID, "null argument");
  if (n < 0) H2OT_FAIL(H2OT_ERR_INVALID, "negative row count");
  if (n_leaves < 1 || n_leaves > kMaxLeaves)
    H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "n_leaves must be in 1.." + std::to_string(kMaxLeaves) + " for the no-allocation entry point");
  NoAllocArgs a{};
  long long bins = 0;
  for (int l = 0; l < n_leaves; l++) {
    const h2ot_hist_desc& s = descs[l];
    if (s.nbin < 1 || s.nbin > 65534) H2OT_FAIL(H2OT_ERR_INVALID, "nbin out of range (1..65534)");
    if (!(s.step > 0) || std::isinf(s.step)) H2OT_FAIL(H2OT_ERR_INVALID, "histogram step must be positive and finite");
    a.d[l].min = s.min;
    a.d[l].step = s.step;
    a.d[l].nbin = s.nbin;
    a.d[l].bin0 = (int32_t)bins;
    bins += s.nbin + 1;
  }
  if (bins > kMaxBins)
    H2OT_FAIL(H2OT_ERR_UNSUPPORTED, "histogram needs " + std::to_string(bins) + " bins (NA buckets included); the fixed scratch of the "
              "no-allocation entry point holds " + std::to_string(kMaxBins) + " — use h2ot_build_level for larger levels");
  long long* scratch = nullptr;
  {
    const int rc = scratch_for_current_device(&scratch);
    if (rc != H2OT_OK) return rc;
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int bitsN = ceil_log2_u64((uint64_t)n + 1);
  const double b = work_abs_bound > 0 ? work_abs_bound : 1.0;
  const int ey = exponent_for_bound(b, bitsN), eyy = exponent_for_bound(b * b, bitsN);
  a.x = d_x;
  a.work = d_work;
  a.leaf = d_leaf_code;
  a.n = n;
  a.n_leaves = n_leaves;
  a.acc = scratch;
  a.out = d_out_vals;
  a.sy = std::ldexp(1.0, ey);
  a.syy = std::ldexp(1.0, eyy);
  a.inv_sy = std::ldexp(1.0, -ey);
  a.inv_syy = std::ldexp(1.0, -eyy);
  int sms = 148;
  {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const long long words = bins * 3;
  const int g_small = (int)std::max<long long>(1, std::min<long long>((words + 255) / 256, (long long)sms * 8));
  const int g_rows = (int)std::max<long long>(1, std::min<long long>((n + 255) / 256, (long long)sms * 8));
  // the scratch still holds the previous call's sums: zero the part this call uses
  noalloc_clear_kernel<<<g_small, 256, 0, st>>>(scratch, words);
  if (n > 0) noalloc_hist_kernel<<<g_rows, 256, 0, st>>>(a);
  noalloc_finalize_kernel<<<g_small, 256, 0, st>>>(scratch, d_out_vals, bins, a.inv_sy, a.inv_syy);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(st));
  return H2OT_OK;
}
