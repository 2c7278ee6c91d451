// svdj_noalloc.cu -- svdj_step_device_noalloc: the split-path round-robin step for callers that
// must not allocate device memory.
//
// Who cannot allocate, and why.  cudaMalloc and cudaFree are not ordinary calls: they synchronise the
// whole device, so a caller that keeps several streams busy (a ring driver that overlaps the exchange
// of one column block with the rotations of another, say) stalls all of them every time a step
// allocates its scratch.  Other callers simply have nothing left to allocate from: the process has
// handed the device's memory to one arena at start-up (a framework's caching pool, a fixed
// partition given to each tenant of a shared GPU) and the driver's allocator will answer "out of
// memory".  And a caller with a latency budget per step cannot absorb a call whose cost depends on
// the state of the driver's heap.  For them the scratch of the split path lives in the library image
// itself: a file-scope __device__ array, present on every GPU that loads libsvdj.so from the moment
// the module is loaded until the process exits.
//
// Consequences the caller must know:
//   * the array is kSplitScratchMaxDoubles doubles (768 KiB) per GPU, whether or not this entry point
//     is ever called; problems whose split path needs more are refused (SVDJ_ERR_UNSUPPORTED);
//   * there is one array per GPU, shared by every call: two calls whose kernels can overlap on the same
//     GPU (different streams, different host threads) race on it.  Serialise them;
//   * the array keeps what the previous call left in it.  Nothing here needs it cleared: the Gram
//     kernel writes every partial triple of the step before the rotation kernel reads any;
//   * the first call on each GPU looks the array's device address up (cudaGetSymbolAddress) and
//     remembers it.  That first call should be made before any stream capture begins.
#include <mutex>

#include "svdj_common.cuh"

namespace svdj {

// pairs * splits * 3 doubles; 98304 doubles = 32768 (pair, split) cells = 768 KiB
static constexpr size_t kSplitScratchMaxDoubles = SVDJ_NOALLOC_MAX_SCRATCH_DOUBLES;

// In host code this name is only a handle for the runtime's symbol table, not a device address:
// a kernel handed `g_split_scratch` directly would fault (or scribble on host memory where the GPU can
// reach it).  The address the kernels get comes from cudaGetSymbolAddress below.
static __device__ __align__(256) double g_split_scratch[kSplitScratchMaxDoubles];

static std::mutex g_setup_mutex;
static double *g_scratch_addr[kMaxDevices] = {};  // per device ordinal, filled on first use

// Once per device, with that device current.
static int scratch_address_for_current_device(double **out) {
  int dev = -1;
  SVDJ_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("svdj_step_device_noalloc: device ordinal %d is outside the table of %d", dev, kMaxDevices);
    return SVDJ_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (g_scratch_addr[dev] == nullptr) {
    double *ptr = nullptr;
    SVDJ_CUDA_TRY(cudaGetSymbolAddress((void **)&ptr, g_split_scratch));
    if (ptr == nullptr) {
      set_error("svdj_step_device_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
      return SVDJ_ERR_CUDA;
    }
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return SVDJ_OK;
}

}  // namespace svdj

using namespace svdj;

extern "C" {

int svdj_step_device_noalloc(size_t m, size_t n, double *dA, size_t lda, double *dV, size_t ldv, int rotate_v,
                             size_t step, double tol, int threshold_mode, unsigned long long *d_counters,
                             void *stream) {
  clear_error();
  int bad = 0;
  if (m == 0 || m > 0x7fffffffu) bad = -1;
  else if (n > 0x7fffffffu) bad = -2;
  else if (dA == nullptr) bad = -3;
  else if (lda < m) bad = -4;
  else if (rotate_v && dV == nullptr) bad = -5;
  else if (rotate_v && ldv < n) bad = -6;
  else if (step >= svdj_num_steps(n)) bad = -8;
  if (bad) {
    set_error("svdj_step_device_noalloc: argument %d is invalid", -bad);
    return bad;
  }
  const size_t need = l1_scratch_doubles(m, n);
  if (need > kSplitScratchMaxDoubles) {
    set_error("svdj_step_device_noalloc: a %zu x %zu step needs %zu doubles of scratch, the built-in array holds %zu; "
              "use svdj_step_device with a caller-provided buffer",
              m, n, need, kSplitScratchMaxDoubles);
    return SVDJ_ERR_UNSUPPORTED;
  }
  double *scratch = nullptr;
  if (int rc = scratch_address_for_current_device(&scratch)) return rc;

  StepArgs a;
  a.A = dA;
  a.lda = lda;
  a.V = dV;
  a.ldv = ldv;
  a.m = (unsigned)m;
  a.n = (unsigned)n;
  a.vrows = (unsigned)n;
  a.rotate_v = rotate_v ? 1 : 0;
  a.tol = tol;
  a.threshold_mode = threshold_mode;
  a.counters = reinterpret_cast<SweepCounters *>(d_counters);
  return l1_launch_split(a, (unsigned)step + 1u, scratch, static_cast<cudaStream_t>(stream), nullptr);
}

}  // extern "C"
