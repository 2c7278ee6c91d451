// svdj_common.cuh -- shared device/host helpers of libsvdj (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <atomic>
#include <cstdlib>
#include <utility>
#include <stddef.h>
#include <stdint.h>

#include "../../include/svdj.h"

namespace svdj {

// ---------------------------------------------------------------------------------------------
// error plumbing: every CUDA call is checked (the reference checks none)
// ---------------------------------------------------------------------------------------------
void set_error(const char *fmt, ...);
void clear_error();  // empties the calling thread's error text (every entry point starts with it)

#define SVDJ_CUDA_TRY(expr)                                                                   \
  do {                                                                                        \
    cudaError_t e__ = (expr);                                                                 \
    if (e__ != cudaSuccess) {                                                                 \
      ::svdj::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      return SVDJ_ERR_CUDA;                                                                   \
    }                                                                                         \
  } while (0)

// ---------------------------------------------------------------------------------------------
// Per-device set-up.  cudaFuncSetAttribute (the opt-in to more than 48 KB of dynamic shared memory) and the SM
// count belong to a DEVICE, not to the process: a process that drives several GPUs (svdj_opts.device, the
// multi-GPU driver in svdj_mg.cu) must repeat the set-up on each of them.  `DeviceOnce` remembers, per device
// ordinal, that `fn` succeeded there; a failure is returned and retried on the next call.
// ---------------------------------------------------------------------------------------------
constexpr int kMaxDevices = 64;
struct DeviceOnce {
  std::atomic<int> done[kMaxDevices];
  int sm_count[kMaxDevices];
  template <class F>
  cudaError_t run(F &&fn, int *sms = nullptr) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    if (!done[dev].load(std::memory_order_acquire)) {
      e = fn();
      if (e != cudaSuccess) return e;
      int n = 0;
      e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess) return e;
      sm_count[dev] = n;
      done[dev].store(1, std::memory_order_release);
    }
    if (sms) *sms = sm_count[dev];
    return cudaSuccess;
  }
};

// ---------------------------------------------------------------------------------------------
// per-sweep counters kept on the device (what lib/JacobiMethods.cu:462 computes and drops)
// ---------------------------------------------------------------------------------------------
struct SweepCounters {
  unsigned long long applied;       // pairs rotated
  unsigned long long offdiag_bits;  // bit pattern of max |alpha|/sqrt(beta*gamma); non-negative doubles order like uint64
};

// ---------------------------------------------------------------------------------------------
// The reference's round-robin ordering (lib/JacobiMethods.cu:250-251 attributes it to Sameh 1971).
// Step k = 1 .. 2*m_ord-1 with m_ord = (n+1)/2; every step holds n/2 disjoint pairs, indexed here by
// idx = 0 .. n/2-1 in the order of the reference's q loop.  Columns are 1-based inside, 0-based out.
// Restates lib/JacobiMethods.cu:278-290 (k < m_ord) and :723-735 (k >= m_ord).
// ---------------------------------------------------------------------------------------------
struct Pair {
  unsigned p, q;
};

__host__ __device__ inline Pair sameh_pair(unsigned n, unsigned k, unsigned idx) {
  const unsigned h = (n + 1u) >> 1;  // m_ordering
  unsigned q, p;
  if (k < h) {
    q = h - k + 1u + idx;  // q runs h-k+1 .. n-k
    if (q <= 2u * h - 2u * k) {
      p = 2u * h - 2u * k + 1u - q;
    } else if (q <= 2u * h - k - 1u) {
      p = 4u * h - 2u * k - q;
    } else {
      p = n;
    }
  } else {
    q = 4u * h - n - k + idx;  // q runs 4h-n-k .. 3h-k-1
    if (q < 2u * h - k + 1u) {
      p = n;
    } else if (q <= 4u * h - 2u * k - 1u) {
      p = 4u * h - 2u * k - q;
    } else {
      p = 6u * h - 2u * k - 1u - q;
    }
  }
  Pair r;
  r.p = p - 1u;
  r.q = q - 1u;
  return r;
}

// Two-level tournament over nb = 2 P wb blocks (P "virtual ranks"), the ordering the multi-GPU ring walks, as a
// function of (step k = 1 .. nb-1, pair index idx = 0 .. nb/2-1).  The blocks form 2P super-blocks of wb blocks.
// Steps 1 .. 2wb-1 (round 0): round-robin inside each pair of super-blocks (r, P + r).  Then rounds 1 .. 2P-2 of
// wb steps each: the circle method moves the super-blocks one place (super-block 0 stays at top[0]; the other
// 2P-1 places form the ring top[1] -> ... -> top[P-1] -> bottom[P-1] -> ... -> bottom[0] -> top[1]) and only the
// pairs across the two super-blocks of a virtual rank are visited (block li of the top one with block
// (li + t) mod wb of the bottom one).  Every block pair exactly once per nb-1 steps, nb/2 disjoint pairs per step.
// The lower block index comes first.
__host__ __device__ inline Pair two_level_pair(unsigned nb, unsigned P, unsigned k, unsigned idx) {
  const unsigned wb = nb / (2u * P);
  const unsigned rk = idx / wb, li = idx - rk * wb;
  Pair r;
  if (k <= 2u * wb - 1u) {
    const Pair lp = sameh_pair(2u * wb, k, li);
    r.p = (lp.p < wb ? rk : P + rk) * wb + (lp.p < wb ? lp.p : lp.p - wb);
    r.q = (lp.q < wb ? rk : P + rk) * wb + (lp.q < wb ? lp.q : lp.q - wb);
  } else {
    const unsigned rd = 1u + (k - 2u * wb) / wb, t = (k - 2u * wb) % wb;
    const unsigned ring = 2u * P - 1u;
    const unsigned sh = ring - rd % ring;
    const unsigned pt = (rk + ring - 1u + sh) % ring;           // ring position the occupant of top[rk] started at
    const unsigned pb = (2u * P - 2u - rk + sh) % ring;         // ... of bottom[rk]
    const unsigned top = (rk == 0u) ? 0u : (pt < P - 1u ? pt + 1u : P + (2u * P - 2u - pt));
    const unsigned bot = pb < P - 1u ? pb + 1u : P + (2u * P - 2u - pb);
    r.p = top * wb + li;
    r.q = bot * wb + (li + t) % wb;
  }
  if (r.p > r.q) {
    const unsigned t2 = r.p;
    r.p = r.q;
    r.q = t2;
  }
  return r;
}

// ---------------------------------------------------------------------------------------------
// Rotation parameters from the Gram triple; restates lib/JacobiMethods.cu:464-477.
// alpha = a_p.a_q, beta = |a_p|^2, gamma = |a_q|^2.  Returns false when the pair is left alone.
// ---------------------------------------------------------------------------------------------
struct Rot {
  double c, s;
};

// floor2: squared norm below which a column counts as numerically zero (0 = no floor).  Without it the
// null columns of a rank-deficient or m < n input keep "rotating" for ever: their direction is rounding
// noise, so their cosines never fall under a relative tolerance.
__host__ __device__ inline bool rotation_from_gram(double alpha, double beta, double gamma, double tol,
                                                   int threshold_mode, Rot &r, double &offdiag,
                                                   double floor2 = 0.0) {
  const double denom = sqrt(beta * gamma);
  const bool live = beta > floor2 && gamma > floor2;
  offdiag = (denom > 0.0 && live) ? fabs(alpha) / denom : 0.0;
  const bool rotate =
      live && ((threshold_mode == SVDJ_THRESH_ABS) ? (fabs(alpha) > tol) : (fabs(alpha) > tol * denom));
  if (!rotate) {
    r.c = 1.0;
    r.s = 0.0;
    return false;
  }
  const double tau = (gamma - beta) / (2.0 * alpha);
  const double root = sqrt(1.0 + tau * tau);
  const double t = 1.0 / (tau >= 0.0 ? tau + root : tau - root);
  r.c = 1.0 / sqrt(1.0 + t * t);
  r.s = t * r.c;
  return true;
}

// ---------------------------------------------------------------------------------------------
// warp / block reductions
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum of three values over the block.  Every thread returns the same bits (xor-butterfly in the warp,
// then each thread adds the per-warp partials in the same order), so every thread derives the same (c,s).
template <int THREADS>
__device__ __forceinline__ void block_sum3(double &a, double &b, double &c, double *smem /* 3*THREADS/32 */) {
  constexpr int WARPS = THREADS / 32;
  a = warp_sum(a);
  b = warp_sum(b);
  c = warp_sum(c);
  if (WARPS == 1) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    smem[warp] = a;
    smem[WARPS + warp] = b;
    smem[2 * WARPS + warp] = c;
  }
  __syncthreads();
  double sa = 0.0, sb = 0.0, sc = 0.0;
#pragma unroll
  for (int w = 0; w < WARPS; ++w) {
    sa += smem[w];
    sb += smem[WARPS + w];
    sc += smem[2 * WARPS + w];
  }
  a = sa;
  b = sb;
  c = sc;
}

__device__ __forceinline__ void publish_pair(SweepCounters *ctr, bool rotated, double offdiag) {
  if (ctr == nullptr) return;
  if (rotated) atomicAdd(&ctr->applied, 1ull);
  atomicMax(&ctr->offdiag_bits, (unsigned long long)__double_as_longlong(offdiag));
}

// ---------------------------------------------------------------------------------------------
// launchers implemented in the .cu files
// ---------------------------------------------------------------------------------------------
struct StepArgs {
  double *A;
  size_t lda;
  double *V;
  size_t ldv;
  unsigned m, n;
  unsigned vrows;  // rows of V to rotate: n, or n padded with zero rows
  int rotate_v;
  double tol;
  int threshold_mode;
  SweepCounters *counters;  // may be null
  double floor2 = 0.0;      // columns with squared norm <= floor2 are treated as zero (never rotated)
};

// Level-1 fused step (svdj_l1.cu).  k = 1-based step of the reference ordering.
// scratch: device buffer of l1_scratch_doubles(m, n) doubles, only needed when l1_uses_split(m, n).
int l1_launch_step(const StepArgs &a, unsigned k, double *scratch, cudaStream_t st, unsigned long long *launches);
int l1_launch_split(const StepArgs &a, unsigned k, double *scratch, cudaStream_t st, unsigned long long *launches);
bool l1_uses_split(size_t m, size_t n);
size_t l1_scratch_doubles(size_t m, size_t n);

// Block-Jacobi step (svdj_block.cu): column blocks of 16, block pairs in the same round-robin ordering.
struct BlockArgs {
  double *A;
  size_t lda;
  double *V;
  size_t ldv;
  unsigned m, n;
  unsigned vrows;
  int rotate_v;
  unsigned nb;  // number of column blocks, ceil(n / 16)
  // second buffers: column c >= split lives at A2 + (c - split) * lda (V2 likewise); split = n when unused
  double *A2;
  double *V2;
  unsigned split;
  int sched;  // 0: round-robin over all nb blocks; 1: bipartite, first half of the blocks against the second
  // 1: order the 32 columns of a block pair by decreasing eigenvalue (fewer sweeps on one GPU).  Ring drivers
  // pass 0: the ordering moves columns from one travelling block to the other, and with the tournament's
  // fixed block 0 the largest column of the last rank then never meets the largest column of rank 0.
  int sort_cols;
  // Optional store targets (null = store where the column was loaded from): block bases for the first and the
  // second buffer.  A ring driver sets them, for the last step of a round, to peer-mapped buffers of the
  // neighbour GPUs.  force_update makes that step rewrite every block pair, rotated or not.
  double *A_st = nullptr, *A2_st = nullptr, *V_st = nullptr, *V2_st = nullptr;
  int force_update = 0;
  int vranks = 1;      // sched 2: number of virtual ranks P of the two-level tournament
  double *gcache = nullptr;  // b = 32: per column block, its 32 x 32 Gram matrix as the last eigen-solve left it
  int gram_mode = 0;   // b = 32 with gcache: 1 = full Gram, refresh the cache; 2 = cross block only, diagonal blocks from the cache
  int cross_only = 0;  // eigen-solve visits only the pairs with one column in each block (set per step by the launchers)
  double tol;
  int threshold_mode;
  int max_inner_sweeps;     // cap on the cyclic sweeps of the 32 x 32 eigen-solve
  int inner_again = 1;      // another inner sweep only if the last one applied at least this many rotations
  SweepCounters *counters;  // may be null
  double floor2 = 0.0;      // columns with squared norm <= floor2 are treated as zero (never rotated)
  // doubles the caller's partial-Gram buffer holds; the launchers refuse a step that needs more (0 = unchecked)
  size_t gpart_capacity = 0;
};
// Optional timing of the update kernel (the dominant, HBM-bound one) of a step: events recorded around it
// and a device counter that receives the number of block pairs it really updated.
// Kernel launch with the programmatic-dependent-launch attribute (on unless SVDJ_PDL=0): the next kernel of the
// stream may be scheduled while this one drains; the kernels order themselves with griddepcontrol.wait
// (svdj_block_dev.cuh).  Measured: 385 -> 372 ms at 4096^2, 80 -> 75 ms at 2048^2.
inline bool pdl_enabled() {
  static const bool on = [] {
    const char *e = std::getenv("SVDJ_PDL");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, unsigned block, size_t smem, cudaStream_t st,
                            Args &&...args) {
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t c = {};
  c.gridDim = grid;
  c.blockDim = dim3(block);
  c.dynamicSmemBytes = smem;
  c.stream = st;
  c.attrs = attr;
  c.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&c, kern, std::forward<Args>(args)...);
}

// Row-sharded driver: every rank holds a slice of the rows of A and V and all the columns.  The Gram matrices of a
// step are the only quantity that needs the other ranks: after K1 the per-pair partial sums are reduced over the
// row splits into `gred` and summed over the ranks by the caller's all-reduce (in place, on the stream); every
// rank then runs the same eigen-solves and updates its own rows.
typedef int (*allreduce_fn)(void *ctx, double *dptr, size_t count, void *stream);
struct RowHook {
  allreduce_fn fn = nullptr;
  void *ctx = nullptr;
  double *gred = nullptr;  // pairs x kGramPart32 doubles
};

struct BlockProbe {
  cudaEvent_t before = nullptr, after = nullptr;
  unsigned long long *active_pairs = nullptr;
};
// A pool of probes a multi-step driver hands out to every 8th step (the events are read back once per sweep)
struct ProbeSet {
  BlockProbe *probes = nullptr;
  int count = 0, used = 0;
  unsigned long long steps_seen = 0;
  const BlockProbe *next() {
    if (probes == nullptr || (steps_seen++ % 8) != 0 || used >= count) return nullptr;
    return &probes[used++];
  }
};
unsigned blk_num_blocks(size_t n);
// scratch of a step over nb column blocks (nb as the launcher will see it, padding included)
void blk_scratch_sizes(size_t m, unsigned nb, size_t *gpart_doubles, size_t *w_doubles, size_t *flag_ints);
// k = 1-based step of the round-robin ordering over the nb blocks
struct RowHook;
int blk_launch_step(const BlockArgs &a, unsigned k, double *gpart, double *wbuf, int *wflag, cudaStream_t st,
                    unsigned long long *launches, const BlockProbe *probe = nullptr, const RowHook *rows = nullptr);

// The same step with column blocks of 32 (svdj_block32.cu); single buffer, aligned layouts only.
unsigned blk32_num_blocks(size_t n);
void blk32_scratch_sizes(size_t m, unsigned nb, size_t *gpart_doubles, size_t *w_doubles, size_t *flag_ints);
bool blk32_supported(const BlockArgs &a);
int blk32_launch_step(const BlockArgs &a, unsigned k, double *gpart, double *wbuf, int *wflag, cudaStream_t st,
                      unsigned long long *launches, const BlockProbe *probe = nullptr, const RowHook *rows = nullptr);
size_t blk32_gred_doubles(unsigned nb);  // all-reduce buffer of the row-sharded driver for nb column blocks
double default_tol(size_t m, int jobu, int jobv);

// One phase of a ring rank on its two resident column blocks (svdj_ring.cu; ws from svdj_ring_ws_create).
// redirect_last: the last step's update kernel stores the blocks at dst* (null = in place) and rewrites every pair.
int ring_phase(void *ws, size_t m, size_t vrows, size_t w, double *dA_left, double *dA_right, size_t lda,
               double *dV_left, double *dV_right, size_t ldv, int rotate_v, int phase, double tol, int inner_sweeps,
               unsigned long long *d_counters, void *stream, double *dstA_left, double *dstA_right,
               double *dstV_left, double *dstV_right, bool redirect_last, unsigned long long *launches,
               ProbeSet *probes = nullptr, int sort_cols = 0);

// Finalisation (svdj_finalize.cu)
int launch_set_identity(double *V, size_t ldv, size_t rows, size_t n, cudaStream_t st, size_t row0 = 0);
int launch_square(double *x, size_t n, cudaStream_t st);
int launch_sqrt(double *x, size_t n, cudaStream_t st);
int launch_column_norms(const double *A, size_t lda, size_t m, size_t n, double *norms, cudaStream_t st);
// out[:, k] = in[:, perm[k]] * (scale ? 1/sigma[k] : 1); columns k >= ncols_scaled copied unscaled
int launch_gather_columns(const double *in, size_t ldin, double *out, size_t ldout, size_t rows, size_t ncols,
                          const unsigned *perm, const double *sigma, size_t ncols_scaled, cudaStream_t st);

}  // namespace svdj
