// svdj_api.cu -- the C ABI (include/svdj.h) and the host-side sweep driver.
//
// Host code is C++ above a plain-C boundary, like the reference's solver functions it replaces
// (lib/JacobiMethods.cu:191 and :1177).  The structure of a solve:
//   V = I on the device                      (lib/JacobiMethods.cu:213-226)
//   repeat sweeps until one rotates nothing  (the reference stops after 1: :234, :248)
//     n-1 round-robin steps, one fused kernel launch each (:252, :697)
//     one 16-byte read of the device counters (the reference computes :462 and drops it)
//   sigma, sort, U = A/sigma on the device   (:1146-1164)
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <numeric>
#include <vector>

#include "svdj_common.cuh"
#include "svdj_host.cuh"
#include "svdj_qr.cuh"

namespace svdj {

static thread_local char g_error[512] = "";

// What opts.precondition < 0 resolves to where no second attempt is possible (see solve_entry)
static constexpr int kAutoPrecondition = 0;

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

void clear_error() { g_error[0] = 0; }

static double now_s() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

Pool g_pool;

// Default threshold of the relative test: LAPACK dgesvj's CTOL -- sqrt(m) when singular vectors are wanted (the
// columns must come out orthogonal to working precision: |U^T U - I| <= 10 n eps needs cosines ~ sqrt(m) eps,
// with m eps a 1,048,576-row input ends at 1e-10), m when only the singular values are.
double default_tol(size_t m, int jobu, int jobv) {
  const double mm = (double)(m ? m : 1);
  const bool vectors = (jobu != SVDJ_NOVEC) || (jobv != SVDJ_NOVEC);
  return (vectors ? std::sqrt(mm) : mm) * 0x1p-53;
}

static svdj_opts resolve_opts(const svdj_opts *in, size_t m, int jobu, int jobv) {
  svdj_opts o;
  svdj_default_opts(&o);
  if (in) o = *in;
  if (o.max_sweeps <= 0) o.max_sweeps = 60;
  if (o.tol <= 0.0) o.tol = (o.threshold_mode == SVDJ_THRESH_ABS) ? 1e-16 : default_tol(m, jobu, jobv);
  return o;
}

static int check_args(int jobu, int jobv, size_t m, size_t n, const double *A, size_t lda, const double *s,
                      const double *V, size_t ldv) {
  if (jobu < SVDJ_ALLVEC || jobu > SVDJ_NOVEC) return -1;
  if (jobv < SVDJ_ALLVEC || jobv > SVDJ_NOVEC) return -2;
  if (m > 0x7fffffffu) return -3;
  if (n > 0x7fffffffu) return -4;
  if (A == nullptr && m * n > 0) return -5;
  if (lda < m) return -6;
  if (s == nullptr && std::min(m, n) > 0) return -7;
  if (jobv != SVDJ_NOVEC) {
    if (V == nullptr && n > 0) return -8;
    if (ldv < n) return -9;
  }
  return 0;
}

// The sweep loop + finalisation on device-resident buffers.
// vrows: rows of V the kernels rotate (n, or n rounded up to even when the buffer has a zero pad row).
// rs != nullptr: row-sharded solve (svdj_dgesvd_rows_device): m and vrows are this rank's rows of A and V,
// v_row0 the first row of V it holds, rs->fn the all-reduce over the ranks
struct RowShard {
  allreduce_fn fn;
  void *ctx;
  size_t v_row0;
};

static int solve_on_device(int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds,
                           double *dV, size_t ldv, size_t vrows, const svdj_opts &o, svdj_stats &st,
                           cudaStream_t stream, const RowShard *rs = nullptr) {
  const size_t mn = std::min(m, n);
  int variant = o.variant;
  if (variant == SVDJ_VARIANT_AUTO) {
    // block Jacobi moves ~13x fewer bytes per sweep; it needs a few block pairs per step to fill the GPU
    variant = (blk_num_blocks(n) >= 4 && o.threshold_mode == SVDJ_THRESH_REL) ? SVDJ_VARIANT_BLOCK : SVDJ_VARIANT_LEVEL1;
  }
  if (variant != SVDJ_VARIANT_LEVEL1 && variant != SVDJ_VARIANT_BLOCK) {
    set_error("unknown variant %d", variant);
    return SVDJ_ERR_UNSUPPORTED;
  }
  if (variant == SVDJ_VARIANT_BLOCK && (blk_num_blocks(n) < 2 || o.threshold_mode != SVDJ_THRESH_REL)) {
    variant = SVDJ_VARIANT_LEVEL1;  // a single block has no partner; the absolute test is the reference's level-1 rule
  }
  st.variant = variant;
  const bool block = (variant == SVDJ_VARIANT_BLOCK);
  if (m == 0 || n == 0) return SVDJ_OK;
  const bool rotate_v = (jobv != SVDJ_NOVEC);

  Event ev0, ev1, ev2;
  if (int rc = ev0.create()) return rc;
  if (int rc = ev1.create()) return rc;
  if (int rc = ev2.create()) return rc;

  DevBuf d_counters, d_scratch, d_norms, d_perm, d_tmp;
  PinBuf h_counters, h_norms;
  if (int rc = d_counters.alloc(sizeof(SweepCounters) * (size_t)o.max_sweeps)) return rc;
  if (int rc = h_counters.alloc(sizeof(SweepCounters))) return rc;
  SVDJ_CUDA_TRY(cudaMemsetAsync(d_counters.p, 0, sizeof(SweepCounters) * (size_t)o.max_sweeps, stream));

  // block width: 16 by default; 32 on request (opts.block == 32 or SVDJ_BLOCK=32) when the layout allows it
  bool wide = false;
  if (block) {
    // 0 = choose: measured on B200 (profiles/r1_block_widths.md) the 64-column block pairs win from n = 2048 up
    // (half the steps and half the HBM traffic per sweep, fewer sweeps), the 32-column ones below
    int want = o.block;
    if (const char *env = std::getenv("SVDJ_BLOCK")) want = std::atoi(env);
    if (want == 0) want = (n >= 2048 || (m >= 8 * n && n >= 128)) ? 32 : 16;  // tall: 0.34 s vs 0.65 s at 1,048,576 x 512
    if (want == 32 && blk32_num_blocks(n) >= 2) {
      BlockArgs probe_args;
      probe_args.A = dA;
      probe_args.lda = lda;
      probe_args.V = dV;
      probe_args.ldv = ldv;
      probe_args.rotate_v = rotate_v ? 1 : 0;
      probe_args.nb = blk32_num_blocks(n);
      probe_args.A2 = nullptr;
      wide = blk32_supported(probe_args);
    }
  }
  // Empty blocks at the end (all their column slots past n: the kernels treat them like the missing columns of a
  // ragged last block) make the number of blocks even, so that every block is in a pair of the sweep's first step
  // (cross-only inner schedule, Gram cache), and from 64 blocks on a multiple of 8, so that the two-level
  // ordering applies whatever n is.  At most 7 extra blocks of 64+: they cost a Gram pass over zeros, no update.
  unsigned nb = wide ? blk32_num_blocks(n) : blk_num_blocks(n);
  if (block && std::getenv("SVDJ_NO_PAD") == nullptr) {
    if (nb >= 64) nb = (nb + 7u) & ~7u;
    else if (nb > 2 && (nb & 1u)) nb += 1u;
  }

  // Scratch of the block path, sized for exactly the kernels and the (padded) block count that will run: the
  // number of row splits of the Gram kernel is not monotonic in the number of pairs, so a size worked out for
  // any other block count can be too small.  The launchers check the capacity again (BlockArgs::gpart_capacity).
  DevBuf d_gpart, d_w, d_wflag, d_gcache;
  size_t gpart_doubles = 0;
  if (block) {
    size_t wd = 0, fl = 0;
    (wide ? blk32_scratch_sizes : blk_scratch_sizes)(m, nb, &gpart_doubles, &wd, &fl);
    if (int rc = d_gpart.alloc(sizeof(double) * gpart_doubles)) return rc;
    if (int rc = d_w.alloc(sizeof(double) * wd)) return rc;
    if (int rc = d_wflag.alloc(sizeof(int) * fl)) return rc;
    // per column block of 32: its Gram matrix as the last eigen-solve left it (8 KiB per block)
    if (wide) {
      if (int rc = d_gcache.alloc(sizeof(double) * 1024 * (size_t)nb)) return rc;
    }
  } else if (l1_uses_split(m, n)) {
    if (int rc = d_scratch.alloc(sizeof(double) * l1_scratch_doubles(m, n))) return rc;
  }

  if (rotate_v) {
    if (int rc = launch_set_identity(dV, ldv, vrows, n, stream, rs ? rs->v_row0 : 0)) return rc;
    st.launches += 1;
  }

  const unsigned steps = (unsigned)svdj_num_steps(block ? nb : n);
  // column pairs looked at per step: n/2 on the level-1 path, 496 (b = 16) or 2016 (b = 32) per block pair
  const unsigned long long pairs_per_step =
      block ? (unsigned long long)(nb / 2) * (wide ? 2016ull : 496ull) : n / 2;

  BlockArgs b;
  b.A = dA;
  b.lda = lda;
  b.V = dV;
  b.ldv = ldv;
  b.m = (unsigned)m;
  b.n = (unsigned)n;
  b.vrows = (unsigned)vrows;
  b.rotate_v = rotate_v ? 1 : 0;
  b.nb = nb;
  b.A2 = nullptr;
  b.V2 = nullptr;
  b.split = (unsigned)n;
  b.sched = 0;
  b.sort_cols = 1;
  b.vranks = 1;
  b.gcache = wide ? d_gcache.as<double>() : nullptr;  // the 64-column kernels drop it where it does not apply
  if (const char *env = std::getenv("SVDJ_GCACHE")) {
    if (env[0] == '0') b.gcache = nullptr;
  }
  // two-level tournament ordering (the ring's ordering with virtual ranks) when the blocks divide evenly
  if (block) {
    // P = 4 virtual ranks where the blocks divide evenly (measured best of 1, 2, 4, 16 at 1024..4096), else 2,
    // else the flat ordering; SVDJ_VRANKS overrides (0 = flat)
    int P = (nb % 8u == 0) ? 4 : (nb % 4u == 0) ? 2 : 0;
    if (const char *env = std::getenv("SVDJ_VRANKS")) P = std::atoi(env);
    if (P >= 2 && nb % (2u * (unsigned)P) == 0 && nb >= 4) {
      b.sched = 2;
      b.vranks = P;
      if (const char *e2 = std::getenv("SVDJ_SORT")) b.sort_cols = std::atoi(e2);
    }
  }
  b.gpart_capacity = gpart_doubles;
  b.tol = o.tol;
  b.threshold_mode = o.threshold_mode;
  // Cyclic sweeps of the small eigen-solve per visit: one below n = 4096, two from there on.  A second inner sweep
  // costs ~14 us per eigen-solve and saves outer sweeps (4096^2: 14 instead of 17, 0.376 -> 0.346 s; 8192^2: 16
  // instead of 19); at 1024 and 2048 the two effects cancel (tools/tune_inner.sh).  Round 1 drew the line at 8192:
  // its eigen-solve loop cost twice as much.  SVDJ_INNER_SWEEPS overrides.
  // Tall inputs (m >= 8 n): the eigen-solve does not grow with m while K1/K3 do, so extra inner sweeps are free
  // and every outer sweep saved counts (1,048,576 x 512: 12 sweeps with one inner sweep, 8 with three).
  b.max_inner_sweeps = (m >= 8 * n) ? 3 : (n >= 4096) ? 2 : 1;
  // ... and the second one only where the first applied at least 64 of its 1024 rotations: in the last outer sweeps
  // it finds nothing to do and costs its 14 us all the same (4096^2: 0.346 -> 0.343 s, same 14 sweeps)
  if (m < 8 * n && b.max_inner_sweeps == 2) b.inner_again = 64;
  if (const char *env = std::getenv("SVDJ_INNER_AGAIN")) b.inner_again = std::atoi(env) > 0 ? std::atoi(env) : 1;
  // row-sharded: every rank must take the same decisions (they all run the same eigen-solves on the same Gram
  // matrices); the local row count differs between ranks, so nothing may depend on it -- tall by definition
  if (rs) b.max_inner_sweeps = 3;
  if (const char *env = std::getenv("SVDJ_INNER_SWEEPS")) {
    const int v = std::atoi(env);
    if (v > 0) b.max_inner_sweeps = v;
  }

  StepArgs a;
  a.A = dA;
  a.lda = lda;
  a.V = dV;
  a.ldv = ldv;
  a.m = (unsigned)m;
  a.n = (unsigned)n;
  a.vrows = (unsigned)vrows;
  a.rotate_v = rotate_v ? 1 : 0;
  a.tol = o.tol;
  a.threshold_mode = o.threshold_mode;

  // probes of the block path's update kernel
  constexpr int kProbes = 256;
  std::vector<Event> probe_events(block ? 2 * kProbes : 0);
  std::vector<BlockProbe> probes(block ? kProbes : 0);
  DevBuf d_probe;
  int probes_used = 0;
  double probe_ms = 0.0;
  unsigned long long probe_launches = 0;
  if (block) {
    if (int rc = d_probe.alloc(sizeof(unsigned long long))) return rc;
    SVDJ_CUDA_TRY(cudaMemsetAsync(d_probe.p, 0, sizeof(unsigned long long), stream));
    for (int i = 0; i < kProbes; ++i) {
      if (int rc = probe_events[2 * i].create()) return rc;
      if (int rc = probe_events[2 * i + 1].create()) return rc;
      probes[i].before = probe_events[2 * i].e;
      probes[i].after = probe_events[2 * i + 1].e;
    }
  }

  // Columns whose norm falls below tol * (largest column norm of the input) are numerically zero: their
  // direction is rounding noise and no relative test on them can ever pass.  They are left alone
  // (rank-deficient inputs, m < n).  One pass over A and n doubles to the host, before the clock starts.
  if (o.threshold_mode == SVDJ_THRESH_REL) {
    if (int rc = d_norms.alloc(sizeof(double) * n)) return rc;
    if (int rc = launch_column_norms(dA, lda, m, n, d_norms.as<double>(), stream)) return rc;
    if (rs) {  // global column norms: sum of the ranks' squares
      if (int rc = launch_square(d_norms.as<double>(), n, stream)) return rc;
      if (int rc = rs->fn(rs->ctx, d_norms.as<double>(), n, stream)) {
        set_error("row-sharded solve: the caller's all-reduce (column norms) returned %d", rc);
        return SVDJ_ERR_CUDA;
      }
      if (int rc = launch_sqrt(d_norms.as<double>(), n, stream)) return rc;
    }
    std::vector<double> hn(n);
    SVDJ_CUDA_TRY(cudaMemcpyAsync(hn.data(), d_norms.p, sizeof(double) * n, cudaMemcpyDeviceToHost, stream));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
    double mx = 0.0;
    for (double v : hn) mx = std::max(mx, v);
    const double fl = o.tol * mx;
    a.floor2 = b.floor2 = fl * fl;
    st.launches += 1;
  }

  RowHook row_hook;
  DevBuf d_gred;
  if (rs) {
    if (!(block && wide)) {
      set_error("row-sharded solve: needs the 64-column block path (n >= 128, 16-byte aligned columns, even ld)");
      return SVDJ_ERR_UNSUPPORTED;
    }
    const size_t gd = blk32_gred_doubles(nb);
    if (int rc = d_gred.alloc(sizeof(double) * gd)) return rc;
    SVDJ_CUDA_TRY(cudaMemsetAsync(d_gred.p, 0, sizeof(double) * gd, stream));
    row_hook.fn = rs->fn;
    row_hook.ctx = rs->ctx;
    row_hook.gred = d_gred.as<double>();
  }

  SVDJ_CUDA_TRY(cudaEventRecord(ev0.e, stream));
  unsigned long long step_launches = 0;
  bool converged = false;
  const bool trace = std::getenv("SVDJ_TRACE") != nullptr;  // per-sweep progress on stderr
  // Off by default: the bound below assumes separated singular values.  Inside a cluster of equal ones a rotation
  // has a large angle however small the cosine, and refills the pairs zeroed before it to FIRST order (found by the
  // clustered-spectrum test: 2e-10 left on U).  SVDJ_EARLY_STOP=1 enables it for inputs known to be separated.
  const char *es_env = std::getenv("SVDJ_EARLY_STOP");
  const bool early_stop = block && es_env != nullptr && es_env[0] == '1';
  for (int sweep = 0; sweep < o.max_sweeps; ++sweep) {
    a.counters = d_counters.as<SweepCounters>() + sweep;
    b.counters = a.counters;
    for (unsigned k = 1; k <= steps; ++k) {
      if (block) {
        // every 8th step carries a probe around its update kernel (events + pairs really updated)
        const BlockProbe *probe = nullptr;
        if (k % 8 == 1 && probes_used < kProbes) {
          probes[probes_used].active_pairs = d_probe.as<unsigned long long>();
          probe = &probes[probes_used++];
        }
        if (rs) {
          if (int rc = blk32_launch_step(b, k, d_gpart.as<double>(), d_w.as<double>(), d_wflag.as<int>(), stream,
                                         &step_launches, probe, &row_hook))
            return rc;
        } else if (int rc = (wide ? blk32_launch_step : blk_launch_step)(b, k, d_gpart.as<double>(), d_w.as<double>(),
                                                                         d_wflag.as<int>(), stream, &step_launches, probe,
                                                                         nullptr))
          return rc;
      } else {
        if (int rc = l1_launch_step(a, k, d_scratch.as<double>(), stream, &step_launches)) return rc;
      }
    }
    SVDJ_CUDA_TRY(cudaMemcpyAsync(h_counters.p, a.counters, sizeof(SweepCounters), cudaMemcpyDeviceToHost, stream));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
    for (int i = 0; i < probes_used; ++i) {
      float ms = 0.f;
      SVDJ_CUDA_TRY(cudaEventElapsedTime(&ms, probes[i].before, probes[i].after));
      probe_ms += ms;
    }
    probe_launches += (unsigned long long)probes_used;
    probes_used = 0;
    const SweepCounters *hc = h_counters.as<SweepCounters>();
    st.sweeps = sweep + 1;
    st.tested += pairs_per_step * steps;
    st.applied += hc->applied;
    double off;
    std::memcpy(&off, &hc->offdiag_bits, sizeof(double));
    st.last_offdiag = off;
    if (trace) {
      std::fprintf(stderr, "[svdj] sweep %2d: rotations %llu, max off-diagonal cosine %.3e (tol %.3e)\n", sweep + 1,
                   (unsigned long long)hc->applied, off, o.tol);
    }
    // Converged: a sweep that rotated nothing (the reference's rule applied to lib/JacobiMethods.cu:462-466), or --
    // block path -- one in which no pair was seen above 2 tol.  tol = sqrt(m) eps sits AT the rounding noise of an
    // m-term inner product, so a finished solve keeps finding a handful of pairs a hair above it (4096^2: "2
    // rotations, max cosine 7.116e-15" against tol 7.105e-15) and would pay a whole extra sweep for them; the pairs
    // above tol were rotated in this sweep anyway.
    if (hc->applied == 0 || (block && o.threshold_mode == SVDJ_THRESH_REL && off <= 2.0 * o.tol)) {
      converged = true;
      break;
    }
    // Quadratic-convergence stop.  Every pair whose cosine exceeded tol was rotated to zero in this sweep; what a
    // pair carries afterwards is what later rotations of its two columns put back: at most (rotations per column)
    // x max_cos^2 <= applied x max_cos^2.  Below tol / 2 the next sweep would rotate nothing, i.e. the reference's
    // stopping condition (a sweep without rotations, lib/JacobiMethods.cu:462-466) already holds and the
    // confirming sweep -- 2-3 % of the solve, a whole extra tournament on several GPUs -- is not run.
    // (4096^2: sweep 16 applies 240 rotations with max cosine 2.9e-8; sweep 17 would apply none.)
    if (early_stop && off > 0.0 && off < 1e-6 && (double)hc->applied * off * off < 0.5 * o.tol &&
        o.threshold_mode == SVDJ_THRESH_REL) {
      converged = true;
      break;
    }
  }
  st.converged = converged ? 1 : 0;
  st.launches += step_launches;
  st.step_kernel_launches = step_launches;
  SVDJ_CUDA_TRY(cudaEventRecord(ev1.e, stream));

  // ---- finalisation -----------------------------------------------------------------------------
  if (d_norms.p == nullptr) {
    if (int rc = d_norms.alloc(sizeof(double) * n)) return rc;
  }
  if (int rc = launch_column_norms(dA, lda, m, n, d_norms.as<double>(), stream)) return rc;
  if (rs) {  // sigma_k^2 = sum over the ranks of the squared norms of their rows of column k
    if (int rc = launch_square(d_norms.as<double>(), n, stream)) return rc;
    if (int rc = rs->fn(rs->ctx, d_norms.as<double>(), n, stream)) {
      set_error("row-sharded solve: the caller's all-reduce (singular values) returned %d", rc);
      return SVDJ_ERR_CUDA;
    }
    if (int rc = launch_sqrt(d_norms.as<double>(), n, stream)) return rc;
  }
  st.launches += 1;
  const bool scale_u = (jobu == SVDJ_ALLVEC || jobu == SVDJ_SOMEVEC);

  if (o.sort_desc) {
    if (int rc = h_norms.alloc(sizeof(double) * n)) return rc;
    SVDJ_CUDA_TRY(cudaMemcpyAsync(h_norms.p, d_norms.p, sizeof(double) * n, cudaMemcpyDeviceToHost, stream));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
    const double *nr = h_norms.as<double>();
    std::vector<unsigned> perm(n);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [nr](unsigned x, unsigned y) { return nr[x] > nr[y]; });
    std::vector<double> sorted(n);
    for (size_t k2 = 0; k2 < n; ++k2) sorted[k2] = nr[perm[k2]];
    if (int rc = d_perm.alloc(sizeof(unsigned) * n)) return rc;
    // sorted sigma goes back into d_norms (the gather reads sigma[k] for output column k)
    SVDJ_CUDA_TRY(cudaMemcpyAsync(d_perm.p, perm.data(), sizeof(unsigned) * n, cudaMemcpyHostToDevice, stream));
    SVDJ_CUDA_TRY(cudaMemcpyAsync(d_norms.p, sorted.data(), sizeof(double) * n, cudaMemcpyHostToDevice, stream));
    const size_t tmp_rows = std::max(m, n);
    if (int rc = d_tmp.alloc(sizeof(double) * tmp_rows * n)) return rc;
    // A: gather + scale into tmp, copy back
    if (int rc = launch_gather_columns(dA, lda, d_tmp.as<double>(), m, m, n, d_perm.as<unsigned>(),
                                       d_norms.as<double>(), scale_u ? mn : 0, stream))
      return rc;
    SVDJ_CUDA_TRY(cudaMemcpy2DAsync(dA, lda * sizeof(double), d_tmp.p, m * sizeof(double), m * sizeof(double), n,
                                    cudaMemcpyDeviceToDevice, stream));
    st.launches += 1;
    if (rotate_v) {
      const size_t vr = rs ? vrows : n;  // rows of V held here
      if (int rc = launch_gather_columns(dV, ldv, d_tmp.as<double>(), vr, vr, n, d_perm.as<unsigned>(), nullptr, 0, stream))
        return rc;
      SVDJ_CUDA_TRY(cudaMemcpy2DAsync(dV, ldv * sizeof(double), d_tmp.p, vr * sizeof(double), vr * sizeof(double), n,
                                      cudaMemcpyDeviceToDevice, stream));
      st.launches += 1;
    }
    // the host vectors must outlive the async uploads
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
  } else if (scale_u) {
    if (int rc = launch_gather_columns(dA, lda, dA, lda, m, n, nullptr, d_norms.as<double>(), mn, stream)) return rc;
    st.launches += 1;
  }
  SVDJ_CUDA_TRY(cudaMemcpyAsync(ds, d_norms.p, sizeof(double) * mn, cudaMemcpyDeviceToDevice, stream));
  SVDJ_CUDA_TRY(cudaEventRecord(ev2.e, stream));
  SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));

  float ms01 = 0.f, ms12 = 0.f;
  SVDJ_CUDA_TRY(cudaEventElapsedTime(&ms01, ev0.e, ev1.e));
  SVDJ_CUDA_TRY(cudaEventElapsedTime(&ms12, ev1.e, ev2.e));
  st.seconds_sweeps = ms01 * 1e-3;
  st.seconds_finalize = ms12 * 1e-3;
  st.step_kernel_ms = step_launches ? (double)ms01 / (double)step_launches : 0.0;
  if (block) {
    // dominant kernel of the block path = the update kernel: mean of the probed launches, and the block
    // pairs those launches updated (each moves 512 (m + n) bytes: 32 columns of A and V read and written)
    unsigned long long active = 0;
    SVDJ_CUDA_TRY(cudaMemcpy(&active, d_probe.p, sizeof(active), cudaMemcpyDeviceToHost));
    st.step_kernel_ms = probe_launches ? probe_ms / (double)probe_launches : 0.0;
    st.step_kernel_launches = probe_launches;
    st.probe_block_pairs = active;
    st.reserved = wide ? 32 : 16;  // block width that ran
  }
  if (o.threshold_mode == SVDJ_THRESH_REL && !converged) return SVDJ_NOT_CONVERGED;
  return SVDJ_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Preconditioned solve (svdj_opts.precondition; Drmac & Veselic).  One-sided Jacobi needs dozens of sweeps when the
// singular values are strongly graded but the columns are not (8192^2 with cond 1e12: more than 60).  Remedy, all of it
// on the device with this library's own kernels (svdj_qr.cu):
//   1. de Rijk pre-sort: columns by decreasing norm (a cheap stand-in for column pivoting),  A P
//   2. A P = Q1 R1 (blocked Householder)                    [steps = 2:  R1^T = Q2 R2 once more]
//   3. one-sided Jacobi on L = R1^T:  L J = W diag(s)        [steps = 2:  on L = R2^T]
//   4. steps = 1:  A P = Q1 R1 = Q1 J s W^T          =>  U = Q1 [J; 0],  V = P W
//      steps = 2:  R1 = R2^T Q2^T = (W s J^T) Q2^T   =>  U = Q1 [W; 0],  V = P Q2 J
// The Jacobi iteration accumulates straight into the buffers the caller gave us where the layout allows it; the
// Householder vectors need one m x n work buffer.  Requires m >= n.
// ---------------------------------------------------------------------------------------------------------
static size_t round16(size_t x) { return (x + 15) & ~size_t(15); }

// `inner` runs the Jacobi iteration on the n x n factor (the single-GPU sweep loop, or the multi-GPU driver's
// collective solve on rank 0).  It is called exactly once, even when the QR part failed (pre_status != 0), because
// a collective inner solve has other ranks waiting for it.
int solve_with_qr(int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds, double *dV, size_t ldv,
                  int qr_steps, const svdj_opts &o, svdj_stats &st, cudaStream_t stream, const InnerSolve &inner) {
  const bool want_u = (jobu != SVDJ_NOVEC), want_v = (jobv != SVDJ_NOVEC);
  Event e0, e1, e2, e3;
  DevBuf d_norms, d_perm, d_Y, d_T1, d_T2, d_qrs, d_L, d_B1, d_J;
  unsigned long long launches = 0;
  const size_t ldy = round16(m);
  const size_t np = qr_num_panels(n);
  // L = R^T: n x n lower triangular, one zero pad row when n is odd so that the vectorised kernels apply
  const size_t rows_l = (n + 1) & ~size_t(1);
  const size_t ldl = round16(rows_l);
  // Where the Jacobi iteration accumulates its rotations J.  steps = 1: J becomes U after Q1 is applied, so it grows
  // in the top n rows of the caller's A (whose contents now live in Y); steps = 2: J becomes V after Q2, in a
  // work buffer.
  double *J = nullptr;
  size_t ldj = 0, jrows = n;
  const bool need_j = (qr_steps >= 2) ? want_v : want_u;
  const bool need_w = (qr_steps >= 2) ? want_u : want_v;

  auto prepare = [&]() -> int {
    if (int rc = e0.create()) return rc;
    if (int rc = e1.create()) return rc;
    if (int rc = e2.create()) return rc;
    if (int rc = e3.create()) return rc;
    SVDJ_CUDA_TRY(cudaEventRecord(e0.e, stream));
    // 1. column norms -> permutation (n doubles to the host and n indices back)
    if (int rc = d_norms.alloc(sizeof(double) * n)) return rc;
    if (int rc = d_perm.alloc(sizeof(unsigned) * n)) return rc;
    if (int rc = launch_column_norms(dA, lda, m, n, d_norms.as<double>(), stream)) return rc;
    std::vector<double> hn(n);
    SVDJ_CUDA_TRY(cudaMemcpyAsync(hn.data(), d_norms.p, sizeof(double) * n, cudaMemcpyDeviceToHost, stream));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
    std::vector<unsigned> perm(n);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&hn](unsigned x, unsigned y) { return hn[x] > hn[y]; });
    SVDJ_CUDA_TRY(cudaMemcpyAsync(d_perm.p, perm.data(), sizeof(unsigned) * n, cudaMemcpyHostToDevice, stream));
    // 2. Y = A P, factored in place
    if (int rc = d_Y.alloc(sizeof(double) * ldy * n)) return rc;
    if (int rc = d_T1.alloc(sizeof(double) * 1024 * np)) return rc;
    if (int rc = d_qrs.alloc(qr_scratch_bytes(n))) return rc;
    if (int rc = launch_gather_columns(dA, lda, d_Y.as<double>(), ldy, m, n, d_perm.as<unsigned>(), nullptr, 0, stream))
      return rc;
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));  // perm (host) must outlive its upload
    launches += 2;
    if (int rc = qr_factor(d_Y.as<double>(), ldy, m, n, d_T1.as<double>(), d_qrs.p, stream, &launches)) return rc;
    // 3. L = R^T
    if (int rc = d_L.alloc(sizeof(double) * ldl * n)) return rc;
    if (qr_steps >= 2) {
      if (int rc = d_B1.alloc(sizeof(double) * ldl * n)) return rc;
      if (int rc = d_T2.alloc(sizeof(double) * 1024 * np)) return rc;
      if (int rc = qr_transpose_r(d_Y.as<double>(), ldy, n, d_B1.as<double>(), ldl, n, stream)) return rc;
      if (int rc = qr_factor(d_B1.as<double>(), ldl, n, n, d_T2.as<double>(), d_qrs.p, stream, &launches)) return rc;
      if (int rc = qr_transpose_r(d_B1.as<double>(), ldl, n, d_L.as<double>(), ldl, rows_l, stream)) return rc;
      launches += 2;
    } else {
      if (int rc = qr_transpose_r(d_Y.as<double>(), ldy, n, d_L.as<double>(), ldl, rows_l, stream)) return rc;
      launches += 1;
    }
    if (qr_steps >= 2) {
      if (need_j) {
        if (int rc = d_J.alloc(sizeof(double) * ldl * n)) return rc;
        J = d_J.as<double>();
        ldj = ldl;
        jrows = rows_l;
      }
    } else if (need_j) {
      SVDJ_CUDA_TRY(cudaMemset2DAsync(dA, lda * sizeof(double), 0, m * sizeof(double), n, stream));
      J = dA;
      ldj = lda;
      jrows = (m >= rows_l) ? rows_l : n;
    }
    SVDJ_CUDA_TRY(cudaEventRecord(e1.e, stream));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));  // a collective inner solve starts from finished data
    return SVDJ_OK;
  };
  const int pre = prepare();

  svdj_opts oi = o;
  oi.precondition = 0;
  const int rc_inner = inner(pre, need_w ? SVDJ_ALLVEC : SVDJ_NOVEC, need_j ? SVDJ_ALLVEC : SVDJ_NOVEC, rows_l, n,
                             d_L.as<double>(), ldl, ds, J, ldj, need_j ? jrows : n, oi, st);
  if (pre != SVDJ_OK) return pre;
  if (rc_inner != SVDJ_OK && rc_inner != SVDJ_NOT_CONVERGED) return rc_inner;
  SVDJ_CUDA_TRY(cudaEventRecord(e2.e, stream));

  // 4. back to the caller's U and V
  void *rsc = static_cast<char *>(d_qrs.p) + 0;  // qr_apply_q reuses the factorisation's scratch (sized for n columns)
  double *W = d_L.as<double>();  // normalised columns of L J: the left singular vectors of L, sigma-sorted
  if (qr_steps >= 2) {
    if (want_v) {  // V = P (Q2 J)
      if (int rc = qr_apply_q(d_B1.as<double>(), ldl, n, n, d_T2.as<double>(), J, ldj, n, rsc, stream, &launches)) return rc;
      if (int rc = qr_scatter_rows(J, ldj, dV, ldv, n, n, d_perm.as<unsigned>(), stream)) return rc;
      launches += 1;
    }
    if (want_u) {  // U = Q1 [W; 0]
      SVDJ_CUDA_TRY(cudaMemset2DAsync(dA, lda * sizeof(double), 0, m * sizeof(double), n, stream));
      SVDJ_CUDA_TRY(cudaMemcpy2DAsync(dA, lda * sizeof(double), W, ldl * sizeof(double), n * sizeof(double), n,
                                      cudaMemcpyDeviceToDevice, stream));
      if (int rc = qr_apply_q(d_Y.as<double>(), ldy, m, n, d_T1.as<double>(), dA, lda, n, rsc, stream, &launches)) return rc;
    }
  } else {
    if (want_v) {  // V = P W
      if (int rc = qr_scatter_rows(W, ldl, dV, ldv, n, n, d_perm.as<unsigned>(), stream)) return rc;
      launches += 1;
    }
    if (want_u) {  // U = Q1 [J; 0], in place
      if (int rc = qr_apply_q(d_Y.as<double>(), ldy, m, n, d_T1.as<double>(), dA, lda, n, rsc, stream, &launches)) return rc;
    }
  }
  SVDJ_CUDA_TRY(cudaEventRecord(e3.e, stream));
  SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
  float ms01 = 0.f, ms23 = 0.f;
  SVDJ_CUDA_TRY(cudaEventElapsedTime(&ms01, e0.e, e1.e));
  SVDJ_CUDA_TRY(cudaEventElapsedTime(&ms23, e2.e, e3.e));
  st.seconds_precond = (ms01 + ms23) * 1e-3;
  st.precondition = qr_steps;
  st.launches += launches;
  return rc_inner;
}

// Which front-end runs.  opts.precondition: 0 none, 1 / 2 QR steps, < 0 the library decides.  SVDJ_PRECOND overrides.
int precondition_steps(const svdj_opts &o, size_t m, size_t n) {
  int want = o.precondition;
  if (const char *env = std::getenv("SVDJ_PRECOND")) want = std::atoi(env);
  const bool possible = m >= n && n >= 64 && qr_supported(m, n) && o.threshold_mode == SVDJ_THRESH_REL;
  if (!possible) return 0;
  if (want < 0) want = kAutoPrecondition;
  return want > 2 ? 2 : want;
}

static const InnerSolve &single_gpu_inner(cudaStream_t stream) {
  static thread_local InnerSolve f;
  f = [stream](int pre, int ju, int jv, size_t mi, size_t ni, double *A, size_t la, double *s, double *V, size_t lv,
               size_t vr, const svdj_opts &oi, svdj_stats &sti) -> int {
    if (pre != SVDJ_OK) return pre;
    return solve_on_device(ju, jv, mi, ni, A, la, s, V, lv, vr, oi, sti, stream);
  };
  return f;
}

// opts.precondition < 0 ("the library decides").  Measured (profiles/r2_precond.md): on dense random inputs the QR
// front-end costs about what the two sweeps it saves are worth (4096^2: 0.389 s with, 0.384 s without), so it is not
// run up front.  What it is indispensable for are strongly graded spectra, which cannot be told from the columns
// beforehand and on which plain Jacobi crawls (8192^2, cond 1e12: not converged after 60 sweeps).  So: sweep the raw
// columns, and if kAutoGiveUpSweeps sweeps -- more than any regular input has needed up to 16384^2 (19) -- have not
// converged, start again from a saved copy of A with two QR steps.  The copy costs one m x n buffer.
static constexpr int kAutoGiveUpSweeps = 26;

static int solve_entry(int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds, double *dV,
                       size_t ldv, size_t vrows, const svdj_opts &o, svdj_stats &st, cudaStream_t stream) {
  if (m == 0 || n == 0) return solve_on_device(jobu, jobv, m, n, dA, lda, ds, dV, ldv, vrows, o, st, stream);
  svdj_opts probe = o;
  probe.precondition = 2;
  const bool qr_possible = precondition_steps(probe, m, n) == 2;
  const bool automatic = o.precondition < 0 && std::getenv("SVDJ_PRECOND") == nullptr;
  if (automatic && qr_possible && o.max_sweeps > kAutoGiveUpSweeps) {
    DevBuf copy;
    if (int rc = copy.alloc(sizeof(double) * m * n)) return rc;
    SVDJ_CUDA_TRY(cudaMemcpy2DAsync(copy.p, m * sizeof(double), dA, lda * sizeof(double), m * sizeof(double), n,
                                    cudaMemcpyDeviceToDevice, stream));
    svdj_opts first = o;
    first.precondition = 0;
    first.max_sweeps = kAutoGiveUpSweeps;
    const int rc1 = solve_on_device(jobu, jobv, m, n, dA, lda, ds, dV, ldv, vrows, first, st, stream);
    if (rc1 != SVDJ_NOT_CONVERGED) return rc1;
    // start again, preconditioned
    const svdj_stats tried = st;
    SVDJ_CUDA_TRY(cudaMemcpy2DAsync(dA, lda * sizeof(double), copy.p, m * sizeof(double), m * sizeof(double), n,
                                    cudaMemcpyDeviceToDevice, stream));
    std::memset(&st, 0, sizeof(st));
    svdj_opts second = o;
    second.precondition = 2;
    const int rc2 = solve_with_qr(jobu, jobv, m, n, dA, lda, ds, dV, ldv, 2, second, st, stream, single_gpu_inner(stream));
    st.sweeps += tried.sweeps;  // the abandoned attempt is part of what this call cost
    st.tested += tried.tested;
    st.applied += tried.applied;
    st.launches += tried.launches;
    st.seconds_sweeps += tried.seconds_sweeps;
    return rc2;
  }
  const int steps = precondition_steps(o, m, n);
  if (steps == 0) return solve_on_device(jobu, jobv, m, n, dA, lda, ds, dV, ldv, vrows, o, st, stream);
  return solve_with_qr(jobu, jobv, m, n, dA, lda, ds, dV, ldv, steps, o, st, stream, single_gpu_inner(stream));
}

}  // namespace svdj

using namespace svdj;

extern "C" {

void svdj_default_opts(svdj_opts *o) {
  if (!o) return;
  o->tol = 0.0;
  o->max_sweeps = 60;
  o->sort_desc = 1;
  o->variant = SVDJ_VARIANT_AUTO;
  o->threshold_mode = SVDJ_THRESH_REL;
  o->block = 0;
  o->device = -1;
  o->precondition = -1;
}

const char *svdj_last_error(void) { return g_error; }

int svdj_release_workspace(void) {
  mg_release_team();
  g_pool.trim();
  return SVDJ_OK;
}

int svdj_version(void) { return 200; }

int svdj_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

size_t svdj_num_steps(size_t n) { return n < 2 ? 0 : 2 * ((n + 1) / 2) - 1; }

size_t svdj_schedule_step(size_t n, size_t step, size_t *pq) {
  if (n < 2 || step >= svdj_num_steps(n) || n > 0x7fffffffu) return 0;
  const size_t pairs = n / 2;
  for (size_t i = 0; i < pairs; ++i) {
    Pair pr = sameh_pair((unsigned)n, (unsigned)step + 1u, (unsigned)i);
    pq[2 * i] = pr.p;
    pq[2 * i + 1] = pr.q;
  }
  return pairs;
}

size_t svdj_block_schedule_step(size_t nb, int vranks, size_t step, size_t *pq) {
  if (nb < 2 || nb > 0x7fffffffu || pq == nullptr || step >= svdj_num_steps(nb)) return 0;
  const bool two_level = vranks >= 2 && nb % (2u * (size_t)vranks) == 0;
  if (vranks >= 2 && !two_level) return 0;
  const size_t pairs = nb / 2;
  for (size_t i = 0; i < pairs; ++i) {
    const Pair pr = two_level ? two_level_pair((unsigned)nb, (unsigned)vranks, (unsigned)step + 1u, (unsigned)i)
                              : sameh_pair((unsigned)nb, (unsigned)step + 1u, (unsigned)i);
    pq[2 * i] = pr.p < pr.q ? pr.p : pr.q;  // the lower block index first, as the kernels take them
    pq[2 * i + 1] = pr.p < pr.q ? pr.q : pr.p;
  }
  return pairs;
}

int svdj_dgesvd_rows_device(int jobu, int jobv, size_t m_local, size_t n, double *dA, size_t lda, double *ds,
                            double *dV, size_t ldv, size_t vrows_local, size_t v_row0, const svdj_opts *opts,
                            svdj_stats *stats, void *stream, svdj_allreduce_fn allreduce, void *ctx) {
  g_error[0] = 0;
  svdj_stats local;
  svdj_stats &st = stats ? *stats : local;
  std::memset(&st, 0, sizeof(st));
  if (m_local == 0 || m_local > 0x7fffffffu) return -3;
  if (n < 128 || n > 0x7fffffffu) return -4;
  if (dA == nullptr) return -5;
  if (lda < m_local) return -6;
  if (ds == nullptr) return -7;
  const bool want_v = (jobv == SVDJ_ALLVEC || jobv == SVDJ_SOMEVEC);
  if (want_v && dV == nullptr) return -8;
  if (want_v && ldv < vrows_local) return -9;
  if (want_v && (vrows_local == 0 || v_row0 + vrows_local > n)) return -10;
  if (opts == nullptr || !(opts->tol > 0.0)) {
    set_error("svdj_dgesvd_rows_device: opts->tol must be set (the relative threshold uses the global row count)");
    return -12;
  }
  if (allreduce == nullptr) return -15;
  const double t0 = now_s();
  svdj_opts o = resolve_opts(opts, m_local, jobu, jobv);
  o.variant = SVDJ_VARIANT_BLOCK;
  o.block = 32;
  o.threshold_mode = SVDJ_THRESH_REL;
  DeviceGuard guard;
  if (int rc = guard.enter(o.device)) return rc;
  RowShard rs{allreduce, ctx, v_row0};
  const int rc = solve_on_device(jobu, jobv, m_local, n, dA, lda, ds, dV, ldv, want_v ? vrows_local : 0, o, st,
                               static_cast<cudaStream_t>(stream), &rs);
  st.seconds_total = now_s() - t0;
  return rc;
}

size_t svdj_step_scratch_doubles(size_t m, size_t n) { return l1_scratch_doubles(m, n); }

size_t svdj_qr_tfactor_doubles(size_t n) { return 1024 * qr_num_panels(n); }

int svdj_qr_factor_device(size_t m, size_t n, double *dY, size_t ld, double *dT, void *stream) {
  clear_error();
  if (m > 0x7fffffffu) return -1;
  if (n > 0x7fffffffu || n > m) return -2;
  if (dY == nullptr && m * n > 0) return -3;
  if (ld < m) return -4;
  if (dT == nullptr && n > 0) return -5;
  if (n == 0) return SVDJ_OK;
  DevBuf scratch;
  if (int rc = scratch.alloc(qr_scratch_bytes(n))) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (int rc = qr_factor(dY, ld, m, n, dT, scratch.p, st, nullptr)) return rc;
  SVDJ_CUDA_TRY(cudaStreamSynchronize(st));  // the scratch goes back to the pool on return
  return SVDJ_OK;
}

int svdj_qr_apply_q_device(size_t m, size_t n, const double *dY, size_t ld, const double *dT, double *dC, size_t ldc,
                           size_t ncols, void *stream) {
  clear_error();
  if (m > 0x7fffffffu) return -1;
  if (n > 0x7fffffffu || n > m) return -2;
  if (dY == nullptr && m * n > 0) return -3;
  if (ld < m) return -4;
  if (dT == nullptr && n > 0) return -5;
  if (dC == nullptr && m * ncols > 0) return -6;
  if (ldc < m) return -7;
  if (ncols > 0x7fffffffu) return -8;
  DevBuf scratch;
  if (int rc = scratch.alloc(qr_scratch_bytes(ncols))) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (int rc = qr_apply_q(dY, ld, m, n, dT, dC, ldc, ncols, scratch.p, st, nullptr)) return rc;
  SVDJ_CUDA_TRY(cudaStreamSynchronize(st));  // the scratch goes back to the pool on return
  return SVDJ_OK;
}

int svdj_dgesvd_device(int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds, double *dV,
                       size_t ldv, const svdj_opts *opts, svdj_stats *stats, void *stream) {
  g_error[0] = 0;
  svdj_stats local;
  svdj_stats &st = stats ? *stats : local;
  std::memset(&st, 0, sizeof(st));
  if (int bad = check_args(jobu, jobv, m, n, dA, lda, ds, dV, ldv)) {
    set_error("svdj_dgesvd_device: argument %d is invalid", -bad);
    return bad;
  }
  const double t0 = now_s();
  svdj_opts o = resolve_opts(opts, m, jobu, jobv);
  DeviceGuard guard;
  if (int rc = guard.enter(o.device)) return rc;
  int rc = solve_entry(jobu, jobv, m, n, dA, lda, ds, dV, ldv, n, o, st, static_cast<cudaStream_t>(stream));
  st.seconds_total = now_s() - t0;
  return rc;
}

int svdj_dgesvd(int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s, double *V, size_t ldv,
                const svdj_opts *opts, svdj_stats *stats) {
  g_error[0] = 0;
  svdj_stats local;
  svdj_stats &st = stats ? *stats : local;
  std::memset(&st, 0, sizeof(st));
  if (int bad = check_args(jobu, jobv, m, n, A, lda, s, V, ldv)) {
    set_error("svdj_dgesvd: argument %d is invalid", -bad);
    return bad;
  }
  const double t0 = now_s();
  svdj_opts o = resolve_opts(opts, m, jobu, jobv);
  DeviceGuard guard;
  if (int rc = guard.enter(o.device)) return rc;
  const size_t mn = std::min(m, n);
  if (m == 0 || n == 0) return SVDJ_OK;

  // Device copies are padded to an even number of rows (a zero row changes neither the Gram triples
  // nor the rotations) and to 128-byte columns, so that the vectorised kernels always apply.
  const bool want_v = (jobv != SVDJ_NOVEC);
  const size_t m_pad = (m + 1) & ~size_t(1);
  const size_t n_pad = (n + 1) & ~size_t(1);
  const size_t lda_d = (m_pad + 15) & ~size_t(15);
  const size_t ldv_d = (n_pad + 15) & ~size_t(15);
  const size_t n_cols = n;

  cudaStream_t stream = nullptr;
  SVDJ_CUDA_TRY(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  struct StreamGuard {
    cudaStream_t s;
    ~StreamGuard() { cudaStreamDestroy(s); }
  } sguard{stream};

  DevBuf dA, dV, ds;
  if (int rc = dA.alloc(sizeof(double) * lda_d * n_cols)) return rc;
  if (int rc = ds.alloc(sizeof(double) * std::max<size_t>(mn, 1))) return rc;
  if (want_v) {
    if (int rc = dV.alloc(sizeof(double) * ldv_d * n_cols)) return rc;
  }

  double t = now_s();
  if (m_pad != m) SVDJ_CUDA_TRY(cudaMemsetAsync(dA.p, 0, sizeof(double) * lda_d * n_cols, stream));
  SVDJ_CUDA_TRY(cudaMemcpy2DAsync(dA.p, lda_d * sizeof(double), A, lda * sizeof(double), m * sizeof(double), n,
                                  cudaMemcpyHostToDevice, stream));
  SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
  st.seconds_h2d = now_s() - t;

  // V is n x n; with n odd the kernels see n_pad rows (the identity init writes the pad row as zero)
  int rc = solve_entry(jobu, jobv, m_pad, n, dA.as<double>(), lda_d, ds.as<double>(), dV.as<double>(), ldv_d, n_pad,
                       o, st, stream);
  if (rc != SVDJ_OK && rc != SVDJ_NOT_CONVERGED) return rc;

  t = now_s();
  SVDJ_CUDA_TRY(cudaMemcpy2DAsync(A, lda * sizeof(double), dA.p, lda_d * sizeof(double), m * sizeof(double), n,
                                  cudaMemcpyDeviceToHost, stream));
  SVDJ_CUDA_TRY(cudaMemcpyAsync(s, ds.p, sizeof(double) * mn, cudaMemcpyDeviceToHost, stream));
  if (want_v) {
    SVDJ_CUDA_TRY(cudaMemcpy2DAsync(V, ldv * sizeof(double), dV.p, ldv_d * sizeof(double), n * sizeof(double), n,
                                    cudaMemcpyDeviceToHost, stream));
  }
  SVDJ_CUDA_TRY(cudaStreamSynchronize(stream));
  st.seconds_d2h = now_s() - t;
  st.seconds_total = now_s() - t0;
  return rc;
}

int svdj_step_device(size_t m, size_t n, double *dA, size_t lda, double *dV, size_t ldv, int rotate_v, size_t step,
                     double tol, int threshold_mode, int path, double *d_scratch, unsigned long long *d_counters,
                     void *stream) {
  g_error[0] = 0;
  if (m == 0 || m > 0x7fffffffu) return -1;
  if (n > 0x7fffffffu) return -2;
  if (dA == nullptr) return -3;
  if (lda < m) return -4;
  if (rotate_v && dV == nullptr) return -5;
  if (rotate_v && ldv < n) return -6;
  if (step >= svdj_num_steps(n)) return -8;
  if (path < SVDJ_PATH_AUTO || path > SVDJ_PATH_SPLIT) return -11;
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned k = (unsigned)step + 1u;
  const bool split = (path == SVDJ_PATH_SPLIT) || (path == SVDJ_PATH_AUTO && l1_uses_split(m, n));
  if (!split) {
    if (l1_uses_split(m, n)) {
      set_error("svdj_step_device: the register path holds at most 8192 rows (m = %zu)", m);
      return SVDJ_ERR_UNSUPPORTED;
    }
    return l1_launch_step(a, k, nullptr, st, nullptr);
  }
  DevBuf own;
  if (d_scratch == nullptr) {
    if (int rc = own.alloc(sizeof(double) * l1_scratch_doubles(m, n))) return rc;
    d_scratch = own.as<double>();
  }
  int rc = l1_launch_split(a, k, d_scratch, st, nullptr);
  if (own.p) SVDJ_CUDA_TRY(cudaStreamSynchronize(st));  // the scratch is freed on return
  return rc;
}

}  // extern "C"
