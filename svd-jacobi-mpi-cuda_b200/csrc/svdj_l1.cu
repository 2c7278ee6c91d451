// svdj_l1.cu -- level-1 sweep step: Gram triple + rotation parameters + Givens update of the A and V
// column pairs, fused so that every column is read once and written once per step.
//
// Replaces, for all n/2 pairs of a round-robin step at once, what the reference does per pair with
// a host loop, 8 cudaMemcpy and 2 launches of a 16-thread kernel (lib/JacobiMethods.cu:450-510;
// kernel :1483-1491).  HBM-bound: 32*(m+n) bytes and 12m+6n flops per pair (0.28 flop/B).
//
// Two code paths:
//   pair-in-registers  one CTA per pair keeps both A columns in registers between the Gram
//                      reduction and the rotation (m <= kRegPathMaxRows): one HBM pass.
//   split              rows of a pair are spread over several CTAs; partial Gram triples go through
//                      a scratch buffer and a second kernel rotates (any m; tall-skinny shapes).
#include "svdj_common.cuh"

namespace svdj {

static constexpr unsigned kRegPathMaxRows = 8192;

// ------------------------------------------------------------------------------------------------
// pair-in-registers kernel
// ------------------------------------------------------------------------------------------------
template <int THREADS, int EPT, bool VEC>
__global__ void __launch_bounds__(THREADS) l1_pair_reg_kernel(const StepArgs a, const unsigned k) {
  __shared__ double red[3 * (THREADS / 32)];
  const Pair pr = sameh_pair(a.n, k, blockIdx.x);
  double *__restrict__ ap = a.A + (size_t)pr.p * a.lda;
  double *__restrict__ aq = a.A + (size_t)pr.q * a.lda;
  const unsigned t = threadIdx.x;
  const unsigned m = a.m;

  double alpha = 0.0, beta = 0.0, gamma = 0.0;
  Rot r;
  double offdiag;
  bool rotate;

  if constexpr (VEC) {
    constexpr int NV = EPT / 2;
    const unsigned m2 = m >> 1;  // VEC requires m even
    double2 xp[NV], xq[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const unsigned i = j * THREADS + t;
      if (i < m2) {
        xp[j] = reinterpret_cast<const double2 *>(ap)[i];
        xq[j] = reinterpret_cast<const double2 *>(aq)[i];
      } else {
        xp[j] = make_double2(0.0, 0.0);
        xq[j] = make_double2(0.0, 0.0);
      }
    }
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      alpha = fma(xp[j].x, xq[j].x, alpha);
      alpha = fma(xp[j].y, xq[j].y, alpha);
      beta = fma(xp[j].x, xp[j].x, beta);
      beta = fma(xp[j].y, xp[j].y, beta);
      gamma = fma(xq[j].x, xq[j].x, gamma);
      gamma = fma(xq[j].y, xq[j].y, gamma);
    }
    block_sum3<THREADS>(alpha, beta, gamma, red);
    rotate = rotation_from_gram(alpha, beta, gamma, a.tol, a.threshold_mode, r, offdiag, a.floor2);
    if (rotate) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const unsigned i = j * THREADS + t;
        if (i < m2) {
          double2 np, nq;
          np.x = r.c * xp[j].x - r.s * xq[j].x;
          np.y = r.c * xp[j].y - r.s * xq[j].y;
          nq.x = r.s * xp[j].x + r.c * xq[j].x;
          nq.y = r.s * xp[j].y + r.c * xq[j].y;
          reinterpret_cast<double2 *>(ap)[i] = np;
          reinterpret_cast<double2 *>(aq)[i] = nq;
        }
      }
    }
  } else {
    double xp[EPT], xq[EPT];
#pragma unroll
    for (int j = 0; j < EPT; ++j) {
      const unsigned i = j * THREADS + t;
      xp[j] = i < m ? ap[i] : 0.0;
      xq[j] = i < m ? aq[i] : 0.0;
    }
#pragma unroll
    for (int j = 0; j < EPT; ++j) {
      alpha = fma(xp[j], xq[j], alpha);
      beta = fma(xp[j], xp[j], beta);
      gamma = fma(xq[j], xq[j], gamma);
    }
    block_sum3<THREADS>(alpha, beta, gamma, red);
    rotate = rotation_from_gram(alpha, beta, gamma, a.tol, a.threshold_mode, r, offdiag, a.floor2);
    if (rotate) {
#pragma unroll
      for (int j = 0; j < EPT; ++j) {
        const unsigned i = j * THREADS + t;
        if (i < m) {
          ap[i] = r.c * xp[j] - r.s * xq[j];
          aq[i] = r.s * xp[j] + r.c * xq[j];
        }
      }
    }
  }

  if (t == 0) publish_pair(a.counters, rotate, offdiag);

  // V columns: pure streaming update with the same (c,s) (lib/JacobiMethods.cu:501-510)
  if (rotate && a.rotate_v) {
    double *__restrict__ vp = a.V + (size_t)pr.p * a.ldv;
    double *__restrict__ vq = a.V + (size_t)pr.q * a.ldv;
    const unsigned n = a.vrows;
    if constexpr (VEC) {
      constexpr int U = 4;
      const unsigned n2 = n >> 1;  // VEC requires n even
      for (unsigned base = 0; base < n2; base += U * THREADS) {
        double2 yp[U], yq[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned i = base + u * THREADS + t;
          if (i < n2) {
            yp[u] = reinterpret_cast<const double2 *>(vp)[i];
            yq[u] = reinterpret_cast<const double2 *>(vq)[i];
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned i = base + u * THREADS + t;
          if (i < n2) {
            double2 np, nq;
            np.x = r.c * yp[u].x - r.s * yq[u].x;
            np.y = r.c * yp[u].y - r.s * yq[u].y;
            nq.x = r.s * yp[u].x + r.c * yq[u].x;
            nq.y = r.s * yp[u].y + r.c * yq[u].y;
            reinterpret_cast<double2 *>(vp)[i] = np;
            reinterpret_cast<double2 *>(vq)[i] = nq;
          }
        }
      }
    } else {
      for (unsigned i = t; i < n; i += THREADS) {
        const double x = vp[i], y = vq[i];
        vp[i] = r.c * x - r.s * y;
        vq[i] = r.s * x + r.c * y;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// split path: partial Gram triples through scratch, then rotation
// ------------------------------------------------------------------------------------------------
static constexpr int kSplitThreads = 256;

// scratch layout: [pair][split][3]
template <bool VEC>
__global__ void __launch_bounds__(kSplitThreads) l1_gram_split_kernel(const StepArgs a, const unsigned k,
                                                                       const unsigned rows_per_split,
                                                                       double *__restrict__ scratch) {
  __shared__ double red[3 * (kSplitThreads / 32)];
  const Pair pr = sameh_pair(a.n, k, blockIdx.x);
  const double *__restrict__ ap = a.A + (size_t)pr.p * a.lda;
  const double *__restrict__ aq = a.A + (size_t)pr.q * a.lda;
  const unsigned r0 = blockIdx.y * rows_per_split;
  const unsigned r1 = min(a.m, r0 + rows_per_split);
  double alpha = 0.0, beta = 0.0, gamma = 0.0;
  if constexpr (VEC) {
    // rows_per_split is even, so r0 is even; m even
    const double2 *p2 = reinterpret_cast<const double2 *>(ap);
    const double2 *q2 = reinterpret_cast<const double2 *>(aq);
    const unsigned i1 = r1 >> 1;
    unsigned i = (r0 >> 1) + threadIdx.x;
    for (; i + 3 * kSplitThreads < i1; i += 4 * kSplitThreads) {
      double2 x[4], y[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        x[u] = p2[i + u * kSplitThreads];
        y[u] = q2[i + u * kSplitThreads];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        alpha = fma(x[u].x, y[u].x, alpha);
        alpha = fma(x[u].y, y[u].y, alpha);
        beta = fma(x[u].x, x[u].x, beta);
        beta = fma(x[u].y, x[u].y, beta);
        gamma = fma(y[u].x, y[u].x, gamma);
        gamma = fma(y[u].y, y[u].y, gamma);
      }
    }
    for (; i < i1; i += kSplitThreads) {
      const double2 x = p2[i], y = q2[i];
      alpha = fma(x.x, y.x, alpha);
      alpha = fma(x.y, y.y, alpha);
      beta = fma(x.x, x.x, beta);
      beta = fma(x.y, x.y, beta);
      gamma = fma(y.x, y.x, gamma);
      gamma = fma(y.y, y.y, gamma);
    }
  } else {
    for (unsigned i = r0 + threadIdx.x; i < r1; i += kSplitThreads) {
      const double x = ap[i], y = aq[i];
      alpha = fma(x, y, alpha);
      beta = fma(x, x, beta);
      gamma = fma(y, y, gamma);
    }
  }
  block_sum3<kSplitThreads>(alpha, beta, gamma, red);
  if (threadIdx.x == 0) {
    double *dst = scratch + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 3;
    dst[0] = alpha;
    dst[1] = beta;
    dst[2] = gamma;
  }
}

template <bool VEC>
__global__ void __launch_bounds__(kSplitThreads) l1_rotate_split_kernel(const StepArgs a, const unsigned k,
                                                                         const unsigned rows_per_split,
                                                                         const unsigned vrows_per_split,
                                                                         const double *__restrict__ scratch) {
  const Pair pr = sameh_pair(a.n, k, blockIdx.x);
  // every CTA of the pair adds the partials in the same order: same (c,s) everywhere
  double alpha = 0.0, beta = 0.0, gamma = 0.0;
  const double *src = scratch + (size_t)blockIdx.x * gridDim.y * 3;
  for (unsigned sidx = 0; sidx < gridDim.y; ++sidx) {
    alpha += src[3 * sidx + 0];
    beta += src[3 * sidx + 1];
    gamma += src[3 * sidx + 2];
  }
  Rot r;
  double offdiag;
  const bool rotate = rotation_from_gram(alpha, beta, gamma, a.tol, a.threshold_mode, r, offdiag, a.floor2);
  if (threadIdx.x == 0 && blockIdx.y == 0) publish_pair(a.counters, rotate, offdiag);
  if (!rotate) return;

  auto rotate_rows = [&](double *__restrict__ xcol, double *__restrict__ ycol, unsigned r0, unsigned r1) {
    if constexpr (VEC) {
      double2 *p2 = reinterpret_cast<double2 *>(xcol);
      double2 *q2 = reinterpret_cast<double2 *>(ycol);
      const unsigned i1 = r1 >> 1;
      unsigned i = (r0 >> 1) + threadIdx.x;
      for (; i + 3 * kSplitThreads < i1; i += 4 * kSplitThreads) {
        double2 x[4], y[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          x[u] = p2[i + u * kSplitThreads];
          y[u] = q2[i + u * kSplitThreads];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          double2 np, nq;
          np.x = r.c * x[u].x - r.s * y[u].x;
          np.y = r.c * x[u].y - r.s * y[u].y;
          nq.x = r.s * x[u].x + r.c * y[u].x;
          nq.y = r.s * x[u].y + r.c * y[u].y;
          p2[i + u * kSplitThreads] = np;
          q2[i + u * kSplitThreads] = nq;
        }
      }
      for (; i < i1; i += kSplitThreads) {
        const double2 x = p2[i], y = q2[i];
        double2 np, nq;
        np.x = r.c * x.x - r.s * y.x;
        np.y = r.c * x.y - r.s * y.y;
        nq.x = r.s * x.x + r.c * y.x;
        nq.y = r.s * x.y + r.c * y.y;
        p2[i] = np;
        q2[i] = nq;
      }
    } else {
      for (unsigned i = r0 + threadIdx.x; i < r1; i += kSplitThreads) {
        const double x = xcol[i], y = ycol[i];
        xcol[i] = r.c * x - r.s * y;
        ycol[i] = r.s * x + r.c * y;
      }
    }
  };

  {
    const unsigned r0 = blockIdx.y * rows_per_split;
    const unsigned r1 = min(a.m, r0 + rows_per_split);
    if (r0 < r1) rotate_rows(a.A + (size_t)pr.p * a.lda, a.A + (size_t)pr.q * a.lda, r0, r1);
  }
  if (a.rotate_v) {
    const unsigned r0 = blockIdx.y * vrows_per_split;
    const unsigned r1 = min(a.vrows, r0 + vrows_per_split);
    if (r0 < r1) rotate_rows(a.V + (size_t)pr.p * a.ldv, a.V + (size_t)pr.q * a.ldv, r0, r1);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static unsigned split_count(size_t m, size_t n) {
  const size_t pairs = n / 2;
  if (pairs == 0) return 1;
  // aim at >= 4 CTAs per SM over 148 SMs, but keep >= 2048 rows per CTA
  size_t want = (4 * 148 + pairs - 1) / pairs;
  size_t cap = (m + 2047) / 2048;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  if (want > 65535) want = 65535;
  return (unsigned)want;
}

bool l1_uses_split(size_t m, size_t n) {
  (void)n;
  return m > kRegPathMaxRows;
}

size_t l1_scratch_doubles(size_t m, size_t n) { return (n / 2) * (size_t)split_count(m, n) * 3; }

static bool vec_ok(const StepArgs &a) {
  const bool a_ok = (a.m % 2 == 0) && (a.lda % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.A) & 15) == 0);
  const bool v_ok = !a.rotate_v ||
                    ((a.vrows % 2 == 0) && (a.ldv % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.V) & 15) == 0));
  return a_ok && v_ok;
}

template <int THREADS, int EPT>
static void launch_reg(const StepArgs &a, unsigned k, unsigned pairs, bool vec, cudaStream_t st) {
  if (vec) {
    l1_pair_reg_kernel<THREADS, EPT, true><<<pairs, THREADS, 0, st>>>(a, k);
  } else {
    l1_pair_reg_kernel<THREADS, EPT, false><<<pairs, THREADS, 0, st>>>(a, k);
  }
}

int l1_launch_split(const StepArgs &a, unsigned k, double *scratch, cudaStream_t st,
                    unsigned long long *launches) {
  const unsigned pairs = a.n / 2;
  if (pairs == 0) return SVDJ_OK;
  if (scratch == nullptr) {
    set_error("l1 split path needs a scratch buffer");
    return SVDJ_ERR_ALLOC;
  }
  const bool vec = vec_ok(a);
  const unsigned splits = split_count(a.m, a.n);
  unsigned rps = (a.m + splits - 1) / splits;
  rps = (rps + 1u) & ~1u;
  unsigned vrps = (a.vrows + splits - 1) / splits;
  vrps = (vrps + 1u) & ~1u;
  dim3 grid(pairs, splits);
  if (vec) {
    l1_gram_split_kernel<true><<<grid, kSplitThreads, 0, st>>>(a, k, rps, scratch);
    l1_rotate_split_kernel<true><<<grid, kSplitThreads, 0, st>>>(a, k, rps, vrps, scratch);
  } else {
    l1_gram_split_kernel<false><<<grid, kSplitThreads, 0, st>>>(a, k, rps, scratch);
    l1_rotate_split_kernel<false><<<grid, kSplitThreads, 0, st>>>(a, k, rps, vrps, scratch);
  }
  if (launches) *launches += 2;
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int l1_launch_step(const StepArgs &a, unsigned k, double *scratch, cudaStream_t st,
                   unsigned long long *launches) {
  const unsigned pairs = a.n / 2;
  if (pairs == 0) return SVDJ_OK;
  if (l1_uses_split(a.m, a.n)) return l1_launch_split(a, k, scratch, st, launches);
  const bool vec = vec_ok(a);
  const unsigned m = a.m;
  if (m <= 256) launch_reg<32, 8>(a, k, pairs, vec, st);
  else if (m <= 512) launch_reg<64, 8>(a, k, pairs, vec, st);
  else if (m <= 1024) launch_reg<128, 8>(a, k, pairs, vec, st);
  else if (m <= 2048) launch_reg<128, 16>(a, k, pairs, vec, st);
  else if (m <= 4096) launch_reg<256, 16>(a, k, pairs, vec, st);
  else launch_reg<512, 16>(a, k, pairs, vec, st);
  if (launches) *launches += 1;
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

}  // namespace svdj
