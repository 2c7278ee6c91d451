/*
 * svdj_oracle.c -- CPU restatement of the reference's one-sided (Hestenes) Jacobi SVD.
 *
 * TEST INFRASTRUCTURE ONLY (see svdj_oracle.h).  Parity status: PINNED against oracle/_ref
 * (the reference's own solver built for the host) on tests/golden/.
 *
 * Each function cites the reference lines it restates (paths under /root/reference).
 * Two deliberate additions over the reference, both switchable so the reference's exact
 * behaviour stays reachable (threshold_mode = ABS, max_sweeps = 1, sort_desc = 0):
 *   - an outer sweep loop with a relative convergence test built from the value the reference
 *     computes and throws away at lib/JacobiMethods.cu:462;
 *   - an optional descending sort of sigma (promised by the doc block at lib/JacobiMethods.cu:22-23,
 *     never done).
 */
#include "svdj_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* lib/global.cuh:14 */
#define IDX_C(i, j, ld) (((j) * (ld)) + (i))

static double wall(void) {
#ifdef _OPENMP
  return omp_get_wtime();
#else
  return 0.0;
#endif
}

size_t oracle_num_steps(size_t n) {
  size_t m_ord = (n + 1) / 2; /* lib/JacobiMethods.cu:227 (m_ordering) */
  return n < 2 ? 0 : 2 * m_ord - 1;
}

/* Partner of q (1-based) at step k (1-based); lib/JacobiMethods.cu:279-290 and :724-735. */
static size_t partner_of(size_t n, size_t m_ord, size_t k, size_t q) {
  size_t p = 0;
  if (k < m_ord) {
    if (m_ord - k + 1 <= q && q <= 2 * m_ord - 2 * k) {
      p = (2 * m_ord - 2 * k + 1) - q;
    } else if (2 * m_ord - 2 * k < q && q <= 2 * m_ord - k - 1) {
      p = (4 * m_ord - 2 * k) - q;
    } else if (2 * m_ord - k - 1 < q) {
      p = n;
    }
  } else {
    if (q < 2 * m_ord - k + 1) {
      p = n;
    } else if (2 * m_ord - k + 1 <= q && q <= 4 * m_ord - 2 * k - 1) {
      p = (4 * m_ord - 2 * k) - q;
    } else if (4 * m_ord - 2 * k - 1 < q) {
      p = (6 * m_ord - 2 * k - 1) - q;
    }
  }
  return p;
}

size_t oracle_schedule_step(size_t n, size_t step, size_t *pq) {
  size_t m_ord = (n + 1) / 2;
  size_t k = step + 1;
  size_t q_lo, q_hi; /* inclusive range of the reference's q loop */
  size_t cnt = 0;
  if (n < 2 || k >= 2 * m_ord) return 0;
  if (k < m_ord) { /* lib/JacobiMethods.cu:278: q = m_ord-k+1 .. n-k */
    q_lo = m_ord - k + 1;
    q_hi = n - k;
  } else { /* lib/JacobiMethods.cu:723: q = 4*m_ord-n-k .. 3*m_ord-k-1 */
    q_lo = 4 * m_ord - n - k;
    q_hi = 3 * m_ord - k - 1;
  }
  for (size_t q = q_lo; q <= q_hi; ++q) {
    size_t p = partner_of(n, m_ord, k, q);
    pq[2 * cnt] = p - 1; /* "Translate to (0,0)", lib/JacobiMethods.cu:293-294 */
    pq[2 * cnt + 1] = q - 1;
    ++cnt;
  }
  return cnt;
}

/* One pair: Gram triple (lib/JacobiMethods.cu:450-459), rotation parameters (:464-477),
 * rotation of the A and V columns (the kernel at :1483-1491, run on the host).
 * Returns 1 if rotated.  *offdiag = |alpha|/sqrt(beta*gamma) (:462). */
static int pair_update(size_t m, size_t n, double *A, size_t lda, double *V, size_t ldv, int rotate_v,
                       size_t p, size_t q, const oracle_opts *o, double *offdiag) {
  double *ap = A + p * lda, *aq = A + q * lda;
  double alpha = 0.0, beta = 0.0, gamma = 0.0;
  for (size_t i = 0; i < m; ++i) {
    double tp = ap[i], tq = aq[i];
    alpha += tp * tq;
    beta += tp * tp;
    gamma += tq * tq;
  }
  double denom = sqrt(beta * gamma);
  *offdiag = denom > 0.0 ? fabs(alpha) / denom : 0.0;

  int rotate;
  if (o->threshold_mode == ORACLE_THRESH_ABS) {
    rotate = fabs(alpha) > o->threshold;
  } else {
    rotate = fabs(alpha) > o->threshold * denom;
  }
  if (!rotate) return 0;

  double tau = (gamma - beta) / (2.0 * alpha);
  double t;
  if (tau >= 0) {
    t = 1.0 / (tau + sqrt(1 + tau * tau));
  } else {
    t = 1.0 / (tau - sqrt(1 + tau * tau));
  }
  double c = 1.0 / sqrt(1 + t * t);
  double s = t * c;

  if (o->fma_rotation) {
    /* what nvcc's default -fmad=true makes of the kernel body */
    for (size_t i = 0; i < m; ++i) {
      double x = ap[i], y = aq[i];
      ap[i] = fma(c, x, -(s * y));
      aq[i] = fma(s, x, c * y);
    }
  } else {
    for (size_t i = 0; i < m; ++i) {
      double x = ap[i];
      ap[i] = c * x - s * aq[i];
      aq[i] = s * x + c * aq[i];
    }
  }
  if (rotate_v) { /* the reference rotates V whatever jobv says (guard commented out at :498) */
    double *vp = V + p * ldv, *vq = V + q * ldv;
    if (o->fma_rotation) {
      for (size_t i = 0; i < n; ++i) {
        double x = vp[i], y = vq[i];
        vp[i] = fma(c, x, -(s * y));
        vq[i] = fma(s, x, c * y);
      }
    } else {
      for (size_t i = 0; i < n; ++i) {
        double x = vp[i];
        vp[i] = c * x - s * vq[i];
        vq[i] = s * x + c * vq[i];
      }
    }
  }
  return 1;
}

void oracle_step(size_t m, size_t n, double *A, size_t lda, double *V, size_t ldv, int rotate_v,
                 size_t step, const oracle_opts *o, oracle_stats *st) {
  size_t npairs = n / 2;
  size_t *pq = (size_t *)malloc(sizeof(size_t) * 2 * (npairs ? npairs : 1));
  npairs = oracle_schedule_step(n, step, pq);
  unsigned long long applied = 0;
  double maxoff = 0.0;
  int nt = o->nthreads;
#ifdef _OPENMP
  if (nt <= 0) nt = omp_get_max_threads();
#else
  nt = 1;
#endif
  /* pairs of a step are disjoint: the reference runs them under `omp parallel for` (:439) */
#pragma omp parallel for num_threads(nt) reduction(+ : applied) reduction(max : maxoff) schedule(static)
  for (long long idx = 0; idx < (long long)npairs; ++idx) {
    double off;
    applied += (unsigned long long)pair_update(m, n, A, lda, V, ldv, rotate_v, pq[2 * idx], pq[2 * idx + 1], o, &off);
    if (off > maxoff) maxoff = off;
  }
  free(pq);
  if (st) {
    st->tested += npairs;
    st->applied += applied;
    if (maxoff > st->last_offdiag) st->last_offdiag = maxoff;
  }
}

int oracle_dgesvd(int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s, double *V,
                  size_t ldv, const oracle_opts *o, oracle_stats *st) {
  oracle_stats local;
  if (!st) st = &local;
  memset(st, 0, sizeof(*st));
  if (jobu < 0 || jobu > 2) return -1;
  if (jobv < 0 || jobv > 2) return -2;
  if (!A) return -5;
  if (lda < m) return -6;
  if (!s) return -7;
  if (jobv != ORACLE_NOVEC && !V) return -8;
  if (jobv != ORACLE_NOVEC && ldv < n) return -9;
  if (!o) return -10;
  size_t mn = m < n ? m : n;

  /* V = I (lib/JacobiMethods.cu:213-226 sets the diagonal and relies on the caller's zero fill) */
  int rotate_v = (jobv != ORACLE_NOVEC);
  if (rotate_v) {
    for (size_t j = 0; j < n; ++j) {
      memset(V + j * ldv, 0, sizeof(double) * n);
      V[IDX_C(j, j, ldv)] = 1.0;
    }
  }

  size_t steps = oracle_num_steps(n);
  double t0 = wall();
  int converged = 0;
  for (int sweep = 0; sweep < o->max_sweeps; ++sweep) { /* lib/JacobiMethods.cu:248 */
    unsigned long long before = st->applied;
    st->last_offdiag = 0.0;
    for (size_t step = 0; step < steps; ++step) { /* :252 and :697 */
      oracle_step(m, n, A, lda, V, ldv, rotate_v, step, o, st);
    }
    st->sweeps = sweep + 1;
    if (o->threshold_mode == ORACLE_THRESH_REL && st->applied == before) {
      converged = 1;
      break;
    }
  }
  st->seconds_sweeps = wall() - t0;
  st->converged = converged;

  /* Sigma: column norms (lib/JacobiMethods.cu:1149-1155) */
  for (size_t k = 0; k < n; ++k) {
    double acc = 0.0;
    for (size_t i = 0; i < m; ++i) acc += A[IDX_C(i, k, lda)] * A[IDX_C(i, k, lda)];
    /* sigma of the columns beyond min(m,n) is not part of the contract; keep it in a scratch slot */
    if (k < mn) s[k] = sqrt(acc);
  }

  if (o->sort_desc) {
    /* selection by full column norms so that m < n inputs push their null columns to the end */
    double *nrm = (double *)malloc(sizeof(double) * n);
    size_t *perm = (size_t *)malloc(sizeof(size_t) * n);
    for (size_t k = 0; k < n; ++k) {
      double acc = 0.0;
      for (size_t i = 0; i < m; ++i) acc += A[IDX_C(i, k, lda)] * A[IDX_C(i, k, lda)];
      nrm[k] = sqrt(acc);
      perm[k] = k;
    }
    /* stable insertion sort, descending */
    for (size_t i = 1; i < n; ++i) {
      size_t pi = perm[i];
      size_t j = i;
      while (j > 0 && nrm[perm[j - 1]] < nrm[pi]) {
        perm[j] = perm[j - 1];
        --j;
      }
      perm[j] = pi;
    }
    double *tmp = (double *)malloc(sizeof(double) * (m > n ? m : n) * n);
    for (size_t k = 0; k < n; ++k) memcpy(tmp + k * m, A + perm[k] * lda, sizeof(double) * m);
    for (size_t k = 0; k < n; ++k) memcpy(A + k * lda, tmp + k * m, sizeof(double) * m);
    if (rotate_v) {
      for (size_t k = 0; k < n; ++k) memcpy(tmp + k * n, V + perm[k] * ldv, sizeof(double) * n);
      for (size_t k = 0; k < n; ++k) memcpy(V + k * ldv, tmp + k * n, sizeof(double) * n);
    }
    for (size_t k = 0; k < mn; ++k) s[k] = nrm[perm[k]];
    free(tmp);
    free(perm);
    free(nrm);
  }

  /* U = A / sigma (lib/JacobiMethods.cu:1158-1164; SomeVec as in :1441-1447).  The reference has no
   * guard for sigma = 0; here such a column is left as it is (zero). */
  if (jobu == ORACLE_ALLVEC || jobu == ORACLE_SOMEVEC) {
    for (size_t k = 0; k < mn; ++k) {
      if (s[k] > 0.0) {
        for (size_t i = 0; i < m; ++i) A[IDX_C(i, k, lda)] = A[IDX_C(i, k, lda)] / s[k];
      }
    }
  }
  if (o->threshold_mode == ORACLE_THRESH_REL && !converged) return 1;
  return 0;
}

double oracle_residual_fro(size_t m, size_t n, const double *A0, const double *U, const double *s,
                           const double *V) {
  /* main.cu:1641-1665: value = sum_k U[i,k] * s[k] * V[j,k] */
  double fro = 0.0;
#pragma omp parallel for reduction(+ : fro) schedule(static)
  for (long long i = 0; i < (long long)m; ++i) {
    for (size_t j = 0; j < n; ++j) {
      double v = 0.0;
      for (size_t k = 0; k < n; ++k) v += U[IDX_C(i, k, m)] * s[k] * V[IDX_C(j, k, n)];
      double d = A0[IDX_C(i, j, m)] - v;
      fro += d * d;
    }
  }
  return sqrt(fro);
}
