/*
 * svdj_oracle.h -- CPU restatement of the reference's one-sided Jacobi SVD.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and there only as the checker or as the timed CPU baseline.
 *
 * Parity status: PINNED.  The restatement is checked bit-for-bit against the reference's own
 * solver (main.cu:440-1423, built for the host by oracle/Makefile into oracle/_ref/) on the
 * committed fixtures in tests/golden/ (tests/test_oracle.py), and against the reference's
 * genuine CUDA path on the GPU box (tests/test_reference_gpu.py).
 *
 * All file:line citations are into /root/reference.
 */
#ifndef SVDJ_ORACLE_H_
#define SVDJ_ORACLE_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* lib/JacobiMethods.cuh:25-29 */
enum { ORACLE_ALLVEC = 0, ORACLE_SOMEVEC = 1, ORACLE_NOVEC = 2 };

/* How a pair decides whether to rotate. */
enum {
  ORACLE_THRESH_ABS = 0, /* |alpha| > thr, thr absolute: the reference's test (lib/JacobiMethods.cu:466, 1271) */
  ORACLE_THRESH_REL = 1  /* |alpha| > tol*sqrt(beta*gamma): the test built from the unused value at :462 */
};

typedef struct {
  int threshold_mode; /* ORACLE_THRESH_* */
  double threshold;   /* thr (ABS: reference uses 1e-16 in the MPI solver, 1e-20 in cuda_dgesvd_kernel) or tol (REL) */
  int max_sweeps;     /* reference: exactly 1 (lib/JacobiMethods.cu:234) */
  int sort_desc;      /* 0 = reference behaviour (unsorted); 1 = sigma descending, U and V columns permuted alike */
  int nthreads;       /* OpenMP threads over the pairs of a step (lib/JacobiMethods.cu:439); <=0: library default */
  int fma_rotation;   /* 1: rotate with fma() the way nvcc contracts the device kernel (lib/JacobiMethods.cu:1483-1491) */
} oracle_opts;

typedef struct {
  int sweeps;                 /* sweeps executed */
  int converged;              /* REL mode: 1 if the last sweep applied no rotation */
  unsigned long long tested;  /* pairs whose Gram triple was computed */
  unsigned long long applied; /* pairs that were rotated */
  double last_offdiag;        /* max over the last sweep of |alpha|/sqrt(beta*gamma) (lib/JacobiMethods.cu:462) */
  double seconds_sweeps;      /* wall time inside the sweep loop */
} oracle_stats;

/* Number of round-robin steps of one sweep: 2*((n+1)/2) - 1 (lib/JacobiMethods.cu:252, 697). */
size_t oracle_num_steps(size_t n);

/* Pairs (p,q), 0-based, of step `step` (0-based; k = step+1) in the order the reference's q loop
 * emits them (lib/JacobiMethods.cu:279-290 for k < m_ord, :724-735 for k >= m_ord).
 * pq must hold 2*(n/2) entries; returns the number of pairs (always n/2). */
size_t oracle_schedule_step(size_t n, size_t step, size_t *pq);

/* One-sided Jacobi SVD with the reference's data contract (lib/JacobiMethods.cuh:44-62):
 * A m x n column-major lda (out: U in place), s[min(m,n)] (out: sigma), V n x n column-major ldv (out: V, not V^T).
 * Unlike the reference, s and V need not be zeroed by the caller: they are overwritten.
 * Returns 0, or 1 if REL mode did not converge within max_sweeps, or <0 for a bad argument. */
int oracle_dgesvd(int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s,
                  double *V, size_t ldv, const oracle_opts *opts, oracle_stats *stats);

/* Exactly one round-robin step on raw buffers (no V init, no finalisation): used to time a bounded
 * sample of a big sweep and to check single steps of the CUDA path. */
void oracle_step(size_t m, size_t n, double *A, size_t lda, double *V, size_t ldv, int rotate_v,
                 size_t step, const oracle_opts *opts, oracle_stats *stats);

/* The reference driver's input: zero matrix whose upper triangle is filled row by row from
 * std::default_random_engine(seed) through uniform_real_distribution(0,1) (main.cu:1445, 1559-1567).
 * Implemented in oracle_input.cpp on the same libstdc++ the reference would use. */
void oracle_fill_reference_input(size_t n, unsigned seed, double *A);

/* ||A0 - U diag(s) V^T||_F the way the driver checks it (main.cu:1641-1665). */
double oracle_residual_fro(size_t m, size_t n, const double *A0, const double *U, const double *s,
                           const double *V);

#ifdef __cplusplus
}
#endif
#endif
