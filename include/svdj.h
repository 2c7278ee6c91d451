/*
 * svdj.h -- C ABI of the B200-native one-sided Jacobi SVD (libsvdj.so).
 *
 * This is the drop-in boundary for the solver that the reference driver times
 * (/root/reference/main.cu:1587 -> omp_mpi_cuda_dgesvd_local_matrices, main.cu:440-1423;
 * single-process twin cuda_dgesvd_kernel, main.cu:163-437; declarations
 * lib/JacobiMethods.cuh:44-62).  Plain pointers and sizes only: no C++ or torch types.
 * The C++ shims with the reference's exact signatures live in svdj_reference_api.hpp.
 *
 * Data contract (same as the reference, lib/JacobiMethods.cuh:31-43, lib/global.cuh:14):
 *   A   m x n, column-major, leading dimension lda >= m.  In: the matrix.  Out: U in place
 *       (A[:,k] / sigma_k), or the un-normalised A*V when jobu == SVDJ_NOVEC.
 *   s   min(m,n) doubles.  Out: singular values.
 *   V   n x n, column-major, ldv >= n.  Out: V (not V^T), so that A = sum_k U[:,k] s[k] V[:,k]^T
 *       (main.cu:1645-1647).  Untouched when jobv == SVDJ_NOVEC.
 * Differences from the reference, all on the safe side: s and V need not be zero-filled by the
 * caller (the reference requires it: lib/JacobiMethods.cu:1151, 213-226); sweeps repeat until
 * convergence instead of exactly one (lib/JacobiMethods.cu:234); sigma comes back sorted in
 * descending order with U and V permuted alike, as the reference's own doc block promises
 * (lib/JacobiMethods.cu:22-23) -- set opts.sort_desc = 0 for the reference's raw order;
 * every CUDA call is checked and failures come back as a status code (the reference ignores them).
 *
 * There is no CPU fallback: without a CUDA device every compute entry point returns
 * SVDJ_ERR_CUDA.
 */
#ifndef SVDJ_H_
#define SVDJ_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* lib/JacobiMethods.cuh:25-29 (enum SVD_OPTIONS) -- same numeric values */
enum { SVDJ_ALLVEC = 0, SVDJ_SOMEVEC = 1, SVDJ_NOVEC = 2 };

/* Which sweep implementation runs. */
enum {
  SVDJ_VARIANT_AUTO = 0,
  SVDJ_VARIANT_LEVEL1 = 1, /* fused Gram + Givens rotation per column pair: HBM-bound (main.cu:450-510 in one pass) */
  SVDJ_VARIANT_BLOCK = 2   /* block Jacobi: panel Gram, on-chip eigen-solve, panel update as FP64 tensor-core tiles */
};

/* How a pair decides whether it is rotated. */
enum {
  SVDJ_THRESH_REL = 0, /* |a_p.a_q| > tol * |a_p| |a_q|   (from the value computed at lib/JacobiMethods.cu:462) */
  SVDJ_THRESH_ABS = 1  /* |a_p.a_q| > tol                 (the reference's test, lib/JacobiMethods.cu:466, 1271) */
};

/* Status codes.  0 = success; negative = argument number (1-based, as in LAPACK xerbla and as the
 * reference's doc block describes, lib/JacobiMethods.cu:180-187); positive = run-time condition. */
enum {
  SVDJ_OK = 0,
  SVDJ_NOT_CONVERGED = 1, /* max_sweeps reached; outputs hold the current iterate */
  SVDJ_ERR_CUDA = 2,      /* a CUDA call failed or no device; see svdj_last_error() */
  SVDJ_ERR_ALLOC = 3,     /* device or host allocation failed */
  SVDJ_ERR_UNSUPPORTED = 4
};

typedef struct svdj_opts {
  double tol;         /* <= 0: default.  REL mode: LAPACK dgesvj's choice -- sqrt(m) * 2^-53 when U or V is wanted (the
                         columns then come out orthogonal to 10 n eps), m * 2^-53 for NoVec/NoVec.  ABS mode: 1e-16 */
  int max_sweeps;     /* <= 0: default 60.  The reference runs exactly 1 (lib/JacobiMethods.cu:234) */
  int sort_desc;      /* 1: sigma descending, U/V columns permuted alike (default); 0: reference's raw column order */
  int variant;        /* SVDJ_VARIANT_* */
  int threshold_mode; /* SVDJ_THRESH_* */
  int block;          /* column-block width of SVDJ_VARIANT_BLOCK: 16 or 32; 0 = library decides (32 from n = 4096) */
  int device;         /* CUDA device ordinal; < 0 = the calling thread's current device */
  int precondition;   /* QR front-end before the sweeps (m >= n >= 64): 0 none; 1: de Rijk column sort, A P = Q R by
                         blocked Householder, Jacobi on R^T; 2: R^T factored once more (strongly graded spectra);
                         < 0 (default): plain sweeps first; if 26 of them have not converged (a strongly graded
                         spectrum), the solve starts again from a saved copy of A with two QR steps.  All on the
                         device, this library's own kernels */
} svdj_opts;

typedef struct svdj_stats {
  int sweeps;                    /* sweeps executed */
  int converged;                 /* 1 if the last sweep rotated nothing (REL mode) */
  int variant;                   /* variant that ran */
  int reserved;                  /* block path: the column-block width that ran (16 or 32) */
  unsigned long long tested;     /* column pairs whose Gram triple was evaluated */
  unsigned long long applied;    /* column pairs rotated */
  unsigned long long launches;   /* kernels launched by this call */
  double last_offdiag;           /* max over the last sweep of |a_p.a_q| / (|a_p||a_q|) */
  double seconds_sweeps;         /* device time of the sweep loop (CUDA events) */
  double seconds_finalize;       /* device time of sigma / U / sort */
  double seconds_h2d;            /* host->device copies (svdj_dgesvd only) */
  double seconds_d2h;            /* device->host copies (svdj_dgesvd only) */
  double seconds_total;          /* wall time of the call */
  double step_kernel_ms;         /* mean device time per launch of the dominant sweep kernel: the fused level-1
                                    step, or the block path's update kernel (probed with events every 8th step) */
  unsigned long long step_kernel_launches; /* launches of that kernel (block path: the probed ones) */
  unsigned long long probe_block_pairs;    /* block path: block pairs updated by the probed launches */
  double seconds_precond;        /* device time of the QR front-end and of applying Q afterwards (0 without it) */
  int precondition;              /* QR steps that ran (0, 1 or 2) */
  int reserved1;
} svdj_stats;

/* Fills *opts with the defaults described above. */
void svdj_default_opts(svdj_opts *opts);

/*
 * Host-pointer entry point: replaces the body of omp_mpi_cuda_dgesvd_local_matrices
 * (main.cu:440-1423) and of cuda_dgesvd_kernel (main.cu:163-437).  Copies A to the device, runs the
 * sweeps there (A and V stay resident in HBM, unlike the per-pair staging at main.cu:727-757),
 * finalises (main.cu:1395-1411) and copies U, s, V back.  opts may be NULL (defaults); stats may be NULL.
 */
int svdj_dgesvd(int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s, double *V,
                size_t ldv, const svdj_opts *opts, svdj_stats *stats);

/*
 * Device-pointer entry point: same contract with A, s, V in device memory of the current (or
 * opts->device) device; nothing crosses PCIe except the per-sweep convergence word.  `stream` is a
 * cudaStream_t (NULL = the legacy default stream).  The reference has no device-resident
 * counterpart (its matrices never live on the GPU: main.cu:727-757); this is the same solver minus
 * the copies.
 */
int svdj_dgesvd_device(int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds,
                       double *dV, size_t ldv, const svdj_opts *opts, svdj_stats *stats, void *stream);

/*
 * One round-robin step (all n/2 disjoint pairs of step `step`, 0-based, of the reference's ordering:
 * main.cu:526-545 for the first half, :971-990 for the second) on device buffers: Gram triple
 * (main.cu:696-706), rotation parameters (:708-724) and the rotation of the A and V columns
 * (jacobi_rotation, main.cu:139-147) fused in one pass.  No V initialisation, no finalisation.
 * path: SVDJ_PATH_AUTO, or force SVDJ_PATH_REG (pair kept in registers, m <= 8192) or SVDJ_PATH_SPLIT
 * (rows of a pair spread over several CTAs; partial Gram triples pass through d_scratch).
 * d_scratch: device buffer of at least svdj_step_scratch_doubles(m, n) doubles; only the split path
 * uses it; NULL makes the call allocate and free one itself.
 * d_counters (device, 2 x unsigned long long, may be NULL): [0] += pairs rotated,
 * [1] = max([1], bit pattern of the largest off-diagonal cosine seen).
 * Used by the tests and by multi-GPU drivers.
 */
enum { SVDJ_PATH_AUTO = 0, SVDJ_PATH_REG = 1, SVDJ_PATH_SPLIT = 2 };
int svdj_step_device(size_t m, size_t n, double *dA, size_t lda, double *dV, size_t ldv,
                     int rotate_v, size_t step, double tol, int threshold_mode, int path,
                     double *d_scratch, unsigned long long *d_counters, void *stream);

/* Doubles of scratch the split path needs for an m x n problem. */
size_t svdj_step_scratch_doubles(size_t m, size_t n);

/*
 * The same round-robin step as svdj_step_device(path = SVDJ_PATH_SPLIT), for callers that must not
 * allocate device memory.  It has no counterpart in the reference (which allocates per call,
 * main.cu:484-494); it exists for the callers of THIS library who cannot call cudaMalloc: drivers that
 * keep several streams busy and would have all of them stalled by the device-wide synchronisation
 * inside cudaMalloc/cudaFree, processes whose device memory was handed to a single arena at start-up so
 * that the driver has nothing left to give, and callers with a fixed latency budget per step.
 * The partial Gram triples go through an array that is part of the library image on every GPU that
 * loads libsvdj.so (SVDJ_NOALLOC_MAX_SCRATCH_DOUBLES doubles = 768 KiB, held until the process exits).
 * The call makes no cudaMalloc/cudaFree and no other allocating call.
 *   - Problems with svdj_step_scratch_doubles(m, n) > SVDJ_NOALLOC_MAX_SCRATCH_DOUBLES are refused
 *     with SVDJ_ERR_UNSUPPORTED and a message in svdj_last_error().
 *   - Every call on a GPU uses the same array: calls whose kernels may overlap on that GPU (two streams,
 *     two threads) race.  The array keeps the previous call's contents; the step overwrites every cell
 *     it later reads, so nothing has to be cleared.
 *   - The first call on each GPU (made with that GPU current) resolves the array's device address and
 *     caches it; make that first call before any stream capture starts.
 * A, V, counters and stream are as for svdj_step_device.
 */
#define SVDJ_NOALLOC_MAX_SCRATCH_DOUBLES 98304
int svdj_step_device_noalloc(size_t m, size_t n, double *dA, size_t lda, double *dV, size_t ldv,
                             int rotate_v, size_t step, double tol, int threshold_mode,
                             unsigned long long *d_counters, void *stream);

/*
 * Per-rank compute of the multi-GPU solver.  These replace what a non-root rank does between the
 * reference's MPI_Recv and MPI_Send (main.cu:769-850: Gram triples, rotation parameters and rotations
 * of the pairs whose columns it was sent) -- except that the rank keeps its columns: two column blocks of
 * w columns each ("left" and "right": A-blocks m x w, V-blocks vrows x w, column-major, in separate
 * allocations so that one can travel to a ring neighbour while the other stays).
 *
 * svdj_ring_ws_create / _destroy: scratch for 2w local columns (w a multiple of 16).
 * svdj_ring_phase_device: SVDJ_PHASE_ALL orthogonalises every pair of the 2w local columns (one local
 *   block-Jacobi sweep); SVDJ_PHASE_CROSS only the pairs with one column in each buffer (what is left to do
 *   when both blocks have already been swept internally in this sweep).  d_counters (2 x unsigned long
 *   long on the device, may be NULL) accumulates [0] += rotations, [1] = max(bits of off-diagonal cosine).
 *   inner_sweeps <= 0 selects the default (2).
 * svdj_ring_finalize_device: sigma = column norms of the local columns and, if scale_u, U = A / sigma in
 *   place (main.cu:1395-1411, on the device and per rank).
 */
enum { SVDJ_PHASE_ALL = 0, SVDJ_PHASE_CROSS = 1 };
int svdj_ring_ws_create(size_t m, size_t ncols_local, void **ws_out);
int svdj_ring_ws_destroy(void *ws);
/* Columns whose norm is <= floor_norm are numerically zero and are never rotated (rank-deficient and m < n
 * inputs: without it their noise keeps every sweep rotating).  The driver passes tol * (largest column norm of the
 * whole input), the same floor the single-GPU solver uses; 0 (the default) switches it off.  The reference has no
 * such guard (it runs one sweep whatever happens, main.cu:482). */
int svdj_ring_ws_set_floor(void *ws, double floor_norm);
int svdj_ring_phase_device(void *ws, size_t m, size_t vrows, size_t w, double *dA_left, double *dA_right,
                           size_t lda, double *dV_left, double *dV_right, size_t ldv, int rotate_v, int phase,
                           double tol, int inner_sweeps, unsigned long long *d_counters, void *stream);
int svdj_ring_finalize_device(void *ws, size_t m, size_t ncols, double *dA, size_t lda, double *d_sigma,
                              int scale_u, void *stream);

/*
 * The block exchange fused into the update kernel.  In the reference the columns a rank has rotated
 * go back to rank 0 by MPI_Send from host memory and are re-packed and re-sent before the next step
 * (main.cu:852-942).  svdj_ring_phase_send_device is svdj_ring_phase_device whose LAST step stores the
 * updated columns of the left and/or right block straight into dst* -- normally buffers of the neighbour
 * GPUs mapped through CUDA IPC, reached by peer stores over NVLink -- instead of local memory (NULL = keep
 * that block local).  That step rewrites every block pair, rotated or not, so the destination holds the
 * complete block; the local copy of a redirected block is stale afterwards.  The caller orders the rounds
 * (a barrier between the end of this call on all ranks and the first use of the received blocks).
 * svdj_ipc_*: device allocations other processes can map, and their 64-byte handles (cudaIpcMemHandle_t).
 */
int svdj_ring_phase_send_device(void *ws, size_t m, size_t vrows, size_t w, double *dA_left, double *dA_right,
                                size_t lda, double *dV_left, double *dV_right, size_t ldv, int phase, double tol,
                                int inner_sweeps, unsigned long long *d_counters, double *dstA_left,
                                double *dstV_left, double *dstA_right, double *dstV_right, void *stream);
int svdj_ipc_alloc(size_t bytes, void **dptr);
int svdj_ipc_free(void *dptr);
int svdj_ipc_export(void *dptr, unsigned char *handle64);
int svdj_ipc_open(const unsigned char *handle64, void **peer_ptr);
int svdj_ipc_close(void *peer_ptr);

/* Row-sharded solve for tall inputs on several GPUs ("Tall-skinny"): every rank -- one process
 * per GPU -- holds m_local rows of A (all n columns, column-major, lda >= m_local) and rows v_row0 ..
 * v_row0 + vrows_local - 1 of V (ldv >= vrows_local); the row slices of the ranks tile A and V.  All ranks call
 * with the same n, opts and job flags.  The only quantity a step needs from the other ranks is the Gram matrices
 * of its block pairs: `allreduce(ctx, dptr, count, stream)` must sum `count` doubles at device pointer `dptr` over
 * all ranks, in place, ordered on `stream` (e.g. ncclAllReduce / torch.distributed.all_reduce), and return 0.
 * Every rank then runs the same eigen-solves and updates its own rows; sigma comes out complete on every rank, U
 * and V as the caller's row slices.  opts->tol must be set (sqrt(m_global) * 2^-53 for the library default).  n >= 128,
 * 16-byte aligned columns and even leading dimensions are required (SVDJ_ERR_UNSUPPORTED otherwise).
 * Replaces, for m >> n, the reference's distribution of a step's pairs over MPI ranks through rank 0
 * (main.cu:583-942): there every rank needs whole columns of m doubles per pair, here 18 KiB per block pair. */
typedef int (*svdj_allreduce_fn)(void *ctx, double *dptr, size_t count, void *stream);
int svdj_dgesvd_rows_device(int jobu, int jobv, size_t m_local, size_t n, double *dA, size_t lda, double *ds,
                            double *dV, size_t ldv, size_t vrows_local, size_t v_row0, const svdj_opts *opts,
                            svdj_stats *stats, void *stream, svdj_allreduce_fn allreduce, void *ctx);

/*
 * The multi-GPU solve as ONE collective call, like the reference's omp_mpi_cuda_dgesvd_local_matrices (main.cu:1587:
 * every rank calls, the matrix is valid on rank 0 only, main.cu:1455; rank and size are taken inside,
 * lib/JacobiMethods.cu:209-211).  The host driver is C++ inside the library (csrc/svdj_mg.cu): 2P column blocks on a
 * ring of P GPUs, the block move fused into the update kernel (peer stores over NVLink), rounds ordered by flags in
 * peer memory, convergence data exchanged through peer mailboxes -- no collective library call inside a solve.
 * It replaces the reference's per-step MPI_Send / MPI_Recv / MPI_Barrier star (main.cu:583-942).
 *
 * svdj_dgesvd_mg: single process, ngpus devices (devices[] = their ordinals, NULL = 0 .. ngpus-1).  Same contract
 *   as svdj_dgesvd (host pointers, U in A, sigma sorted).  The ranks are threads inside the call; their buffers and
 *   peer mappings are kept between calls (svdj_release_workspace frees them).  ngpus = 1 is svdj_dgesvd.
 *
 * svdj_mg_rank_*: one rank per process / GPU (e.g. under torchrun or mpirun).  The communicator needs one
 *   operation, used at set-up only: allgather(ctx, send, recv, bytes) gathers `bytes` bytes of host memory from
 *   every rank into recv in rank order (recv holds size * bytes) and returns 0 (MPI_Allgather on MPI_BYTE).  The
 *   ranks then map each other's buffers through CUDA IPC.  svdj_mg_rank_create is collective and uses the calling
 *   thread's current device; max_m x max_n bounds the problems the handle can solve.
 *   svdj_mg_rank_dgesvd (host pointers) and svdj_mg_rank_dgesvd_device (device pointers on rank 0's GPU) are
 *   collective: every rank calls with the same sizes, jobs and options; only rank 0's A, s, V are used, the other
 *   ranks may pass NULL.  A rank that waits more than 60 s
 *   (SVDJ_MG_TIMEOUT_S) for a peer gives up with SVDJ_ERR_CUDA instead of hanging its GPU.
 *   opts->precondition = 1 / 2 works here too: rank 0 factors on its GPU, the ring iterates on the n x n triangular
 *   factor, rank 0 applies Q; "library decides" (< 0) means no front-end in the multi-GPU entry points.
 * svdj_mg_plan: host-side copy of the tournament the driver walks, for tests: after `round` moves, which block
 *   (0 .. 2P-1) rank r holds as top / bottom and in which of its four buffers, and the transfers of the next move as
 *   (src rank, src buffer, dst rank, dst buffer) quadruples (at most 2P-2).  It stands where the reference's
 *   per-step rank assignment and packing stood (main.cu:526-580).
 */
typedef int (*svdj_allgather_fn)(void *ctx, const void *send, void *recv, size_t bytes);
typedef struct svdj_mg_comm {
  int rank;
  int size;
  svdj_allgather_fn allgather;
  void *ctx;
} svdj_mg_comm;
int svdj_dgesvd_mg(int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s, double *V, size_t ldv,
                   const svdj_opts *opts, svdj_stats *stats, int ngpus, const int *devices);
int svdj_mg_rank_create(const svdj_mg_comm *comm, size_t max_m, size_t max_n, void **handle);
int svdj_mg_rank_destroy(void *handle);
int svdj_mg_plan(int P, int round, int *top_block, int *bottom_block, int *top_buf, int *bottom_buf, int *transfers,
                 int *ntransfers);
int svdj_mg_rank_dgesvd(void *handle, int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s,
                        double *V, size_t ldv, const svdj_opts *opts, svdj_stats *stats);
int svdj_mg_rank_dgesvd_device(void *handle, int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda,
                               double *ds, double *dV, size_t ldv, const svdj_opts *opts, svdj_stats *stats);

/*
 * The QR front-end's building blocks on device buffers (what svdj_opts.precondition runs before and after the
 * sweeps).  The reference has no factorisation step: its doc block names tall inputs as the use-case
 * (lib/JacobiMethods.cu:12-14) and then sweeps the raw columns (main.cu:478-1384).
 * svdj_qr_factor_device: blocked Householder QR of Y (m x n, m >= n, column-major, ld >= m) in place, LAPACK dgeqrf's
 *   storage: R on and above the diagonal, the Householder vectors below it (unit diagonal implied); dT receives one
 *   32 x 32 upper-triangular factor per panel of 32 columns, svdj_qr_tfactor_doubles(n) doubles in all
 *   (H_1 .. H_32 = I - V T V^T).  m <= 65536.
 * svdj_qr_apply_q_device: C (m x ncols, ldc >= m) <- Q C with the Q of a previous svdj_qr_factor_device.
 */
size_t svdj_qr_tfactor_doubles(size_t n);
int svdj_qr_factor_device(size_t m, size_t n, double *dY, size_t ld, double *dT, void *stream);
int svdj_qr_apply_q_device(size_t m, size_t n, const double *dY, size_t ld, const double *dT, double *dC, size_t ldc,
                           size_t ncols, void *stream);

/* Number of steps of one sweep for n columns: 2*((n+1)/2) - 1 (main.cu:478, 498, 943). */
size_t svdj_num_steps(size_t n);

/* Host-side copy of the schedule the kernels use: writes the n/2 pairs (p,q), 0-based, of `step`
 * into pq[0..2*(n/2)) and returns their count (main.cu:526-545, 971-990). */
size_t svdj_schedule_step(size_t n, size_t step, size_t *pq);

/* Host-side copy of the schedule the block path uses over its nb column blocks: the nb/2 block pairs (p < q for
 * the two-level ordering), 0-based, of `step` (0-based, < svdj_num_steps(nb)).  vranks <= 1: the reference's
 * round-robin applied to block indices (main.cu:526-545); vranks = P >= 2 (nb divisible by 2P): the two-level
 * tournament the multi-GPU ring walks, which the single-GPU solver uses with P = 4.  The
 * reference has no block level; this replaces the role of its per-step pair lists (main.cu:971-990) for tests.
 * Returns the number of pairs written, 0 for invalid arguments. */
size_t svdj_block_schedule_step(size_t nb, int vranks, size_t step, size_t *pq);

/* The solver keeps its work buffers (device and pinned) in a process-wide pool between calls, because
 * cudaMalloc / cudaFree synchronise the whole device and showed up as 100 ms stalls inside solves; the
 * reference allocates and frees its staging vectors on every call (main.cu:484-494, 1385-1392).
 * svdj_release_workspace() returns every idle pooled buffer to the driver. */
int svdj_release_workspace(void);

/* CUDA devices visible to this process (0 without a driver or device). */
int svdj_device_count(void);

/* Text of the most recent error on the calling thread ("" if none). */
const char *svdj_last_error(void);

/* Library version: major*10000 + minor*100 + patch. */
int svdj_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SVDJ_H_ */
