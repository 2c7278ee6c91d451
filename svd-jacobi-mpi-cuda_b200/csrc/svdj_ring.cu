// svdj_ring.cu -- entry points for drivers that spread the columns over several GPUs.
//
// The reference distributes the pairs of a step over MPI ranks through a star: rank 0 packs and sends
// every rank its columns of A and V, receives them back and barriers, every step
// (lib/JacobiMethods.cu:335-432, 607-694).  Here each rank (one process per GPU) keeps two column
// blocks resident -- A-block m x w and V-block n x w each -- orthogonalises them against each other with
// the block-Jacobi kernels, and passes blocks to its ring neighbours between phases (the driver is
// svd-jacobi-mpi-cuda_b200/ring.py: torch.distributed / NCCL for the exchange, or peer stores fused into
// the update kernel).  This file is the per-rank compute: a "phase" = all round-robin steps over the
// blocks of the two buffers (SVDJ_PHASE_ALL: every pair of local columns; SVDJ_PHASE_CROSS: only pairs
// with one column in each buffer).
#include <cstdlib>
#include <cstring>
#include <new>

#include "svdj_common.cuh"

namespace svdj {

struct RingWorkspace {
  size_t m = 0, ncols = 0;
  double *gpart = nullptr;
  double *wbuf = nullptr;
  int *wflag = nullptr;
  double *norms = nullptr;
  double *gcache = nullptr;  // 64-column kernels: Gram matrix of every local column block of 32
  size_t gpart_doubles = 0;  // capacity of gpart (checked by the launchers)
  double floor2 = 0.0;       // squared norm below which a column is numerically zero (svdj_ring_ws_set_floor)
};

static void ws_free(RingWorkspace *w) {
  if (!w) return;
  if (w->gpart) cudaFree(w->gpart);
  if (w->wbuf) cudaFree(w->wbuf);
  if (w->wflag) cudaFree(w->wflag);
  if (w->norms) cudaFree(w->norms);
  if (w->gcache) cudaFree(w->gcache);
  delete w;
}

}  // namespace svdj

using namespace svdj;

extern "C" {

int svdj_ring_ws_create(size_t m, size_t ncols_local, void **ws_out) {
  if (ws_out == nullptr) return -3;
  *ws_out = nullptr;
  if (m == 0 || m > 0x7fffffffu) return -1;
  if (ncols_local == 0 || ncols_local % 32 != 0 || ncols_local > 0x7fffffffu) return -2;
  RingWorkspace *w = new (std::nothrow) RingWorkspace();
  if (!w) return SVDJ_ERR_ALLOC;
  w->m = m;
  w->ncols = ncols_local;
  size_t gp = 0, wd = 0, fl = 0, gp2 = 0, wd2 = 0, fl2 = 0;
  blk_scratch_sizes(m, blk_num_blocks(ncols_local), &gp, &wd, &fl);
  blk32_scratch_sizes(m, blk32_num_blocks(ncols_local), &gp2, &wd2, &fl2);  // either block width
  gp = gp > gp2 ? gp : gp2;
  w->gpart_doubles = gp;
  wd = wd > wd2 ? wd : wd2;
  fl = fl > fl2 ? fl : fl2;
  cudaError_t e = cudaMalloc(&w->gpart, sizeof(double) * gp);
  if (e == cudaSuccess) e = cudaMalloc(&w->gcache, sizeof(double) * 1024 * (ncols_local / 32 + 1));
  if (e == cudaSuccess) e = cudaMalloc(&w->wbuf, sizeof(double) * wd);
  if (e == cudaSuccess) e = cudaMalloc(&w->wflag, sizeof(int) * fl);
  if (e == cudaSuccess) e = cudaMalloc(&w->norms, sizeof(double) * ncols_local);
  if (e != cudaSuccess) {
    set_error("svdj_ring_ws_create: cudaMalloc -> %s", cudaGetErrorString(e));
    cudaGetLastError();
    ws_free(w);
    return SVDJ_ERR_ALLOC;
  }
  *ws_out = w;
  return SVDJ_OK;
}

int svdj_ring_ws_set_floor(void *ws, double floor_norm) {
  RingWorkspace *W = static_cast<RingWorkspace *>(ws);
  if (W == nullptr) return -1;
  if (!(floor_norm >= 0.0)) return -2;
  W->floor2 = floor_norm * floor_norm;
  return SVDJ_OK;
}

int svdj_ring_ws_destroy(void *ws) {
  ws_free(static_cast<RingWorkspace *>(ws));
  return SVDJ_OK;
}

}  // extern "C"

namespace svdj {
int ring_phase(void *ws, size_t m, size_t vrows, size_t w, double *dA_left, double *dA_right, size_t lda,
               double *dV_left, double *dV_right, size_t ldv, int rotate_v, int phase, double tol, int inner_sweeps,
               unsigned long long *d_counters, void *stream, double *dstA_left, double *dstA_right,
               double *dstV_left, double *dstV_right, bool redirect_last, unsigned long long *launches,
               ProbeSet *probes, int sort_cols) {
  RingWorkspace *W = static_cast<RingWorkspace *>(ws);
  if (W == nullptr) return -1;
  if (m == 0 || m != W->m) return -2;
  if (w == 0 || w % 16 != 0 || 2 * w != W->ncols) return -4;
  if (dA_left == nullptr || dA_right == nullptr) return -5;
  if (lda < m) return -7;
  if (rotate_v && (dV_left == nullptr || dV_right == nullptr)) return -8;
  if (rotate_v && ldv < vrows) return -10;
  if (phase != SVDJ_PHASE_ALL && phase != SVDJ_PHASE_CROSS) return -12;
  BlockArgs b;
  b.A = dA_left;
  b.lda = lda;
  b.V = dV_left;
  b.ldv = ldv;
  b.m = (unsigned)m;
  b.n = (unsigned)(2 * w);
  b.vrows = (unsigned)vrows;
  b.rotate_v = rotate_v ? 1 : 0;
  // 64-column block pairs where the local blocks divide into column blocks of 32 (SVDJ_BLOCK=16 forces the narrow ones)
  bool wide = (w % 32 == 0);
  if (const char *env = std::getenv("SVDJ_BLOCK")) wide = wide && std::atoi(env) != 16;
  b.nb = wide ? blk32_num_blocks(2 * w) : blk_num_blocks(2 * w);
  b.A2 = dA_right;
  b.V2 = dV_right;
  b.split = (unsigned)w;
  b.sched = (phase == SVDJ_PHASE_CROSS) ? 1 : 0;
  // ordering the columns of a block pair by size is only sound when "left" is the block with the lower global
  // index in every call (the multi-GPU driver sees to that); the Python ring driver leaves it off
  b.sort_cols = sort_cols;
  b.tol = tol;
  b.threshold_mode = SVDJ_THRESH_REL;
  b.max_inner_sweeps = inner_sweeps > 0 ? inner_sweeps : 2;
  b.counters = reinterpret_cast<SweepCounters *>(d_counters);
  b.floor2 = W->floor2;
  b.gpart_capacity = W->gpart_doubles;
  if (wide) {
    wide = blk32_supported(b);
    if (!wide) b.nb = blk_num_blocks(2 * w);
  }
  // the cache is refreshed in the first step of every phase (the blocks may just have arrived from a neighbour)
  b.gcache = wide ? W->gcache : nullptr;
  if (const char *env = std::getenv("SVDJ_GCACHE")) {
    if (env[0] == '0') b.gcache = nullptr;
  }
  const unsigned steps = (phase == SVDJ_PHASE_CROSS) ? b.nb / 2 : (unsigned)svdj_num_steps(b.nb);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  for (unsigned k = 1; k <= steps; ++k) {
    if (redirect_last && k == steps) {
      // last step of the round: the update kernel stores the travelling blocks where their next owner wants
      // them (peer-mapped memory of a neighbour GPU) and rewrites every pair, rotated or not
      b.A_st = dstA_left ? dstA_left : dA_left;
      b.A2_st = dstA_right ? dstA_right : dA_right;
      b.V_st = dstV_left ? dstV_left : dV_left;
      b.V2_st = dstV_right ? dstV_right : dV_right;
      b.force_update = 1;
    }
    if (int rc = (wide ? blk32_launch_step : blk_launch_step)(b, k, W->gpart, W->wbuf, W->wflag, st, launches,
                                                                    probes ? probes->next() : nullptr, nullptr))
      return rc;
  }
  return SVDJ_OK;
}
}  // namespace svdj

extern "C" {

int svdj_ring_phase_device(void *ws, size_t m, size_t vrows, size_t w, double *dA_left, double *dA_right, size_t lda,
                           double *dV_left, double *dV_right, size_t ldv, int rotate_v, int phase, double tol,
                           int inner_sweeps, unsigned long long *d_counters, void *stream) {
  return ring_phase(ws, m, vrows, w, dA_left, dA_right, lda, dV_left, dV_right, ldv, rotate_v, phase, tol,
                    inner_sweeps, d_counters, stream, nullptr, nullptr, nullptr, nullptr, false, nullptr);
}

int svdj_ring_phase_send_device(void *ws, size_t m, size_t vrows, size_t w, double *dA_left, double *dA_right,
                                size_t lda, double *dV_left, double *dV_right, size_t ldv, int phase, double tol,
                                int inner_sweeps, unsigned long long *d_counters, double *dstA_left,
                                double *dstV_left, double *dstA_right, double *dstV_right, void *stream) {
  return ring_phase(ws, m, vrows, w, dA_left, dA_right, lda, dV_left, dV_right, ldv, 1, phase, tol, inner_sweeps,
                    d_counters, stream, dstA_left, dstA_right, dstV_left, dstV_right, true, nullptr);
}

// ---- device memory that other processes can map (CUDA IPC) ------------------------------------------------
int svdj_ipc_alloc(size_t bytes, void **dptr) {
  if (dptr == nullptr) return -2;
  *dptr = nullptr;
  cudaError_t e = cudaMalloc(dptr, bytes ? bytes : 8);
  if (e != cudaSuccess) {
    set_error("svdj_ipc_alloc: cudaMalloc(%zu) -> %s", bytes, cudaGetErrorString(e));
    cudaGetLastError();
    return SVDJ_ERR_ALLOC;
  }
  SVDJ_CUDA_TRY(cudaMemset(*dptr, 0, bytes ? bytes : 8));
  return SVDJ_OK;
}

int svdj_ipc_free(void *dptr) {
  if (dptr) SVDJ_CUDA_TRY(cudaFree(dptr));
  return SVDJ_OK;
}

int svdj_ipc_export(void *dptr, unsigned char *handle64) {
  if (dptr == nullptr) return -1;
  if (handle64 == nullptr) return -2;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  SVDJ_CUDA_TRY(cudaIpcGetMemHandle(&h, dptr));
  std::memcpy(handle64, &h, 64);
  return SVDJ_OK;
}

int svdj_ipc_open(const unsigned char *handle64, void **peer_ptr) {
  if (handle64 == nullptr) return -1;
  if (peer_ptr == nullptr) return -2;
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle64, 64);
  SVDJ_CUDA_TRY(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return SVDJ_OK;
}

int svdj_ipc_close(void *peer_ptr) {
  if (peer_ptr) SVDJ_CUDA_TRY(cudaIpcCloseMemHandle(peer_ptr));
  return SVDJ_OK;
}

int svdj_ring_finalize_device(void *ws, size_t m, size_t ncols, double *dA, size_t lda, double *d_sigma, int scale_u,
                              void *stream) {
  RingWorkspace *W = static_cast<RingWorkspace *>(ws);
  if (W == nullptr) return -1;
  if (ncols > W->ncols) return -3;
  if (dA == nullptr) return -4;
  if (lda < m) return -5;
  if (d_sigma == nullptr) return -6;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (int rc = launch_column_norms(dA, lda, m, ncols, d_sigma, st)) return rc;
  if (scale_u) {
    if (int rc = launch_gather_columns(dA, lda, dA, lda, m, ncols, nullptr, d_sigma, ncols, st)) return rc;
  }
  return SVDJ_OK;
}

}  // extern "C"
