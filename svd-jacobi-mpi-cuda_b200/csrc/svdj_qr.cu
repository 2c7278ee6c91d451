// svdj_qr.cu -- blocked Householder QR and the application of Q, hand-written for sm_100a: the preconditioning
// front-end of the Jacobi solve (svdj_opts.precondition, driver: solve_with_qr in svdj_api.cu).
//
// The reference's doc block names the use-case ("M-by-N with m >> n", lib/JacobiMethods.cu:12-14) but runs plain
// one-sided Jacobi on whatever it is given.  Strongly graded spectra then need many dozens of sweeps; Drmac and
// Veselic's remedy is A P = Q R first and Jacobi on R^T.  This file is the QR part:
//
//   qr_factor   Y (m x n, column-major) -> Householder vectors below the diagonal, R on and above it, one 32 x 32
//               triangular factor T per panel of 32 columns (compact WY: H_1 .. H_32 = I - V T V^T).
//               Per panel two kernels:
//                 qr_panel_kernel   the m_j x 32 panel lives in REGISTERS, one row per thread (two for tall inputs),
//                                   spread over up to 148 co-resident CTAs; every column step is one fused pass
//                                   (products of the pivot column with all 32 columns: the trailing dots AND the
//                                   entries of V^T V that T needs), one deterministic grid-wide reduction
//                                   (partials through global memory, fixed summation order, one grid barrier),
//                                   and the rank-1 update in registers.  HBM sees the panel once in, once out.
//                 qr_block_reflector_kernel<true>   C <- (I - V T^T V^T) C on the trailing columns, one CTA per
//                                   strip of 32 columns: W = V^T C as FP64 DMMA tiles over the CTA's rows, M = T^T W
//                                   in shared memory, C -= V M as DMMA tiles.  5.3 flop/B at b = 32: at the
//                                   HBM / FP64-tensor ridge of a B200 (5.7).
//   qr_apply_q  C <- Q C with the panels in reverse order (qr_block_reflector_kernel<false>).
#include <algorithm>
#include <cstdlib>

#include "svdj_block_dev.cuh"
#include "svdj_qr.cuh"

namespace svdj {

static constexpr int kQB = 32;          // panel width
static constexpr int kPanelThreads = 256;

// ---------------------------------------------------------------------------------------------------------
// grid barrier for the co-resident CTAs of the panel kernel (cooperative launch guarantees co-residency)
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void grid_barrier(unsigned *bar, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    while (ld_acquire_u32(bar) < target) {
    }
  }
  __syncthreads();
}

// Thread-block cluster primitives (sm_90+): the panels of up to 8192 rows run as ONE cluster of <= 16 CTAs, whose
// hardware barrier and distributed shared memory replace the global-memory barrier and partials (5.5 -> ~1.5 us per
// column step).
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// store v at the address of `local_ptr` (this CTA's shared memory) in CTA `rank` of the cluster
__device__ __forceinline__ void st_cluster_f64(const void *local_ptr, unsigned rank, double v) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(local_ptr);
  unsigned ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(ra), "d"(v) : "memory");
}

// Sum over the warp of 32 values per lane; lane j ends with the total of value j (31 shuffles of doubles instead
// of 160 for 32 butterflies).
__device__ __forceinline__ double warp_transpose_sum(double (&p)[32], unsigned lane) {
  double q16[16];
  {
    const bool hi = (lane & 16u) != 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const double send = hi ? p[k] : p[k + 16];
      const double keep = hi ? p[k + 16] : p[k];
      q16[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  double q8[8];
  {
    const bool hi = (lane & 8u) != 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const double send = hi ? q16[k] : q16[k + 8];
      const double keep = hi ? q16[k + 8] : q16[k];
      q8[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  double q4[4];
  {
    const bool hi = (lane & 4u) != 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const double send = hi ? q8[k] : q8[k + 4];
      const double keep = hi ? q8[k + 4] : q8[k];
      q4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
  }
  double q2[2];
  {
    const bool hi = (lane & 2u) != 0;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const double send = hi ? q4[k] : q4[k + 2];
      const double keep = hi ? q4[k + 2] : q4[k];
      q2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
  }
  const bool hi = (lane & 1u) != 0;
  const double send = hi ? q2[0] : q2[1];
  const double keep = hi ? q2[1] : q2[0];
  return keep + __shfl_xor_sync(0xffffffffu, send, 1);  // lane j holds value index j (bits 4..0 = the halves kept)
}

// ---------------------------------------------------------------------------------------------------------
// Panel factorisation.  Rows j0 .. m-1, columns j0 .. j0+pcols-1 of Y.  Row i (relative to j0) belongs to
// register set rr = i / (G*256) of thread (i % 256) of CTA (i / 256) % G: rows 0..31 -- the pivot rows -- are
// lanes 0..31 of warp 0 of CTA 0.
//
// Column step c (LAPACK dlarfg / dlarf, restated for one fused pass):
//   x = column c below the pivot;  S_j = sum_i x_i row_i[j] for all 32 j   (S_c = |x|^2; j > c: trailing dots;
//                                                                          j < c: x . v_j, needed for T)
//   alpha = pivot[c];  beta = -sign(alpha) hypot(alpha, |x|);  tau = (beta - alpha) / beta;  v = x / (alpha - beta)
//   columns j > c:  w_j = pivot[j] + S_j / (alpha - beta);  row_i[j] -= tau w_j v_i;  pivot[j] -= tau w_j
//   T[a][c] = -tau * sum_{k=a}^{c-1} T[a][k] z_k,  z_k = v_k . v_c = pivot[k] + S_k / (alpha - beta)   (a < c)
// ---------------------------------------------------------------------------------------------------------
// The register array of a row is ROTATED by one position after every column step (position j <- position j+1,
// position 31 <- the finished Householder column): the pivot column is always position 0, the trailing columns
// positions 1 .. 31-c, the finished columns v_a positions 32-c+a.  Every step then runs the same code with
// compile-time register indices -- a run-time loop of ~600 instructions instead of 32 unrolled copies (the unrolled
// version was a 300 KB kernel that ran out of the instruction cache: 6 us per column step even on a single CTA) --
// and after 32 steps the array is back in column order.
template <int R, bool CL>
__global__ void __launch_bounds__(kPanelThreads) qr_panel_kernel(double *__restrict__ Y, const size_t ld,
                                                                 const unsigned m, const unsigned j0,
                                                                 const unsigned pcols, double *__restrict__ Tout,
                                                                 double *__restrict__ part,   // [2][G][32]
                                                                 double *__restrict__ pivot,  // [2][32]
                                                                 unsigned *__restrict__ bar) {
  __shared__ double red[8][32];
  __shared__ double coef[32];   // by position: tau * w for the trailing columns, 0 elsewhere
  __shared__ double scal[2];    // scale = 1 / (alpha - beta), beta
  __shared__ double Z[32][33];  // CTA 0: z_k of every step (column c of Z), then T
  __shared__ double taus[32];
  // cluster variant: every CTA receives all CTAs' partial sums and the pivot row in its own shared memory
  __shared__ double cpart[CL ? 2 : 1][CL ? 16 : 1][32];
  __shared__ double cpiv[CL ? 2 : 1][32];
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned G = gridDim.x, cta = blockIdx.x;  // CL: the grid is one cluster, blockIdx.x = rank in it
  const unsigned mj = m - j0;

  double row[R][32];
  unsigned ridx[R];
#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    ridx[rr] = (unsigned)rr * G * kPanelThreads + cta * kPanelThreads + tid;
    const bool ok = ridx[rr] < mj;
    const double *src = Y + (size_t)j0 * ld + j0 + ridx[rr];
#pragma unroll
    for (int j = 0; j < 32; ++j) row[rr][j] = (ok && (unsigned)j < pcols) ? src[(size_t)j * ld] : 0.0;
  }

#pragma unroll 1
  for (unsigned c = 0; c < pcols; ++c) {  // grid-uniform trip count
    // products of the pivot column (position 0, rows below the pivot) with all 32 positions
    double p[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) p[j] = 0.0;
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const double x = (ridx[rr] > c && ridx[rr] < mj) ? row[rr][0] : 0.0;
#pragma unroll
      for (int j = 0; j < 32; ++j) p[j] = fma(x, row[rr][j], p[j]);
    }
    const double ws = warp_transpose_sum(p, lane);
    red[warp][lane] = ws;
    __syncthreads();
    if (warp == 0) {
      double s = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][lane];
      if (CL) {
        for (unsigned g = 0; g < G; ++g) st_cluster_f64(&cpart[c & 1][cta][lane], g, s);
        if (cta == 0) {  // the pivot row of this step is lane c's register set 0: through `coef` to one value per lane
          if (lane == c) {
#pragma unroll
            for (int j = 0; j < 32; ++j) coef[j] = row[0][j];
          }
          __syncwarp();
          const double pvj = coef[lane];
          for (unsigned g = 0; g < G; ++g) st_cluster_f64(&cpiv[c & 1][lane], g, pvj);
        }
      } else {
        part[((size_t)(c & 1) * G + cta) * 32 + lane] = s;
        if (cta == 0 && lane == c) {  // the pivot row of this step: register set 0 of this thread
#pragma unroll
          for (int j = 0; j < 32; ++j) pivot[(c & 1) * 32 + j] = row[0][j];
        }
      }
    }
    if (CL) cluster_sync_all();
    else grid_barrier(bar, (c + 1) * G);
    {
      double s = 0.0;
      if (CL) {
        for (unsigned g = warp; g < G; g += 8) s += cpart[c & 1][g][lane];
      } else {
        for (unsigned g = warp; g < G; g += 8) s += __ldcg(part + ((size_t)(c & 1) * G + g) * 32 + lane);
      }
      red[warp][lane] = s;
    }
    __syncthreads();
    if (warp == 0) {
      double S = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) S += red[w][lane];
      const double pv = CL ? cpiv[c & 1][lane] : __ldcg(pivot + (c & 1) * 32 + lane);
      const double alpha = __shfl_sync(0xffffffffu, pv, 0);  // position 0 = the pivot column
      const double xn2 = __shfl_sync(0xffffffffu, S, 0);
      double beta = alpha, tau = 0.0, scale = 0.0;
      if (xn2 > 0.0) {
        const double nrm = sqrt(fma(alpha, alpha, xn2));
        beta = alpha >= 0.0 ? -nrm : nrm;
        tau = (beta - alpha) / beta;
        scale = 1.0 / (alpha - beta);
      }
      const double wj = fma(S, scale, pv);  // trailing position: v . y;  finished position 32-c+a: v_a . v
      coef[lane] = (lane >= 1u && lane <= 31u - c) ? tau * wj : 0.0;
      if (lane == 0) {
        scal[0] = scale;
        scal[1] = beta;
      }
      if (cta == 0) {
        if (lane >= 32u - c) Z[lane - (32u - c)][c] = wj;  // a = lane - (32 - c) < c
        if (lane == 0) taus[c] = tau;
      }
    }
    __syncthreads();
    const double scale = scal[0], beta = scal[1];
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      double v = row[rr][0];  // rows above the pivot keep their R entry
      if (ridx[rr] < mj) {
        if (ridx[rr] > c) {
          v = v * scale;
#pragma unroll
          for (int j = 1; j < 32; ++j) row[rr][j] = fma(-coef[j], v, row[rr][j]);
        } else if (ridx[rr] == c) {
          v = beta;
#pragma unroll
          for (int j = 1; j < 32; ++j) row[rr][j] -= coef[j];
        }
      }
      // rotate: position j <- j + 1, position 31 <- this step's column
#pragma unroll
      for (int j = 0; j < 31; ++j) row[rr][j] = row[rr][j + 1];
      row[rr][31] = v;
    }
  }

  // after pcols rotations position j holds column (j + pcols) mod 32 (pcols = 32: column j)
#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    if (ridx[rr] < mj) {
      double *dst = Y + (size_t)j0 * ld + j0 + ridx[rr];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const unsigned col = ((unsigned)j + pcols) & 31u;
        if (col < pcols) dst[(size_t)col * ld] = row[rr][j];
      }
    }
  }

  if (CL) cluster_sync_all();  // nobody leaves while a peer may still store into its shared memory
  if (cta == 0 && warp == 0) {
    // T, column by column (forward recurrence of LAPACK dlarft); lane a owns row a:
    //   T[c][c] = tau_c,  T[a][c] = -tau_c sum_{k=a}^{c-1} T[a][k] z_k(c)   (a < c; Z[k][c] = v_k . v_c)
    // (T[a][k], a <= k, is kept at Z[k][a]: the lower triangle of Z is free)
    __syncwarp();
    for (int c = 0; c < 32; ++c) {
      const double tau = ((unsigned)c < pcols) ? taus[c] : 0.0;
      double t = 0.0;
      if ((int)lane < c) {
        for (int k = (int)lane; k < c; ++k) t = fma(Z[k][lane], Z[k][c], t);
        t *= -tau;
      } else if ((int)lane == c) {
        t = tau;
      }
      __syncwarp();
      if ((int)lane <= c) Z[c][lane] = t;  // T[lane][c]
      __syncwarp();
    }
    // column-major T[a + 32 c] = Z[c][a] for a <= c, zero below the diagonal
    for (int c = 0; c < 32; ++c) Tout[32 * c + lane] = ((int)lane <= c) ? Z[c][lane] : 0.0;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Block reflector  C <- (I - V op(T) V^T) C  on rows j0.. of the columns c0 .. c0+ncols-1 of Cmat, in three kernels
// tiled like a GEMM so that the panel V is re-read once per 64 columns of C instead of once per 16:
//   qr_vtc_kernel      Wpart[split] = V^T C over a slice of the rows, for a tile of 64 columns: 32 DMMA per 12 loads
//   qr_tw_kernel       M = -op(T) sum_split Wpart    (op(T) = T^T: Q^T C, the trailing update; T: Q C, applying Q)
//   qr_vm_kernel       C += V M on 128-row x 64-column tiles: 128 DMMA per 48 loads and 32 stores
// V = the unit lower trapezoidal panel stored in Y (rows j0.., columns j0 .. j0+pcols-1).  The first version (one CTA
// per strip of 16 / 32 columns doing all three steps) re-read V from L2 for every strip: 8.6 TFLOP/s at 8192^2.
// ---------------------------------------------------------------------------------------------------------
static constexpr int kRefThreads = 256;
static constexpr unsigned kRefCols = 64;    // columns of C per CTA
static constexpr unsigned kVmRows = 128;    // rows of C per CTA of qr_vm_kernel

__device__ __forceinline__ double panel_v(const double *__restrict__ Yp, size_t ldy, unsigned i, unsigned k,
                                          unsigned mj, unsigned pcols) {
  // V(i, k): rows and columns relative to the panel's corner
  if (i >= mj || k >= pcols || i < k) return 0.0;
  if (i == k) return 1.0;
  return Yp[(size_t)k * ldy + i];
}

__global__ void __launch_bounds__(kRefThreads, 1) qr_vtc_kernel(const double *__restrict__ Y, const size_t ldy,
                                                                const unsigned m, const unsigned j0, const unsigned pcols,
                                                                const double *__restrict__ Cmat, const size_t ldc,
                                                                const unsigned c0, const unsigned ncols,
                                                                const unsigned rows_per_split,  // a multiple of 4
                                                                double *__restrict__ wpart) {  // [split][ncols_pad][32]
  __shared__ double buf[32 * 64];  // the CTA's 32 x 64 result, fragment order: [(ti * 8 + tj) * 64 + 2 lane + e]
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned g = lane >> 2, tg = lane & 3u;
  const unsigned mj = m - j0;
  const unsigned cs = c0 + kRefCols * blockIdx.x;
  const unsigned sw = min(kRefCols, c0 + ncols - cs);
  const unsigned rb = blockIdx.y * rows_per_split, re = min(mj, rb + rows_per_split);
  const double *Yp = Y + (size_t)j0 * ldy + j0;
  const double *Cp = Cmat + (size_t)cs * ldc + j0;

  double acc[4][8][2];
#pragma unroll
  for (int ti = 0; ti < 4; ++ti)
#pragma unroll
    for (int tj = 0; tj < 8; ++tj) acc[ti][tj][0] = acc[ti][tj][1] = 0.0;
  const unsigned ks0 = rb >> 2, ks1 = (re + 3u) >> 2;
  for (unsigned ks = ks0 + warp; ks < ks1; ks += 8) {
    const unsigned ri = 4u * ks + tg;
    const bool ok = ri < re;
    double a[4], b[8];
    if (4u * ks < 32u) {  // the panel's triangle: unit diagonal, zeros above
#pragma unroll
      for (int ti = 0; ti < 4; ++ti) a[ti] = ok ? panel_v(Yp, ldy, ri, 8u * ti + g, mj, pcols) : 0.0;
    } else {
#pragma unroll
      for (int ti = 0; ti < 4; ++ti) a[ti] = (ok && 8u * ti + g < pcols) ? Yp[(size_t)(8u * ti + g) * ldy + ri] : 0.0;
    }
#pragma unroll
    for (int tj = 0; tj < 8; ++tj) b[tj] = (ok && 8u * tj + g < sw) ? Cp[(size_t)(8u * tj + g) * ldc + ri] : 0.0;
#pragma unroll
    for (int ti = 0; ti < 4; ++ti)
#pragma unroll
      for (int tj = 0; tj < 8; ++tj) dmma884(acc[ti][tj][0], acc[ti][tj][1], a[ti], b[tj]);
  }
  // the 8 warps' partial results, summed in warp order (deterministic)
  for (unsigned w = 0; w < 8; ++w) {
    if (warp == w) {
#pragma unroll
      for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = 0; tj < 8; ++tj) {
          double *d = buf + (ti * 8 + tj) * 64 + 2 * lane;
          if (w == 0) {
            d[0] = acc[ti][tj][0];
            d[1] = acc[ti][tj][1];
          } else {
            d[0] += acc[ti][tj][0];
            d[1] += acc[ti][tj][1];
          }
        }
    }
    __syncthreads();
  }
  // out: wpart[split][column][k], k = row of W (panel column) -- 32 contiguous doubles per column of C
  const unsigned ncols_pad = (ncols + kRefCols - 1) / kRefCols * kRefCols;
  double *out = wpart + ((size_t)blockIdx.y * ncols_pad + (size_t)kRefCols * blockIdx.x) * 32;
  for (unsigned idx = tid; idx < 32 * 64; idx += kRefThreads) {
    const unsigned k = idx & 31u, n = idx >> 5;                       // W[k][n]
    const unsigned ti = k >> 3, gg = k & 7u, tj = n >> 3, e = n & 1u, t4 = (n & 7u) >> 1;
    out[idx] = buf[(ti * 8 + tj) * 64 + 2 * (gg * 4 + t4) + e];
  }
}

// The same product with the tiles staged through shared memory by cp.async (3 stages of 64 rows x (32 + 64)
// columns, 52 KB each): the loads run two tiles ahead of the DMMAs instead of stalling every k-step on global
// memory (the direct version streams C at 1.7 TB/s).  Needs 16-byte aligned columns (even leading dimensions).
static constexpr int kVtRows = 64;
static constexpr int kVtStride = kVtRows + 4;
static constexpr int kVtCols = 32 + (int)kRefCols;
static constexpr int kVtStageDoubles = kVtCols * kVtStride;
static constexpr int kVtStages = 3;
static constexpr size_t kVtSmem = sizeof(double) * kVtStages * kVtStageDoubles;

__global__ void __launch_bounds__(kRefThreads, 1) qr_vtc_staged_kernel(const double *__restrict__ Y, const size_t ldy,
                                                                       const unsigned m, const unsigned j0,
                                                                       const unsigned pcols,
                                                                       const double *__restrict__ Cmat, const size_t ldc,
                                                                       const unsigned c0, const unsigned ncols,
                                                                       const unsigned rows_per_split,  // multiple of 64
                                                                       double *__restrict__ wpart) {
  extern __shared__ __align__(16) double tiles[];  // kVtStages x 96 columns x kVtStride; then the 32 x 64 result
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned g = lane >> 2, tg = lane & 3u;
  const unsigned mj = m - j0;
  const unsigned cs = c0 + kRefCols * blockIdx.x;
  const unsigned sw = min(kRefCols, c0 + ncols - cs);
  const unsigned rb = blockIdx.y * rows_per_split, re = min(mj, rb + rows_per_split);
  const double *Yp = Y + (size_t)j0 * ldy + j0;
  const double *Cp = Cmat + (size_t)cs * ldc + j0;
  const unsigned ntiles = rb < re ? (re - rb + kVtRows - 1) / kVtRows : 0;

  const double *cptr[12];  // this thread copies columns warp + 8 i of the 96 (nullptr: beyond the panel / the strip)
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    const unsigned col = warp + 8u * i;
    if (col < 32u) cptr[i] = col < pcols ? Yp + (size_t)col * ldy : nullptr;
    else cptr[i] = (col - 32u) < sw ? Cp + (size_t)(col - 32u) * ldc : nullptr;
  }
  auto issue_tile = [&](unsigned t, int stage) {
    const unsigned row = rb + t * kVtRows + 2 * lane;
    double *dst = tiles + stage * kVtStageDoubles + 2 * lane;
    const int nb = (row < re) ? ((row + 1 < re) ? 16 : 8) : 0;
#pragma unroll
    for (int i = 0; i < 12; ++i) {
      const bool ok = cptr[i] != nullptr && nb != 0;
      cp_async16(dst + (warp + 8 * i) * kVtStride, ok ? cptr[i] + row : Y, ok ? nb : 0);
    }
  };
  unsigned issued = 0;
#pragma unroll
  for (int st = 0; st < kVtStages - 1; ++st) {
    if (issued < ntiles) {
      issue_tile(issued, st);
      ++issued;
    }
    cp_async_commit();
  }
  double acc[4][8][2];
#pragma unroll
  for (int ti = 0; ti < 4; ++ti)
#pragma unroll
    for (int tj = 0; tj < 8; ++tj) acc[ti][tj][0] = acc[ti][tj][1] = 0.0;

  for (unsigned t = 0; t < ntiles; ++t) {
    cp_async_wait<kVtStages - 2>();
    __syncthreads();
    if (issued < ntiles) {
      issue_tile(issued, (int)(issued % kVtStages));
      ++issued;
    }
    cp_async_commit();
    double *tile = tiles + (t % kVtStages) * kVtStageDoubles;
    if (rb + t * kVtRows < 32u) {  // the panel's triangle lives in this tile: unit diagonal, zeros above
      for (unsigned idx = tid; idx < 32 * 32; idx += kRefThreads) {
        const unsigned k = idx >> 5, i = idx & 31u;  // V(i, k), rows relative to the panel's corner
        const unsigned r = i - (rb + t * kVtRows);   // rb + t*64 is 0 here
        if (i < k) tile[k * kVtStride + r] = 0.0;
        else if (i == k) tile[k * kVtStride + r] = (k < pcols && i < mj) ? 1.0 : 0.0;
      }
      __syncthreads();
    }
    const double *tp = tile + g * kVtStride + 8 * warp + tg;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      double a[4], b[8];
#pragma unroll
      for (int ti = 0; ti < 4; ++ti) a[ti] = tp[(8 * ti) * kVtStride + 4 * ks];
#pragma unroll
      for (int tj = 0; tj < 8; ++tj) b[tj] = tp[(32 + 8 * tj) * kVtStride + 4 * ks];
#pragma unroll
      for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = 0; tj < 8; ++tj) dmma884(acc[ti][tj][0], acc[ti][tj][1], a[ti], b[tj]);
    }
  }
  cp_async_wait<0>();
  __syncthreads();
  double *buf = tiles;  // the 8 warps' partial results, summed in warp order (deterministic)
  for (unsigned w = 0; w < 8; ++w) {
    if (warp == w) {
#pragma unroll
      for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = 0; tj < 8; ++tj) {
          double *d = buf + (ti * 8 + tj) * 64 + 2 * lane;
          if (w == 0) {
            d[0] = acc[ti][tj][0];
            d[1] = acc[ti][tj][1];
          } else {
            d[0] += acc[ti][tj][0];
            d[1] += acc[ti][tj][1];
          }
        }
    }
    __syncthreads();
  }
  const unsigned ncols_pad = (ncols + kRefCols - 1) / kRefCols * kRefCols;
  double *out = wpart + ((size_t)blockIdx.y * ncols_pad + (size_t)kRefCols * blockIdx.x) * 32;
  for (unsigned idx = tid; idx < 32 * 64; idx += kRefThreads) {
    const unsigned k = idx & 31u, n = idx >> 5;
    const unsigned ti = k >> 3, gg = k & 7u, tj = n >> 3, e = n & 1u, t4 = (n & 7u) >> 1;
    out[idx] = buf[(ti * 8 + tj) * 64 + 2 * (gg * 4 + t4) + e];
  }
}

// M[:, n] = -op(T) (sum over the splits of Wpart[:, n]); one thread per entry, T in shared memory
template <bool TRANS_T>
__global__ void __launch_bounds__(256) qr_tw_kernel(const double *__restrict__ T, const double *__restrict__ wpart,
                                                    const unsigned splits, const unsigned ncols_pad,
                                                    double *__restrict__ M) {  // [ncols_pad][32]
  __shared__ double Ts[32][33];
  __shared__ double Ws[8][33];
  const unsigned tid = threadIdx.x;
  for (unsigned i = tid; i < 1024; i += 256) Ts[i & 31][i >> 5] = T[i];  // Ts[a][c] = T[a + 32 c]
  const unsigned n0 = 8u * blockIdx.x, nl = tid >> 5, k = tid & 31u;      // 8 columns per CTA
  double s = 0.0;
  for (unsigned sp = 0; sp < splits; ++sp) s += wpart[((size_t)sp * ncols_pad + n0 + nl) * 32 + k];
  Ws[nl][k] = s;
  __syncthreads();
  double r = 0.0;
  if (TRANS_T) {
    for (unsigned a = 0; a <= k; ++a) r = fma(Ts[a][k], Ws[nl][a], r);
  } else {
    for (unsigned a = k; a < 32; ++a) r = fma(Ts[k][a], Ws[nl][a], r);
  }
  M[(size_t)(n0 + nl) * 32 + k] = -r;
}

__global__ void __launch_bounds__(kRefThreads, 2) qr_vm_kernel(const double *__restrict__ Y, const size_t ldy,
                                                               const unsigned m, const unsigned j0, const unsigned pcols,
                                                               const double *__restrict__ M, double *__restrict__ Cmat,
                                                               const size_t ldc, const unsigned c0, const unsigned ncols) {
  __shared__ double Ms[kRefCols][36];  // Ms[n][k] = (-M)[k][n]
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned g = lane >> 2, tg = lane & 3u;
  const unsigned mj = m - j0;
  const unsigned cs = c0 + kRefCols * blockIdx.x;
  const unsigned sw = min(kRefCols, c0 + ncols - cs);
  const double *Yp = Y + (size_t)j0 * ldy + j0;
  double *Cp = Cmat + (size_t)cs * ldc + j0;
  const double *Mt = M + (size_t)kRefCols * blockIdx.x * 32;
  for (unsigned idx = tid; idx < kRefCols * 32; idx += kRefThreads) Ms[idx >> 5][idx & 31u] = Mt[idx];
  __syncthreads();
#pragma unroll 1
  for (unsigned h = 0; h < 2; ++h) {
    const unsigned r0 = kVmRows * blockIdx.y + 16u * warp + 8u * h;
    if (r0 >= mj) break;
    const unsigned ri = r0 + g;
    const bool ok = ri < mj;
    double a[8];
    if (r0 < 32u) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) a[ks] = panel_v(Yp, ldy, ri, 4u * ks + tg, mj, pcols);
    } else {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) a[ks] = (ok && 4u * ks + tg < pcols) ? Yp[(size_t)(4u * ks + tg) * ldy + ri] : 0.0;
    }
    double d[8][2];
#pragma unroll
    for (int tj = 0; tj < 8; ++tj) {
      const unsigned col = 8u * tj + 2u * tg;
      d[tj][0] = (ok && col < sw) ? Cp[(size_t)col * ldc + ri] : 0.0;
      d[tj][1] = (ok && col + 1 < sw) ? Cp[(size_t)(col + 1) * ldc + ri] : 0.0;
    }
#pragma unroll
    for (int tj = 0; tj < 8; ++tj)
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) dmma884(d[tj][0], d[tj][1], a[ks], Ms[8 * tj + g][4 * ks + tg]);
#pragma unroll
    for (int tj = 0; tj < 8; ++tj) {
      const unsigned col = 8u * tj + 2u * tg;
      if (ok && col < sw) Cp[(size_t)col * ldc + ri] = d[tj][0];
      if (ok && col + 1 < sw) Cp[(size_t)(col + 1) * ldc + ri] = d[tj][1];
    }
  }
}

// L = R^T: out (rows_out x n, column-major, ldo) gets out[i][j] = R[j][i] for j <= i < n and zero elsewhere
// (rows n .. rows_out-1 are zero pad rows), R = the upper triangle of Y.
__global__ void __launch_bounds__(256) qr_transpose_r_kernel(const double *__restrict__ Y, size_t ldy, unsigned n,
                                                             double *__restrict__ out, size_t ldo, unsigned rows_out) {
  __shared__ double tile[32][33];
  const unsigned bi = blockIdx.x, bj = blockIdx.y;  // out tile: rows 32 bi.., columns 32 bj..
  const unsigned tx = threadIdx.x & 31u, ty = threadIdx.x >> 5;
  // read R[32 bj + r][32 bi + c] (row of R = column of out)
  if (bi >= bj) {
    for (unsigned c = ty; c < 32; c += 8) {
      const unsigned rr = 32u * bj + tx, cc = 32u * bi + c;  // R[rr][cc]: column cc of Y, row rr
      tile[c][tx] = (rr < n && cc < n && rr <= cc) ? Y[(size_t)cc * ldy + rr] : 0.0;
    }
  }
  __syncthreads();
  for (unsigned c = ty; c < 32; c += 8) {
    const unsigned i = 32u * bi + tx, j = 32u * bj + c;  // out[i][j] = R[j][i] = tile[i - 32 bi][j - 32 bj]
    if (i < rows_out && j < n) out[(size_t)j * ldo + i] = (bi >= bj) ? tile[tx][c] : 0.0;
  }
}

// out[perm[i]][k] = in[i][k]: undoes the column pre-sort on the rows of V
__global__ void __launch_bounds__(256) qr_scatter_rows_kernel(const double *__restrict__ in, size_t ldi,
                                                              double *__restrict__ out, size_t ldo, unsigned rows,
                                                              const unsigned *__restrict__ perm) {
  const double *ci = in + (size_t)blockIdx.x * ldi;
  double *co = out + (size_t)blockIdx.x * ldo;
  for (unsigned i = threadIdx.x; i < rows; i += 256) co[perm[i]] = ci[i];
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static constexpr unsigned kMaxPanelCtas = 128;
// strips of 32 columns only when they alone give every SM its two CTAs (148 x 2 x 32 columns)
static constexpr unsigned kWideStripMinCols = 148 * 2 * 32;  // co-resident by a wide margin on 148 SMs

bool qr_supported(size_t m, size_t n) { return n >= 1 && m >= n && m <= (size_t)kMaxPanelCtas * kPanelThreads * 2; }

size_t qr_num_panels(size_t n) { return (n + kQB - 1) / kQB; }

size_t qr_reflector_scratch_doubles(size_t ncols);
static size_t panel_scratch_bytes(size_t n) {
  // part [2][G][32] + pivot [2][32] doubles, one barrier word per panel; rounded to 256 bytes
  const size_t b = sizeof(double) * (2 * (size_t)kMaxPanelCtas * 32 + 64) + sizeof(unsigned) * (qr_num_panels(n) + 2);
  return (b + 255) & ~size_t(255);
}
// scratch of qr_factor on n columns, also enough for qr_apply_q on up to n columns of C
size_t qr_scratch_bytes(size_t n) { return panel_scratch_bytes(n) + sizeof(double) * qr_reflector_scratch_doubles(n); }

static int reflector_setup() {
  static DeviceOnce once;
  SVDJ_CUDA_TRY(once.run([] {
    cudaError_t e = cudaFuncSetAttribute(qr_vtc_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVtSmem);
    // clusters of 16 CTAs are beyond the portable size of 8
    if (e == cudaSuccess) e = cudaFuncSetAttribute(qr_panel_kernel<1, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(qr_panel_kernel<2, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    return e;
  }));
  return SVDJ_OK;
}

// row slices of qr_vtc_kernel: enough CTAs for two per SM, every slice a whole number of 64-row groups
static unsigned vtc_splits(size_t mj, size_t ncols) {
  const size_t tiles = (ncols + kRefCols - 1) / kRefCols;
  size_t want = (2 * 148 + tiles - 1) / tiles;
  const size_t cap = (mj + 255) / 256;  // at least 256 rows per slice
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  if (want > 64) want = 64;
  return (unsigned)want;
}

// C <- (I - V op(T) V^T) C; wpart / mbuf: scratch of qr_reflector_scratch_doubles(ncols)
template <bool TRANS_T>
static int apply_block_reflector(const double *Y, size_t ldy, unsigned m, unsigned j0, unsigned pcols, const double *T,
                                 double *C, size_t ldc, unsigned c0, unsigned ncols, double *scratch, cudaStream_t st,
                                 unsigned long long *launches) {
  const size_t mj = m - j0;
  const unsigned tiles = (ncols + kRefCols - 1) / kRefCols;
  const unsigned ncols_pad = tiles * kRefCols;
  const unsigned splits = vtc_splits(mj, ncols);
  unsigned rps = (unsigned)((mj + splits - 1) / splits);
  rps = (rps + 63u) & ~63u;
  double *mbuf = scratch;                               // [ncols_pad][32]
  double *wpart = scratch + (size_t)ncols_pad * 32;     // [splits][ncols_pad][32]
  const bool aligned = (ldy % 2 == 0) && (ldc % 2 == 0) && ((reinterpret_cast<uintptr_t>(Y) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(C) & 15) == 0) && (j0 % 2 == 0);
  if (aligned)
    qr_vtc_staged_kernel<<<dim3(tiles, splits), kRefThreads, kVtSmem, st>>>(Y, ldy, m, j0, pcols, C, ldc, c0, ncols, rps,
                                                                            wpart);
  else
    qr_vtc_kernel<<<dim3(tiles, splits), kRefThreads, 0, st>>>(Y, ldy, m, j0, pcols, C, ldc, c0, ncols, rps, wpart);
  qr_tw_kernel<TRANS_T><<<ncols_pad / 8, 256, 0, st>>>(T, wpart, splits, ncols_pad, mbuf);
  qr_vm_kernel<<<dim3(tiles, (unsigned)((mj + kVmRows - 1) / kVmRows)), kRefThreads, 0, st>>>(Y, ldy, m, j0, pcols, mbuf,
                                                                                            C, ldc, c0, ncols);
  if (launches) *launches += 3;
  return SVDJ_OK;
}

size_t qr_reflector_scratch_doubles(size_t ncols) {
  const size_t pad = (ncols + kRefCols - 1) / kRefCols * kRefCols;
  return pad * 32 * (1 + 64);
}

int qr_factor(double *Y, size_t ld, size_t m, size_t n, double *Tbuf, void *scratch, cudaStream_t st,
              unsigned long long *launches) {
  if (!qr_supported(m, n)) {
    set_error("qr_factor: %zu x %zu is outside the supported range (n <= m <= %u)", m, n,
              kMaxPanelCtas * kPanelThreads * 2);
    return SVDJ_ERR_UNSUPPORTED;
  }
  if (int rc = reflector_setup()) return rc;
  // SVDJ_QR_CLUSTER=0 keeps every panel on the global-memory barrier (for comparisons and as a fallback)
  static const bool cluster_env = [] {
    const char *e = std::getenv("SVDJ_QR_CLUSTER");
    return !(e != nullptr && e[0] == '0');
  }();
  const bool use_cluster = cluster_env;
  double *part = static_cast<double *>(scratch);
  double *pivot = part + 2 * (size_t)kMaxPanelCtas * 32;
  unsigned *bar = reinterpret_cast<unsigned *>(pivot + 64);
  double *rscratch = reinterpret_cast<double *>(static_cast<char *>(scratch) + panel_scratch_bytes(n));
  const size_t np = qr_num_panels(n);
  SVDJ_CUDA_TRY(cudaMemsetAsync(bar, 0, sizeof(unsigned) * np, st));
  for (size_t p = 0; p < np; ++p) {
    unsigned j0 = (unsigned)(p * kQB);
    unsigned pcols = (unsigned)std::min<size_t>(kQB, n - j0);
    unsigned mu = (unsigned)m;
    const size_t mj = m - j0;
    double *Tp = Tbuf + p * 1024;
    unsigned *barp = bar + p;
    if (use_cluster && mj <= (size_t)16 * kPanelThreads * 2) {
      // one cluster: rows over 1, 2, 4, 8 or 16 CTAs, one row per thread up to 4096 rows, two above
      const int R = (mj > (size_t)16 * kPanelThreads) ? 2 : 1;
      unsigned G = 1;
      while ((size_t)G * kPanelThreads * R < mj) G *= 2;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = G;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(G);
      cfg.blockDim = dim3(kPanelThreads);
      cfg.dynamicSmemBytes = 0;
      cfg.stream = st;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      if (R == 1)
        SVDJ_CUDA_TRY(cudaLaunchKernelEx(&cfg, qr_panel_kernel<1, true>, Y, ld, mu, j0, pcols, Tp, part, pivot, barp));
      else
        SVDJ_CUDA_TRY(cudaLaunchKernelEx(&cfg, qr_panel_kernel<2, true>, Y, ld, mu, j0, pcols, Tp, part, pivot, barp));
    } else {
      const int R = (mj > (size_t)kMaxPanelCtas * kPanelThreads) ? 2 : 1;
      unsigned G = (unsigned)((mj + (size_t)kPanelThreads * R - 1) / ((size_t)kPanelThreads * R));
      if (G < 1) G = 1;
      void *args[] = {&Y, &ld, &mu, &j0, &pcols, &Tp, &part, &pivot, &barp};
      if (R == 1)
        SVDJ_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)qr_panel_kernel<1, false>, dim3(G), dim3(kPanelThreads), args, 0, st));
      else
        SVDJ_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)qr_panel_kernel<2, false>, dim3(G), dim3(kPanelThreads), args, 0, st));
    }
    if (launches) *launches += 1;
    const size_t c0 = (size_t)j0 + pcols;
    if (c0 < n) {
      const unsigned ncols = (unsigned)(n - c0);
      if (int rc = apply_block_reflector<true>(Y, ld, mu, j0, pcols, Tp, Y, ld, (unsigned)c0, ncols, rscratch, st, launches))
        return rc;
    }
  }
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int qr_apply_q(const double *Y, size_t ld, size_t m, size_t n, const double *Tbuf, double *C, size_t ldc, size_t ncols,
               void *scratch, cudaStream_t st, unsigned long long *launches) {
  if (ncols == 0 || n == 0) return SVDJ_OK;
  if (int rc = reflector_setup()) return rc;
  const size_t np = qr_num_panels(n);
  for (size_t p = np; p-- > 0;) {
    const unsigned j0 = (unsigned)(p * kQB);
    const unsigned pcols = (unsigned)std::min<size_t>(kQB, n - j0);
    if (int rc = apply_block_reflector<false>(Y, ld, (unsigned)m, j0, pcols, Tbuf + p * 1024, C, ldc, 0u, (unsigned)ncols,
                                              static_cast<double *>(scratch), st, launches))
      return rc;
  }
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int qr_transpose_r(const double *Y, size_t ldy, size_t n, double *out, size_t ldo, size_t rows_out, cudaStream_t st) {
  const unsigned tb = (unsigned)((rows_out + 31) / 32), tc = (unsigned)((n + 31) / 32);
  qr_transpose_r_kernel<<<dim3(tb, tc), 256, 0, st>>>(Y, ldy, (unsigned)n, out, ldo, (unsigned)rows_out);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int qr_scatter_rows(const double *in, size_t ldi, double *out, size_t ldo, size_t rows, size_t ncols,
                    const unsigned *perm, cudaStream_t st) {
  if (ncols == 0 || rows == 0) return SVDJ_OK;
  qr_scatter_rows_kernel<<<(unsigned)ncols, 256, 0, st>>>(in, ldi, out, ldo, (unsigned)rows, perm);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

}  // namespace svdj
