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
template <int R>
__global__ void __launch_bounds__(kPanelThreads) qr_panel_kernel(double *__restrict__ Y, const size_t ld,
                                                                 const unsigned m, const unsigned j0,
                                                                 const unsigned pcols, double *__restrict__ Tout,
                                                                 double *__restrict__ part,   // [2][G][32]
                                                                 double *__restrict__ pivot,  // [2][32]
                                                                 unsigned *__restrict__ bar) {
  __shared__ double red[8][32];
  __shared__ double coef[32];   // tau * w_j for the columns right of the pivot
  __shared__ double scal[2];    // scale = 1 / (alpha - beta), beta
  __shared__ double Z[32][33];  // CTA 0: z_k of every step (column c of Z), then T
  __shared__ double taus[32];
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned G = gridDim.x, cta = blockIdx.x;
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

#pragma unroll
  for (int c = 0; c < 32; ++c) {
    if ((unsigned)c < pcols) {  // grid-uniform
      // products of the pivot column (rows below the pivot) with all 32 columns
      double p[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) p[j] = 0.0;
#pragma unroll
      for (int rr = 0; rr < R; ++rr) {
        const double x = (ridx[rr] > (unsigned)c && ridx[rr] < mj) ? row[rr][c] : 0.0;
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
        part[((size_t)(c & 1) * G + cta) * 32 + lane] = s;
        if (cta == 0 && (int)lane == c) {  // the pivot row of this step: register set 0 of this thread
#pragma unroll
          for (int j = 0; j < 32; ++j) pivot[(c & 1) * 32 + j] = row[0][j];
        }
      }
      grid_barrier(bar, (unsigned)(c + 1) * G);
      {
        double s = 0.0;
        for (unsigned g = warp; g < G; g += 8) s += __ldcg(part + ((size_t)(c & 1) * G + g) * 32 + lane);
        red[warp][lane] = s;
      }
      __syncthreads();
      if (warp == 0) {
        double S = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) S += red[w][lane];
        const double pv = __ldcg(pivot + (c & 1) * 32 + lane);
        const double alpha = __shfl_sync(0xffffffffu, pv, c);
        const double xn2 = __shfl_sync(0xffffffffu, S, c);
        double beta = alpha, tau = 0.0, scale = 0.0;
        if (xn2 > 0.0) {
          const double nrm = sqrt(fma(alpha, alpha, xn2));
          beta = alpha >= 0.0 ? -nrm : nrm;
          tau = (beta - alpha) / beta;
          scale = 1.0 / (alpha - beta);
        }
        const double wj = fma(S, scale, pv);  // j > c: v . y_j;  j < c: v_j . v
        coef[lane] = ((int)lane > c) ? tau * wj : 0.0;
        if (lane == 0) {
          scal[0] = scale;
          scal[1] = beta;
        }
        if (cta == 0) {
          if ((int)lane < c) Z[lane][c] = wj;
          if (lane == 0) taus[c] = tau;
        }
      }
      __syncthreads();
      const double scale = scal[0], beta = scal[1];
#pragma unroll
      for (int rr = 0; rr < R; ++rr) {
        if (ridx[rr] < mj) {
          if (ridx[rr] > (unsigned)c) {
            const double v = row[rr][c] * scale;
            row[rr][c] = v;
#pragma unroll
            for (int j = c + 1; j < 32; ++j) row[rr][j] = fma(-coef[j], v, row[rr][j]);
          } else if (ridx[rr] == (unsigned)c) {
            row[rr][c] = beta;
#pragma unroll
            for (int j = c + 1; j < 32; ++j) row[rr][j] -= coef[j];
          }
        }
      }
    }
  }

#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    if (ridx[rr] < mj) {
      double *dst = Y + (size_t)j0 * ld + j0 + ridx[rr];
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if ((unsigned)j < pcols) dst[(size_t)j * ld] = row[rr][j];
    }
  }

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
// Block reflector on a strip of 32 columns per CTA:
//   TRANS_T = true :  C <- (I - V T^T V^T) C   (Q^T C: the trailing update of the factorisation)
//   TRANS_T = false:  C <- (I - V T   V^T) C   (Q C:  applying Q)
// V = the unit lower trapezoidal panel stored in Y (rows j0.., columns j0 .. j0+pcols-1), C = rows j0.. of
// columns c0 + 32 blockIdx.x ... of Cmat.  W = V^T C and C -= V M run as FP64 DMMA tiles (mma.sync m8n8k4).
// ---------------------------------------------------------------------------------------------------------
static constexpr int kRefThreads = 256;
static constexpr size_t kRefSmem = sizeof(double) * 8 * 1024;  // the 8 warps' partial W

__device__ __forceinline__ double panel_v(const double *__restrict__ Yp, size_t ldy, unsigned i, unsigned k,
                                          unsigned mj, unsigned pcols) {
  // V(i, k): rows and columns relative to the panel's corner
  if (i >= mj || k >= pcols || i < k) return 0.0;
  if (i == k) return 1.0;
  return Yp[(size_t)k * ldy + i];
}

template <bool TRANS_T>
__global__ void __launch_bounds__(kRefThreads) qr_block_reflector_kernel(const double *__restrict__ Y, const size_t ldy,
                                                                         const unsigned m, const unsigned j0,
                                                                         const unsigned pcols,
                                                                         const double *__restrict__ T,
                                                                         double *__restrict__ Cmat, const size_t ldc,
                                                                         const unsigned c0, const unsigned ncols) {
  extern __shared__ __align__(16) double red[];  // [8 warps][16 tiles][64]
  __shared__ double Ts[32][33];
  __shared__ double Ws[32][33];
  __shared__ double Ms[32][36];  // Ms[n][k] = -M[k][n]
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned g = lane >> 2, tg = lane & 3u;
  const unsigned mj = m - j0;
  const unsigned cs = c0 + 32u * blockIdx.x;
  const unsigned sw = min(32u, c0 + ncols - cs);
  const double *Yp = Y + (size_t)j0 * ldy + j0;
  double *Cp = Cmat + (size_t)cs * ldc + j0;

  for (unsigned i = tid; i < 1024; i += kRefThreads) Ts[i & 31][i >> 5] = T[i];  // Ts[a][c] = T[a + 32 c]

  // ---- pass 1: W = V^T C ----
  double acc[4][4][2];
#pragma unroll
  for (int ti = 0; ti < 4; ++ti)
#pragma unroll
    for (int tj = 0; tj < 4; ++tj) acc[ti][tj][0] = acc[ti][tj][1] = 0.0;
  const unsigned nks = (mj + 3u) >> 2;
  for (unsigned ks = warp; ks < nks; ks += 8) {
    const unsigned ri = 4u * ks + tg;
    const bool ok = ri < mj;
    double a[4], b[4];
    if (4u * ks < 32u) {
#pragma unroll
      for (int ti = 0; ti < 4; ++ti) a[ti] = panel_v(Yp, ldy, ri, 8u * ti + g, mj, pcols);
    } else {
#pragma unroll
      for (int ti = 0; ti < 4; ++ti) a[ti] = (ok && 8u * ti + g < pcols) ? Yp[(size_t)(8u * ti + g) * ldy + ri] : 0.0;
    }
#pragma unroll
    for (int tj = 0; tj < 4; ++tj) b[tj] = (ok && 8u * tj + g < sw) ? Cp[(size_t)(8u * tj + g) * ldc + ri] : 0.0;
#pragma unroll
    for (int ti = 0; ti < 4; ++ti)
#pragma unroll
      for (int tj = 0; tj < 4; ++tj) dmma884(acc[ti][tj][0], acc[ti][tj][1], a[ti], b[tj]);
  }
#pragma unroll
  for (int ti = 0; ti < 4; ++ti)
#pragma unroll
    for (int tj = 0; tj < 4; ++tj) {
      red[(warp * 16 + ti * 4 + tj) * 64 + 2 * lane + 0] = acc[ti][tj][0];
      red[(warp * 16 + ti * 4 + tj) * 64 + 2 * lane + 1] = acc[ti][tj][1];
    }
  __syncthreads();
  for (unsigned idx = tid; idx < 1024; idx += kRefThreads) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w * 1024 + idx];
    const unsigned tile = idx >> 6, l = (idx & 63u) >> 1, e = idx & 1u;
    Ws[8u * (tile >> 2) + (l >> 2)][8u * (tile & 3u) + 2u * (l & 3u) + e] = s;
  }
  __syncthreads();
  // ---- M = T^T W (factor) or T W (apply), stored negated and transposed ----
  {
    const unsigned k = tid >> 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const unsigned n = 4u * (tid & 7u) + q;
      double s = 0.0;
      if (TRANS_T) {
        for (unsigned a = 0; a <= k; ++a) s = fma(Ts[a][k], Ws[a][n], s);
      } else {
        for (unsigned a = k; a < 32; ++a) s = fma(Ts[k][a], Ws[a][n], s);
      }
      Ms[n][k] = -s;
    }
  }
  __syncthreads();
  // ---- pass 2: C += V (-M) ----
  const unsigned nrb = (mj + 7u) >> 3;
  for (unsigned rb = warp; rb < nrb; rb += 8) {
    const unsigned ri = 8u * rb + g;
    const bool ok = ri < mj;
    double a[8];
    if (8u * rb < 32u) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) a[ks] = panel_v(Yp, ldy, ri, 4u * ks + tg, mj, pcols);
    } else {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) a[ks] = (ok && 4u * ks + tg < pcols) ? Yp[(size_t)(4u * ks + tg) * ldy + ri] : 0.0;
    }
    double d[4][2];
#pragma unroll
    for (int tj = 0; tj < 4; ++tj) {
      const unsigned col = 8u * tj + 2u * tg;
      d[tj][0] = (ok && col < sw) ? Cp[(size_t)col * ldc + ri] : 0.0;
      d[tj][1] = (ok && col + 1 < sw) ? Cp[(size_t)(col + 1) * ldc + ri] : 0.0;
    }
#pragma unroll
    for (int tj = 0; tj < 4; ++tj)
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) dmma884(d[tj][0], d[tj][1], a[ks], Ms[8 * tj + g][4 * ks + tg]);
#pragma unroll
    for (int tj = 0; tj < 4; ++tj) {
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
static constexpr unsigned kMaxPanelCtas = 128;  // co-resident by a wide margin on 148 SMs

bool qr_supported(size_t m, size_t n) { return n >= 1 && m >= n && m <= (size_t)kMaxPanelCtas * kPanelThreads * 2; }

size_t qr_num_panels(size_t n) { return (n + kQB - 1) / kQB; }

size_t qr_scratch_bytes(size_t n) {
  // part [2][G][32] + pivot [2][32] doubles, one barrier word per panel
  return sizeof(double) * (2 * (size_t)kMaxPanelCtas * 32 + 64) + sizeof(unsigned) * (qr_num_panels(n) + 2);
}

static int reflector_setup() {
  static DeviceOnce once;
  SVDJ_CUDA_TRY(once.run([] {
    cudaError_t e = cudaFuncSetAttribute(qr_block_reflector_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kRefSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(qr_block_reflector_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)kRefSmem);
    return e;
  }));
  return SVDJ_OK;
}

int qr_factor(double *Y, size_t ld, size_t m, size_t n, double *Tbuf, void *scratch, cudaStream_t st,
              unsigned long long *launches) {
  if (!qr_supported(m, n)) {
    set_error("qr_factor: %zu x %zu is outside the supported range (n <= m <= %u)", m, n,
              kMaxPanelCtas * kPanelThreads * 2);
    return SVDJ_ERR_UNSUPPORTED;
  }
  if (int rc = reflector_setup()) return rc;
  double *part = static_cast<double *>(scratch);
  double *pivot = part + 2 * (size_t)kMaxPanelCtas * 32;
  unsigned *bar = reinterpret_cast<unsigned *>(pivot + 64);
  const size_t np = qr_num_panels(n);
  SVDJ_CUDA_TRY(cudaMemsetAsync(bar, 0, sizeof(unsigned) * np, st));
  for (size_t p = 0; p < np; ++p) {
    unsigned j0 = (unsigned)(p * kQB);
    unsigned pcols = (unsigned)std::min<size_t>(kQB, n - j0);
    unsigned mu = (unsigned)m;
    const size_t mj = m - j0;
    const int R = (mj > (size_t)kMaxPanelCtas * kPanelThreads) ? 2 : 1;
    unsigned G = (unsigned)((mj + (size_t)kPanelThreads * R - 1) / ((size_t)kPanelThreads * R));
    if (G < 1) G = 1;
    double *Tp = Tbuf + p * 1024;
    unsigned *barp = bar + p;
    void *args[] = {&Y, &ld, &mu, &j0, &pcols, &Tp, &part, &pivot, &barp};
    if (R == 1)
      SVDJ_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)qr_panel_kernel<1>, dim3(G), dim3(kPanelThreads), args, 0, st));
    else
      SVDJ_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)qr_panel_kernel<2>, dim3(G), dim3(kPanelThreads), args, 0, st));
    if (launches) *launches += 1;
    const size_t c0 = (size_t)j0 + pcols;
    if (c0 < n) {
      const unsigned ncols = (unsigned)(n - c0);
      qr_block_reflector_kernel<true><<<(ncols + 31) / 32, kRefThreads, kRefSmem, st>>>(Y, ld, mu, j0, pcols, Tp, Y, ld,
                                                                                       (unsigned)c0, ncols);
      if (launches) *launches += 1;
    }
  }
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int qr_apply_q(const double *Y, size_t ld, size_t m, size_t n, const double *Tbuf, double *C, size_t ldc, size_t ncols,
               cudaStream_t st, unsigned long long *launches) {
  if (ncols == 0 || n == 0) return SVDJ_OK;
  if (int rc = reflector_setup()) return rc;
  const size_t np = qr_num_panels(n);
  for (size_t p = np; p-- > 0;) {
    const unsigned j0 = (unsigned)(p * kQB);
    const unsigned pcols = (unsigned)std::min<size_t>(kQB, n - j0);
    qr_block_reflector_kernel<false><<<(unsigned)((ncols + 31) / 32), kRefThreads, kRefSmem, st>>>(
        Y, ld, (unsigned)m, j0, pcols, Tbuf + p * 1024, C, ldc, 0u, (unsigned)ncols);
    if (launches) *launches += 1;
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
