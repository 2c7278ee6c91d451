// svdj_qr.cuh -- blocked Householder QR / apply-Q launchers (svdj_qr.cu), used by the preconditioned solve
#pragma once

#include "svdj_common.cuh"

namespace svdj {

bool qr_supported(size_t m, size_t n);     // n <= m <= 65536 rows (the panel lives in registers across <= 128 CTAs)
size_t qr_num_panels(size_t n);            // panels of 32 columns
size_t qr_scratch_bytes(size_t n);         // device scratch of qr_factor on n columns / of qr_apply_q on <= n columns
// Y (m x n, ld) -> Householder vectors below the diagonal, R on and above; Tbuf: qr_num_panels(n) x 1024 doubles
int qr_factor(double *Y, size_t ld, size_t m, size_t n, double *Tbuf, void *scratch, cudaStream_t st,
              unsigned long long *launches);
// C (m x ncols, ldc) <- Q C with Q = H_1 ... H_n from qr_factor
int qr_apply_q(const double *Y, size_t ld, size_t m, size_t n, const double *Tbuf, double *C, size_t ldc, size_t ncols,
               void *scratch, cudaStream_t st, unsigned long long *launches);
// out (rows_out x n, ldo) = R^T, zero above the diagonal and in the pad rows n .. rows_out-1
int qr_transpose_r(const double *Y, size_t ldy, size_t n, double *out, size_t ldo, size_t rows_out, cudaStream_t st);
// out[perm[i]][k] = in[i][k] for i < rows, k < ncols
int qr_scatter_rows(const double *in, size_t ldi, double *out, size_t ldo, size_t rows, size_t ncols,
                    const unsigned *perm, cudaStream_t st);

}  // namespace svdj
