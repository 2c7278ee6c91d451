// svdj_finalize.cu -- the steps around the sweeps: V = I, sigma = column norms, U = A / sigma,
// descending order.  Restates lib/JacobiMethods.cu:213-226 (V init) and :1146-1164 (Sigma, U), which
// the reference runs on the host; here they stay on the device so that A and V never leave HBM.
#include "svdj_common.cuh"

namespace svdj {

// rows row0 .. row0 + rows - 1 of the identity (row0 != 0: a rank of the row-sharded driver holds a slice of V)
__global__ void set_identity_kernel(double *__restrict__ V, size_t ldv, unsigned rows, unsigned row0) {
  const unsigned col = blockIdx.x;
  double *c = V + (size_t)col * ldv;
  for (unsigned i = threadIdx.x; i < rows; i += blockDim.x) c[i] = (row0 + i == col) ? 1.0 : 0.0;
}

__global__ void square_kernel(double *__restrict__ x, unsigned n) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = x[i] * x[i];
}
__global__ void sqrt_kernel(double *__restrict__ x, unsigned n) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = sqrt(x[i]);
}

__global__ void __launch_bounds__(256) column_norms_kernel(const double *__restrict__ A, size_t lda, unsigned m,
                                                           double *__restrict__ norms) {
  __shared__ double red[3 * 8];
  const double *c = A + (size_t)blockIdx.x * lda;
  double acc = 0.0, z0 = 0.0, z1 = 0.0;
  for (unsigned i = threadIdx.x; i < m; i += 256) {
    const double x = c[i];
    acc = fma(x, x, acc);
  }
  block_sum3<256>(acc, z0, z1, red);
  if (threadIdx.x == 0) norms[blockIdx.x] = sqrt(acc);
}

// out[:, k] = in[:, perm ? perm[k] : k] / (k < ncols_scaled && sigma[k] > 0 ? sigma[k] : 1)
__global__ void gather_columns_kernel(const double *__restrict__ in, size_t ldin, double *__restrict__ out,
                                      size_t ldout, unsigned rows, const unsigned *__restrict__ perm,
                                      const double *__restrict__ sigma, unsigned ncols_scaled) {
  const unsigned k = blockIdx.x;
  const unsigned src = perm ? perm[k] : k;
  const double *ci = in + (size_t)src * ldin;
  double *co = out + (size_t)k * ldout;
  double d = 1.0;
  bool scale = false;
  if (k < ncols_scaled) {
    d = sigma[k];
    scale = d > 0.0;  // the reference divides unguarded (lib/JacobiMethods.cu:1162); a null column stays null here
  }
  if (scale) {
    for (unsigned i = threadIdx.x; i < rows; i += blockDim.x) co[i] = ci[i] / d;
  } else {
    for (unsigned i = threadIdx.x; i < rows; i += blockDim.x) co[i] = ci[i];
  }
}

int launch_square(double *x, size_t n, cudaStream_t st) {
  if (n == 0) return SVDJ_OK;
  square_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, (unsigned)n);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}
int launch_sqrt(double *x, size_t n, cudaStream_t st) {
  if (n == 0) return SVDJ_OK;
  sqrt_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, (unsigned)n);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int launch_set_identity(double *V, size_t ldv, size_t rows, size_t n, cudaStream_t st, size_t row0) {
  if (n == 0) return SVDJ_OK;
  set_identity_kernel<<<(unsigned)n, 256, 0, st>>>(V, ldv, (unsigned)rows, (unsigned)row0);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int launch_column_norms(const double *A, size_t lda, size_t m, size_t n, double *norms, cudaStream_t st) {
  if (n == 0) return SVDJ_OK;
  column_norms_kernel<<<(unsigned)n, 256, 0, st>>>(A, lda, (unsigned)m, norms);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

int launch_gather_columns(const double *in, size_t ldin, double *out, size_t ldout, size_t rows, size_t ncols,
                          const unsigned *perm, const double *sigma, size_t ncols_scaled, cudaStream_t st) {
  if (ncols == 0 || rows == 0) return SVDJ_OK;
  gather_columns_kernel<<<(unsigned)ncols, 256, 0, st>>>(in, ldin, out, ldout, (unsigned)rows, perm, sigma,
                                                         (unsigned)ncols_scaled);
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

}  // namespace svdj
