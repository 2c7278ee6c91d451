// svdj_reference_api.hpp -- the reference's solver interface, implemented on libsvdj.
//
// Drop-in for the two functions the reference driver calls (/root/reference/main.cu:1497 and :1587):
//
//   void cuda_dgesvd_kernel(SVD_OPTIONS, SVD_OPTIONS, size_t m, size_t n, Matrix &A, size_t lda,
//                           Matrix &s, Matrix &V, size_t ldv);                 // main.cu:163
//   void omp_mpi_cuda_dgesvd_local_matrices(SVD_OPTIONS, SVD_OPTIONS, size_t m, size_t n,
//                           MatrixMPI &A, size_t lda, MatrixMPI &s, MatrixMPI &V, size_t ldv);  // main.cu:440
//
// both at global scope (the copies main.cu calls) and in namespace Thesis (lib/JacobiMethods.cuh:44-62).
// Same argument meaning: A column-major m x n in / U out in place, s = singular values, V = right
// singular vectors (not transposed).  Error behaviour: the reference returns void and ignores every
// failure; these throw std::runtime_error with the library's message instead of continuing silently.
// SVDJ_NOT_CONVERGED returns the current iterate (the reference always returns after its one sweep) but says
// so on stderr, and svdj_shim::last_status() reports the status of the calling thread's last solve.
//
// Several GPUs: the reference's MPI entry point is one collective call with the data on rank 0 (main.cu:1455,
// 1587).  Here omp_mpi_cuda_dgesvd_local_matrices runs the in-library multi-GPU driver (svdj_dgesvd_mg) on
// SVDJ_NGPUS devices of the calling process (environment; default 1, "all" = every visible GPU).
//
// Types: a translation unit that already has the reference's Matrix / MatrixMPI / SVD_OPTIONS
// (e.g. main.cu itself, which defines them at main.cu:17-69 and :152-160) defines
// SVDJ_HAVE_REFERENCE_TYPES before including this header; otherwise layout-compatible ones are
// provided here: fields in the order {width, height, elements}, constructor taking (height, width).
#pragma once

#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

#include "svdj.h"

#ifndef SVDJ_HAVE_REFERENCE_TYPES
// Non-owning-or-owning host matrix with manual release, as the MPI solver uses (main.cu:17-52).
struct MatrixMPI {
  unsigned long width = 0;
  unsigned long height = 0;
  double *elements = nullptr;
  MatrixMPI() = default;
  MatrixMPI(unsigned long rows, unsigned long cols) : width(cols), height(rows), elements(new double[rows * cols]) {}
  void free() {
    delete[] elements;
    elements = nullptr;
  }
};

// Owning host matrix (main.cu:54-69).
struct Matrix {
  unsigned long width;
  unsigned long height;
  double *elements;
  Matrix(unsigned long rows, unsigned long cols) : width(cols), height(rows), elements(new double[rows * cols]) {}
  Matrix(const Matrix &) = delete;
  Matrix &operator=(const Matrix &) = delete;
  ~Matrix() { delete[] elements; }
};

// main.cu:152-156 / lib/JacobiMethods.cuh:25-29
enum SVD_OPTIONS { AllVec = SVDJ_ALLVEC, SomeVec = SVDJ_SOMEVEC, NoVec = SVDJ_NOVEC };
#endif  // SVDJ_HAVE_REFERENCE_TYPES

namespace svdj_shim {

inline int &status_slot() {
  static thread_local int status = SVDJ_OK;
  return status;
}
// Status of the calling thread's most recent solve through this header (SVDJ_OK, SVDJ_NOT_CONVERGED, ...).
inline int last_status() { return status_slot(); }

inline void finish(int rc) {
  status_slot() = rc;
  if (rc == SVDJ_OK) return;
  if (rc == SVDJ_NOT_CONVERGED) {
    std::fprintf(stderr, "libsvdj: warning: max_sweeps reached, the returned factors are the current iterate, not a "
                         "converged SVD (svdj_shim::last_status() == SVDJ_NOT_CONVERGED)\n");
    return;
  }
  throw std::runtime_error("libsvdj: status " + std::to_string(rc) + ": " + svdj_last_error());
}

inline void solve(int jobu, int jobv, std::size_t m, std::size_t n, double *A, std::size_t lda, double *s,
                  double *V, std::size_t ldv, const svdj_opts *opts = nullptr, svdj_stats *stats = nullptr) {
  finish(svdj_dgesvd(jobu, jobv, m, n, A, lda, s, V, ldv, opts, stats));
}

// SVDJ_NGPUS: how many GPUs of this process the distributed entry point uses
inline int ngpus_from_env() {
  const char *e = std::getenv("SVDJ_NGPUS");
  if (e == nullptr || *e == 0) return 1;
  if (!std::strcmp(e, "all")) return 0;
  const int v = std::atoi(e);
  return v < 1 ? 1 : v;
}

inline void solve_mg(int jobu, int jobv, std::size_t m, std::size_t n, double *A, std::size_t lda, double *s,
                     double *V, std::size_t ldv, int ngpus, const svdj_opts *opts = nullptr,
                     svdj_stats *stats = nullptr) {
  if (ngpus == 0) ngpus = svdj_device_count();
  if (ngpus <= 1) return solve(jobu, jobv, m, n, A, lda, s, V, ldv, opts, stats);
  finish(svdj_dgesvd_mg(jobu, jobv, m, n, A, lda, s, V, ldv, opts, stats, ngpus, nullptr));
}

}  // namespace svdj_shim

// main.cu:163.  The reference's single-process entry point.
inline void cuda_dgesvd_kernel(SVD_OPTIONS jobu, SVD_OPTIONS jobv, std::size_t m, std::size_t n, Matrix &A,
                               std::size_t lda, Matrix &s, Matrix &V, std::size_t ldv) {
  svdj_shim::solve(static_cast<int>(jobu), static_cast<int>(jobv), m, n, A.elements, lda, s.elements, V.elements, ldv);
}

// main.cu:440.  In the reference this is a collective over MPI_COMM_WORLD in which only rank 0 holds
// A, s and V (the other ranks pass default-constructed matrices, main.cu:1455).  Here the whole solve
// runs on the GPU(s) of the calling process -- SVDJ_NGPUS of them, through the in-library multi-GPU driver --
// so a call without data (elements == nullptr) returns at once.
inline void omp_mpi_cuda_dgesvd_local_matrices(SVD_OPTIONS jobu, SVD_OPTIONS jobv, std::size_t m, std::size_t n,
                                               MatrixMPI &A, std::size_t lda, MatrixMPI &s, MatrixMPI &V,
                                               std::size_t ldv) {
  if (A.elements == nullptr) return;
  svdj_shim::solve_mg(static_cast<int>(jobu), static_cast<int>(jobv), m, n, A.elements, lda, s.elements, V.elements,
                      ldv, svdj_shim::ngpus_from_env());
}

namespace Thesis {
using ::AllVec;
using ::NoVec;
using ::SomeVec;
using ::SVD_OPTIONS;
inline void cuda_dgesvd_kernel(SVD_OPTIONS jobu, SVD_OPTIONS jobv, std::size_t m, std::size_t n, Matrix &A,
                               std::size_t lda, Matrix &s, Matrix &V, std::size_t ldv) {
  ::cuda_dgesvd_kernel(jobu, jobv, m, n, A, lda, s, V, ldv);
}
inline void omp_mpi_cuda_dgesvd_local_matrices(SVD_OPTIONS jobu, SVD_OPTIONS jobv, std::size_t m, std::size_t n,
                                               MatrixMPI &A, std::size_t lda, MatrixMPI &s, MatrixMPI &V,
                                               std::size_t ldv) {
  ::omp_mpi_cuda_dgesvd_local_matrices(jobu, jobv, m, n, A, lda, s, V, ldv);
}
}  // namespace Thesis
