// svd_jacobi <n> [--no-check] [--seed S] [--dense] [--gpus N]
//
// Driver with the reference's workflow (main.cu:1426-1676): build the n x n input (upper triangle filled
// row by row from std::default_random_engine(1000000), main.cu:1445 and :1559-1567), time the solve
// through the reference's own entry point (here: the shim in include/svdj_reference_api.hpp), print the
// reference's lines ("Dimensions, ...", "SVD MPI+OMP time with U,V calculation: t", "||A-USVt||_F: r",
// main.cu:1583, :1638, :1665), write reporte-dimension-<n>-time-<dd-mm-YYYY-HH-MM-SS>.txt (main.cu:1667),
// and add one JSON line with what the reference never reported (sweeps, rotations, accuracy).
//
// Differences by design: no MPI_Init (one process drives the GPUs: --gpus N or SVDJ_NGPUS, through the in-library
// multi-GPU driver, where the reference uses mpiexec -n P), no "Test 1" on a random unseeded matrix
// (main.cu:1461-1534), thread count is not forced to 36 (main.cu:1431; the report's "Number of threads" line is kept).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <omp.h>
#include <random>
#include <sstream>
#include <string>
#include <vector>

#include "svdj_reference_api.hpp"

static double wall() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char **argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: %s <n> [--no-check] [--seed S] [--dense] [--gpus N]\n", argv[0]);
    return 2;
  }
  const std::size_t n = std::stoul(argv[1]);
  bool check = true, dense = false;
  unsigned seed = 1000000;
  int gpus = svdj_shim::ngpus_from_env();
  for (int i = 2; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--no-check")) check = false;
    else if (!std::strcmp(argv[i], "--dense")) dense = true;
    else if (!std::strcmp(argv[i], "--seed") && i + 1 < argc) seed = (unsigned)std::stoul(argv[++i]);
    else if (!std::strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = std::atoi(argv[++i]);
  }
  const std::size_t height = n, width = n;

  std::time_t t = std::time(nullptr);
  std::tm tmv = *std::localtime(&t);
  std::ostringstream stamp;
  stamp << std::put_time(&tmv, "%d-%m-%Y-%H-%M-%S");
  std::stringstream report;

  MatrixMPI A(height, width), V(width, width), s(1, width), A_copy(height, width);
  std::fill_n(A.elements, height * width, 0.0);
  std::fill_n(V.elements, width * width, 0.0);
  std::fill_n(s.elements, width, 0.0);
  std::default_random_engine engine(seed);
  std::uniform_real_distribution<double> dist(0.0, 1.0);
  for (std::size_t row = 0; row < n; ++row)
    for (std::size_t col = dense ? 0 : row; col < n; ++col) A.elements[col * height + row] = dist(engine);
  std::copy(A.elements, A.elements + height * width, A_copy.elements);

  // main.cu:1581 (omp_get_num_threads() outside a parallel region: 1 in the reference too)
  report << "Number of threads: " << omp_get_num_threads() << '\n';
  report << "Dimensions, height: " << height << ", width: " << width << "\n";
  std::cout << "Dimensions, height: " << height << ", width: " << width << "\n";

  svdj_stats st;
  std::memset(&st, 0, sizeof(st));
  const double ti = wall();
  int rc = 0;
  try {
    // same call as main.cu:1587-1595, plus the stats the reference has no channel for
    svdj_shim::solve_mg(AllVec, AllVec, height, width, A.elements, height, s.elements, V.elements, width, gpus, nullptr,
                        &st);
  } catch (const std::exception &e) {
    std::cerr << e.what() << "\n";
    rc = 1;
  }
  const double time = wall() - ti;
  if (rc) return rc;
  report << "SVD MPI+OMP time with U,V calculation: " << time << "\n";
  std::cout << "SVD MPI+OMP time with U,V calculation: " << time << "\n";

  double fro = -1.0;
  if (check) {
    // ||A - U S V^T||_F as at main.cu:1641-1665
    double acc = 0.0;
#pragma omp parallel for reduction(+ : acc) schedule(static)
    for (long long i = 0; i < (long long)height; ++i) {
      std::vector<double> rowu(width);
      for (std::size_t k = 0; k < width; ++k) rowu[k] = A.elements[k * height + i] * s.elements[k];
      for (std::size_t j = 0; j < width; ++j) {
        double v = 0.0;
        for (std::size_t k = 0; k < width; ++k) v += rowu[k] * V.elements[k * width + j];
        const double d = A_copy.elements[j * height + i] - v;
        acc += d * d;
      }
    }
    fro = std::sqrt(acc);
    report << "||A-USVt||_F: " << fro << "\n";
    std::cout << "||A-USVt||_F: " << fro << "\n";
  }
  const std::string name = "reporte-dimension-" + std::to_string(height) + "-time-" + stamp.str() + ".txt";
  std::ofstream file(name, std::ofstream::out | std::ofstream::trunc);
  file << report.rdbuf();
  file.close();

  std::printf("{\"n\": %zu, \"seconds\": %.6f, \"sweeps\": %d, \"converged\": %d, \"rotations_applied\": %llu, "
              "\"pairs_tested\": %llu, \"last_offdiag\": %.3e, \"seconds_sweeps\": %.6f, \"residual_fro\": %.3e, "
              "\"gpus\": %d, \"report\": \"%s\"}\n",
              n, time, st.sweeps, st.converged, st.applied, st.tested, st.last_offdiag, st.seconds_sweeps, fro,
              gpus, name.c_str());
  A.free();
  V.free();
  s.free();
  A_copy.free();
  return 0;
}
