// oracle_input.cpp -- the reference driver's input matrix, regenerated on the same libstdc++.
//
// TEST INFRASTRUCTURE ONLY (see svdj_oracle.h).
//
// Restates main.cu:1445 (seed 1000000) and main.cu:1559-1567: an n x n zero matrix whose upper
// triangle (row index <= column index) is filled row by row with uniform(0,1) draws from
// std::default_random_engine.  The engine and distribution are library types, so the only way to
// get the same bits is to use the same library; that is why this one file is C++.
#include <cstddef>
#include <random>

#include "svdj_oracle.h"

extern "C" void oracle_fill_reference_input(size_t n, unsigned seed, double *A) {
  for (size_t i = 0; i < n * n; ++i) A[i] = 0.0;
  std::default_random_engine engine(seed);
  std::uniform_real_distribution<double> dist(0.0, 1.0);
  for (size_t row = 0; row < n; ++row) {
    for (size_t col = row; col < n; ++col) {
      A[col * n + row] = dist(engine);  // iteratorC(row, col, n), lib/global.cuh:14
    }
  }
}
