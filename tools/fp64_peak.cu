// fp64_peak.cu -- measures what the B200 gives for FP64: DFMA issue rate and mma.sync DMMA shapes.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double (&d)[4], const double (&a)[4], const double (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double (&d)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
               : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int MODE>
__global__ void __launch_bounds__(256) k(double *out, int iters, double seed) {
  double acc[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = seed * (i + j);
  double a8[8], b4[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a8[i] = seed + threadIdx.x * 1e-9 + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) b4[i] = seed - threadIdx.x * 1e-9 - i;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {  // DFMA: 64 independent fmas
#pragma unroll
      for (int i = 0; i < 16; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma(acc[i][j], a8[j], b4[j]);
    } else if (MODE == 1) {  // m8n8k4: 16 independent accumulators
#pragma unroll
      for (int i = 0; i < 16; ++i) dmma884(acc[i][0], acc[i][1], a8[i & 7], b4[i & 3]);
    } else if (MODE == 2) {  // m16n8k8
      double a4[4] = {a8[0], a8[1], a8[2], a8[3]};
      double b2[2] = {b4[0], b4[1]};
#pragma unroll
      for (int i = 0; i < 16; ++i) dmma1688(acc[i], a4, b2);
    } else if (MODE == 4) {  // even warps DMMA m8n8k4, odd warps DFMA: do the two pipes run side by side?
      if ((threadIdx.x >> 5) & 1) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fma(acc[i][j], a8[j], b4[j]);
      } else {
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
          for (int i = 0; i < 16; ++i) dmma884(acc[i][0], acc[i][1], a8[i & 7], b4[i & 3]);
      }
    } else if (MODE == 5) {  // every warp interleaves 16 DMMA with 32 DFMA
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        dmma884(acc[i][0], acc[i][1], a8[i & 7], b4[i & 3]);
        acc[i][2] = fma(acc[i][2], a8[i & 7], b4[i & 3]);
        acc[i][3] = fma(acc[i][3], a8[(i + 1) & 7], b4[i & 3]);
      }
    } else {  // m16n8k16
#pragma unroll
      for (int i = 0; i < 16; ++i) dmma16816(acc[i], a8, b4);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += acc[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double flops_per_warp_iter) {
  const int blocks = 148 * 4, threads = 256, iters = 4096;
  double *out;
  cudaMalloc(&out, sizeof(double) * blocks * threads);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<MODE><<<blocks, threads>>>(out, 16, 1.0);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, iters, 1.0);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double warps = (double)blocks * threads / 32;
  double tf = warps * iters * flops_per_warp_iter / (ms * 1e-3) / 1e12;
  printf("%-12s %8.3f ms  %7.2f TFLOP/s  (%s)\n", name, ms, tf, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

int main() {
  run<0>("DFMA", 64.0 * 32 * 2);
  run<1>("DMMA m8n8k4", 16.0 * 8 * 8 * 4 * 2);
  run<2>("DMMA m16n8k8", 16.0 * 16 * 8 * 8 * 2);
  run<3>("DMMA m16n8k16", 16.0 * 16 * 8 * 16 * 2);
  // mode 4: per iteration an even warp does 32 DMMA (16384 flop), an odd warp 64 DFMA (4096 flop): mean per warp
  run<4>("DMMA|DFMA warps", 0.5 * (32.0 * 512 + 64.0 * 64));
  run<5>("DMMA+DFMA mixed", 16.0 * 512 + 32.0 * 64);
  return 0;
}
