// DMMA (mma.sync.m8n8k4.f64) rate against resident warps per SM and independent accumulators per warp.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
template <int NACC>
__global__ void k(double *out, int iters, double seed) {
  double acc[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i][0] = acc[i][1] = seed * i;
  double a = seed + threadIdx.x * 1e-9, b = seed - threadIdx.x * 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma884(acc[i][0], acc[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += acc[i][0] + acc[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
void run(int warps_per_sm) {
  const int threads = 32 * warps_per_sm, blocks = 148, iters = 8192 * 16 / NACC;
  double *out; cudaMalloc(&out, sizeof(double) * blocks * threads);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<NACC><<<blocks, threads>>>(out, 16, 1.0); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<NACC><<<blocks, threads>>>(out, iters, 1.0); cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double tf = (double)blocks * warps_per_sm * iters * NACC * 512.0 / (ms * 1e-3) / 1e12;
  printf("warps/SM %2d  acc/warp %2d: %6.2f TFLOP/s\n", warps_per_sm, NACC, tf);
  cudaFree(out);
}
int main() {
  for (int w : {4, 8, 16, 32}) { run<16>(w); run<8>(w); run<4>(w); run<2>(w); }
  return 0;
}
