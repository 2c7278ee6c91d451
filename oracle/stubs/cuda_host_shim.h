/* Host stand-ins for the handful of CUDA runtime names /root/reference/main.cu uses, so that the
 * reference solver can be built with g++ and run on CPU cores: "device" buffers are heap buffers,
 * cudaMemcpy is memcpy, and a kernel launch becomes a loop that runs the reference's own kernel
 * body once per thread index.  Force-included by oracle/Makefile.  TEST INFRASTRUCTURE ONLY. */
#ifndef SVDJ_CUDA_HOST_SHIM_H_
#define SVDJ_CUDA_HOST_SHIM_H_
#include <cmath>
#include <cstdlib>
#include <cstring>
#define __global__
#define __host__
#define __device__
struct shim_uint3 { unsigned x = 1, y = 1, z = 1; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
  dim3(double x_) : x((unsigned)x_), y(1), z(1) {}
};
static thread_local shim_uint3 blockDim, blockIdx, threadIdx;
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
template <class T> static inline int cudaMalloc(T **p, size_t bytes) { *p = (T *)std::malloc(bytes); return 0; }
static inline int cudaFree(void *p) { std::free(p); return 0; }
static inline int cudaMemcpy(void *d, const void *s, size_t bytes, cudaMemcpyKind) { std::memcpy(d, s, bytes); return 0; }
static inline int cudaMemset(void *d, int v, size_t bytes) { std::memset(d, v, bytes); return 0; }
/* jacobi_rotation<<<grid, block>>>(n, x, y, c, s) is rewritten by the Makefile into
 * SHIM_LAUNCH(jacobi_rotation, grid, block, n, x, y, c, s). */
#define SHIM_LAUNCH(kernel, grid, block, ...)                                   \
  do {                                                                          \
    dim3 g_ = (grid);                                                           \
    unsigned b_ = (unsigned)(block);                                            \
    blockDim.x = b_;                                                            \
    for (unsigned bi_ = 0; bi_ < g_.x; ++bi_) {                                 \
      blockIdx.x = bi_;                                                         \
      for (unsigned ti_ = 0; ti_ < b_; ++ti_) {                                 \
        threadIdx.x = ti_;                                                      \
        kernel(__VA_ARGS__);                                                    \
      }                                                                         \
    }                                                                           \
  } while (0)
#endif
