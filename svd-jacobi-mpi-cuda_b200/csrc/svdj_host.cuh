// svdj_host.cuh -- host-side helpers shared by the drivers (svdj_api.cu, svdj_mg.cu): pooled buffers, RAII wrappers
#pragma once

#include <functional>
#include <mutex>
#include <vector>

#include "svdj_common.cuh"

namespace svdj {

// Work buffers come from a small process-wide pool (per device for device memory): cudaMalloc, cudaFree and
// cudaMallocHost are device-wide synchronising driver calls that can stall for ~100 ms behind any other driver
// client (a monitoring query is enough), which used to land inside the solve.  A buffer returns to the pool when
// its RAII holder dies; svdj_release_workspace() gives everything idle back to the driver.
struct Pool {
  struct Entry {
    void *p;
    size_t bytes;
    int device;  // -1 for pinned host memory
    bool busy;
  };
  std::mutex mu;
  std::vector<Entry> entries;

  void *acquire(size_t bytes, int device, cudaError_t *err) {
    std::lock_guard<std::mutex> lock(mu);
    Entry *best = nullptr;
    for (auto &e : entries) {
      if (!e.busy && e.device == device && e.bytes >= bytes && e.bytes <= 2 * bytes + 4096 &&
          (best == nullptr || e.bytes < best->bytes))
        best = &e;
    }
    if (best) {
      best->busy = true;
      *err = cudaSuccess;
      return best->p;
    }
    void *p = nullptr;
    *err = (device < 0) ? cudaMallocHost(&p, bytes) : cudaMalloc(&p, bytes);
    if (*err != cudaSuccess) return nullptr;
    entries.push_back({p, bytes, device, true});
    return p;
  }
  void release(void *p) {
    std::lock_guard<std::mutex> lock(mu);
    for (auto &e : entries)
      if (e.p == p) e.busy = false;
  }
  void trim() {
    std::lock_guard<std::mutex> lock(mu);
    std::vector<Entry> keep;
    int cur = -1;
    cudaGetDevice(&cur);
    for (auto &e : entries) {
      if (e.busy) {
        keep.push_back(e);
      } else if (e.device < 0) {
        cudaFreeHost(e.p);
      } else {
        cudaSetDevice(e.device);
        cudaFree(e.p);
      }
    }
    if (cur >= 0) cudaSetDevice(cur);
    entries.swap(keep);
  }
};
extern Pool g_pool;  // svdj_api.cu

// RAII for device / pinned buffers and events so that every early return releases them
struct DevBuf {
  void *p = nullptr;
  ~DevBuf() {
    if (p) g_pool.release(p);
  }
  int alloc(size_t bytes) {
    if (bytes == 0) bytes = 8;
    int dev = 0;
    SVDJ_CUDA_TRY(cudaGetDevice(&dev));
    cudaError_t e;
    p = g_pool.acquire(bytes, dev, &e);
    if (p == nullptr) {
      set_error("cudaMalloc(%zu bytes) -> %s", bytes, cudaGetErrorString(e));
      cudaGetLastError();
      return SVDJ_ERR_ALLOC;
    }
    return SVDJ_OK;
  }
  template <class T>
  T *as() const {
    return static_cast<T *>(p);
  }
};

struct PinBuf {
  void *p = nullptr;
  ~PinBuf() {
    if (p) g_pool.release(p);
  }
  int alloc(size_t bytes) {
    cudaError_t e;
    p = g_pool.acquire(bytes ? bytes : 8, -1, &e);
    if (p == nullptr) {
      set_error("cudaMallocHost(%zu bytes) -> %s", bytes, cudaGetErrorString(e));
      cudaGetLastError();
      return SVDJ_ERR_ALLOC;
    }
    return SVDJ_OK;
  }
  template <class T>
  T *as() const {
    return static_cast<T *>(p);
  }
};

struct Event {
  cudaEvent_t e = nullptr;
  ~Event() {
    if (e) cudaEventDestroy(e);
  }
  int create() {
    SVDJ_CUDA_TRY(cudaEventCreate(&e));
    return SVDJ_OK;
  }
};

struct DeviceGuard {
  int prev = -1;
  bool active = false;
  int enter(int dev) {
    if (dev < 0) return SVDJ_OK;
    SVDJ_CUDA_TRY(cudaGetDevice(&prev));
    if (prev != dev) {
      SVDJ_CUDA_TRY(cudaSetDevice(dev));
      active = true;
    }
    return SVDJ_OK;
  }
  ~DeviceGuard() {
    if (active) cudaSetDevice(prev);
  }
};

// The QR-preconditioned solve (svdj_api.cu).  inner(pre_status, jobu, jobv, m, n, A, lda, s, V, ldv, vrows, opts,
// stats) runs the Jacobi iteration on the triangular factor; it is called exactly once.
typedef std::function<int(int, int, int, size_t, size_t, double *, size_t, double *, double *, size_t, size_t,
                          const svdj_opts &, svdj_stats &)>
    InnerSolve;
int solve_with_qr(int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds, double *dV, size_t ldv,
                  int qr_steps, const svdj_opts &o, svdj_stats &st, cudaStream_t stream, const InnerSolve &inner);
// QR steps that opts ask for on an m x n problem (0 when the front-end does not apply)
int precondition_steps(const svdj_opts &o, size_t m, size_t n);

void mg_release_team();  // svdj_mg.cu: tears down the cached ranks of svdj_dgesvd_mg

}  // namespace svdj
