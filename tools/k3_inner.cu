// k3_inner.cu -- ceiling of the update kernel's inner loop: the exact LDS + DMMA pattern of blk32_update_kernel on
// shared-memory-resident data (no global traffic, no tile changes), 2 CTAs of 8 warps per SM.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/k3_inner tools/k3_inner.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
constexpr int kStride = 68, kTile = 64 * kStride;

// MODE 0: as in the kernel (2 LDS.64 + 2 LDS.128 per 8 DMMA); 1: B fragments hoisted (registers), A from smem;
// 2: everything in registers (pure DMMA with 8 accumulators); 3: like 0 with a __syncthreads() every 16 k-steps
template <int MODE>
__global__ void __launch_bounds__(256, 2) k(double *out, int tiles) {
  extern __shared__ __align__(16) double sm[];
  double *tile_s = sm, *Wsm = sm + 2 * kTile;
  for (int i = threadIdx.x; i < 2 * kTile + 4096; i += 256) sm[i] = 1e-3 * (i % 97);
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3, rowgroup = warp & 3, nhalf = warp >> 2;
  double d[2][4][2];
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) d[0][nt][0] = d[0][nt][1] = d[1][nt][0] = d[1][nt][1] = 0.0;
  const double2 *wl = reinterpret_cast<const double2 *>(Wsm + nhalf * 16 * 128 + 2 * lane);
  for (int t = 0; t < tiles; ++t) {
    const double *tile = tile_s + (t & 1) * kTile + 16 * rowgroup + g;
#pragma unroll 8
    for (int ks = 0; ks < 16; ++ks) {
      double a0, a1;
      double2 b01, b23;
      if (MODE == 2) {
        a0 = 1.0 + ks; a1 = 2.0 + ks; b01 = make_double2(0.5, 0.25); b23 = make_double2(0.125, 0.75);
      } else {
        a0 = tile[(4 * ks + tg) * kStride];
        a1 = tile[(4 * ks + tg) * kStride + 8];
        if (MODE == 1) { b01 = make_double2(0.5, 0.25); b23 = make_double2(0.125, 0.75); }
        else { b01 = wl[ks * 64]; b23 = wl[ks * 64 + 32]; }
      }
      dmma884(d[0][0][0], d[0][0][1], a0, b01.x);
      dmma884(d[1][0][0], d[1][0][1], a1, b01.x);
      dmma884(d[0][1][0], d[0][1][1], a0, b01.y);
      dmma884(d[1][1][0], d[1][1][1], a1, b01.y);
      dmma884(d[0][2][0], d[0][2][1], a0, b23.x);
      dmma884(d[1][2][0], d[1][2][1], a1, b23.x);
      dmma884(d[0][3][0], d[0][3][1], a0, b23.y);
      dmma884(d[1][3][0], d[1][3][1], a1, b23.y);
    }
    if (MODE == 3) __syncthreads();
  }
  double s = 0;
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) s += d[0][nt][0] + d[0][nt][1] + d[1][nt][0] + d[1][nt][1];
  out[blockIdx.x * 256 + threadIdx.x] = s;
}

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
// Streaming variant: tiles of 64 rows x 64 columns of a column-major matrix with `ld` rows come in through the
// kernel's 2-stage cp.async pipeline (LOAD), results go back with the kernel's store pattern (STORE).
__device__ unsigned g_sm_slot[256];
__device__ unsigned g_smid_of_block[1024];
template <bool LOAD, int STORE>
__global__ void __launch_bounds__(256, 2) ks(double *X, size_t ld, int tiles_per_cta, int phase_mode, unsigned phase_ns) {
  extern __shared__ __align__(16) double sm[];
  double *tile_s = sm, *Wsm = sm + 2 * kTile;
  for (int i = threadIdx.x; i < 2 * kTile + 4096; i += 256) sm[i] = 1e-3 * (i % 97);
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3, rowgroup = warp & 3, nhalf = warp >> 2;
  const double2 *wl = reinterpret_cast<const double2 *>(Wsm + nhalf * 16 * 128 + 2 * lane);
  {
    __shared__ unsigned slot_s;
    if (threadIdx.x == 0) {
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      g_smid_of_block[blockIdx.x] = smid;
      slot_s = (phase_mode == 3) ? (atomicAdd(&g_sm_slot[smid], 1u) & 1u)
               : (phase_mode == 2) ? (blockIdx.x & 1u) : (phase_mode == 1) ? (blockIdx.x >= gridDim.x / 2) : 0u;
    }
    __syncthreads();
    if (slot_s && phase_ns) __nanosleep(phase_ns);
  }
  // CTA b owns a contiguous range of 64-row tiles of a (ld x 64 x panels) matrix: tile t -> panel t / (ld/64)
  const size_t tpp = ld / 64;
  const size_t t0 = (size_t)blockIdx.x * tiles_per_cta;
  auto issue = [&](size_t t, int stage) {
    size_t panel = t / tpp, r0 = (t % tpp) * 64;
    if (phase_mode == 9) { panel &= 1; r0 &= 1023; }  // loads from an L2-resident window
    double *dst = tile_s + stage * kTile + 2 * lane;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const unsigned slot = warp + 8 * i;
      cp_async16(dst + slot * kStride, X + (panel * 64 + slot) * ld + r0 + 2 * lane);
    }
  };
  if (LOAD) issue(t0, 0);
  asm volatile("cp.async.commit_group;\n" ::: "memory");
  for (int t = 0; t < tiles_per_cta; ++t) {
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();
    if (LOAD && t + 1 < tiles_per_cta) issue(t0 + t + 1, (t + 1) & 1);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    const double *tile = tile_s + (t & 1) * kTile + 16 * rowgroup + g;
    double d[2][4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) d[0][nt][0] = d[0][nt][1] = d[1][nt][0] = d[1][nt][1] = 0.0;
#pragma unroll 8
    for (int ks2 = 0; ks2 < 16; ++ks2) {
      const double a0 = tile[(4 * ks2 + tg) * kStride], a1 = tile[(4 * ks2 + tg) * kStride + 8];
      const double2 b01 = wl[ks2 * 64], b23 = wl[ks2 * 64 + 32];
      if (STORE == 4) {  // transposed roles: D[n][row], a lane ends up with two consecutive rows of one column
        dmma884(d[0][0][0], d[0][0][1], b01.x, a0);
        dmma884(d[1][0][0], d[1][0][1], b01.x, a1);
        dmma884(d[0][1][0], d[0][1][1], b01.y, a0);
        dmma884(d[1][1][0], d[1][1][1], b01.y, a1);
        dmma884(d[0][2][0], d[0][2][1], b23.x, a0);
        dmma884(d[1][2][0], d[1][2][1], b23.x, a1);
        dmma884(d[0][3][0], d[0][3][1], b23.y, a0);
        dmma884(d[1][3][0], d[1][3][1], b23.y, a1);
      } else {
      dmma884(d[0][0][0], d[0][0][1], a0, b01.x);
      dmma884(d[1][0][0], d[1][0][1], a1, b01.x);
      dmma884(d[0][1][0], d[0][1][1], a0, b01.y);
      dmma884(d[1][1][0], d[1][1][1], a1, b01.y);
      dmma884(d[0][2][0], d[0][2][1], a0, b23.x);
      dmma884(d[1][2][0], d[1][2][1], a1, b23.x);
      dmma884(d[0][3][0], d[0][3][1], a0, b23.y);
      dmma884(d[1][3][0], d[1][3][1], a1, b23.y);
      }
    }
    const size_t tt = t0 + t, panel = tt / tpp, r0 = (tt % tpp) * 64 + 16 * rowgroup;
    if (STORE == 1 || STORE == 2 || (STORE == 0 && ld == 12345)) {  // STORE 0: never taken, but keeps the arithmetic alive
      const size_t pn = (STORE == 2) ? (panel & 1) : panel;   // 2: everything lands in two panels (L2-resident)
      const size_t rr = (STORE == 2) ? (r0 & 1023) : r0;
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const unsigned s0 = 32 * nhalf + 8 * nt + 2 * tg;
          X[(pn * 64 + s0) * ld + rr + 8 * h + g] = d[h][nt][0];
          X[(pn * 64 + s0 + 1) * ld + rr + 8 * h + g] = d[h][nt][1];
        }
    } else if (STORE == 4) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const unsigned col = 32 * nhalf + 8 * nt + g;
          *reinterpret_cast<double2 *>(X + (panel * 64 + col) * ld + r0 + 8 * h + 2 * tg) = make_double2(d[h][nt][0], d[h][nt][1]);
        }
    } else if (STORE == 3) {
      // stage the 64 x 64 result in the tile buffer just consumed (everyone must be done reading it), then write
      // it out as 512-byte column pieces with 16-byte stores
      __syncthreads();
      double *stg = tile_s + (t & 1) * kTile;
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const unsigned s0 = 32 * nhalf + 8 * nt + 2 * tg;
          stg[s0 * kStride + 16 * rowgroup + 8 * h + g] = d[h][nt][0];
          stg[(s0 + 1) * kStride + 16 * rowgroup + 8 * h + g] = d[h][nt][1];
        }
      __syncthreads();
      const size_t rt = (tt % tpp) * 64;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const unsigned slot = warp + 8 * i;
        const double2 v = *reinterpret_cast<const double2 *>(stg + slot * kStride + 2 * lane);
        *reinterpret_cast<double2 *>(X + (panel * 64 + slot) * ld + rt + 2 * lane) = v;
      }
      __syncthreads();  // the staging buffer is the next tile's landing zone... (issued after this barrier in a real kernel)
    } else {
      asm volatile("" ::"d"(d[0][0][0]), "d"(d[0][1][0]), "d"(d[0][2][0]), "d"(d[0][3][0]), "d"(d[1][0][0]), "d"(d[1][1][0]), "d"(d[1][2][0]), "d"(d[1][3][0]));
      asm volatile("" ::"d"(d[0][0][1]), "d"(d[0][1][1]), "d"(d[0][2][1]), "d"(d[0][3][1]), "d"(d[1][0][1]), "d"(d[1][1][1]), "d"(d[1][2][1]), "d"(d[1][3][1]));
    }
  }
}

// Variant: warp tile 32 rows x 16 output columns (2 row groups x 4 column quarters), the warp's W fragments
// (16 k-steps x 2 n-tiles = 32 doubles... per lane) held in registers: only A fragments come from shared memory.
template <bool LOAD, bool STORE>
__global__ void __launch_bounds__(256, 2) ks2(double *X, size_t ld, int tiles_per_cta) {
  extern __shared__ __align__(16) double sm[];
  double *tile_s = sm, *Wsm = sm + 2 * kTile;
  for (int i = threadIdx.x; i < 2 * kTile + 4096; i += 256) sm[i] = 1e-3 * (i % 97);
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3, rowgroup = warp & 1, nq = warp >> 1;
  double wf[16][2];
#pragma unroll
  for (int ks2_ = 0; ks2_ < 16; ++ks2_) {
    wf[ks2_][0] = Wsm[(16 * nq + g) * 64 + 4 * ks2_ + tg];
    wf[ks2_][1] = Wsm[(16 * nq + 8 + g) * 64 + 4 * ks2_ + tg];
  }
  const size_t tpp = ld / 64;
  const size_t t0 = (size_t)blockIdx.x * tiles_per_cta;
  auto issue = [&](size_t t, int stage) {
    const size_t panel = t / tpp, r0 = (t % tpp) * 64;
    double *dst = tile_s + stage * kTile + 2 * lane;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const unsigned slot = warp + 8 * i;
      cp_async16(dst + slot * kStride, X + (panel * 64 + slot) * ld + r0 + 2 * lane);
    }
  };
  if (LOAD) issue(t0, 0);
  asm volatile("cp.async.commit_group;\n" ::: "memory");
  for (int t = 0; t < tiles_per_cta; ++t) {
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();
    if (LOAD && t + 1 < tiles_per_cta) issue(t0 + t + 1, (t + 1) & 1);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    const double *tile = tile_s + (t & 1) * kTile + 32 * rowgroup + g;
    double d[4][2][2];
#pragma unroll
    for (int h = 0; h < 4; ++h) d[h][0][0] = d[h][0][1] = d[h][1][0] = d[h][1][1] = 0.0;
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
      double af[4];
#pragma unroll
      for (int h = 0; h < 4; ++h) af[h] = tile[(4 * k2 + tg) * kStride + 8 * h];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        dmma884(d[h][0][0], d[h][0][1], af[h], wf[k2][0]);
        dmma884(d[h][1][0], d[h][1][1], af[h], wf[k2][1]);
      }
    }
    const size_t tt = t0 + t, panel = tt / tpp, r0 = (tt % tpp) * 64 + 32 * rowgroup;
    if (STORE || ld == 12345) {
#pragma unroll
      for (int h = 0; h < 4; ++h)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const unsigned s0 = 16 * nq + 8 * nt + 2 * tg;
          X[(panel * 64 + s0) * ld + r0 + 8 * h + g] = d[h][nt][0];
          X[(panel * 64 + s0 + 1) * ld + r0 + 8 * h + g] = d[h][nt][1];
        }
    }
  }
}

template <bool LOAD, bool STORE>
void run_stream2(const char *name) {
  const size_t ld = 8192, panels = 64;
  const int blocks = 296, tpc = (int)(ld / 64 * panels / blocks);
  const size_t smem = sizeof(double) * (2 * kTile + 4096);
  cudaFuncSetAttribute(ks2<LOAD, STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  double *X;
  cudaMalloc(&X, sizeof(double) * ld * 64 * panels);
  cudaMemset(X, 0, sizeof(double) * ld * 64 * panels);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  ks2<LOAD, STORE><<<blocks, 256, smem>>>(X, ld, tpc);
  cudaDeviceSynchronize();
  float best = 1e9f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    ks2<LOAD, STORE><<<blocks, 256, smem>>>(X, ld, tpc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
  }
  const double flops = (double)blocks * tpc * 8 * 16 * 8 * 512.0;
  printf("%-44s %8.1f us  %6.2f TFLOP/s  (%s)\n", name, best * 1e3, flops / (best * 1e-3) / 1e12,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(X);
}

// ---- bulk-async (TMA) variant: tile loads as 64 x cp.async.bulk (512 B column pieces, completion on an mbarrier),
// results written back into the consumed stage and stored with 64 x cp.async.bulk shared->global: no LDGSTS and no
// STG instructions in the warps' instruction streams, 16 STS per thread and tile instead.
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ unsigned mbar_try_wait(unsigned long long *bar, unsigned parity) {
  unsigned ok;
  asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok;
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ int g_tma_timeout;

__global__ void __launch_bounds__(256, 2) ks_tma(double *X, size_t ld, int tiles_per_cta) {
  extern __shared__ __align__(128) double sm[];
  double *tile_s = sm, *Wsm = sm + 2 * kTile;
  __shared__ __align__(8) unsigned long long full[2];
  for (int i = threadIdx.x; i < 2 * kTile + 4096; i += 256) sm[i] = 1e-3 * (i % 97);
  if (threadIdx.x == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the generic-proxy initialisation of the stages
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3, rowgroup = warp & 3, nhalf = warp >> 2;
  const double2 *wl = reinterpret_cast<const double2 *>(Wsm + nhalf * 16 * 128 + 2 * lane);
  const size_t tpp = ld / 64;
  const size_t t0 = (size_t)blockIdx.x * tiles_per_cta;
  auto issue = [&](size_t t, int stage) {  // threads 0..63: one 512-byte column piece each
    if (threadIdx.x < 64) {
      const size_t panel = t / tpp, r0 = (t % tpp) * 64;
      if (threadIdx.x == 0) mbar_expect_tx(&full[stage], 64 * 512);
      bulk_g2s(tile_s + stage * kTile + threadIdx.x * kStride, X + (panel * 64 + threadIdx.x) * ld + r0, 512, &full[stage]);
    }
  };
  issue(t0, 0);
  bool dead = false;
  for (int t = 0; t < tiles_per_cta && !dead; ++t) {
    const int stage = t & 1;
    if (t + 1 < tiles_per_cta) {
      // the other stage still feeds the bulk store of tile t-1 (this thread's own column piece): wait for its reads
      if (threadIdx.x < 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      issue(t0 + t + 1, stage ^ 1);
    }
    {
      const unsigned parity = (unsigned)(t >> 1) & 1u;
      int spins = 0;
      while (!mbar_try_wait(&full[stage], parity)) {
        if (++spins > (1 << 22)) { dead = true; g_tma_timeout = 1; break; }
      }
    }
    const double *tile = tile_s + stage * kTile + 16 * rowgroup + g;
    double d[2][4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) d[0][nt][0] = d[0][nt][1] = d[1][nt][0] = d[1][nt][1] = 0.0;
#pragma unroll 8
    for (int ks2 = 0; ks2 < 16; ++ks2) {
      const double a0 = tile[(4 * ks2 + tg) * kStride], a1 = tile[(4 * ks2 + tg) * kStride + 8];
      const double2 b01 = wl[ks2 * 64], b23 = wl[ks2 * 64 + 32];
      dmma884(d[0][0][0], d[0][0][1], a0, b01.x);
      dmma884(d[1][0][0], d[1][0][1], a1, b01.x);
      dmma884(d[0][1][0], d[0][1][1], a0, b01.y);
      dmma884(d[1][1][0], d[1][1][1], a1, b01.y);
      dmma884(d[0][2][0], d[0][2][1], a0, b23.x);
      dmma884(d[1][2][0], d[1][2][1], a1, b23.x);
      dmma884(d[0][3][0], d[0][3][1], a0, b23.y);
      dmma884(d[1][3][0], d[1][3][1], a1, b23.y);
    }
    __syncthreads();  // everyone has read the input tile: it becomes the output tile
    double *stg = tile_s + stage * kTile;
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const unsigned s0 = 32 * nhalf + 8 * nt + 2 * tg;
        stg[s0 * kStride + 16 * rowgroup + 8 * h + g] = d[h][nt][0];
        stg[(s0 + 1) * kStride + 16 * rowgroup + 8 * h + g] = d[h][nt][1];
      }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 64) {
      const size_t tt = t0 + t, panel = tt / tpp, r0 = (tt % tpp) * 64;
      bulk_s2g(X + (panel * 64 + threadIdx.x) * ld + r0, stg + threadIdx.x * kStride, 512);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (threadIdx.x < 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__global__ void __launch_bounds__(256, 2) ks_tma_wreg(double *X, size_t ld, int tiles_per_cta) {
  extern __shared__ __align__(128) double sm[];
  double *tile_s = sm, *Wsm = sm + 2 * kTile;
  __shared__ __align__(8) unsigned long long full[2];
  for (int i = threadIdx.x; i < 2 * kTile + 4096; i += 256) sm[i] = 1e-3 * (i % 97);
  if (threadIdx.x == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the generic-proxy initialisation of the stages
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3, rowgroup = warp & 1, nq = warp >> 1;
  double wf[16][2];
#pragma unroll
  for (int k2 = 0; k2 < 16; ++k2) {
    wf[k2][0] = Wsm[(16 * nq + g) * 64 + 4 * k2 + tg];
    wf[k2][1] = Wsm[(16 * nq + 8 + g) * 64 + 4 * k2 + tg];
  }
  const size_t tpp = ld / 64;
  const size_t t0 = (size_t)blockIdx.x * tiles_per_cta;
  auto issue = [&](size_t t, int stage) {  // threads 0..63: one 512-byte column piece each
    if (threadIdx.x < 64) {
      const size_t panel = t / tpp, r0 = (t % tpp) * 64;
      if (threadIdx.x == 0) mbar_expect_tx(&full[stage], 64 * 512);
      bulk_g2s(tile_s + stage * kTile + threadIdx.x * kStride, X + (panel * 64 + threadIdx.x) * ld + r0, 512, &full[stage]);
    }
  };
  issue(t0, 0);
  bool dead = false;
  for (int t = 0; t < tiles_per_cta && !dead; ++t) {
    const int stage = t & 1;
    if (t + 1 < tiles_per_cta) {
      // the other stage still feeds the bulk store of tile t-1 (this thread's own column piece): wait for its reads
      if (threadIdx.x < 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      issue(t0 + t + 1, stage ^ 1);
    }
    {
      const unsigned parity = (unsigned)(t >> 1) & 1u;
      int spins = 0;
      while (!mbar_try_wait(&full[stage], parity)) {
        if (++spins > (1 << 22)) { dead = true; g_tma_timeout = 1; break; }
      }
    }
    const double *tile = tile_s + stage * kTile + 32 * rowgroup + g;
    double d[4][2][2];
#pragma unroll
    for (int h = 0; h < 4; ++h) d[h][0][0] = d[h][0][1] = d[h][1][0] = d[h][1][1] = 0.0;
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
      double af[4];
#pragma unroll
      for (int h = 0; h < 4; ++h) af[h] = tile[(4 * k2 + tg) * kStride + 8 * h];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        dmma884(d[h][0][0], d[h][0][1], af[h], wf[k2][0]);
        dmma884(d[h][1][0], d[h][1][1], af[h], wf[k2][1]);
      }
    }
    __syncthreads();  // everyone has read the input tile: it becomes the output tile
    double *stg = tile_s + stage * kTile;
#pragma unroll
    for (int h = 0; h < 4; ++h)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const unsigned s0 = 16 * nq + 8 * nt + 2 * tg;
        stg[s0 * kStride + 32 * rowgroup + 8 * h + g] = d[h][nt][0];
        stg[(s0 + 1) * kStride + 32 * rowgroup + 8 * h + g] = d[h][nt][1];
      }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 64) {
      const size_t tt = t0 + t, panel = tt / tpp, r0 = (tt % tpp) * 64;
      bulk_s2g(X + (panel * 64 + threadIdx.x) * ld + r0, stg + threadIdx.x * kStride, 512);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (threadIdx.x < 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

void run_tma(const char *name, void (*kern)(double *, size_t, int) = ks_tma) {
  const size_t ld = 8192, panels = 64;
  const int blocks = 296, tpc = (int)(ld / 64 * panels / blocks);
  const size_t smem = sizeof(double) * (2 * kTile + 4096);
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  double *X;
  cudaMalloc(&X, sizeof(double) * ld * 64 * panels);
  cudaMemset(X, 0, sizeof(double) * ld * 64 * panels);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  kern<<<blocks, 256, smem>>>(X, ld, tpc);
  cudaError_t err = cudaDeviceSynchronize();
  int flag = 0;
  cudaMemcpyFromSymbol(&flag, g_tma_timeout, sizeof(int));
  if (err != cudaSuccess || flag) {
    printf("%-44s FAILED (%s, mbarrier timeout flag %d)\n", name, cudaGetErrorString(err), flag);
    return;
  }
  float best = 1e9f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    kern<<<blocks, 256, smem>>>(X, ld, tpc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
  }
  const double flops = (double)blocks * tpc * 8 * 16 * 8 * 512.0;
  printf("%-44s %8.1f us  %6.2f TFLOP/s  (%s)\n", name, best * 1e3, flops / (best * 1e-3) / 1e12,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(X);
}

template <bool LOAD, int STORE>
void run_stream(const char *name, int phase_mode = 0, unsigned phase_ns = 0) {
  const size_t ld = 8192, panels = 64;                 // 8192 x 4096 doubles = 268 MB, as A and V of the 4096^2 case
  const int blocks = 296, tpc = (int)(ld / 64 * panels / blocks);  // 27 tiles per CTA
  const size_t smem = sizeof(double) * (2 * kTile + 4096);
  cudaFuncSetAttribute(ks<LOAD, STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  double *X;
  cudaMalloc(&X, sizeof(double) * ld * 64 * panels);
  cudaMemset(X, 0, sizeof(double) * ld * 64 * panels);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  ks<LOAD, STORE><<<blocks, 256, smem>>>(X, ld, tpc, phase_mode, phase_ns);
  cudaDeviceSynchronize();
  float best = 1e9f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    ks<LOAD, STORE><<<blocks, 256, smem>>>(X, ld, tpc, phase_mode, phase_ns);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
  }
  const double flops = (double)blocks * tpc * 8 * 16 * 8 * 512.0;
  printf("%-44s %8.1f us  %6.2f TFLOP/s  (%s)\n", name, best * 1e3, flops / (best * 1e-3) / 1e12,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(X);
}

template <int MODE>
void run(const char *name) {
  const int blocks = 148 * 2, tiles = 2000;
  const size_t smem = sizeof(double) * (2 * kTile + 4096);
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  double *out;
  cudaMalloc(&out, sizeof(double) * blocks * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<MODE><<<blocks, 256, smem>>>(out, 10);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, 256, smem>>>(out, tiles);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double flops = (double)blocks * 8 * tiles * 16 * 8 * 512.0;
  printf("%-44s %8.3f ms  %6.2f TFLOP/s  (%s)\n", name, ms, flops / (ms * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

int main() {
  run<0>("kernel pattern: 2 LDS.64 + 2 LDS.128 / 8 DMMA");
  run<1>("A fragments from smem, B in registers");
  run<2>("all operands in registers");
  run<3>("kernel pattern + __syncthreads per tile");
  run_stream<false, 0>("27 tiles per CTA, no global traffic");
  run_stream<true, 0>("+ cp.async tile loads (268 MB)");
  run_stream<false, 1>("+ result stores (268 MB to HBM)");
  run_stream<false, 2>("+ result stores into an L2-resident window");
  run_stream<true, 1>("+ loads and stores (the update kernel)");
  run_stream<true, 2>("+ loads, stores into an L2-resident window");
  run_stream<false, 3>("+ stores staged through smem, 16 B coalesced");
  run_stream<true, 0>("+ loads from an L2-resident window, no stores", 9, 0);
  run_stream<true, 2>("+ loads and stores, both L2-resident windows", 9, 0);
  run_stream<false, 4>("+ stores, transposed roles, 16 B per lane");
  run_stream<true, 4>("+ loads and stores, transposed roles, 16 B");
  run_tma("bulk-async (TMA) loads and stores via smem");
  run_tma("TMA + W in registers (32x16 warp tiles)", ks_tma_wreg);
  run_stream2<false, false>("W in registers, 32x16 warp tiles: no traffic");
  run_stream2<true, false>("W in registers: + loads");
  run_stream2<false, true>("W in registers: + stores");
  run_stream2<true, true>("W in registers: + loads and stores");
  return 0;
}
