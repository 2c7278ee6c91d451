// svdj_block_dev.cuh -- device helpers shared by the block-Jacobi kernels (b = 16: svdj_block.cu, b = 32: svdj_block32.cu)
#pragma once

#include "svdj_common.cuh"

namespace svdj {

static constexpr unsigned kNoCol = 0xffffffffu;

// FP64 tensor-core tile: D(8x8) += A(8x4) B(4x8).  tcgen05 has no FP64 kind; this is the sm_100a FP64 tensor path.
__device__ __forceinline__ void dmma884(double &d0, double &d1, const double a, const double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}


// Same rotation as lib/JacobiMethods.cu:464-477 (tau, t = sgn(tau)/(|tau|+sqrt(1+tau^2)), c = 1/sqrt(1+t^2),
// s = t c), evaluated without divisions through cos(2 theta) = |zeta| / h, zeta = gamma - beta, h = hypot(zeta, 2 alpha):
//   c = sqrt((1 + |zeta|/h) / 2),  s = sgn(zeta) * alpha / (h c).
// Two rsqrt instead of three divisions and two square roots on the critical path of every step.
__device__ __forceinline__ void fast_rotation(double alpha, double beta, double gamma, double &c, double &s) {
  const double zeta = gamma - beta;
  const double a2 = 2.0 * alpha;
  // scale to avoid under/overflow of the squares (Gram entries span the square of the column norms' range)
  // (a power of two built from the exponent bits of the larger magnitude: a handful of integer instructions
  // where ilogb/scalbn cost ~60 with branches, on the critical path of every step)
  const double big = fmax(fabs(zeta), fabs(a2));
  const int ex = (__double2hiint(big) >> 20) & 0x7ff;
  const double scale = __hiloint2double((2046 - ex) << 20, 0);  // 2^(1023 - ex): max(|z|, |y|) lands in [1, 2)
  const double z = zeta * scale, y = a2 * scale;
  const double rh = rsqrt(z * z + y * y);
  const double cc = 0.5 + 0.5 * fabs(z) * rh;
  const double rcc = rsqrt(cc);
  c = cc * rcc;
  s = (zeta >= 0.0 ? 0.5 : -0.5) * y * rh * rcc;
}


// Sum of `splits` partial values src[sp * stride], sp = 0 .. splits-1, added in that order.  The loads of a group of
// eight are issued before the first add: written as `s += src[...]` in a loop every add stalls on its own load and
// the loads of a thread run one L2 latency after the other (3 us of the eigen-solve's set-up at 4 splits, ~10 at 16).
__device__ __forceinline__ double sum_partials(const double *__restrict__ src, size_t stride, unsigned splits) {
  double s = 0.0;
  for (unsigned sp = 0; sp < splits; sp += 8) {
    double v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = (sp + u < splits) ? src[(size_t)(sp + u) * stride] : 0.0;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (sp + u < splits) s += v[u];
  }
  return s;
}

// Programmatic dependent launch (PDL): a kernel launched with the programmatic-serialization attribute may start
// while its predecessor in the stream is still draining; griddep_wait() returns once the predecessor has completed
// and its memory is visible, griddep_launch() lets the successor's CTAs be scheduled as soon as SM resources free
// up.  Everything before griddep_wait() must not touch global memory the predecessor reads or writes.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src, int src_bytes) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}


// Columns live in up to two buffers (multi-GPU drivers keep the two column blocks that a rank holds in
// separate allocations so that one can travel while the other stays): column c < split is in the first.
__device__ __forceinline__ double *col_ptr(const BlockArgs &a, bool is_a, unsigned c) {
  // A2 / V2 arrive pre-biased by -split * ld (blk_launch_step), so both halves index with c itself
  double *base = is_a ? (c < a.split ? a.A : a.A2) : (c < a.split ? a.V : a.V2);
  return base + (size_t)c * (is_a ? a.lda : a.ldv);
}

// Where an updated column is written.  Normally the place it was read from; in the last step of a ring
// round the drivers point the stores of a travelling column block at the neighbour GPU's receive buffer
// (a peer mapping over NVLink), so that the block exchange rides on the update kernel's own stores.
__device__ __forceinline__ double *col_ptr_store(const BlockArgs &a, bool is_a, unsigned c) {
  double *base = is_a ? (c < a.split ? a.A_st : a.A2_st) : (c < a.split ? a.V_st : a.V2_st);
  return base + (size_t)c * (is_a ? a.lda : a.ldv);
}

// Block pair idx of step k (1-based).
//   sched 0: the reference's round-robin over all nb blocks (every pair once per nb-1 steps);
//   sched 1: bipartite, block i of the first half with block (i + k - 1) mod (nb/2) of the second half (every
//            cross pair once per nb/2 steps) -- the CROSS phase of a ring rank;
//   sched 2: two-level tournament on one GPU, the ordering the multi-GPU ring walks with P = vranks ranks: the
//            blocks form 2P super-blocks of wb = nb/(2P) blocks; steps 1..2wb-1 sweep every pair of super-blocks
//            (r, P+r) internally (round 0), then rounds 1..2P-2 of wb steps each rotate only the pairs across
//            the two super-blocks a (virtual) rank holds after the circle-method move.  Same nb-1 steps and
//            nb/2 pairs per step as sched 0, every pair once per sweep; it needs fewer sweeps.
__device__ __forceinline__ Pair block_pair(const BlockArgs &a, unsigned k, unsigned idx) {
  if (a.sched == 0) {
    Pair r = sameh_pair(a.nb, k, idx);
    if (r.p > r.q) {  // the lower block index first: the eigen-solve orders the columns of a pair towards it
      const unsigned t = r.p;
      r.p = r.q;
      r.q = t;
    }
    return r;
  }
  if (a.sched == 1) {
    const unsigned half = a.nb >> 1;
    Pair r;
    r.p = idx;
    r.q = half + (idx + k - 1u) % half;
    return r;
  }
  return two_level_pair(a.nb, (unsigned)a.vranks, k, idx);
}

}  // namespace svdj
