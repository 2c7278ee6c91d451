// svdj_block32.cu -- the block-Jacobi step with column blocks of 32 (block pairs of 64 columns).
//
// Same three kernels as svdj_block.cu (Gram, eigen-solve, update), twice as wide.  Per sweep the flops are
// the same (10 n^3 at m = n) but there are half as many steps, so half the HBM traffic (a step always touches
// all of A and V once), half the launches and half the eigen-solve latencies; the update kernel moves from the
// HBM roofline to the FP64 tensor roofline (0.31 b flop/B = 10 at b = 32, ridge 5.7).  Aligned layouts only
// (even leading dimensions, 16-byte aligned bases): the callers fall back to the b = 16 kernels otherwise.
#include <cstdlib>

#include "svdj_block_dev.cuh"

namespace svdj {

static constexpr int kB32 = 32;
static constexpr int kD64 = 64;                 // columns of a block pair
static constexpr int kTG = 8;                   // 8 x 8 tiles per side
static constexpr int kGramTiles32 = 36;         // upper triangle of the 8 x 8 grid of tiles
static constexpr int kGramPart32 = kGramTiles32 * 64;
static constexpr int kMaxPairs32 = 2048;

__device__ __forceinline__ unsigned slot_col32(unsigned I, unsigned J, unsigned s, unsigned n) {
  const unsigned c = (s < kB32) ? I * kB32 + s : J * kB32 + (s - kB32);
  return c < n ? c : kNoCol;
}

// ------------------------------------------------------------------------------------------------
// K1: partial Gram matrices of the 64-column panels.  grid (block pairs, row splits), 4 warps.  Tiles of
// 64 rows x 64 columns come in through a 3-stage cp.async pipeline.  The 36 output tiles do not fit one
// warp's registers, so warps (0,2) accumulate the even-numbered tiles and warps (1,3) the odd ones, each
// over half of the tile's rows.
// ------------------------------------------------------------------------------------------------
static constexpr int kG32Rows = 64;
static constexpr int kG32Stride = kG32Rows + 4;
static constexpr int kG32TileDoubles = kD64 * kG32Stride;
static constexpr int kG32Stages = 3;

template <int GSEL>
__device__ __forceinline__ void gram32_ksteps(const double *__restrict__ tile, double (&acc)[18][2]) {
  // tile points at (row 32 rh + tg, column g) of the stage; 8 k-steps of 4 rows
#pragma unroll 2
  for (int ks = 0; ks < 8; ++ks) {
    double f[kTG];
#pragma unroll
    for (int gi = 0; gi < kTG; ++gi) f[gi] = tile[(8 * gi) * kG32Stride + 4 * ks];
    int t = 0;
#pragma unroll
    for (int ti = 0; ti < kTG; ++ti)
#pragma unroll
      for (int tj = ti; tj < kTG; ++tj) {
        if ((t & 1) == GSEL) dmma884(acc[t >> 1][0], acc[t >> 1][1], f[ti], f[tj]);
        ++t;
      }
  }
}

// CROSS: only the 16 tiles of X_p^T X_q (the cross block of the Gram matrix); the two diagonal blocks then come
// from the per-block cache the eigen-solve maintains (BlockArgs::gcache).  Every warp accumulates all 16 tiles over
// a quarter of the tile's rows.  Output layout: tile (ti, tj) at gpart[(4 ti + tj) * 64 ...], same fragment order.
template <bool CROSS>
__global__ void __launch_bounds__(128) blk32_gram_kernel(const BlockArgs a, const unsigned k,
                                                         const unsigned rows_per_split,
                                                         double *__restrict__ gpart) {
  extern __shared__ __align__(16) double tiles[];  // kG32Stages x 64 columns x kG32Stride; reused for the reduction
  const Pair bp = block_pair(a, k, blockIdx.x);
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3;
  const unsigned r0 = blockIdx.y * rows_per_split;  // a multiple of 64
  const unsigned r1 = min(a.m, r0 + rows_per_split);
  const unsigned ntiles = r0 < r1 ? (r1 - r0 + kG32Rows - 1) / kG32Rows : 0;

  const double *cptr[16];  // this thread copies column slots warp + 4 i (nullptr: past the last column)
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const unsigned c = slot_col32(bp.p, bp.q, warp + 4 * i, a.n);
    cptr[i] = (c != kNoCol) ? col_ptr(a, true, c) : nullptr;
  }
  griddep_launch();
  griddep_wait();  // the columns come from the previous step's update kernel
  auto issue_tile = [&](unsigned t, int stage) {
    const unsigned row = r0 + t * kG32Rows + 2 * lane;
    double *dst = tiles + stage * kG32TileDoubles + 2 * lane;
    const int rb = (row < r1) ? ((row + 1 < r1) ? 16 : 8) : 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const bool ok = cptr[i] != nullptr && rb != 0;
      cp_async16(dst + (warp + 4 * i) * kG32Stride, ok ? cptr[i] + row : a.A, ok ? rb : 0);
    }
  };

  unsigned issued = 0;
#pragma unroll
  for (int s = 0; s < kG32Stages - 1; ++s) {
    if (issued < ntiles) {
      issue_tile(issued, s);
      ++issued;
    }
    cp_async_commit();
  }
  constexpr int kAcc = CROSS ? 16 : 18;
  double acc[kAcc][2];
#pragma unroll
  for (int t = 0; t < kAcc; ++t) acc[t][0] = acc[t][1] = 0.0;
  const unsigned gsel = warp & 1, rh = warp >> 1;

  for (unsigned t = 0; t < ntiles; ++t) {
    cp_async_wait<kG32Stages - 2>();
    __syncthreads();
    if (issued < ntiles) {
      issue_tile(issued, (int)(issued % kG32Stages));
      ++issued;
    }
    cp_async_commit();
    if constexpr (CROSS) {
      const double *tile = tiles + (t % kG32Stages) * kG32TileDoubles + g * kG32Stride + 16 * warp + tg;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        double f[kTG];
#pragma unroll
        for (int gi = 0; gi < kTG; ++gi) f[gi] = tile[(8 * gi) * kG32Stride + 4 * ks];
#pragma unroll
        for (int ti = 0; ti < 4; ++ti)
#pragma unroll
          for (int tj = 0; tj < 4; ++tj) dmma884(acc[4 * ti + tj][0], acc[4 * ti + tj][1], f[ti], f[4 + tj]);
      }
    } else {
      const double *tile = tiles + (t % kG32Stages) * kG32TileDoubles + g * kG32Stride + 32 * rh + tg;
      if (gsel == 0) gram32_ksteps<0>(tile, acc);
      else gram32_ksteps<1>(tile, acc);
    }
  }
  cp_async_wait<0>();
  __syncthreads();  // the tile buffers become the reduction buffer: [4 warps][kAcc tiles][64]
  double *red = tiles;
#pragma unroll
  for (int t = 0; t < kAcc; ++t) {
    red[(warp * kAcc + t) * 64 + lane * 2 + 0] = acc[t][0];
    red[(warp * kAcc + t) * 64 + lane * 2 + 1] = acc[t][1];
  }
  __syncthreads();
  double *dst = gpart + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * kGramPart32;
  if constexpr (CROSS) {
    for (unsigned i = threadIdx.x; i < 16 * 64; i += 128)
      dst[i] = (red[i] + red[16 * 64 + i]) + (red[2 * 16 * 64 + i] + red[3 * 16 * 64 + i]);
  } else {
    for (unsigned i = threadIdx.x; i < kGramPart32; i += 128) {
      const unsigned t = i >> 6, e = i & 63;
      const unsigned gs = t & 1, lt = t >> 1;  // tile t lives in warps gs and gs + 2 as local tile lt
      dst[i] = red[((gs)*18 + lt) * 64 + e] + red[((gs + 2) * 18 + lt) * 64 + e];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K2: eigen-decomposition of the 64 x 64 Gram matrix, cyclic two-sided Jacobi in shared memory, 512 threads:
// warp j1, lane j2 owns the 2 x 2 block G[{p1,q1}][{p2,q2}]; each warp computes the 32 rotations of a step in
// its 32 lanes (lane = pair index) and fetches its own row rotation by shuffle; one barrier per step.
// ------------------------------------------------------------------------------------------------
static constexpr int kE32Stride = kD64 + 2;  // odd (stride - 1): conflict-free rotation inputs, see svdj_block.cu
static constexpr size_t kE32Smem = sizeof(double) * 5 * kD64 * kE32Stride;
static constexpr int kE32Threads = 512;  // 16 warps: 128 registers per thread for the 4 x 4 register blocks of the CROSS loop

__device__ __forceinline__ void circle_pair64(int t, int j, int &p, int &q) {
  int x, y;
  if (j == 0) {
    x = t;
    y = kD64 - 1;
  } else {
    x = t + j;
    x = x >= kD64 - 1 ? x - (kD64 - 1) : x;
    y = t - j;
    y = y < 0 ? y + (kD64 - 1) : y;
  }
  p = min(x, y);
  q = max(x, y);
}

__global__ void __launch_bounds__(kE32Threads) blk32_eig_kernel(const BlockArgs a, const unsigned k, const unsigned splits,
                                                         const double *__restrict__ gpart, double *__restrict__ wbuf,
                                                         int *__restrict__ wflag) {
  extern __shared__ __align__(16) double sm[];
  double(*G0)[kE32Stride] = reinterpret_cast<double(*)[kE32Stride]>(sm);
  double(*G1)[kE32Stride] = reinterpret_cast<double(*)[kE32Stride]>(sm + kD64 * kE32Stride);
  double(*W)[kE32Stride] = reinterpret_cast<double(*)[kE32Stride]>(sm + 2 * kD64 * kE32Stride);
  double(*T)[kE32Stride] = reinterpret_cast<double(*)[kE32Stride]>(sm + 3 * kD64 * kE32Stride);
  double(*W1)[kE32Stride] = reinterpret_cast<double(*)[kE32Stride]>(sm + 4 * kD64 * kE32Stride);
  __shared__ unsigned long long fresh_max;
  __shared__ int rank_of[kD64];
  __shared__ int valid[kD64];
  __shared__ double rot_c[32], rot_s[32], rin_al[32], rin_be[32], rin_ga[32];
  __shared__ int sweep_rot_s;
  const unsigned tid = threadIdx.x, lane = tid & 31;
  const unsigned pair = blockIdx.x;

  griddep_launch();
  griddep_wait();  // the partial Gram matrices come from K1
  const double *src = gpart + (size_t)pair * splits * kGramPart32;
  const bool cached = a.gcache != nullptr && a.gram_mode == 2;
  if (cached) {
    // cross block from this step's partial sums (16 tiles), the two diagonal blocks from the cache: they are the
    // rotated Gram blocks the eigen-solves of the blocks' previous pairs left behind (refreshed from the columns
    // themselves in the first step of every sweep, which is also the only step that tests intra-block pairs)
    const Pair bp = block_pair(a, k, pair);
    for (unsigned i = tid; i < 16 * 64; i += kE32Threads) {
      const double s = sum_partials(src + i, kGramPart32, splits);
      const int t = i >> 6, l = (i & 63) >> 1, e = i & 1;
      const int r = 8 * (t >> 2) + (l >> 2), c = 32 + 8 * (t & 3) + 2 * (l & 3) + e;
      G0[r][c] = s;
      G0[c][r] = s;
    }
    const double *cp = a.gcache + (size_t)bp.p * (kB32 * kB32), *cq = a.gcache + (size_t)bp.q * (kB32 * kB32);
    for (unsigned i = tid; i < kB32 * kB32; i += kE32Threads) {
      const int r = i >> 5, c = i & 31;
      G0[r][c] = cp[i];
      G0[32 + r][32 + c] = cq[i];
    }
  } else {
    for (unsigned i = tid; i < kGramPart32; i += kE32Threads) {
      const double s = sum_partials(src + i, kGramPart32, splits);
      const int t = i >> 6, l = (i & 63) >> 1, e = i & 1;
      int ti = 0, rem = t;  // t -> (ti, tj): rows of 8, 7, ..., 1 tiles
      while (rem >= kTG - ti) {
        rem -= kTG - ti;
        ++ti;
      }
      const int tj = ti + rem;
      const int r = 8 * ti + (l >> 2), c = 8 * tj + 2 * (l & 3) + e;
      G0[r][c] = s;
      if (ti != tj) G0[c][r] = s;
    }
  }
  const bool cross = a.cross_only != 0;
  if (!cross)  // the CROSS loop keeps W in registers and writes it out at the end
    for (unsigned i = tid; i < kD64 * kD64; i += kE32Threads) W[i >> 6][i & 63] = ((i >> 6) == (i & 63)) ? 1.0 : 0.0;
  if (tid == 0) fresh_max = 0ull;
  __syncthreads();

  // One pass over the pairs this schedule visits (CROSS: one pair per thread): is there anything to rotate (the
  // same test the steps apply), and the largest cosine of the fresh matrix (what lib/JacobiMethods.cu:462
  // computes per pair), as max of alpha^2 / (beta gamma) with one square root per warp.
  const double tol2 = a.tol * a.tol;
  int any = 0;
  {
    double mx2 = 0.0;
    auto look = [&](int r, int c) {
      const double al = G0[r][c], be = G0[r][r], ga = G0[c][c];
      if (be > a.floor2 && ga > a.floor2) {
        any |= (al * al > tol2 * be * ga) ? 1 : 0;
        mx2 = fmax(mx2, al * al / (be * ga));
      }
    };
    if (cross) {
      for (unsigned i = tid; i < kB32 * kB32; i += kE32Threads) look((int)(i >> 5), 32 + (int)(i & 31));
    } else {
      for (unsigned i = tid; i < kD64 * kD64; i += kE32Threads)
        if ((i >> 6) < (i & 63)) look((int)(i >> 6), (int)(i & 63));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx2 = fmax(mx2, __shfl_xor_sync(0xffffffffu, mx2, o));
    if (lane == 0) atomicMax(&fresh_max, (unsigned long long)__double_as_longlong(sqrt(mx2)));
  }

  // Systolic layout.  Rows and columns of G (and the columns of W) are kept in SLOT order: slot j and slot 32 + j
  // always hold the two indices of pair j, so thread (warp w, lane l) always works on G[{w, 32+w}][{l, 32+l}] --
  // fixed, conflict-free addresses (with the circle-method indices themselves half of all shared-memory
  // wavefronts were bank conflicts).  The round-robin move of the indices is folded into the write-back: each
  // updated value is stored at the slot its index occupies in the next step (top 0 stays, top j -> top j+1,
  // top 31 -> bottom 31, bottom j -> bottom j-1, bottom 0 -> top 1), into the other buffer.  After the 63 steps
  // of a sweep every index is back in its slot, so slot order = index order at every sweep boundary.
  // Two schedules.  FULL: all 2016 pairs of the 64 indices in 63 steps (circle method).  CROSS: only the 1024
  // pairs with one index in each column block, in 32 steps (the bottom slots rotate, the top slots stay).  A
  // sweep uses FULL for its first step, which orthogonalises every block internally, and CROSS afterwards:
  // every column pair is then rotated once per sweep, as in the reference's ordering, instead of re-rotating
  // the 992 intra-block pairs in each of the sweep's steps.
  const int nsteps = cross ? 32 : kD64 - 1;
  auto next_slot = [cross](int p) {
    if (cross) return p < 32 ? p : 32 + ((p - 32 + 31) & 31);
    return p == 0 ? 0 : (p < 31 ? p + 1 : (p == 31 ? 63 : (p == 32 ? 1 : p - 1)));
  };
  int total_rot = 0;
  int cur = 0;
  const int w1 = tid >> 5, l2 = lane;
  const int nc_t = next_slot(l2), nc_b = next_slot(32 + l2);
  if (cross && tid < 32) {  // rotation inputs of the first CROSS step
    rin_al[tid] = G0[tid][32 + tid];
    rin_be[tid] = G0[tid][tid];
    rin_ga[tid] = G0[32 + tid][32 + tid];
  }
  any = __syncthreads_or(any);
  if (cross && any) {
    // CROSS schedule with the matrices in registers, on 9 of the 32 warps.  In this schedule the top slots stay and
    // the bottom slots move down by one per step.  256 worker threads (warps 0..7) hold G and W as 4 x 4 register
    // blocks: thread (R, C) owns the row pairs {2R, 2R+1} x the column pairs {2C, 2C+1} of G (each pair = one top and
    // one bottom slot) and rows {2R, 2R+1, 32+2R, 32+2R+1} of W over the same column slots.  Half of every shift
    // then stays inside the thread (pair 2R+1 -> pair 2R); the other half comes from the next lane (columns:
    // shuffle) or the next thread row (rows: shared memory).  Against one 2 x 2 block per thread on 1024 threads this
    // halves the shared-memory traffic and the barrier width, and gives every thread four independent FMA chains.
    // Warp 8 does nothing but the 32 rotations of the NEXT step (lane = pair): they depend only on the G update of
    // this step, so they run while the workers do this step's W update and shifts (software pipeline).
    // A step:  workers: update the blocks that feed the next rotations, post them, ARRIVE (no wait)  ->  warp 8
    //          starts the rotations of step t+1 while the workers finish the G update, hand the bottom rows on
    //          (worker-only barrier), shift and update W; one barrier of all nine warps ends the step.
    __shared__ double2 rot_cs[32];
    __shared__ int total_rot_s;
    // named barriers: 1 = all 9 warps ("the rotations of the next step are in place"); 2 = the workers ARRIVE (do
    // not wait) once the next rotation inputs are posted, the rotation warp waits on it; 3 = the workers among
    // themselves (the bottom rows have been handed on)
    auto bar_all = [] { asm volatile("bar.sync 1, 288;" ::: "memory"); };
    auto bar_rin_arrive = [] { asm volatile("bar.arrive 2, 288;" ::: "memory"); };
    auto bar_rin_wait = [] { asm volatile("bar.sync 2, 288;" ::: "memory"); };
    auto bar_workers = [] { asm volatile("bar.sync 3, 256;" ::: "memory"); };
    double(*Xbt)[32] = reinterpret_cast<double(*)[32]>(&G1[0][0]);           // [16 thread rows][32 column pairs]
    double(*Xbb)[32] = reinterpret_cast<double(*)[32]>(&G1[0][0] + 16 * 32);
    if (tid >= 256 && tid < 288) {
      // ---- the rotation warp ----
      auto rotations = [&]() -> int {
        const double al = rin_al[lane], be = rin_be[lane], ga = rin_ga[lane];
        const bool rot = al * al > tol2 * be * ga && be > a.floor2 && ga > a.floor2;
        double c = 1.0, sn = 0.0;
        if (rot) fast_rotation(al, be, ga, c, sn);
        rot_cs[lane] = make_double2(c, sn);
        return __popc(__ballot_sync(0xffffffffu, rot));
      };
      int pending = rotations();
      int tot = 0;
      bar_all();  // rotations of step 0 are in place
      for (int sw = 0; sw < a.max_inner_sweeps; ++sw) {
        int sweep_rot = 0;
        for (int t = 0; t < 32; ++t) {
          sweep_rot += pending;
          bar_rin_wait();            // the workers have posted rin_* of the next step (and read rot_cs of this one)
          pending = rotations();
          bar_all();
        }
        if (lane == 0) sweep_rot_s = sweep_rot;
        tot += sweep_rot;
        bar_all();
        if (sweep_rot < a.inner_again) break;
      }
      if (lane == 0) total_rot_s = tot;
    } else if (tid < 256) {
      // ---- the workers ----
      const int R = (int)(tid >> 4), C = (int)(tid & 15);
      const int upl = (int)(lane & 16u) | ((C + 1) & 15);  // lane of thread (R, C + 1)
      const int Rdn = (R + 15) & 15;                        // the thread row that takes over our pair 2R
      double gtt[2][2], gtb[2][2], gbt[2][2], gbb[2][2];
      double wt[4][2], wb[4][2];                            // W rows {2R, 2R+1, 32+2R, 32+2R+1} x column pairs
#pragma unroll
      for (int ia = 0; ia < 2; ++ia)
#pragma unroll
        for (int ib = 0; ib < 2; ++ib) {
          const int ra = 2 * R + ia, cb = 2 * C + ib;
          gtt[ia][ib] = G0[ra][cb];
          gtb[ia][ib] = G0[ra][32 + cb];
          gbt[ia][ib] = G0[32 + ra][cb];
          gbb[ia][ib] = G0[32 + ra][32 + cb];
        }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int ib = 0; ib < 2; ++ib) {
          const int wrow = 2 * R + (r & 1) + 32 * (r >> 1), cb = 2 * C + ib;
          wt[r][ib] = (wrow == cb) ? 1.0 : 0.0;
          wb[r][ib] = (wrow == 32 + cb) ? 1.0 : 0.0;
        }
      // which of this thread's four blocks feed the next step's rotations (pivot blocks and their right neighbours):
      // they are on the critical path rotation -> G update -> rotation, so they are updated and posted first
      bar_all();
      for (int sw = 0; sw < a.max_inner_sweeps; ++sw) {
        for (int t = 0; t < 32; ++t) {
          double2 rc[2], rr[2];
          rc[0] = rot_cs[2 * C];
          rc[1] = rot_cs[2 * C + 1];
          rr[0] = rot_cs[2 * R];
          rr[1] = rot_cs[2 * R + 1];
          double ntt[2][2], ntb[2][2], nbt[2][2], nbb[2][2];
          auto update_block = [&](int ia, int ib) {
            const double c2 = rc[ib].x, s2 = rc[ib].y, c1 = rr[ia].x, s1 = rr[ia].y;
            const double h11 = c2 * gtt[ia][ib] - s2 * gtb[ia][ib], h12 = s2 * gtt[ia][ib] + c2 * gtb[ia][ib];
            const double h21 = c2 * gbt[ia][ib] - s2 * gbb[ia][ib], h22 = s2 * gbt[ia][ib] + c2 * gbb[ia][ib];
            ntt[ia][ib] = c1 * h11 - s1 * h21;
            nbt[ia][ib] = s1 * h11 + c1 * h21;
            ntb[ia][ib] = c1 * h12 - s1 * h22;
            nbb[ia][ib] = s1 * h12 + c1 * h22;
            const int ra = 2 * R + ia, cb = 2 * C + ib;
            if (ra == cb) {
              if (s1 != 0.0) {
                ntb[ia][ib] = 0.0;  // the rotated pair is orthogonal by construction
                nbt[ia][ib] = 0.0;
              }
              rin_be[ra] = ntt[ia][ib];              // G[j][j] of the next step, j = ra
              rin_ga[(ra + 31) & 31] = nbb[ia][ib];  // G[32+j][32+j] of the next step, j = ra - 1
            }
            if (cb == ((ra + 1) & 31)) rin_al[ra] = ntb[ia][ib];  // G[j][32+j] of the next step, j = ra
          };
          // blocks (0,0), (1,1), (0,1) hold every rotation input a thread with R == C or C == R + 1 can own
          // ((1,0) of thread (R, R + 1) is pair (2R+1, 2R+2): posted from the late group of THAT thread only when
          // C == R + 1, so it goes first there) -- simplest correct rule: everything but one block first
          update_block(0, 0);
          update_block(1, 1);
          update_block(0, 1);
          if (C == ((R + 1) & 15)) update_block(1, 0);
          bar_rin_arrive();  // rin_* of the next step are complete; rot_cs of this step is in registers
          if (C != ((R + 1) & 15)) update_block(1, 0);
          // the bottom row of our pair 2R goes to thread row R - 1 (as its pair 2R - 1), bottom columns already
          // shifted: column pair cb lands on cb - 1
          *reinterpret_cast<double2 *>(&Xbt[Rdn][2 * C]) = make_double2(nbt[0][0], nbt[0][1]);
          Xbb[Rdn][(2 * C + 31) & 31] = nbb[0][0];
          Xbb[Rdn][2 * C] = nbb[0][1];
          // ---- shifts and the W update, under the rotation warp's latency chain ----
          // top rows: bottom column (b) <- column pair + 1
#pragma unroll
          for (int ia = 0; ia < 2; ++ia) {
            gtt[ia][0] = ntt[ia][0];
            gtt[ia][1] = ntt[ia][1];
            gtb[ia][0] = ntb[ia][1];
            gtb[ia][1] = __shfl_sync(0xffffffffu, ntb[ia][0], upl);
          }
          // bottom row of pair 2R <- bottom row of pair 2R + 1 (own), with the column shift
          gbt[0][0] = nbt[1][0];
          gbt[0][1] = nbt[1][1];
          gbb[0][0] = nbb[1][1];
          gbb[0][1] = __shfl_sync(0xffffffffu, nbb[1][0], upl);
          // W: columns rotate with the column pairs; the bottom column shifts like G's
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            double nwb[2];
#pragma unroll
            for (int ib = 0; ib < 2; ++ib) {
              const double c2 = rc[ib].x, s2 = rc[ib].y;
              nwb[ib] = s2 * wt[r][ib] + c2 * wb[r][ib];
              wt[r][ib] = c2 * wt[r][ib] - s2 * wb[r][ib];
            }
            wb[r][0] = nwb[1];
            wb[r][1] = __shfl_sync(0xffffffffu, nwb[0], upl);
          }
          bar_workers();  // every thread row has handed its bottom row on
          // bottom row of pair 2R + 1 <- thread row R + 1
          {
            const double2 xt = *reinterpret_cast<const double2 *>(&Xbt[R][2 * C]);
            const double2 xb = *reinterpret_cast<const double2 *>(&Xbb[R][2 * C]);
            gbt[1][0] = xt.x;
            gbt[1][1] = xt.y;
            gbb[1][0] = xb.x;
            gbb[1][1] = xb.y;
          }
          bar_all();  // rot_cs of the next step is in place; X may be rewritten
        }
        bar_all();
        if (sweep_rot_s < a.inner_again) break;
      }
#pragma unroll
      for (int ia = 0; ia < 2; ++ia)
#pragma unroll
        for (int ib = 0; ib < 2; ++ib) {
          const int ra = 2 * R + ia, cb = 2 * C + ib;
          G0[ra][cb] = gtt[ia][ib];
          G0[ra][32 + cb] = gtb[ia][ib];
          G0[32 + ra][cb] = gbt[ia][ib];
          G0[32 + ra][32 + cb] = gbb[ia][ib];
        }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int ib = 0; ib < 2; ++ib) {
          const int wrow = 2 * R + (r & 1) + 32 * (r >> 1), cb = 2 * C + ib;
          W[wrow][cb] = wt[r][ib];
          W[wrow][32 + cb] = wb[r][ib];
        }
    }
    __syncthreads();
    total_rot += total_rot_s;
  }
  if (cross && !any && a.force_update) {  // W = I still moves the columns (last step of a ring round)
    for (unsigned i = tid; i < kD64 * kD64; i += kE32Threads) W[i >> 6][i & 63] = ((i >> 6) == (i & 63)) ? 1.0 : 0.0;
    __syncthreads();
  }
  for (int sw = 0; sw < ((any && !cross) ? a.max_inner_sweeps : 0); ++sw) {
    int sweep_rot = 0;  // counted by warp 0
    for (int t = 0; t < nsteps; ++t) {
      double(*Gc)[kE32Stride] = cur ? G1 : G0;
      double(*Gn)[kE32Stride] = cur ? G0 : G1;
      double(*Wc)[kE32Stride] = cur ? W1 : W;
      double(*Wn)[kE32Stride] = cur ? W : W1;
      // every thread works on two 2 x 2 blocks: row pairs warp and warp + 16 (the kernel runs 16 warps)
      double g11[2], g12[2], g21[2], g22[2], wp0[2], wq0[2], wp1[2], wq1[2];
#pragma unroll
      for (int vv = 0; vv < 2; ++vv) {
        const int wv = w1 + 16 * vv;
        g11[vv] = Gc[wv][l2];
        g12[vv] = Gc[wv][32 + l2];
        g21[vv] = Gc[32 + wv][l2];
        g22[vv] = Gc[32 + wv][32 + l2];
        wp0[vv] = Wc[wv][l2];
        wq0[vv] = Wc[wv][32 + l2];
        wp1[vv] = Wc[wv + 32][l2];
        wq1[vv] = Wc[wv + 32][32 + l2];
      }
      if (w1 == 0) {
        // the 32 rotations of the step, once per CTA (one per lane); with every warp computing them the kernel
        // was bound by instruction issue and the FP64 pipe (~half of its instructions were this arithmetic)
        const double al = Gc[l2][32 + l2], be = Gc[l2][l2], ga = Gc[32 + l2][32 + l2];
        const bool rot = al * al > tol2 * be * ga && be > a.floor2 && ga > a.floor2;
        double c = 1.0, sn = 0.0;
        if (rot) fast_rotation(al, be, ga, c, sn);
        rot_c[l2] = c;
        rot_s[l2] = sn;
        sweep_rot += __popc(__ballot_sync(0xffffffffu, rot));
      }
      __syncthreads();
      const double c2 = rot_c[l2], s2 = rot_s[l2];
#pragma unroll
      for (int vv = 0; vv < 2; ++vv) {
        const int wv = w1 + 16 * vv;
        const int nr_t = next_slot(wv), nr_b = next_slot(32 + wv);
        const double c1 = rot_c[wv], s1 = rot_s[wv];
        const double h11 = c2 * g11[vv] - s2 * g12[vv], h12 = s2 * g11[vv] + c2 * g12[vv];
        const double h21 = c2 * g21[vv] - s2 * g22[vv], h22 = s2 * g21[vv] + c2 * g22[vv];
        double n11 = c1 * h11 - s1 * h21, n21 = s1 * h11 + c1 * h21;
        double n12 = c1 * h12 - s1 * h22, n22 = s1 * h12 + c1 * h22;
        if (wv == l2 && s1 != 0.0) {
          n12 = 0.0;  // the rotated pair is orthogonal by construction
          n21 = 0.0;
        }
        Gn[nr_t][nc_t] = n11;
        Gn[nr_t][nc_b] = n12;
        Gn[nr_b][nc_t] = n21;
        Gn[nr_b][nc_b] = n22;
        Wn[wv][nc_t] = c2 * wp0[vv] - s2 * wq0[vv];
        Wn[wv][nc_b] = s2 * wp0[vv] + c2 * wq0[vv];
        Wn[wv + 32][nc_t] = c2 * wp1[vv] - s2 * wq1[vv];
        Wn[wv + 32][nc_b] = s2 * wp1[vv] + c2 * wq1[vv];
      }
      cur ^= 1;
      __syncthreads();
    }
    if (tid == 0) sweep_rot_s = sweep_rot;
    __syncthreads();
    const int sr = sweep_rot_s;
    total_rot += sr;
    if (sr == 0) break;
  }
  if (cur) {  // an odd number of steps was taken: the live copies are G1 / W1; bring W back (G is only read)
    for (unsigned i = tid; i < kD64 * kD64; i += kE32Threads) W[i >> 6][i & 63] = W1[i >> 6][i & 63];
    __syncthreads();
  }

  if (total_rot > 0 || a.force_update) {
    double(*Gc)[kE32Stride] = cur ? G1 : G0;
    // Newton-Schulz: W <- W (1.5 I - 0.5 W^T W).  Both 64 x 64 x 64 products run as DMMA tiles: warp w owns the
    // 8 x 16 piece (tile row w >> 2, tile columns 2 (w & 3), +1) of the result.  (As scalar FMAs they were 45 % of
    // this kernel's time, bound by the shared-memory loads: 5 loads per 4 FMAs.)
    const int g = lane >> 2, tg = lane & 3;
    // 16 warps, 64 output tiles: warp w owns the 2 x 2 tiles (2 (w >> 2) .. +1) x (2 (w & 3) .. +1): two A- and two
    // B-fragments feed four DMMAs per k-step
    const int ti0 = 2 * (w1 >> 2), tj0 = 2 * (w1 & 3);
    {
      double d[2][2][2];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) d[i][j][0] = d[i][j][1] = 0.0;
#pragma unroll 4
      for (int ks = 0; ks < 16; ++ks) {
        const double a0 = W[4 * ks + tg][8 * ti0 + g], a1 = W[4 * ks + tg][8 * ti0 + 8 + g];
        const double b0 = W[4 * ks + tg][8 * tj0 + g], b1 = W[4 * ks + tg][8 * tj0 + 8 + g];
        dmma884(d[0][0][0], d[0][0][1], a0, b0);
        dmma884(d[0][1][0], d[0][1][1], a0, b1);
        dmma884(d[1][0][0], d[1][0][1], a1, b0);
        dmma884(d[1][1][0], d[1][1][1], a1, b1);
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int r = 8 * (ti0 + i) + g, c = 8 * (tj0 + j) + 2 * tg;
          T[r][c] = (r == c ? 1.5 : 0.0) - 0.5 * d[i][j][0];
          T[r][c + 1] = (r == c + 1 ? 1.5 : 0.0) - 0.5 * d[i][j][1];
        }
    }
    if (tid < kD64) {
      const Pair bp = block_pair(a, k, pair);
      valid[tid] = slot_col32(bp.p, bp.q, tid, a.n) != kNoCol;
    }
    __syncthreads();
    {
      // output slot of every column: eigenvalues descending over the valid slots (single GPU only).  Warp w ranks
      // columns w and 32 + w: lane j compares them with columns j and 32 + j, two ballots each.
      const double dj0 = Gc[lane][lane], dj1 = Gc[32 + lane][32 + lane];
      const bool v0 = valid[lane] != 0, v1 = valid[32 + lane] != 0;
      const unsigned vm0 = __ballot_sync(0xffffffffu, v0), vm1 = __ballot_sync(0xffffffffu, v1);
#pragma unroll
      for (int hsel = 0; hsel < 4; ++hsel) {
        const int i = w1 + 16 * hsel;  // 16 warps rank 4 columns each
        const double di = Gc[i][i];
        const unsigned b0 = __ballot_sync(0xffffffffu, v0 && (dj0 > di || (dj0 == di && (int)lane < i)));
        const unsigned b1 = __ballot_sync(0xffffffffu, v1 && (dj1 > di || (dj1 == di && 32 + (int)lane < i)));
        if (lane == 0) {
          int out = i;
          if (a.sort_cols && valid[i]) {
            const int rk = __popc(b0) + __popc(b1);
            const int n0 = __popc(vm0);  // the rk-th valid slot (ragged last block: not every slot is one)
            out = rk < n0 ? (int)__fns(vm0, 0, rk + 1) : 32 + (int)__fns(vm1, 0, rk - n0 + 1);
          }
          rank_of[i] = out;
        }
      }
    }
    __syncthreads();
    if (a.gcache != nullptr) {  // the rotated diagonal blocks, in the order the columns are written back in
      const Pair bp = block_pair(a, k, pair);
      for (unsigned i = tid; i < kD64 * kD64; i += kE32Threads) {
        const int orow = rank_of[i >> 6], ocol = rank_of[i & 63];
        if ((orow < 32) == (ocol < 32))
          a.gcache[(size_t)(orow < 32 ? bp.p : bp.q) * (kB32 * kB32) + (orow & 31) * kB32 + (ocol & 31)] =
              Gc[i >> 6][i & 63];
      }
    }
    double *wdst = wbuf + (size_t)pair * kD64 * kD64;  // column-major: wdst[c * 64 + r]
    {
      double u[2][2][2];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) u[i][j][0] = u[i][j][1] = 0.0;
#pragma unroll 4
      for (int ks = 0; ks < 16; ++ks) {
        const double a0 = W[8 * ti0 + g][4 * ks + tg], a1 = W[8 * ti0 + 8 + g][4 * ks + tg];
        const double b0 = T[4 * ks + tg][8 * tj0 + g], b1 = T[4 * ks + tg][8 * tj0 + 8 + g];
        dmma884(u[0][0][0], u[0][0][1], a0, b0);
        dmma884(u[0][1][0], u[0][1][1], a0, b1);
        dmma884(u[1][0][0], u[1][0][1], a1, b0);
        dmma884(u[1][1][0], u[1][1][1], a1, b1);
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int r = 8 * (ti0 + i) + g, c = 8 * (tj0 + j) + 2 * tg;
          wdst[rank_of[c] * kD64 + r] = u[i][j][0];
          wdst[rank_of[c + 1] * kD64 + r] = u[i][j][1];
        }
    }
  }
  else if (a.gcache != nullptr && a.gram_mode == 1) {  // refresh step, nothing rotated: the fresh diagonal blocks
    const Pair bp = block_pair(a, k, pair);
    double(*Gc)[kE32Stride] = cur ? G1 : G0;
    for (unsigned i = tid; i < kB32 * kB32; i += kE32Threads) {
      const int r = i >> 5, c = i & 31;
      a.gcache[(size_t)bp.p * (kB32 * kB32) + i] = Gc[r][c];
      a.gcache[(size_t)bp.q * (kB32 * kB32) + i] = Gc[32 + r][32 + c];
    }
  }
  if (tid == 0) {
    wflag[pair] = (total_rot > 0 || a.force_update) ? 1 : 0;
    if (a.counters) {
      if (total_rot > 0) atomicAdd(&a.counters->applied, (unsigned long long)total_rot);
      atomicMax(&a.counters->offdiag_bits, fresh_max);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K3: X <- X W, Y <- Y W for the 64-column panels.  Two persistent CTAs of 8 warps per SM walk the flat list of
// 64-row x 64-column tiles of the block pairs that rotated; tiles arrive through a 2-stage cp.async pipeline
// (512 contiguous bytes per warp-level copy); W (64 x 64) sits in shared memory in fragment order and feeds the
// DMMA B-fragments.  Warp w owns rows 16 (w & 3) .. +15 of the tile and output columns 32 (w >> 2) .. +31: each
// B-fragment serves the two 8-row halves.  What bounds it (0.67-0.74 of the DMMA peak): profiles/r1_k3_inner.md --
// the loop itself runs at the peak from shared memory; the tile loads and the result stores each cost ~15 % on the
// SM side, and a launch of this size pays ~15 us of ramp and tail.
// ------------------------------------------------------------------------------------------------
static constexpr int kU32Rows = 64;
static constexpr int kU32Stride = kU32Rows + 4;
static constexpr int kU32TileDoubles = kD64 * kU32Stride;
static constexpr int kU32WStride = kD64 + 4;
static constexpr int kU32Threads = 256;      // 8 warps: 4 row groups of 16 rows x 2 halves of the output columns
static constexpr int kU32CtasPerSm = 2;      // two independent CTAs per SM: their barriers and tile waits interleave
static constexpr size_t kU32Smem = sizeof(double) * (2 * kU32TileDoubles + kD64 * kU32WStride);

__global__ void __launch_bounds__(kU32Threads, kU32CtasPerSm) blk32_update_kernel(const BlockArgs a, const unsigned k,
                                                                     const double *__restrict__ wbuf,
                                                                     const int *__restrict__ wflag,
                                                                     const unsigned pairs,
                                                                     unsigned long long *__restrict__ active_out) {
  extern __shared__ __align__(16) double sm[];
  double *tiles = sm;                          // 2 stages
  double *Wsm = sm + 2 * kU32TileDoubles;      // Wsm[c * kU32WStride + k] = W[k][c]
  __shared__ unsigned short active[kMaxPairs32];
  __shared__ unsigned n_active_s;
  __shared__ double *coltab[2][4][kD64];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3;

  griddep_launch();
  griddep_wait();  // W and the flags come from K2
  if (warp == 0) {
    unsigned cnt = 0;
    for (unsigned base = 0; base < pairs; base += 32) {
      const unsigned p = base + lane;
      const bool f = p < pairs && wflag[p] != 0;
      const unsigned mask = __ballot_sync(0xffffffffu, f);
      if (f) active[cnt + __popc(mask & ((1u << lane) - 1u))] = (unsigned short)p;
      cnt += __popc(mask);
    }
    if (lane == 0) {
      n_active_s = cnt;
      if (active_out != nullptr && blockIdx.x == 0) *active_out += cnt;
    }
  }
  __syncthreads();
  const unsigned n_active = n_active_s;
  const unsigned tpa = (a.m + kU32Rows - 1) / kU32Rows;
  const unsigned tpv = a.rotate_v ? (a.vrows + kU32Rows - 1) / kU32Rows : 0;
  const unsigned tpp = tpa + tpv;
  const unsigned long long total = (unsigned long long)n_active * tpp;
  const unsigned t0 = (unsigned)(total * blockIdx.x / gridDim.x);
  const unsigned t1 = (unsigned)(total * (blockIdx.x + 1) / gridDim.x);
  if (t0 >= t1) return;

  // Column pointers of a block pair, worked out once per pair (schedule, ragged-block and split/peer selection
  // are dozens of integer instructions per column; per tile they cost more issue slots than the DMMAs).  Two
  // sets: the loads run one tile ahead of the arithmetic and may already be in the next pair.
  auto fill_table = [&](unsigned ap) {
    const Pair bp = block_pair(a, k, active[ap]);
    const unsigned set = ap & 1u;
    for (unsigned i = threadIdx.x; i < 4 * kD64; i += kU32Threads) {
      const unsigned which = i >> 6, slot = i & 63u;  // 0: A load, 1: V load, 2: A store, 3: V store
      const unsigned col = slot_col32(bp.p, bp.q, slot, a.n);
      double *ptr = nullptr;
      if (col != kNoCol) ptr = (which < 2) ? col_ptr(a, which == 0, col) : col_ptr_store(a, which == 2, col);
      coltab[set][which][slot] = ptr;
    }
  };

  unsigned issue_ap = 0xffffffffu, issue_next = 0;  // pair of the tile being issued, first tile of the next pair
  auto issue_tile = [&](unsigned tau, int stage) {
    if (tau >= issue_next || issue_ap == 0xffffffffu) {  // CTA-uniform
      issue_ap = tau / tpp;
      issue_next = (issue_ap + 1) * tpp;
      fill_table(issue_ap);
      __syncthreads();
    }
    const unsigned t = tau - issue_ap * tpp;
    const bool isa = t < tpa;
    const unsigned limit = isa ? a.m : a.vrows;
    const unsigned rin = 2 * lane;  // one warp covers the 64 rows of one column (512 contiguous bytes)
    const unsigned row = (isa ? t : t - tpa) * kU32Rows + rin;
    double *dst = tiles + stage * kU32TileDoubles + rin;
    double *const *tab = coltab[issue_ap & 1u][isa ? 0 : 1];
    const int rb = (row < limit) ? ((row + 1 < limit) ? 16 : 8) : 0;
    const double *base[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) base[i] = tab[warp + 8 * i];  // all loads first: their latencies overlap
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const bool ok = base[i] != nullptr && rb != 0;
      cp_async16(dst + (warp + 8 * i) * kU32Stride, ok ? base[i] + row : a.A, ok ? rb : 0);
    }
  };

  // W of a block pair into shared memory, fragment-major: per k-step and half of the output columns, 32 lanes x
  // (b0, b1) then 32 lanes x (b2, b3): each of the two 128-bit loads of a warp reads 512 contiguous bytes
  auto load_w = [&](unsigned ap) {
    const double *w = wbuf + (size_t)active[ap] * kD64 * kD64;
    for (unsigned i = threadIdx.x; i < kD64 * kD64; i += kU32Threads) {
      const unsigned lo = i & 1, ln = (i >> 1) & 31, hi = (i >> 6) & 1, ks = (i >> 7) & 15, nh = i >> 11;
      Wsm[i] = w[(32 * nh + 8 * (2 * hi + lo) + (ln >> 2)) * kD64 + 4 * ks + (ln & 3)];
    }
  };

  unsigned issued = t0;
  issue_tile(issued, 0);
  ++issued;
  cp_async_commit();

  const unsigned rowgroup = warp & 3, nhalf = warp >> 2;
  // the first pair's W while the first tile is in flight (the barrier at the top of the loop publishes it)
  unsigned cur_ap = t0 / tpp, cur_next = (cur_ap + 1) * tpp;
  load_w(cur_ap);
  for (unsigned tau = t0; tau < t1; ++tau) {
    cp_async_wait<0>();
    __syncthreads();  // tile tau has landed; everyone is done with the other stage and with the previous W
    if (issued < t1) {
      issue_tile(issued, (int)((issued - t0) & 1));
      ++issued;
    }
    cp_async_commit();

    if (tau >= cur_next) {  // CTA-uniform: a new block pair, load its W (everyone is past the barrier above)
      cur_ap = tau / tpp;
      cur_next = (cur_ap + 1) * tpp;
      load_w(cur_ap);
      __syncthreads();
    }
    const unsigned t = tau - cur_ap * tpp;
    const bool isa = t < tpa;
    const unsigned limit = isa ? a.m : a.vrows;
    const unsigned row0 = (isa ? t : t - tpa) * kU32Rows + 16 * rowgroup;
    const double *tile = tiles + ((tau - t0) & 1) * kU32TileDoubles + 16 * rowgroup + g;
    // The product is formed transposed -- D' = W^T X^T: W's fragment goes in as the A operand, the tile's as B (the
    // registers are the same ones either way) -- so that the two accumulator values of a thread are two consecutive
    // rows of one column: d[h][nt][0..1] = out[row0 + 8h + 2tg, +1][32 nhalf + 8 nt + g], one 16-byte store.
    double d[2][4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) d[0][nt][0] = d[0][nt][1] = d[1][nt][0] = d[1][nt][1] = 0.0;
    const double2 *wl = reinterpret_cast<const double2 *>(Wsm + nhalf * 16 * 128 + 2 * lane);
#pragma unroll 8
    for (int ks = 0; ks < 16; ++ks) {
      const double a0 = tile[(4 * ks + tg) * kU32Stride];
      const double a1 = tile[(4 * ks + tg) * kU32Stride + 8];
      const double2 b01 = wl[ks * 64], b23 = wl[ks * 64 + 32];
      dmma884(d[0][0][0], d[0][0][1], b01.x, a0);
      dmma884(d[1][0][0], d[1][0][1], b01.x, a1);
      dmma884(d[0][1][0], d[0][1][1], b01.y, a0);
      dmma884(d[1][1][0], d[1][1][1], b01.y, a1);
      dmma884(d[0][2][0], d[0][2][1], b23.x, a0);
      dmma884(d[1][2][0], d[1][2][1], b23.x, a1);
      dmma884(d[0][3][0], d[0][3][1], b23.y, a0);
      dmma884(d[1][3][0], d[1][3][1], b23.y, a1);
    }
    double *const *stab = coltab[cur_ap & 1u][isa ? 2 : 3] + 32 * nhalf + g;
    double *cst[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) cst[nt] = stab[8 * nt];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const unsigned row = row0 + 8 * h + 2 * tg;
      if (row + 1 < limit) {
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
          if (cst[nt] != nullptr)  // (explicitly global: the pointer comes out of a table, the compiler cannot tell)
            asm volatile("st.global.v2.f64 [%0], {%1, %2};" ::"l"(cst[nt] + row), "d"(d[h][nt][0]), "d"(d[h][nt][1])
                         : "memory");
      } else if (row < limit) {  // odd row count: the last row on its own
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
          if (cst[nt] != nullptr) cst[nt][row] = d[h][nt][0];
      }
    }
  }
  cp_async_wait<0>();
}

// Sum of the row splits' partial Gram matrices (row-sharded driver: what goes into the all-reduce over the ranks)
__global__ void __launch_bounds__(256) blk32_gram_reduce_kernel(const double *__restrict__ gpart, const unsigned splits,
                                                                const unsigned count, double *__restrict__ gred) {
  griddep_launch();
  griddep_wait();
  const unsigned pair = blockIdx.x;
  const double *src = gpart + (size_t)pair * splits * kGramPart32;
  for (unsigned i = threadIdx.x; i < count; i += 256) {
    const double s = sum_partials(src + i, kGramPart32, splits);
    gred[(size_t)pair * kGramPart32 + i] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static bool blk_full_inner() {  // SVDJ_FULL_INNER=1: always the full schedule (for comparisons)
  const char *e = std::getenv("SVDJ_FULL_INNER");
  return e != nullptr && e[0] == '1';
}

unsigned blk32_num_blocks(size_t n) { return (unsigned)((n + kB32 - 1) / kB32); }

// Row splits of the Gram kernel: as many as fit in ONE wave of resident CTAs (2 per SM), rounded so that every
// split is a whole number of 64-row tiles.  (Two waves, the first version, left the second wave 16 % full at
// 4096 x 4096 and gave the eigen-solve 10 partial sums to add instead of 4.)  SVDJ_GRAM_WAVES overrides.
static unsigned gram32_splits(size_t m, unsigned pairs) {
  static const int waves = [] {
    const char *e = std::getenv("SVDJ_GRAM_WAVES");
    const int w = e ? std::atoi(e) : 1;
    return w >= 1 ? w : 1;
  }();
  const size_t tiles = (m + kG32Rows - 1) / kG32Rows;
  size_t want = ((size_t)waves * 2 * 148) / (pairs ? pairs : 1);
  if (want < 1) want = 1;
  // few pairs (a rank of a multi-GPU ring): every split is one more partial matrix the eigen-solve has to add up
  // (37 dependent loads per thread at 8 pairs), so beyond 16 splits a split must keep at least 4096 rows
  const size_t cap = tiles / 64 > 16 ? tiles / 64 : 16;
  if (want > cap) want = cap;
  if (want > tiles) want = tiles;
  if (want > 1024) want = 1024;
  const size_t tiles_per_split = (tiles + want - 1) / want;
  return (unsigned)((tiles + tiles_per_split - 1) / tiles_per_split);
}

void blk32_scratch_sizes(size_t m, unsigned nb, size_t *gpart_doubles, size_t *w_doubles, size_t *flag_ints) {
  const unsigned pairs = nb / 2;
  *gpart_doubles = (size_t)pairs * gram32_splits(m, pairs) * kGramPart32;
  *w_doubles = (size_t)pairs * kD64 * kD64;
  *flag_ints = pairs;
}

bool blk32_supported(const BlockArgs &a) {
  const bool aligned = (a.lda % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.A) & 15) == 0) &&
                       (!a.rotate_v || ((a.ldv % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.V) & 15) == 0)));
  const bool aligned2 = a.A2 == nullptr || (((reinterpret_cast<uintptr_t>(a.A2) & 15) == 0) &&
                                           (!a.rotate_v || ((reinterpret_cast<uintptr_t>(a.V2) & 15) == 0)));
  return aligned && aligned2 && (a.nb / 2) <= (unsigned)kMaxPairs32 && a.nb >= 2;
}

size_t blk32_gred_doubles(unsigned nb) { return (size_t)(nb / 2) * kGramPart32; }

int blk32_launch_step(const BlockArgs &a_in, unsigned k, double *gpart, double *wbuf, int *wflag, cudaStream_t st,
                      unsigned long long *launches, const BlockProbe *probe, const RowHook *rows) {
  const unsigned pairs = a_in.nb / 2;
  if (pairs == 0) return SVDJ_OK;
  BlockArgs a = a_in;
  if (a.A2 == nullptr || a.split >= a.n) {  // single buffer
    a.split = a.n;
    a.A2 = a.A;
    a.V2 = a.V;
  } else {  // pre-bias the second buffers so that kernels index every column with its own number
    a.A2 = a.A2 - (size_t)a.split * a.lda;
    if (a.V2 != nullptr) a.V2 = a.V2 - (size_t)a.split * a.ldv;
  }
  // store targets: the load buffers unless the caller redirected them (already as plain block bases)
  a.A_st = a_in.A_st ? a_in.A_st : a_in.A;
  a.V_st = a_in.V_st ? a_in.V_st : a_in.V;
  if (a_in.A2 == nullptr || a_in.split >= a_in.n) {
    a.A2_st = a.A_st;
    a.V2_st = a.V_st;
  } else {
    a.A2_st = (a_in.A2_st ? a_in.A2_st : a_in.A2) - (size_t)a.split * a.lda;
    double *v2 = a_in.V2_st ? a_in.V2_st : a_in.V2;
    a.V2_st = v2 ? v2 - (size_t)a.split * a.ldv : nullptr;
  }
  // the update kernel stores two rows at a time
  for (const double *ptr : {a.A_st, a.A2_st, a.rotate_v ? a.V_st : nullptr, a.rotate_v ? a.V2_st : nullptr})
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) {
      set_error("block step: a store target is not 16-byte aligned");
      return -5;
    }
  // first step of a sweep or ring phase: full eigen-solve (every block is in exactly one pair when their number is
  // even); later steps, and every step of the ring's bipartite CROSS phase: cross pairs only
  a.cross_only = (!blk_full_inner() && (a.sched == 1 || (k > 1 && a.nb % 2 == 0))) ? 1 : 0;
  // The Gram cache needs every block in a pair of the first step, which refreshes it
  if (a.gcache != nullptr && (a.nb % 2 != 0 || blk_full_inner())) a.gcache = nullptr;
  a.gram_mode = (a.gcache == nullptr) ? 0 : (k == 1 ? 1 : 2);

  static DeviceOnce once;  // the shared-memory opt-in is per device
  int sm_count = 148;
  SVDJ_CUDA_TRY(once.run(
      [] {
        cudaError_t e = cudaFuncSetAttribute(blk32_gram_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(sizeof(double) * kG32Stages * kG32TileDoubles));
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(blk32_gram_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(sizeof(double) * kG32Stages * kG32TileDoubles));
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(blk32_eig_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kE32Smem);
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(blk32_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kU32Smem);
        return e;
      },
      &sm_count));
  const unsigned splits = gram32_splits(a.m, pairs);
  if (a.gpart_capacity != 0 && (size_t)pairs * splits * kGramPart32 > a.gpart_capacity) {
    set_error("block step: %u pairs x %u row splits need %zu doubles of partial Gram matrices, the buffer holds %zu",
              pairs, splits, (size_t)pairs * splits * kGramPart32, a.gpart_capacity);
    return SVDJ_ERR_ALLOC;
  }
  unsigned rps = (a.m + splits - 1) / splits;
  rps = (rps + kG32Rows - 1) / kG32Rows * kG32Rows;
  // SVDJ_PROBE_KERNEL=1|2 moves the probe events from the update kernel to K1 or K2 (profiling aid)
  static const int probe_kernel = [] {
    const char *e = std::getenv("SVDJ_PROBE_KERNEL");
    return e ? std::atoi(e) : 3;
  }();
  if (probe && probe_kernel == 1) SVDJ_CUDA_TRY(cudaEventRecord(probe->before, st));
  if (a.gram_mode == 2)
    SVDJ_CUDA_TRY(launch_k(blk32_gram_kernel<true>, dim3(pairs, splits), 128, sizeof(double) * kG32Stages * kG32TileDoubles, st, a, k, rps, gpart));
  else
    SVDJ_CUDA_TRY(launch_k(blk32_gram_kernel<false>, dim3(pairs, splits), 128, sizeof(double) * kG32Stages * kG32TileDoubles, st, a, k, rps, gpart));
  if (probe && probe_kernel == 1) SVDJ_CUDA_TRY(cudaEventRecord(probe->after, st));
  const double *gsrc = gpart;
  unsigned gsplits = splits;
  if (rows != nullptr) {  // row-sharded: this rank's rows only so far; sum over the splits, then over the ranks
    const unsigned count = (a.gram_mode == 2) ? 16u * 64u : (unsigned)kGramPart32;
    SVDJ_CUDA_TRY(launch_k(blk32_gram_reduce_kernel, dim3(pairs), 256, 0, st, (const double *)gpart, splits, count, rows->gred));
    if (int rc = rows->fn(rows->ctx, rows->gred, (size_t)pairs * kGramPart32, st)) {
      set_error("row-sharded step: the caller's all-reduce returned %d", rc);
      return SVDJ_ERR_CUDA;
    }
    if (launches) *launches += 1;
    gsrc = rows->gred;
    gsplits = 1;
  }
  if (probe && probe_kernel == 2) SVDJ_CUDA_TRY(cudaEventRecord(probe->before, st));
  SVDJ_CUDA_TRY(launch_k(blk32_eig_kernel, dim3(pairs), kE32Threads, kE32Smem, st, a, k, gsplits, gsrc, wbuf, wflag));
  if (probe && probe_kernel == 2) SVDJ_CUDA_TRY(cudaEventRecord(probe->after, st));
  if (probe && probe_kernel == 3) SVDJ_CUDA_TRY(cudaEventRecord(probe->before, st));
  SVDJ_CUDA_TRY(launch_k(blk32_update_kernel, dim3(sm_count * kU32CtasPerSm), kU32Threads, kU32Smem, st, a, k, wbuf, wflag, pairs,
                         probe ? probe->active_pairs : nullptr));
  if (probe && probe_kernel == 3) SVDJ_CUDA_TRY(cudaEventRecord(probe->after, st));
  if (launches) *launches += 3;
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

}  // namespace svdj
