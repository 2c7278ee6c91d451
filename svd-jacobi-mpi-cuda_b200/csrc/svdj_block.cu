// svdj_block.cu -- block one-sided Jacobi step: the same orthogonalisation as the level-1 path, done
// for 2b = 32 columns at a time so that the work becomes dense FP64 contractions.
//
// A round-robin step of the reference rotates n/2 column pairs, each from its own Gram triple
// (lib/JacobiMethods.cu:450-510).  Here a step takes nb/2 disjoint pairs of column BLOCKS (b = 16
// columns each, blocks paired by the same round-robin ordering, lib/JacobiMethods.cu:278-290) and for
// every block pair X = [A_I A_J] (m x 32), Y = [V_I V_J] (n x 32):
//   K1 blk_gram_kernel    G = X^T X            FP64 tensor-core tiles (mma.sync m8n8k4 = DMMA), HBM/tensor ridge
//   K2 blk_eig_kernel     G = W L W^T          cyclic two-sided Jacobi on the 32 x 32 Gram matrix in shared
//                                              memory: the reference's rotation formulas (:464-477) applied to
//                                              Gram entries instead of freshly summed dot products
//   K3 blk_update_kernel  X <- X W, Y <- Y W   FP64 tensor-core tiles, HBM-bound
// One block pair covers b(2b-1) = 496 column pairs for 16b(3m+2n) bytes of traffic, against 32(m+n) per
// pair on the level-1 path: ~0.8 b = 13x fewer bytes per sweep; flops rise to 10 n^3 per sweep (m = n).
// tcgen05 has no FP64 kind, so mma.sync DMMA is the sm_100a FP64 tensor path (profiles/r1_fp64_peaks.md).
#include <cstdlib>

#include "svdj_block_dev.cuh"

namespace svdj {

static constexpr int kBW = 16;                 // block width b
static constexpr int kD = 2 * kBW;             // columns of a block pair
static constexpr int kGramTiles = 10;          // upper triangle of the 4 x 4 grid of 8 x 8 tiles
static constexpr int kGramPart = kGramTiles * 64;

// Global column held by slot s (0..31) of block pair (I, J); kNoCol past the last column (ragged last block:
// such a slot behaves as a zero column and is never written).
__device__ __forceinline__ unsigned slot_col(unsigned I, unsigned J, unsigned s, unsigned n) {
  const unsigned c = (s < kBW) ? I * kBW + s : J * kBW + (s - kBW);
  return c < n ? c : kNoCol;
}

// ------------------------------------------------------------------------------------------------
// K1: partial Gram matrices.  grid (block pairs, row splits); each warp walks 4-row k-steps of its
// rows.  For G = X^T X the A fragment of tile row ti and the B fragment of tile column tj are the same
// 32 numbers X[k0 + lane%4][8t + lane/4], so 4 loads feed the 10 DMMAs of a k-step.
// gpart layout: [pair][split][tile t][lane][e], tile t = (ti<=tj) in row-major order of the upper triangle.
// ------------------------------------------------------------------------------------------------
template <int WARPS, int UNROLL>
__global__ void __launch_bounds__(WARPS * 32) blk_gram_kernel(const BlockArgs a, const unsigned k,
                                                              const unsigned rows_per_split,
                                                              double *__restrict__ gpart) {
  griddep_launch();
  griddep_wait();  // everything this kernel reads comes from the kernel before it in the stream
  __shared__ double red[WARPS][kGramPart];
  const Pair bp = block_pair(a, k, blockIdx.x);
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3;
  const double *cp[4];
  bool cv[4];
#pragma unroll
  for (int gi = 0; gi < 4; ++gi) {
    const unsigned c = slot_col(bp.p, bp.q, 8 * gi + g, a.n);
    cv[gi] = (c != kNoCol);
    cp[gi] = col_ptr(a, true, cv[gi] ? c : 0);
  }
  const unsigned r0 = blockIdx.y * rows_per_split;
  const unsigned r1 = min(a.m, r0 + rows_per_split);
  double acc[kGramTiles][2];
#pragma unroll
  for (int t = 0; t < kGramTiles; ++t) acc[t][0] = acc[t][1] = 0.0;

  for (unsigned kr = r0 + 4 * warp; kr < r1; kr += 4 * WARPS * UNROLL) {
    double f[UNROLL][4];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const unsigned row = kr + 4 * WARPS * u + tg;
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) f[u][gi] = (cv[gi] && row < r1) ? cp[gi][row] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      int t = 0;
#pragma unroll
      for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = ti; tj < 4; ++tj) {
          dmma884(acc[t][0], acc[t][1], f[u][ti], f[u][tj]);
          ++t;
        }
    }
  }
#pragma unroll
  for (int t = 0; t < kGramTiles; ++t) {
    red[warp][t * 64 + lane * 2 + 0] = acc[t][0];
    red[warp][t * 64 + lane * 2 + 1] = acc[t][1];
  }
  __syncthreads();
  double *dst = gpart + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * kGramPart;
  for (unsigned i = threadIdx.x; i < kGramPart; i += WARPS * 32) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) s += red[w][i];
    dst[i] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// K2: eigen-decomposition of the 32 x 32 Gram matrix by cyclic two-sided Jacobi, all in shared memory.
// One CTA of 256 threads per block pair.  A parallel step holds 16 disjoint index pairs (circle
// method).  Every warp computes the step's 16 rotations itself (lanes j and j+16 both do pair j), so no
// broadcast through shared memory is needed; thread (j1, j2) owns the 2 x 2 block G[{p1,q1}][{p2,q2}]
// and applies J1^T . J2 to it, reading the old matrix and writing the other buffer: one barrier per
// step.  W accumulates the rotations, is pulled back to orthogonality to O(eps) by one Newton-Schulz
// step (otherwise the defect of thousands of products adds up linearly in V), and its columns are
// ordered by decreasing eigenvalue, which keeps the larger columns in the first block of the pair
// (de Rijk's ordering at block level: ~25 % fewer outer sweeps).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void circle_pair(int t, int j, int &p, int &q) {
  // 31 rounds over 32 players: player 31 stays, the others rotate
  int x, y;
  if (j == 0) {
    x = t;
    y = kD - 1;
  } else {
    x = t + j;
    x = x >= kD - 1 ? x - (kD - 1) : x;
    y = t - j;
    y = y < 0 ? y + (kD - 1) : y;
  }
  p = min(x, y);
  q = max(x, y);
}

__global__ void __launch_bounds__(256) blk_eig_kernel(const BlockArgs a, const unsigned k, const unsigned splits,
                                                      const double *__restrict__ gpart, double *__restrict__ wbuf,
                                                      int *__restrict__ wflag) {
  griddep_launch();
  griddep_wait();  // everything this kernel reads comes from the kernel before it in the stream
  // Row stride kD + 2: the rotation inputs G[p][q] of a step satisfy p + q = const (circle method), so their
  // addresses are (stride - 1) p + const; an odd (stride - 1) spreads them over the banks (stride kD + 1 put
  // all sixteen pairs on one bank), and the diagonal, (stride + 1) p, stays conflict-free as well.
  __shared__ double Gb[2][kD][kD + 2];
  __shared__ double Wb[2][kD][kD + 2];
  __shared__ double T[kD][kD + 2];
  __shared__ unsigned long long fresh_max;
  __shared__ int rank_of[kD];
  const unsigned tid = threadIdx.x, lane = tid & 31;
  const unsigned pair = blockIdx.x;

  // sum the partial Gram matrices in a fixed order (deterministic) and unpack the DMMA tile layout
  const double *src = gpart + (size_t)pair * splits * kGramPart;
  for (unsigned i = tid; i < kGramPart; i += 256) {
    double s = 0.0;
    s = sum_partials(src + i, kGramPart, splits);
    const int t = i >> 6, l = (i & 63) >> 1, e = i & 1;
    int ti = 0, rem = t;  // t -> (ti, tj) of the upper triangle: rows of 4, 3, 2, 1 tiles
    while (rem >= 4 - ti) {
      rem -= 4 - ti;
      ++ti;
    }
    const int tj = ti + rem;
    const int r = 8 * ti + (l >> 2), c = 8 * tj + 2 * (l & 3) + e;
    Gb[0][r][c] = s;
    if (ti != tj) Gb[0][c][r] = s;
  }
  for (unsigned i = tid; i < kD * kD; i += 256) Wb[0][i >> 5][i & 31] = ((i >> 5) == (i & 31)) ? 1.0 : 0.0;
  if (tid == 0) fresh_max = 0ull;
  __syncthreads();

  // largest off-diagonal cosine of the fresh Gram matrix (what lib/JacobiMethods.cu:462 computes per pair)
  {
    double mx = 0.0;
    for (unsigned i = tid; i < kD * kD; i += 256) {
      const int r = i >> 5, c = i & 31;
      if (r < c) {
        const double grr = Gb[0][r][r], gcc = Gb[0][c][c];
        const double den = sqrt(grr * gcc);
        const double off = (den > 0.0 && grr > a.floor2 && gcc > a.floor2) ? fabs(Gb[0][r][c]) / den : 0.0;
        mx = fmax(mx, off);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) atomicMax(&fresh_max, (unsigned long long)__double_as_longlong(mx));
  }

  // Systolic layout: rows and columns of G (and the columns of W) are kept in SLOT order.  Slot j and slot
  // 16 + j always hold the two indices of pair j, so thread (j1, j2) always works on G[{j1, 16+j1}][{j2, 16+j2}]:
  // fixed, conflict-free addresses (indexing by the circle-method indices themselves put the sixteen rotation
  // inputs of a step on one bank).  The round-robin move of the indices is folded into the write-back: every
  // updated value goes to the slot its index occupies in the next step, in the other buffer; after a whole
  // sweep every index is back in its own slot, so slot order = index order at sweep boundaries.
  // FULL schedule: all 496 pairs in 31 steps (top 0 stays, top j -> top j+1, top 15 -> bottom 15, bottom j ->
  // bottom j-1, bottom 0 -> top 1).  CROSS: the 256 pairs with one index in each column block, 16 steps (bottom
  // slots rotate, top slots stay) -- used after the first step of a sweep, so that a column pair is rotated once
  // per sweep as in the reference's ordering, not re-rotated with its own block in every step.
  const bool cross = a.cross_only != 0;
  const int nsteps = cross ? kBW : kD - 1;
  auto next_slot = [cross](int p) {
    if (cross) return p < kBW ? p : kBW + ((p - kBW + kBW - 1) & (kBW - 1));
    return p == 0 ? 0 : (p < kBW - 1 ? p + 1 : (p == kBW - 1 ? kD - 1 : (p == kBW ? 1 : p - 1)));
  };
  int total_rot = 0;
  int cur = 0;
  const int j1 = tid >> 4, j2 = tid & 15;
  const double tol2 = a.tol * a.tol;
  const int nr_t = next_slot(j1), nr_b = next_slot(kBW + j1), nc_t = next_slot(j2), nc_b = next_slot(kBW + j2);
  int any = 0;  // anything to rotate among the pairs this schedule visits?  (same test as the steps)
  for (unsigned i = tid; i < kD * kD; i += 256) {
    const int r = i >> 5, c = i & 31;
    if (r < c && (!cross || (r < kBW && c >= kBW))) {
      const double al = Gb[0][r][c], be = Gb[0][r][r], ga = Gb[0][c][c];
      any |= (al * al > tol2 * be * ga && be > a.floor2 && ga > a.floor2) ? 1 : 0;
    }
  }
  any = __syncthreads_or(any);
  for (int sw = 0; sw < (any ? a.max_inner_sweeps : 0); ++sw) {
    int sweep_rot = 0;
    for (int t = 0; t < nsteps; ++t) {
      // the step's 16 rotations, computed by every warp (lane & 15 = pair index); all loads first
      const double al = Gb[cur][j2][kBW + j2], be = Gb[cur][j2][j2], ga = Gb[cur][kBW + j2][kBW + j2];
      const double g11 = Gb[cur][j1][j2], g12 = Gb[cur][j1][kBW + j2];
      const double g21 = Gb[cur][kBW + j1][j2], g22 = Gb[cur][kBW + j1][kBW + j2];
      const double wp0 = Wb[cur][j1][j2], wq0 = Wb[cur][j1][kBW + j2];
      const double wp1 = Wb[cur][j1 + 16][j2], wq1 = Wb[cur][j1 + 16][kBW + j2];
      // |alpha| > tol sqrt(beta gamma) (lib/JacobiMethods.cu:462/466), and neither column numerically zero
      const bool rot = al * al > tol2 * be * ga && be > a.floor2 && ga > a.floor2;
      double c2 = 1.0, s2 = 0.0;
      if (rot) fast_rotation(al, be, ga, c2, s2);
      sweep_rot += __popc(__ballot_sync(0xffffffffu, rot) & 0xffffu);
      const double c1 = __shfl_sync(0xffffffffu, c2, j1), s1 = __shfl_sync(0xffffffffu, s2, j1);
      // columns: col_p' = c col_p - s col_q, col_q' = s col_p + c col_q (the kernel at :1483-1491)
      const double h11 = c2 * g11 - s2 * g12, h12 = s2 * g11 + c2 * g12;
      const double h21 = c2 * g21 - s2 * g22, h22 = s2 * g21 + c2 * g22;
      // rows: the transpose of the same rotation
      double n11 = c1 * h11 - s1 * h21, n21 = s1 * h11 + c1 * h21;
      double n12 = c1 * h12 - s1 * h22, n22 = s1 * h12 + c1 * h22;
      if (j1 == j2 && s1 != 0.0) {
        n12 = 0.0;  // the rotated pair is orthogonal by construction
        n21 = 0.0;
      }
      Gb[cur ^ 1][nr_t][nc_t] = n11;
      Gb[cur ^ 1][nr_t][nc_b] = n12;
      Gb[cur ^ 1][nr_b][nc_t] = n21;
      Gb[cur ^ 1][nr_b][nc_b] = n22;
      Wb[cur ^ 1][j1][nc_t] = c2 * wp0 - s2 * wq0;  // rows j1 and j1 + 16 of W, pair j2
      Wb[cur ^ 1][j1][nc_b] = s2 * wp0 + c2 * wq0;
      Wb[cur ^ 1][j1 + 16][nc_t] = c2 * wp1 - s2 * wq1;
      Wb[cur ^ 1][j1 + 16][nc_b] = s2 * wp1 + c2 * wq1;
      cur ^= 1;
      __syncthreads();
    }
    total_rot += sweep_rot;
    if (sweep_rot == 0) break;
  }
  double(*W)[kD + 2] = Wb[cur];  // the live copy

  if (total_rot > 0 || a.force_update) {
    // Newton-Schulz: W <- W (1.5 I - 0.5 W^T W).  Both 32 x 32 x 32 products run as DMMA tiles: warp w owns the
    // 8 x 16 piece (tile row w >> 1, tile columns 2 (w & 1), +1) of the result (as scalar FMAs they were bound
    // by shared-memory loads, 5 per 4 FMAs).
    const int g = lane >> 2, tg = lane & 3;
    const int ti = (tid >> 5) >> 1, tj = 2 * ((tid >> 5) & 1);
    {
      double d0a = 0.0, d0b = 0.0, d1a = 0.0, d1b = 0.0;
#pragma unroll
      for (int ks = 0; ks < kD / 4; ++ks) {
        const double af = W[4 * ks + tg][8 * ti + g];
        dmma884(d0a, d0b, af, W[4 * ks + tg][8 * tj + g]);
        dmma884(d1a, d1b, af, W[4 * ks + tg][8 * tj + 8 + g]);
      }
      const int r = 8 * ti + g, c = 8 * tj + 2 * tg;
      T[r][c] = (r == c ? 1.5 : 0.0) - 0.5 * d0a;
      T[r][c + 1] = (r == c + 1 ? 1.5 : 0.0) - 0.5 * d0b;
      T[r][c + 8] = (r == c + 8 ? 1.5 : 0.0) - 0.5 * d1a;
      T[r][c + 9] = (r == c + 9 ? 1.5 : 0.0) - 0.5 * d1b;
    }
    // Output slot of W's column i: the slots that hold real columns, in order, receive the eigenvalues in
    // descending order (stable).  Slots past the last column of a ragged block keep their (zero) column.
    if (tid < kD) {
      const Pair bp = block_pair(a, k, pair);
      const unsigned vmask = __ballot_sync(0xffffffffu, slot_col(bp.p, bp.q, tid, a.n) != kNoCol);
      int out = tid;
      if (a.sort_cols && ((vmask >> tid) & 1u)) {
        const double di = Gb[cur][tid][tid];
        int rk = 0;
        for (int j = 0; j < kD; ++j) {
          const double dj = Gb[cur][j][j];
          rk += (((vmask >> j) & 1u) && (dj > di || (dj == di && j < (int)tid))) ? 1 : 0;
        }
        out = __fns(vmask, 0, rk + 1);  // position of the (rk+1)-th valid slot
      }
      rank_of[tid] = out;
    }
    __syncthreads();
    double u0a = 0.0, u0b = 0.0, u1a = 0.0, u1b = 0.0;
#pragma unroll
    for (int ks = 0; ks < kD / 4; ++ks) {
      const double af = W[8 * ti + g][4 * ks + tg];
      dmma884(u0a, u0b, af, T[4 * ks + tg][8 * tj + g]);
      dmma884(u1a, u1b, af, T[4 * ks + tg][8 * tj + 8 + g]);
    }
    double *wdst = wbuf + (size_t)pair * kD * kD;  // column-major: wdst[c*32 + r]
    const int r = 8 * ti + g, c = 8 * tj + 2 * tg;
    wdst[rank_of[c] * kD + r] = u0a;
    wdst[rank_of[c + 1] * kD + r] = u0b;
    wdst[rank_of[c + 8] * kD + r] = u1a;
    wdst[rank_of[c + 9] * kD + r] = u1b;
  }
  if (tid == 0) {
    wflag[pair] = (total_rot > 0 || a.force_update) ? 1 : 0;  // force_update: W = I still moves the columns
    if (a.counters) {
      if (total_rot > 0) atomicAdd(&a.counters->applied, (unsigned long long)total_rot);
      atomicMax(&a.counters->offdiag_bits, fresh_max);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K3: X <- X W and Y <- Y W.  Each warp owns 8-row strips of the block pair's 32 columns (rows of A first,
// then rows of V), reads a strip as 8 DMMA A-fragments, multiplies by W held in registers as 32
// B-fragments, and writes the strip back in place (a strip depends only on itself).
// ------------------------------------------------------------------------------------------------
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) blk_update_kernel(const BlockArgs a, const unsigned k,
                                                                const double *__restrict__ wbuf,
                                                                const int *__restrict__ wflag,
                                                                unsigned long long *__restrict__ active_out) {
  griddep_launch();
  griddep_wait();  // everything this kernel reads comes from the kernel before it in the stream
  if (active_out != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
    unsigned cnt = 0;
    for (unsigned p = 0; p < gridDim.x; ++p) cnt += wflag[p] != 0;
    *active_out += cnt;
  }
  if (!wflag[blockIdx.x]) return;  // nothing was rotated in this block pair: W = I
  const Pair bp = block_pair(a, k, blockIdx.x);
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3;

  double wf[8][4];  // wf[ks][nt] = W[4 ks + tg][8 nt + g]
  {
    const double *w = wbuf + (size_t)blockIdx.x * kD * kD;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) wf[ks][nt] = w[(8 * nt + g) * kD + 4 * ks + tg];
  }
  unsigned lc[8];     // load columns: slot 4 ks + tg
  unsigned sc[4][2];  // store columns: slot 8 nt + 2 tg + e
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) lc[ks] = slot_col(bp.p, bp.q, 4 * ks + tg, a.n);
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    sc[nt][0] = slot_col(bp.p, bp.q, 8 * nt + 2 * tg, a.n);
    sc[nt][1] = slot_col(bp.p, bp.q, 8 * nt + 2 * tg + 1, a.n);
  }

  const unsigned strips_a = (a.m + 7) >> 3;
  const unsigned strips_v = a.rotate_v ? (a.vrows + 7) >> 3 : 0;
  const unsigned total = strips_a + strips_v;
  const unsigned stride = gridDim.y * WARPS;

  auto load_strip = [&](unsigned s, double (&af)[8]) {
    const bool isa = s < strips_a;
    const unsigned limit = isa ? a.m : a.vrows;
    const unsigned row = (isa ? s : s - strips_a) * 8 + g;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)
      af[ks] = (row < limit && lc[ks] != kNoCol) ? col_ptr(a, isa, lc[ks])[row] : 0.0;
  };

  unsigned s = blockIdx.y * WARPS + warp;
  double cur[8], nxt[8];
  if (s < total) load_strip(s, cur);
  for (; s < total; s += stride) {
    const unsigned sn = s + stride;
    if (sn < total) load_strip(sn, nxt);
    double d[4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) d[nt][0] = d[nt][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) dmma884(d[nt][0], d[nt][1], cur[ks], wf[ks][nt]);
    {
      const bool isa = s < strips_a;
      const unsigned limit = isa ? a.m : a.vrows;
      const unsigned row = (isa ? s : s - strips_a) * 8 + g;
      if (row < limit) {
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          if (sc[nt][0] != kNoCol) col_ptr_store(a, isa, sc[nt][0])[row] = d[nt][0];
          if (sc[nt][1] != kNoCol) col_ptr_store(a, isa, sc[nt][1])[row] = d[nt][1];
        }
      }
    }
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) cur[ks] = nxt[ks];
  }
}

// ------------------------------------------------------------------------------------------------
// K3, staged: the same update for aligned layouts (even leading dimensions, 16-byte aligned bases).
// Persistent CTAs (4 per SM) walk a flat list of 64-row x 32-column tiles of the block pairs that
// rotated (compacted from wflag), so the grid is a multiple of the SM count and nothing is spent on
// pairs whose W is the identity.  Tiles are brought into shared memory by a 3-stage cp.async
// pipeline: every warp-level copy is one column's 512 contiguous bytes, no registers are tied up by
// loads in flight, and ~100 KB per SM are outstanding -- the plain kernel above is latency-bound
// at 17 % occupancy (profiles/r1_block_4096.md).  Column stride 68 doubles makes the DMMA
// A-fragment reads (4 columns x 8 rows per warp) bank-conflict free.
// ------------------------------------------------------------------------------------------------
static constexpr int kTileRows = 64;
static constexpr int kTileStride = kTileRows + 4;
static constexpr int kTileDoubles = kD * kTileStride;
static constexpr int kUpdStages = 3;
static constexpr int kUpdCtasPerSm = 4;
static constexpr int kMaxCompactPairs = 2048;

template <int STAGES>  // STAGES <= 3 (coltab sets)
__global__ void __launch_bounds__(128, kUpdCtasPerSm) blk_update_staged_kernel(const BlockArgs a, const unsigned k,
                                                                  const double *__restrict__ wbuf,
                                                                  const int *__restrict__ wflag,
                                                                  const unsigned pairs,
                                                                  unsigned long long *__restrict__ active_out) {
  griddep_launch();
  griddep_wait();  // everything this kernel reads comes from the kernel before it in the stream
  extern __shared__ __align__(16) double tiles[];  // STAGES x 32 columns x kTileStride
  __shared__ unsigned short active[kMaxCompactPairs];
  __shared__ unsigned n_active_s;
  __shared__ double *coltab[3][4][kD];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3;

  if (warp == 0) {  // compact the block pairs that rotated
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
  const unsigned tpa = (a.m + kTileRows - 1) / kTileRows;
  const unsigned tpv = a.rotate_v ? (a.vrows + kTileRows - 1) / kTileRows : 0;
  const unsigned tpp = tpa + tpv;
  const unsigned long long total = (unsigned long long)n_active * tpp;
  const unsigned t0 = (unsigned)(total * blockIdx.x / gridDim.x);
  const unsigned t1 = (unsigned)(total * (blockIdx.x + 1) / gridDim.x);
  if (t0 >= t1) return;

  // Column pointers of a block pair, worked out once per pair instead of once per tile and column (schedule,
  // ragged-block and split/peer selection are dozens of integer instructions each).  Three sets: the loads run
  // up to STAGES - 1 = 2 tiles ahead of the arithmetic, which for short pairs can be two pairs ahead.
  auto fill_table = [&](unsigned ap) {
    const Pair bp = block_pair(a, k, active[ap]);
    const unsigned set = ap % 3u;
    for (unsigned i = threadIdx.x; i < 4 * kD; i += 128) {
      const unsigned which = i >> 5, slot = i & 31u;  // 0: A load, 1: V load, 2: A store, 3: V store
      const unsigned col = slot_col(bp.p, bp.q, slot, a.n);
      double *ptr = nullptr;
      if (col != kNoCol) ptr = (which < 2) ? col_ptr(a, which == 0, col) : col_ptr_store(a, which == 2, col);
      coltab[set][which][slot] = ptr;
    }
  };

  unsigned issue_ap = 0xffffffffu, issue_next = 0;
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
    const unsigned row = (isa ? t : t - tpa) * kTileRows + 2 * lane;
    double *dst = tiles + stage * kTileDoubles + 2 * lane;
    double *const *tab = coltab[issue_ap % 3u][isa ? 0 : 1];
    const int rb = (row < limit) ? ((row + 1 < limit) ? 16 : 8) : 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const unsigned slot = warp + 4 * i;
      const double *base = tab[slot];
      const bool ok = base != nullptr && rb != 0;
      cp_async16(dst + slot * kTileStride, ok ? base + row : a.A, ok ? rb : 0);
    }
  };

  unsigned issued = t0;
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (issued < t1) {
      issue_tile(issued, s);
      ++issued;
    }
    cp_async_commit();
  }

  double wf[8][4];
  unsigned cur_ap = 0xffffffffu, cur_next = 0;
  for (unsigned tau = t0; tau < t1; ++tau) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();  // tile tau has landed for everyone; everyone is done with the stage refilled next
    if (issued < t1) {
      issue_tile(issued, (int)((issued - t0) % STAGES));
      ++issued;
    }
    cp_async_commit();

    if (tau >= cur_next || cur_ap == 0xffffffffu) {  // new block pair: its W as B fragments
      cur_ap = tau / tpp;
      cur_next = (cur_ap + 1) * tpp;
      const double *w = wbuf + (size_t)active[cur_ap] * kD * kD;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) wf[ks][nt] = w[(8 * nt + g) * kD + 4 * ks + tg];
    }
    const unsigned t = tau - cur_ap * tpp;
    const bool isa = t < tpa;
    const unsigned limit = isa ? a.m : a.vrows;
    const unsigned row0 = (isa ? t : t - tpa) * kTileRows + 16 * warp;
    const double *tile = tiles + ((tau - t0) % STAGES) * kTileDoubles + 16 * warp + g;
    double *const *stab = coltab[cur_ap % 3u][isa ? 2 : 3] + 2 * tg;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      double af[8];
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) af[ks] = tile[(4 * ks + tg) * kTileStride + 8 * h];
      double d[4][2];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) d[nt][0] = d[nt][1] = 0.0;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) dmma884(d[nt][0], d[nt][1], af[ks], wf[ks][nt]);
      const unsigned row = row0 + 8 * h + g;
      if (row < limit) {
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          double *const c0 = stab[8 * nt], *const c1 = stab[8 * nt + 1];
          if (c0 != nullptr) c0[row] = d[nt][0];
          if (c1 != nullptr) c1[row] = d[nt][1];
        }
      }
    }
  }
  cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------------
// K1, staged: the same partial Gram matrices for aligned layouts.  grid (block pairs, row splits) as
// above, but the 64-row x 32-column tiles of a CTA's row range come in through the 3-stage cp.async
// pipeline of the update kernel (512 contiguous bytes per warp-level copy) instead of 8-byte fragment
// loads straight from global memory, which left the plain kernel at half of the HBM rate.
// ------------------------------------------------------------------------------------------------
template <int STAGES>
__global__ void __launch_bounds__(128) blk_gram_staged_kernel(const BlockArgs a, const unsigned k,
                                                              const unsigned rows_per_split,
                                                              double *__restrict__ gpart) {
  griddep_launch();
  griddep_wait();  // everything this kernel reads comes from the kernel before it in the stream
  extern __shared__ __align__(16) double tiles[];  // STAGES x 32 columns x kTileStride; reused for the reduction
  const Pair bp = block_pair(a, k, blockIdx.x);
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned g = lane >> 2, tg = lane & 3;
  const unsigned r0 = blockIdx.y * rows_per_split;  // a multiple of 64
  const unsigned r1 = min(a.m, r0 + rows_per_split);
  const unsigned ntiles = r0 < r1 ? (r1 - r0 + kTileRows - 1) / kTileRows : 0;

  const double *cptr[8];  // this thread copies column slots warp + 4 i
  bool cval[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const unsigned c = slot_col(bp.p, bp.q, warp + 4 * i, a.n);
    cval[i] = (c != kNoCol);
    cptr[i] = col_ptr(a, true, cval[i] ? c : 0);
  }
  auto issue_tile = [&](unsigned t, int stage) {
    const unsigned row = r0 + t * kTileRows + 2 * lane;
    double *dst = tiles + stage * kTileDoubles + 2 * lane;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      int bytes = 0;
      const double *src = a.A;
      if (cval[i] && row < r1) {
        bytes = (row + 1 < r1) ? 16 : 8;
        src = cptr[i] + row;
      }
      cp_async16(dst + (warp + 4 * i) * kTileStride, src, bytes);
    }
  };

  unsigned issued = 0;
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (issued < ntiles) {
      issue_tile(issued, s);
      ++issued;
    }
    cp_async_commit();
  }
  double acc[kGramTiles][2];
#pragma unroll
  for (int t = 0; t < kGramTiles; ++t) acc[t][0] = acc[t][1] = 0.0;

  for (unsigned t = 0; t < ntiles; ++t) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    if (issued < ntiles) {
      issue_tile(issued, (int)(issued % STAGES));
      ++issued;
    }
    cp_async_commit();
    // rows 16 warp .. 16 warp + 15 of the tile: four k-steps of 4 rows; rows past r1 were zero-filled
    const double *tile = tiles + (t % STAGES) * kTileDoubles + 16 * warp + tg;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      double f[4];
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) f[gi] = tile[(8 * gi + g) * kTileStride + 4 * ks];
      int tt = 0;
#pragma unroll
      for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = ti; tj < 4; ++tj) {
          dmma884(acc[tt][0], acc[tt][1], f[ti], f[tj]);
          ++tt;
        }
    }
  }
  cp_async_wait<0>();
  __syncthreads();  // the tile buffers become the reduction buffer
  double *red = tiles;  // [4 warps][kGramPart]
#pragma unroll
  for (int t = 0; t < kGramTiles; ++t) {
    red[warp * kGramPart + t * 64 + lane * 2 + 0] = acc[t][0];
    red[warp * kGramPart + t * 64 + lane * 2 + 1] = acc[t][1];
  }
  __syncthreads();
  double *dst = gpart + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * kGramPart;
  for (unsigned i = threadIdx.x; i < kGramPart; i += 128) {
    dst[i] = red[i] + red[kGramPart + i] + red[2 * kGramPart + i] + red[3 * kGramPart + i];
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static constexpr int kGramWarps = 4, kGramUnroll = 4, kUpdWarps = 4;

// SVDJ_BLOCK_PLAIN=1 forces the unstaged kernels (used by the tests to cover both code paths)
static bool blk_force_plain_update() {
  const char *e = std::getenv("SVDJ_BLOCK_PLAIN");
  return e != nullptr && e[0] == '1';
}

unsigned blk_num_blocks(size_t n) { return (unsigned)((n + kBW - 1) / kBW); }

static unsigned gram_splits(size_t m, unsigned pairs) {
  // about 8 CTAs per SM, at least 64 rows (one unrolled iteration) per CTA
  size_t want = (8 * 148 + pairs - 1) / (pairs ? pairs : 1);
  size_t cap = (m + 63) / 64;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  if (want > 1024) want = 1024;
  return (unsigned)want;
}

void blk_scratch_sizes(size_t m, unsigned nb, size_t *gpart_doubles, size_t *w_doubles, size_t *flag_ints) {
  const unsigned pairs = nb / 2;
  *gpart_doubles = (size_t)pairs * gram_splits(m, pairs) * kGramPart;
  *w_doubles = (size_t)pairs * kD * kD;
  *flag_ints = pairs;
}

int blk_launch_step(const BlockArgs &a_in, unsigned k, double *gpart, double *wbuf, int *wflag, cudaStream_t st,
                    unsigned long long *launches, const BlockProbe *probe, const RowHook *rows) {
  if (rows != nullptr) return SVDJ_ERR_UNSUPPORTED;  // the row-sharded driver uses the 64-column kernels
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
  // first step of a phase or sweep: full eigen-solve (every block sits in exactly one pair when their number is
  // even); afterwards only the cross pairs.  The bipartite schedule of the ring's CROSS phase never needs the
  // intra-block pairs (its blocks were swept internally in the round before).
  {
    const char *e = std::getenv("SVDJ_FULL_INNER");
    const bool always_full = e != nullptr && e[0] == '1';
    a.cross_only = (!always_full && (a.sched == 1 || (k > 1 && a.nb % 2 == 0))) ? 1 : 0;
  }
  const unsigned splits = gram_splits(a.m, pairs);
  if (a.gpart_capacity != 0 && (size_t)pairs * splits * kGramPart > a.gpart_capacity) {
    set_error("block step: %u pairs x %u row splits need %zu doubles of partial Gram matrices, the buffer holds %zu",
              pairs, splits, (size_t)pairs * splits * kGramPart, a.gpart_capacity);
    return SVDJ_ERR_ALLOC;
  }
  unsigned rps = (a.m + splits - 1) / splits;
  rps = (rps + 3u) & ~3u;  // k-steps are 4 rows
  const bool aligned = (a.lda % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.A) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(a.A2) & 15) == 0) &&
                       (!a.rotate_v || ((a.ldv % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.V) & 15) == 0) &&
                                        ((reinterpret_cast<uintptr_t>(a.V2) & 15) == 0)));
  const bool st_aligned = ((reinterpret_cast<uintptr_t>(a.A_st) | reinterpret_cast<uintptr_t>(a.A2_st) |
                            reinterpret_cast<uintptr_t>(a.V_st) | reinterpret_cast<uintptr_t>(a.V2_st)) & 7) == 0;
  (void)st_aligned;
  const bool staged = aligned && pairs <= (unsigned)kMaxCompactPairs && !blk_force_plain_update();
  constexpr size_t kStageSmem = sizeof(double) * kUpdStages * kTileDoubles;
  static DeviceOnce once;  // the shared-memory opt-in is per device
  int sm_count = 148;
  SVDJ_CUDA_TRY(once.run(
      [] {
        cudaError_t e = cudaFuncSetAttribute(blk_gram_staged_kernel<kUpdStages>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStageSmem);
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(blk_update_staged_kernel<kUpdStages>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)kStageSmem);
        return e;
      },
      &sm_count));
  if (staged) {
    rps = (rps + kTileRows - 1) / kTileRows * kTileRows;  // whole tiles per split
    SVDJ_CUDA_TRY(launch_k(blk_gram_staged_kernel<kUpdStages>, dim3(pairs, splits), 128, kStageSmem, st, a, k, rps, gpart));
  } else {
    SVDJ_CUDA_TRY(launch_k(blk_gram_kernel<kGramWarps, kGramUnroll>, dim3(pairs, splits), kGramWarps * 32, 0, st, a, k, rps, gpart));
  }
  SVDJ_CUDA_TRY(launch_k(blk_eig_kernel, dim3(pairs), 256, 0, st, a, k, splits, gpart, wbuf, wflag));
  const unsigned strips = ((a.m + 7) >> 3) + (a.rotate_v ? ((a.vrows + 7) >> 3) : 0);
  unsigned usplits = (8 * 148 + pairs - 1) / pairs;
  const unsigned ucap = (strips + kUpdWarps * 4 - 1) / (kUpdWarps * 4);  // >= 4 strips per warp
  if (usplits > ucap) usplits = ucap;
  if (usplits < 1) usplits = 1;
  if (staged) {
    constexpr size_t smem = kStageSmem;
    if (probe) SVDJ_CUDA_TRY(cudaEventRecord(probe->before, st));
    SVDJ_CUDA_TRY(launch_k(blk_update_staged_kernel<kUpdStages>, dim3(sm_count * kUpdCtasPerSm), 128, smem, st, a, k, wbuf,
                           wflag, pairs, probe ? probe->active_pairs : nullptr));
    if (probe) SVDJ_CUDA_TRY(cudaEventRecord(probe->after, st));
  } else {
    if (probe) SVDJ_CUDA_TRY(cudaEventRecord(probe->before, st));
    SVDJ_CUDA_TRY(launch_k(blk_update_kernel<kUpdWarps>, dim3(pairs, usplits), kUpdWarps * 32, 0, st, a, k, wbuf, wflag,
                           probe ? probe->active_pairs : nullptr));
    if (probe) SVDJ_CUDA_TRY(cudaEventRecord(probe->after, st));
  }
  if (launches) *launches += 3;
  SVDJ_CUDA_TRY(cudaGetLastError());
  return SVDJ_OK;
}

}  // namespace svdj
