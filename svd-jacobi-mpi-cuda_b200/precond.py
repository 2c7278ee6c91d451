"""QR-preconditioned front-end ("next" row 3) -- optional, outside the hot path.

One-sided Jacobi needs many sweeps on matrices whose singular values are strongly graded but whose columns
are not (e.g. A = U diag(sigma) V^T with Haar U, V and cond 1e12: 47 sweeps at n = 512 with the reference's
ordering).  Drmac and Veselic's remedy is to factor A = Q R first and run Jacobi on L = R^T: the same problem,
but the columns of L are graded like sigma, and the sweep count drops (13 at n = 512; measured in DESIGN.md).

    A = Q R,   R^T J = W diag(s)   (one-sided Jacobi on R^T: J accumulates the rotations, W = normalised columns)
    =>  A = (Q J) diag(s) W^T,     U = Q J,  V = W.

The QR factorisation and the product Q J are plain library calls here (torch.linalg.qr = cuSOLVER geqrf/orgqr,
torch.matmul = cuBLAS DGEMM); the Jacobi solve in the middle is libsvdj, unchanged.  A hand-written TSQR would
replace the two library calls; until then this front-end is an option, not the default, and bench.py does not
use it.
"""
from __future__ import annotations

import torch


def dgesvd_preconditioned(A: torch.Tensor, opts=None, qr_steps: int = 1):
    """A: (m, n) float64 CUDA tensor, m >= n (any strides).  Returns (U (m, n), s (n), V (n, n), stats) with
    A = U diag(s) V^T, s descending.  qr_steps = 2 factors R^T once more (helps for cond >~ 1e10)."""
    from . import dgesvd_device, AllVec

    if A.dtype != torch.float64 or not A.is_cuda:
        raise TypeError("A must be a float64 CUDA tensor")
    m, n = A.shape
    if m < n:
        raise ValueError("m >= n expected")
    Q, R = torch.linalg.qr(A, mode="reduced")  # Q (m, n), R (n, n) upper triangular
    Q2 = None
    if qr_steps >= 2:
        # R^T = Q2 R2  =>  R = R2^T Q2^T  =>  A = Q R2^T Q2^T; Jacobi runs on L = R2^T
        Q2, R2 = torch.linalg.qr(R.T.contiguous())
        Lcm = R2.contiguous()    # row j of R2 = column j of R2^T: the column-major bytes of L = R2^T
    else:
        Lcm = R.contiguous()     # row j of R = column j of R^T: the column-major bytes of L = R^T
    ds = torch.empty(n, dtype=torch.float64, device=A.device)
    dJ = torch.empty(n, n, dtype=torch.float64, device=A.device)
    st = dgesvd_device(AllVec, AllVec, n, n, Lcm.data_ptr(), n, ds.data_ptr(), dJ.data_ptr(), n, opts,
                       stream=torch.cuda.current_stream().cuda_stream)
    W = Lcm.T  # normalised columns of L J (left singular vectors of L), (n, n)
    J = dJ.T   # right singular vectors of L, (n, n)
    if qr_steps >= 2:
        # L = R2^T = W s J^T  =>  A = Q (W s J^T) Q2^T = (Q W) s (Q2 J)^T
        U = Q @ W
        V = Q2 @ J
    else:
        # L = R^T = W s J^T  =>  R = J s W^T  =>  A = (Q J) s W^T
        U = Q @ J
        V = W
    return U, ds, V, st
