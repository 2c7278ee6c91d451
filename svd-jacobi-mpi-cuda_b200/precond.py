"""QR-preconditioned solve for torch tensors: a thin call of the library option ``svdj_opts.precondition``.

The front-end itself (de Rijk column sort, blocked Householder QR, Jacobi on R^T, application of Q) is CUDA code
inside libsvdj.so (csrc/svdj_qr.cu, solve_with_qr in csrc/svdj_api.cu); nothing here calls a library QR or GEMM.
The reference names the use-case (tall or ill-conditioned inputs, lib/JacobiMethods.cu:12-14) but has no such step.
"""
from __future__ import annotations

import torch


def dgesvd_preconditioned(A: torch.Tensor, opts=None, qr_steps: int = 1):
    """A: (m, n) float64 CUDA tensor, m >= n (any strides).  Returns (U (m, n), s (n), V (n, n), stats) with
    A = U diag(s) V^T, s descending.  qr_steps = 2 factors R^T once more (strongly graded spectra)."""
    from . import AllVec, default_opts, dgesvd_device

    if A.dtype != torch.float64 or not A.is_cuda:
        raise TypeError("A must be a float64 CUDA tensor")
    m, n = A.shape
    if m < n:
        raise ValueError("m >= n expected")
    if qr_steps not in (1, 2):
        raise ValueError("qr_steps is 1 or 2")
    o = default_opts() if opts is None else opts
    o.precondition = qr_steps
    dA = A.t().contiguous()                      # (n, m): the column-major bytes of the m x n matrix
    ds = torch.empty(n, dtype=torch.float64, device=A.device)
    dV = torch.empty(n, n, dtype=torch.float64, device=A.device)
    with torch.cuda.device(A.device):
        st = dgesvd_device(AllVec, AllVec, m, n, dA.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n, o,
                           stream=torch.cuda.current_stream().cuda_stream)
    return dA.t(), ds, dV.t(), st
