"""Row-sharded solve for tall inputs on several GPUs ("Tall-skinny").

One process per GPU.  Rank r holds a slice of the rows of A (all n columns) and a slice of the rows of V.  A step
of the block path needs the other ranks only for the Gram matrices of its block pairs (18 KiB per pair): the
library reduces its partial sums over the local row splits, this module's callback sums them over the ranks with
one ``torch.distributed.all_reduce`` on the compute stream, and every rank then runs the same small eigen-solves
and updates its own rows.  No column ever travels -- the reference moves whole columns of m doubles through rank 0
for every pair of every step (main.cu:583-942), which is what makes its 1,048,576 x 512 case impractical.

    solver = RowShardSolver(m, n, device, group)        # every rank
    solver.fill_local(A_rows)                           # this rank's rows, as an (n, m_local) tensor
    stats = solver.solve()
    sigma, U_rows, V_rows = solver.sigma, solver.U_local, solver.V_local
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.distributed as dist

from . import AllreduceFn, AllVec, Stats, default_opts, lib
from .ring import _DeviceMemory

def row_partition(m: int, parts: int, multiple: int = 64) -> list[tuple[int, int]]:
    """[(first row, row count)] per rank: contiguous slices, whole multiples of `multiple` rows except the last
    (the kernels stage 64-row tiles), as even as that allows; every rank gets at least one row when m >= parts."""
    if parts < 1 or m < parts:
        raise ValueError("need at least one row per rank")
    chunks = -(-m // multiple)
    out, r0 = [], 0
    for r in range(parts):
        c = chunks // parts + (1 if r < chunks % parts else 0)
        cnt = min(c * multiple, m - r0)
        if r == parts - 1:
            cnt = m - r0
        out.append((r0, cnt))
        r0 += cnt
    if any(cnt <= 0 for _, cnt in out):      # very short inputs: fall back to plain even slices
        base, out, r0 = m // parts, [], 0
        for r in range(parts):
            cnt = base + (1 if r < m % parts else 0)
            out.append((r0, cnt))
            r0 += cnt
    return out


class LibBackend:
    """The product path: libsvdj's row-sharded entry point on this rank's GPU (there is no CPU fallback)."""

    def solve(self, solver: "RowShardSolver", allow_not_converged: bool) -> Stats:
        st = Stats()
        opts = default_opts(tol=solver.tol, max_sweeps=solver.max_sweeps)
        stream = torch.cuda.current_stream(solver.device).cuda_stream
        with torch.cuda.device(solver.device):
            rc = lib.svdj_dgesvd_rows_device(AllVec, AllVec, solver.m_local, solver.n, solver.A.data_ptr(), solver.lda,
                                             solver.sigma.data_ptr(), solver.V.data_ptr(), solver.ldv, solver.v_local,
                                             solver.v_row0, C.byref(opts), C.byref(st), C.c_void_p(stream),
                                             solver._cb, None)
        st.rc = rc
        if rc < 0 or rc > 1 or (rc == 1 and not allow_not_converged):
            raise RuntimeError(f"svdj_dgesvd_rows_device -> {rc}: {lib.svdj_last_error().decode()}")
        return st


class RowShardSolver:
    def __init__(self, m: int, n: int, device, group=None, tol: Optional[float] = None, max_sweeps: int = 60,
                 backend=None):
        self.m, self.n, self.device, self.group = m, n, torch.device(device), group
        self.P = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.row0, self.m_local = row_partition(m, self.P)[self.rank]
        self.v_row0, self.v_local = row_partition(n, self.P, multiple=1)[self.rank]
        self.tol = tol if tol is not None else m ** 0.5 * 2.0 ** -53  # dgesvj CTOL when vectors are wanted
        self.max_sweeps = max_sweeps
        self.lda = (self.m_local + 15) // 16 * 16           # 128-byte columns: aligned, even
        self.ldv = (self.v_local + 15) // 16 * 16
        self.A = torch.zeros(n, self.lda, dtype=torch.float64, device=self.device)   # row j = column j
        self.V = torch.zeros(n, self.ldv, dtype=torch.float64, device=self.device)
        self.sigma = torch.zeros(n, dtype=torch.float64, device=self.device)
        self.allreduce_calls = 0
        self.allreduce_bytes = 0
        self._cb = AllreduceFn(self._allreduce)             # keep the trampoline alive
        self.backend = backend if backend is not None else LibBackend()   # tests pass a numpy stand-in (gloo)

    def allreduce_tensor(self, t: torch.Tensor) -> None:
        """Sum over the ranks, in place (on torch's current stream for CUDA tensors = the solve's stream)."""
        if self.P > 1:
            dist.all_reduce(t, group=self.group)
        self.allreduce_calls += 1
        self.allreduce_bytes += t.numel() * t.element_size()

    # called by libsvdj between the Gram kernel and the eigen-solves of every step (and for the column norms)
    def _allreduce(self, ctx, dptr, count, stream) -> int:
        try:
            self.allreduce_tensor(torch.as_tensor(_DeviceMemory(int(dptr), int(count)), device=self.device))
            return 0
        except Exception:                                    # never let an exception cross the C frame
            import traceback
            traceback.print_exc()
            return 1

    def fill_local(self, A_rows: torch.Tensor) -> None:
        """A_rows: (n, m_local) -- this rank's rows of the column-major matrix (row j = column j)."""
        if tuple(A_rows.shape) != (self.n, self.m_local):
            raise ValueError(f"expected {(self.n, self.m_local)}, got {tuple(A_rows.shape)}")
        self.A.zero_()
        self.A[:, : self.m_local].copy_(A_rows)

    def solve(self, allow_not_converged: bool = False):
        return self.backend.solve(self, allow_not_converged)

    @property
    def U_local(self) -> torch.Tensor:
        """(n, m_local): this rank's rows of U (row k = left singular vector k)."""
        return self.A[:, : self.m_local]

    @property
    def V_local(self) -> torch.Tensor:
        """(n, v_local): this rank's rows of V."""
        return self.V[:, : self.v_local]
