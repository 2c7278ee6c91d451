"""One rank per process: binding of libsvdj's in-library multi-GPU driver (csrc/svdj_mg.cu, svdj_mg_rank_*).

The driver itself -- tournament, buffer roles, fused block exchange, flags, convergence -- is C++ inside libsvdj.so.
This module only supplies the one operation the library asks of a communicator, an all-gather of a few hundred host
bytes at set-up, through ``torch.distributed`` (NCCL under torchrun, gloo in CPU-only tests of the binding), and
passes pointers.  It replaces the reference's ``mpiexec -n P`` launch of ``omp_mpi_cuda_dgesvd_local_matrices``
(build/runSVDMPICUDA.slurm:24, main.cu:1587): every rank calls, only rank 0 holds A, s and V (main.cu:1455).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import AllgatherFn, MgComm, Opts, Stats, SvdjError, _check, last_error, lib


class MgRank:
    def __init__(self, max_m: int, max_n: int, group=None, device=None):
        self.group = group
        self.size = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self._cb = AllgatherFn(self._allgather)      # keep the trampoline alive
        comm = MgComm(self.rank, self.size, self._cb, None)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            rc = lib.svdj_mg_rank_create(C.byref(comm), max_m, max_n, C.byref(h))
        if rc != 0:
            raise SvdjError(rc, last_error())
        self.handle = h

    def _allgather(self, ctx, send, recv, nbytes) -> int:
        try:
            nbytes = int(nbytes)
            mine = torch.frombuffer(bytearray(C.string_at(send, nbytes)), dtype=torch.uint8)
            backend = dist.get_backend(self.group)
            if backend == "nccl":
                mine = mine.to(self.device)
            out = [torch.empty_like(mine) for _ in range(self.size)]
            dist.all_gather(out, mine, group=self.group)
            blob = b"".join(bytes(t.cpu().numpy().tobytes()) for t in out)
            C.memmove(recv, blob, nbytes * self.size)
            return 0
        except Exception:                                # never let an exception cross the C frame
            import traceback
            traceback.print_exc()
            return 1

    def dgesvd_device(self, jobu, jobv, m, n, dA: int = 0, lda: int = 0, ds: int = 0, dV: int = 0, ldv: int = 0,
                      opts: Opts | None = None, allow_not_converged=False) -> Stats:
        """Collective.  Rank 0 passes device addresses on its GPU (A m x n column-major in, U out in place; s; V);
        the other ranks pass nothing."""
        st = Stats()
        with torch.cuda.device(self.device):
            rc = lib.svdj_mg_rank_dgesvd_device(self.handle, jobu, jobv, m, n, dA or None, lda, ds or None, dV or None,
                                                ldv, None if opts is None else C.byref(opts), C.byref(st))
        _check(rc, allow_not_converged)
        st.rc = rc
        return st

    def dgesvd(self, jobu, jobv, m, n, A: int = 0, lda: int = 0, s: int = 0, V: int = 0, ldv: int = 0,
               opts: Opts | None = None, allow_not_converged=False) -> Stats:
        """Collective, host addresses on rank 0 (the reference's contract: main.cu:1455, 1548-1556)."""
        st = Stats()
        with torch.cuda.device(self.device):
            rc = lib.svdj_mg_rank_dgesvd(self.handle, jobu, jobv, m, n, A or None, lda, s or None, V or None, ldv,
                                         None if opts is None else C.byref(opts), C.byref(st))
        _check(rc, allow_not_converged)
        st.rc = rc
        return st

    def close(self):
        if self.handle:
            lib.svdj_mg_rank_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
