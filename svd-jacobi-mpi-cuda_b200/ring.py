"""Multi-GPU driver: column blocks on a ring, one process per GPU.

Replaces the reference's MPI star (``/root/reference/main.cu:583-942``: every round-robin step rank 0 packs
and ``MPI_Send``s each rank its columns of A and V, ``MPI_Recv``s them back and barriers).  Here the n
columns are cut into 2P column blocks of w = n/(2P) columns.  Rank r keeps two of them resident on its
GPU (A-block m x w and the matching V-block n x w, packed in one allocation so that a block travels as
one message).  A sweep is a round-robin tournament of the 2P blocks (circle method, 2P-1 rounds):

    round 0      every rank orthogonalises ALL pairs of its 2w local columns  (SVDJ_PHASE_ALL)
    rounds 1..   only the pairs across its two blocks                         (SVDJ_PHASE_CROSS)
    after each round the blocks move one place round the ring (block 0 stays on rank 0): each rank
    sends at most two blocks to its neighbours and receives at most two.

so every pair of columns meets exactly once per sweep, as in the reference's ordering, and the only
collective is one all-reduce of the rotation count per sweep (the reference barriers every step).

Two ways to move the blocks:

* ``exchange="nccl"``: ``torch.distributed`` point-to-point operations batched per round (grouped
  ncclSend/ncclRecv over NVLink on GPUs; gloo in the CPU tests).
* ``exchange="fused"``: no send/recv at all.  Every rank maps its neighbours' block buffers through CUDA
  IPC; the update kernel of the LAST step of a round stores the travelling blocks straight into the
  buffers their next owner will read (peer stores over NVLink, tile by tile while the kernel is still
  multiplying), and a one-element all-reduce orders the rounds.

The per-rank compute is libsvdj's block-Jacobi phase (``svdj_ring_phase_device``).  The compute backend is
injected so that the host logic (tournament, buffer roles, exchange, convergence) can be tested on CPU with
a stand-in backend; the product backend is :class:`GpuBackend` and there is no CPU fallback in this module.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

PHASE_ALL, PHASE_CROSS = 0, 1


# ----------------------------------------------------------------------------------------------------
# tournament of the 2P column blocks (pure host logic)
# ----------------------------------------------------------------------------------------------------
def initial_arrangement(P: int) -> Tuple[List[int], List[int]]:
    """Block ids held by rank r at round 0: top[r] = r, bottom[r] = P + r."""
    return list(range(P)), list(range(P, 2 * P))


def rotate(top: List[int], bottom: List[int]) -> Tuple[List[int], List[int]]:
    """One move of the circle method: top[0] stays; the other 2P-1 places shift by one round the ring
    top[1] -> top[2] -> ... -> top[P-1] -> bottom[P-1] -> ... -> bottom[0] -> top[1]."""
    P = len(top)
    if P == 1:
        return list(top), list(bottom)
    new_top = [top[0], bottom[0]] + top[1:P - 1]
    new_bottom = bottom[1:] + [top[P - 1]]
    return new_top, new_bottom


def quadratic_stop(applied: int, max_cos: float, tol: float) -> bool:
    """True when the sweep just finished leaves nothing for the next one to do (same rule as csrc/svdj_api.cu).

    Every pair whose cosine exceeded tol was rotated to zero during the sweep; what is left on a pair afterwards
    is what later rotations of its two columns put back, at most (rotations per column) x max_cos^2 <=
    applied x max_cos^2.  If that is below tol/2 the next sweep would find every pair under the threshold and
    rotate nothing -- the reference's stopping condition (lib/JacobiMethods.cu:462-466 yields no rotation in a
    whole sweep) already holds, so the confirming sweep is not run."""
    return 0.0 < max_cos < 1e-6 and applied * max_cos * max_cos < 0.5 * tol  # 0.0: the backend does not track it


def tournament_rounds(P: int) -> List[Tuple[List[int], List[int]]]:
    """The 2P-1 arrangements of one sweep."""
    top, bottom = initial_arrangement(P)
    out = []
    for _ in range(2 * P - 1):
        out.append((list(top), list(bottom)))
        top, bottom = rotate(top, bottom)
    return out


def initial_roles(P: int) -> List[Dict]:
    """Every rank owns four block buffers 0..3: which one is its top, its bottom, and which are spare."""
    return [{"top": 0, "bottom": 1, "spare": [2, 3]} for _ in range(P)]


def advance_roles(P: int, roles: List[Dict]):
    """Buffer roles after one rotation, for every rank, and the transfers that realise it:
    (src_rank, src_buffer, dst_rank, dst_buffer).  Deterministic, so every rank computes the same table and
    a sender knows which buffer of its neighbour is free.  Mirrors :func:`rotate`."""
    if P == 1:
        return [dict(r, spare=list(r["spare"])) for r in roles], []
    new, transfers = [], []
    for r in range(P):
        spare = list(roles[r]["spare"])
        if r == 0:
            nb = spare.pop()
            transfers.append((1, roles[1]["bottom"], 0, nb))          # new bottom[0] = old bottom[1]
            nt = roles[0]["top"]                                       # block 0 never moves
        else:
            nt = spare.pop()
            src = (r - 1, roles[r - 1]["top"]) if r >= 2 else (0, roles[0]["bottom"])
            transfers.append((src[0], src[1], r, nt))                  # new top[r] = old top[r-1] (old bottom[0] for r = 1)
            if r < P - 1:
                nb = spare.pop()
                transfers.append((r + 1, roles[r + 1]["bottom"], r, nb))  # new bottom[r] = old bottom[r+1]
            else:
                nb = roles[r]["top"]                                   # the last rank's top turns the corner
        new.append({"top": nt, "bottom": nb, "spare": [i for i in range(4) if i != nt and i != nb]})
    return new, transfers


# ----------------------------------------------------------------------------------------------------
# compute backend on the GPU (the product path)
# ----------------------------------------------------------------------------------------------------
class GpuBackend:
    """Per-rank compute through the C ABI of libsvdj.so (block-Jacobi phases on two resident column blocks)."""

    def __init__(self, m: int, n: int, w: int, lda: int, ldv: int, tol: float, inner_sweeps: int = 0):
        from . import lib, last_error, SvdjError  # the ctypes binding; fails loudly if libsvdj.so is missing

        self.lib, self._last_error, self._err = lib, last_error, SvdjError
        self.m, self.n, self.w, self.lda, self.ldv, self.tol, self.inner = m, n, w, lda, ldv, tol, inner_sweeps
        ws = C.c_void_p()
        rc = lib.svdj_ring_ws_create(m, 2 * w, C.byref(ws))
        if rc != 0:
            raise SvdjError(rc, last_error())
        self.ws = ws
        self.launches = 0  # kernels launched through this backend (3 per block step, 1-2 per finalize)

    def close(self):
        if self.ws:
            self.lib.svdj_ring_ws_destroy(self.ws)
            self.ws = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _count(self, phase):
        # three kernels per step; the phase runs 64-column kernels (blocks of 32) when the travelling blocks divide
        # into them, as csrc/svdj_ring.cu decides, else the 32-column ones (blocks of 16)
        width = 32 if (self.w % 32 == 0 and os.environ.get("SVDJ_BLOCK") != "16") else 16
        nb = 2 * self.w // width
        self.launches += 3 * (nb // 2 if phase == PHASE_CROSS else int(self.lib.svdj_num_steps(nb)))

    def set_floor(self, floor_norm: float) -> None:
        """Columns with a norm <= floor_norm are numerically zero and are never rotated (svdj_ring_ws_set_floor)."""
        rc = self.lib.svdj_ring_ws_set_floor(self.ws, float(floor_norm))
        if rc != 0:
            raise self._err(rc, self._last_error())

    def phase(self, top, bottom, phase: int, counters: torch.Tensor) -> None:
        """top/bottom: Block objects; counters: int64[2] on the device."""
        stream = torch.cuda.current_stream().cuda_stream
        rc = self.lib.svdj_ring_phase_device(self.ws, self.m, self.n, self.w, top.A.data_ptr(), bottom.A.data_ptr(),
                                             self.lda, top.V.data_ptr(), bottom.V.data_ptr(), self.ldv, 1, phase,
                                             self.tol, self.inner, counters.data_ptr(), stream)
        if rc != 0:
            raise self._err(rc, self._last_error())
        self._count(phase)

    def phase_send(self, top, bottom, phase: int, counters: torch.Tensor, dst_top: Optional[int],
                   dst_bottom: Optional[int]) -> None:
        """Like :meth:`phase`, but the last step stores the updated top / bottom block at the given device
        addresses (base of a peer-mapped block buffer: A-part first, V-part at + w*lda doubles); None = local."""
        stream = torch.cuda.current_stream().cuda_stream
        off = self.w * self.lda * 8

        def split(p):
            return (p, p + off) if p else (None, None)

        ta, tv = split(dst_top)
        ba, bv = split(dst_bottom)
        rc = self.lib.svdj_ring_phase_send_device(self.ws, self.m, self.n, self.w, top.A.data_ptr(),
                                                  bottom.A.data_ptr(), self.lda, top.V.data_ptr(),
                                                  bottom.V.data_ptr(), self.ldv, phase, self.tol, self.inner,
                                                  counters.data_ptr(), ta, tv, ba, bv, stream)
        if rc != 0:
            raise self._err(rc, self._last_error())
        self._count(phase)

    def finalize(self, blk, scale_u: bool = True) -> torch.Tensor:
        sigma = torch.empty(self.w, dtype=torch.float64, device=blk.A.device)
        stream = torch.cuda.current_stream().cuda_stream
        rc = self.lib.svdj_ring_finalize_device(self.ws, self.m, self.w, blk.A.data_ptr(), self.lda, sigma.data_ptr(),
                                                1 if scale_u else 0, stream)
        if rc != 0:
            raise self._err(rc, self._last_error())
        self.launches += 2 if scale_u else 1
        return sigma


class _DeviceMemory:
    """Lets torch wrap a raw device allocation (``torch.as_tensor`` reads ``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, doubles: int):
        self.__cuda_array_interface__ = {"shape": (doubles,), "typestr": "<f8", "data": (ptr, False), "version": 3}


class Block:
    """One column block: A-part (w columns of m rows) and V-part (w columns of n rows) in one allocation,
    column-major (row j of the 2-D views is column j).  With ``ipc=True`` the allocation comes from
    ``svdj_ipc_alloc`` (plain cudaMalloc) so that a neighbour process can map it."""

    def __init__(self, w: int, m: int, n: int, lda: int, ldv: int, device, ipc: bool = False):
        doubles = w * (lda + ldv)
        self.raw_ptr = None
        if ipc:
            from . import lib, last_error, SvdjError
            p = C.c_void_p()
            rc = lib.svdj_ipc_alloc(doubles * 8, C.byref(p))
            if rc != 0:
                raise SvdjError(rc, last_error())
            self.raw_ptr = p.value
            self._keep = _DeviceMemory(self.raw_ptr, doubles)
            self.buf = torch.as_tensor(self._keep, device=device)
        else:
            self.buf = torch.zeros(doubles, dtype=torch.float64, device=device)
        self.A = self.buf[: w * lda].view(w, lda)
        self.V = self.buf[w * lda:].view(w, ldv)
        self.block_id = -1


# ----------------------------------------------------------------------------------------------------
# the ring solver
# ----------------------------------------------------------------------------------------------------
class RingSolver:
    def __init__(self, m: int, n: int, device, group=None, tol: Optional[float] = None, max_sweeps: int = 60,
                 backend=None, inner_sweeps: int = 0, exchange: str = "nccl"):
        self.group = group
        self.P = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        P = self.P
        if n % (2 * P) != 0 or (n // (2 * P)) % 16 != 0:
            raise ValueError(f"n = {n} must be a multiple of 32 * world_size = {32 * P} (column blocks of 16)")
        if exchange not in ("nccl", "fused"):
            raise ValueError("exchange must be 'nccl' or 'fused'")
        self.exchange = exchange if P > 1 else "nccl"
        self.m, self.n, self.w = m, n, n // (2 * P)
        self.lda = (m + 15) // 16 * 16   # 128-byte columns
        self.ldv = (n + 15) // 16 * 16
        self.device = device
        self.tol = tol if tol is not None else m ** 0.5 * 2.0 ** -53  # dgesvj CTOL when vectors are wanted
        self.max_sweeps = max_sweeps
        # the quadratic-convergence stop assumes separated singular values (csrc/svdj_api.cu); off unless asked for
        self.early_stop = os.environ.get("SVDJ_EARLY_STOP", "0") == "1"
        self.backend = backend if backend is not None else GpuBackend(m, n, self.w, self.lda, self.ldv, self.tol,
                                                                    inner_sweeps)
        fused = self.exchange == "fused"
        self.blocks = [Block(self.w, m, n, self.lda, self.ldv, device, ipc=fused) for _ in range(4)]
        self.roles = initial_roles(P)
        t, b = initial_arrangement(P)
        self.top.block_id, self.bottom.block_id = t[self.rank], b[self.rank]
        self.counters = torch.zeros(2, dtype=torch.int64, device=device)
        self.exchanged_bytes = 0
        self._arrangements = tournament_rounds(P)
        self._round = 0
        self.peer: Dict[int, List[int]] = {}
        if fused:
            self._map_neighbours()

    # -- the two active blocks -------------------------------------------------------------------------
    @property
    def top(self) -> Block:
        return self.blocks[self.roles[self.rank]["top"]]

    @property
    def bottom(self) -> Block:
        return self.blocks[self.roles[self.rank]["bottom"]]

    def _map_neighbours(self) -> None:
        """Exchange the IPC handles of the four block buffers and map those of the two ring neighbours."""
        from . import lib, last_error, SvdjError
        mine = torch.zeros(4, 64, dtype=torch.uint8)
        for i, blk in enumerate(self.blocks):
            h = C.create_string_buffer(64)
            rc = lib.svdj_ipc_export(blk.raw_ptr, h)
            if rc != 0:
                raise SvdjError(rc, last_error())
            mine[i] = torch.frombuffer(bytearray(h.raw), dtype=torch.uint8)
        mine = mine.to(self.device)
        everyone = [torch.empty_like(mine) for _ in range(self.P)]
        dist.all_gather(everyone, mine, group=self.group)
        for nb in {(self.rank - 1) % self.P, (self.rank + 1) % self.P} - {self.rank}:
            ptrs = []
            for i in range(4):
                raw = bytes(everyone[nb][i].cpu().numpy().tobytes())
                p = C.c_void_p()
                rc = lib.svdj_ipc_open(raw, C.byref(p))
                if rc != 0:
                    raise SvdjError(rc, last_error())
                ptrs.append(p.value)
            self.peer[nb] = ptrs

    # -- data in ----------------------------------------------------------------------------------------
    def fill_local(self, column_source: Callable[[int, int], torch.Tensor]) -> None:
        """column_source(col0, count) -> (count, m) tensor with the columns [col0, col0+count) of A.
        Resets the tournament to round 0."""
        self.roles = initial_roles(self.P)
        self._round = 0
        t, b = initial_arrangement(self.P)
        self.top.block_id, self.bottom.block_id = t[self.rank], b[self.rank]
        for blk in (self.top, self.bottom):
            c0 = blk.block_id * self.w
            blk.A.zero_()
            blk.A[:, : self.m].copy_(column_source(c0, self.w))
            blk.V.zero_()
            idx = torch.arange(self.w, device=self.device)
            blk.V[idx, c0 + idx] = 1.0  # V = I (main.cu:462-474), this block's columns of it
        if self.exchange == "fused":
            self._fence()  # nobody writes into a neighbour before every rank has finished (re)initialising

    def scatter_from_root(self, A_root: Optional[torch.Tensor]) -> None:
        """Rank 0 holds A as an (n, m) tensor (column-major m x n); every rank receives its two blocks."""
        if self.P == 1:
            self.fill_local(lambda c0, cnt: A_root[c0:c0 + cnt])
            return
        ops, keep = [], []
        t, b = initial_arrangement(self.P)
        if self.rank == 0:
            for r in range(1, self.P):
                for bid in (t[r], b[r]):
                    piece = A_root[bid * self.w:(bid + 1) * self.w].contiguous()
                    keep.append(piece)
                    ops.append(dist.P2POp(dist.isend, piece, r, self.group))
        else:
            recv = [torch.empty(self.w, self.m, dtype=torch.float64, device=self.device) for _ in range(2)]
            for piece in recv:
                ops.append(dist.P2POp(dist.irecv, piece, 0, self.group))
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        if self.rank == 0:
            self.fill_local(lambda c0, cnt: A_root[c0:c0 + cnt])
        else:
            pieces = {t[self.rank] * self.w: recv[0], b[self.rank] * self.w: recv[1]}
            self.fill_local(lambda c0, cnt: pieces[c0])

    # -- one rotation of the blocks ---------------------------------------------------------------------
    def _fence(self) -> None:
        """Orders the rounds of the fused exchange: a one-element all-reduce on the compute stream."""
        token = torch.zeros(1, dtype=torch.int32, device=self.device)
        dist.all_reduce(token, group=self.group)

    def _exchange(self) -> None:
        """NCCL / gloo exchange: send and receive the travelling blocks, then swap buffer roles."""
        P, r = self.P, self.rank
        new_roles, transfers = advance_roles(P, self.roles)
        if P > 1:
            ops = []
            for (sr, sb, dr, db) in transfers:
                if sr == r:
                    ops.append(dist.P2POp(dist.isend, self.blocks[sb].buf, dr, self.group))
                    self.exchanged_bytes += self.blocks[sb].buf.numel() * 8
            for (sr, sb, dr, db) in transfers:
                if dr == r:
                    ops.append(dist.P2POp(dist.irecv, self.blocks[db].buf, sr, self.group))
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        self.roles = new_roles

    def _phase_and_move(self, phase: int) -> None:
        """One round: the local phase, then the rotation of the blocks (separate or fused into the phase)."""
        if self.exchange == "nccl":
            self.backend.phase(self.top, self.bottom, phase, self.counters)
            self._exchange()
        else:
            P, r = self.P, self.rank
            new_roles, transfers = advance_roles(P, self.roles)
            dst = {"top": None, "bottom": None}
            for (sr, sb, dr, db) in transfers:
                if sr == r:
                    which = "top" if sb == self.roles[r]["top"] else "bottom"
                    dst[which] = self.peer[dr][db]
                    self.exchanged_bytes += self.blocks[sb].buf.numel() * 8
            self.backend.phase_send(self.top, self.bottom, phase, self.counters, dst["top"], dst["bottom"])
            self._fence()  # every rank's last step has finished writing before anybody reads a received block
            self.roles = new_roles
        self._round += 1
        t, b = self._arrangements[self._round % len(self._arrangements)]
        self.top.block_id, self.bottom.block_id = t[self.rank], b[self.rank]

    # kept for the tests of the host logic: one rotation without compute
    def _set_ids(self, round_index: int) -> None:
        t, b = self._arrangements[round_index % len(self._arrangements)]
        self.top.block_id, self.bottom.block_id = t[self.rank], b[self.rank]

    # -- the solve ---------------------------------------------------------------------------------------
    def _set_zero_column_floor(self) -> None:
        """tol x (largest column norm of the whole input), as in the single-GPU driver: without it the null columns of
        a rank-deficient or m < n input keep every sweep rotating (their cosines are rounding noise)."""
        if not hasattr(self.backend, "set_floor"):
            return
        local = torch.stack([blk.A[:, : self.m].norm(dim=1).max() for blk in (self.top, self.bottom)]).max().reshape(1)
        if self.P > 1:
            dist.all_reduce(local, op=dist.ReduceOp.MAX, group=self.group)
        self.backend.set_floor(self.tol * float(local.item()))

    def solve(self) -> dict:
        self._set_zero_column_floor()
        rounds = len(self._arrangements)
        sweeps, converged, applied_total, last_off = 0, False, 0, 0.0
        for sweep in range(self.max_sweeps):
            self.counters.zero_()
            for rd in range(rounds):
                self._phase_and_move(PHASE_ALL if rd == 0 else PHASE_CROSS)
            # one collective per sweep: rotations (sum) and largest off-diagonal cosine (max of the bit patterns)
            if self.P > 1:
                dist.all_reduce(self.counters[0:1], op=dist.ReduceOp.SUM, group=self.group)
                dist.all_reduce(self.counters[1:2], op=dist.ReduceOp.MAX, group=self.group)
            host = self.counters.cpu()
            applied = int(host[0])
            last_off = float(host[1:2].view(torch.float64)[0])
            applied_total += applied
            sweeps = sweep + 1
            # (0 < last_off <= 2 tol: nothing seen above the rounding noise of the inner products, csrc/svdj_api.cu)
            if applied == 0 or 0.0 < last_off <= 2.0 * self.tol or (
                    self.early_stop and quadratic_stop(applied, last_off, self.tol)):
                converged = True
                break
        return {"sweeps": sweeps, "converged": converged, "applied": applied_total, "last_offdiag": last_off,
                "exchanged_bytes": self.exchanged_bytes}

    # -- data out ----------------------------------------------------------------------------------------
    def finalize(self, scale_u: bool = True):
        """sigma of the local columns (and U = A / sigma in place).  Returns (sigma_all_sorted, order) where
        order[k] = global position (block_id * w + j) of the k-th largest singular value; both on every rank."""
        s_top = self.backend.finalize(self.top, scale_u)
        s_bot = self.backend.finalize(self.bottom, scale_u)
        local = torch.cat([s_top, s_bot])
        ids = torch.cat([self.top.block_id * self.w + torch.arange(self.w, device=self.device),
                         self.bottom.block_id * self.w + torch.arange(self.w, device=self.device)])
        if self.P > 1:
            all_s = [torch.empty_like(local) for _ in range(self.P)]
            all_i = [torch.empty_like(ids) for _ in range(self.P)]
            dist.all_gather(all_s, local, group=self.group)
            dist.all_gather(all_i, ids, group=self.group)
            local, ids = torch.cat(all_s), torch.cat(all_i)
        order = torch.argsort(local, descending=True, stable=True)
        return local[order], ids[order]

    def gather_to_root(self, order_ids: torch.Tensor):
        """Rank 0 receives every block and returns (U, V) as (n, m) and (n, n) tensors whose row k is the column
        of the k-th largest singular value; other ranks return (None, None)."""
        n, m, w = self.n, self.m, self.w
        if self.rank == 0:
            U_all = torch.empty(n, self.lda, dtype=torch.float64, device=self.device)
            V_all = torch.empty(n, self.ldv, dtype=torch.float64, device=self.device)
        ops = []
        if self.P > 1:
            # everyone tells rank 0 which blocks it holds (ids changed during the tournament)
            mine = torch.tensor([self.top.block_id, self.bottom.block_id], dtype=torch.int64, device=self.device)
            held = [torch.empty_like(mine) for _ in range(self.P)]
            dist.all_gather(held, mine, group=self.group)
            if self.rank == 0:
                for r in range(1, self.P):
                    for slot in range(2):
                        bid = int(held[r][slot])
                        ops.append(dist.P2POp(dist.irecv, U_all[bid * w:(bid + 1) * w], r, self.group))
                        ops.append(dist.P2POp(dist.irecv, V_all[bid * w:(bid + 1) * w], r, self.group))
            else:
                for blk in (self.top, self.bottom):
                    ops.append(dist.P2POp(dist.isend, blk.A, 0, self.group))
                    ops.append(dist.P2POp(dist.isend, blk.V, 0, self.group))
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        if self.rank != 0:
            return None, None
        for blk in (self.top, self.bottom):
            U_all[blk.block_id * w:(blk.block_id + 1) * w].copy_(blk.A)
            V_all[blk.block_id * w:(blk.block_id + 1) * w].copy_(blk.V)
        return U_all[order_ids, :m], V_all[order_ids, :n]
