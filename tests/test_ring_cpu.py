"""CPU tests of the multi-GPU host logic (svd-jacobi-mpi-cuda_b200/ring.py) with gloo, world_size 2 and 3.

The product's per-rank compute is libsvdj's GPU phase; here a numpy stand-in backend (test-only, the oracle's
formulas: lib/JacobiMethods.cu:450-477, 1483-1491) takes its place so that the tournament of column blocks,
the block exchange, the per-sweep all-reduce and the scatter/gather can be checked without a GPU.
"""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H


def _ring():
    return importlib.import_module("svd-jacobi-mpi-cuda_b200.ring")


@pytest.mark.parametrize("P", [1, 2, 3, 4, 5, 8])
def test_tournament_meets_every_block_pair_exactly_once(P):
    ring = _ring()
    rounds = ring.tournament_rounds(P)
    assert len(rounds) == 2 * P - 1
    seen = set()
    for top, bottom in rounds:
        assert sorted(top + bottom) == list(range(2 * P)), "every block is held by exactly one rank"
        assert top[0] == 0, "block 0 never moves"
        for x, y in zip(top, bottom):
            key = (min(x, y), max(x, y))
            assert key not in seen
            seen.add(key)
    assert len(seen) == P * (2 * P - 1)
    # the rotation has period 2P-1: a sweep ends in the arrangement it started from
    t, b = ring.initial_arrangement(P)
    for _ in range(2 * P - 1):
        t, b = ring.rotate(t, b)
    assert (t, b) == ring.initial_arrangement(P)


class NumpyBackend:
    """Stand-in for GpuBackend: same interface, level-1 rotations in numpy on CPU tensors."""

    def __init__(self, m, n, w, tol):
        self.m, self.n, self.w, self.tol = m, n, w, tol

    def phase(self, top, bottom, phase, counters):
        m, n, w = self.m, self.n, self.w
        A = [top.A.numpy(), bottom.A.numpy()]
        V = [top.V.numpy(), bottom.V.numpy()]

        def col(c):
            return (A[c // w][c % w, :m], V[c // w][c % w, :n])

        if phase == 0:
            steps = [H.oracle_schedule(2 * w, st) for st in range(H.oracle_lib().oracle_num_steps(2 * w))]
        else:
            steps = [np.array([(i, w + (i + t) % w) for i in range(w)]) for t in range(w)]
        applied, mx = 0, 0.0
        for pairs in steps:
            for p, q in pairs:
                ap, vp = col(int(p))
                aq, vq = col(int(q))
                alpha, beta, gamma = float(ap @ aq), float(ap @ ap), float(aq @ aq)
                den = np.sqrt(beta * gamma)
                if den > 0:
                    mx = max(mx, abs(alpha) / den)
                if not abs(alpha) > self.tol * den:
                    continue
                tau = (gamma - beta) / (2.0 * alpha)
                t = 1.0 / (tau + np.sqrt(1 + tau * tau)) if tau >= 0 else 1.0 / (tau - np.sqrt(1 + tau * tau))
                c = 1.0 / np.sqrt(1 + t * t)
                s = t * c
                for x, y in ((ap, aq), (vp, vq)):
                    tmp = x.copy()
                    x[:] = c * tmp - s * y
                    y[:] = s * tmp + c * y
                applied += 1
        counters[0] += applied
        bits = int(np.array([mx]).view(np.int64)[0])
        counters[1] = max(int(counters[1]), bits)

    def finalize(self, blk, scale_u=True):
        a = blk.A.numpy()[:, : self.m]
        sigma = np.linalg.norm(a, axis=1)
        if scale_u:
            nz = sigma > 0
            a[nz] /= sigma[nz, None]
        return torch.from_numpy(sigma.copy())


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, m, n, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ring = _ring()
        A0 = H.uniform_matrix(m, n, seed=m + n)
        w = n // (2 * world)
        solver = ring.RingSolver(m, n, torch.device("cpu"), backend=NumpyBackend(m, n, w, m * 2.0 ** -53))
        # rank 0 owns the matrix (the reference's contract: only the root's A, s, V are valid, main.cu:1455)
        A_root = torch.from_numpy(np.ascontiguousarray(A0.T)) if rank == 0 else None
        solver.scatter_from_root(A_root)
        # every rank must have received exactly its two blocks
        for blk in (solver.top, solver.bottom):
            c0 = blk.block_id * w
            assert np.array_equal(blk.A.numpy()[:, :m], A0.T[c0:c0 + w])
        # the physical exchange follows the tournament table: stamp, rotate a whole sweep, compare
        arrangements = ring.tournament_rounds(world)
        solver._arrangements = arrangements
        for rd in range(len(arrangements)):
            t, b = arrangements[rd]
            assert solver.top.block_id == t[rank] and solver.bottom.block_id == b[rank]
            assert np.array_equal(solver.top.A.numpy()[:, :m], A0.T[t[rank] * w:(t[rank] + 1) * w])
            assert np.array_equal(solver.bottom.A.numpy()[:, :m], A0.T[b[rank] * w:(b[rank] + 1) * w])
            solver._exchange()
            solver._set_ids(rd + 1)
        stats = solver.solve()
        assert stats["converged"], stats
        sigma, order = solver.finalize()
        U, V = solver.gather_to_root(order)
        if rank == 0:
            np.savez(out_path, U=U.numpy().T, s=sigma.numpy(), V=V.numpy().T, sweeps=stats["sweeps"],
                     exchanged=stats["exchanged_bytes"])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,m,n", [(2, 80, 64), (3, 100, 96)])
def test_ring_solver_with_gloo(tmp_path, world, m, n):
    out = str(tmp_path / "result.npz")
    mp.spawn(_worker, args=(world, _free_port(), m, n, out), nprocs=world, join=True)
    r = np.load(out)
    A0 = H.uniform_matrix(m, n, seed=m + n)
    U, s, V = r["U"], r["s"], r["V"]
    assert np.all(np.diff(s) <= 0)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    assert int(r["sweeps"]) <= 20 and int(r["exchanged"]) > 0
    # same singular values as the single-process oracle
    Uo, so, Vo, ost, rc = H.oracle_svd(A0)
    assert H.sigma_error_normwise(s, so) <= 1e-12


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
def test_buffer_roles_follow_the_tournament(P):
    """advance_roles() is what both exchanges (send/recv and peer stores) are driven by: every transfer must
    land in a spare buffer of the receiver and the blocks must end up where rotate() says, round after round."""
    ring = _ring()
    roles = ring.initial_roles(P)
    top, bot = ring.initial_arrangement(P)
    held = [{roles[q]["top"]: top[q], roles[q]["bottom"]: bot[q]} for q in range(P)]  # buffer -> block id
    for rd in range(3 * (2 * P - 1)):
        new, transfers = ring.advance_roles(P, roles)
        arrived = [dict() for _ in range(P)]
        for (sr, sb, dr, db) in transfers:
            assert abs(sr - dr) == 1, "only ring neighbours talk"
            assert db in roles[dr]["spare"], "a block may only be written into a spare buffer"
            assert sb in (roles[sr]["top"], roles[sr]["bottom"])
            arrived[dr][db] = held[sr][sb]
        top, bot = ring.rotate(top, bot)
        for q in range(P):
            for role in ("top", "bottom"):
                if new[q][role] not in arrived[q]:
                    arrived[q][new[q][role]] = held[q][new[q][role]]  # stayed in place (or turned the corner)
            assert arrived[q][new[q]["top"]] == top[q] and arrived[q][new[q]["bottom"]] == bot[q]
            assert sorted([new[q]["top"], new[q]["bottom"]] + new[q]["spare"]) == [0, 1, 2, 3]
        roles = new
        held = [{new[q]["top"]: top[q], new[q]["bottom"]: bot[q]} for q in range(P)]


def test_quadratic_stop_rule():
    """The confirming sweep is skipped only when what the sweep can have left behind is below tol/2."""
    from importlib import import_module
    ring = import_module("svd-jacobi-mpi-cuda_b200.ring")
    tol = 4096 * 2.0 ** -53
    assert ring.quadratic_stop(240, 2.87e-8, tol)           # the 4096^2 case: 240 x 8.2e-16 = 2.0e-13 < 2.3e-13
    assert not ring.quadratic_stop(2657, 1.2e-2, tol)       # still far from converged
    assert not ring.quadratic_stop(2153, 1.8e-7, 20000 * 2.0 ** -53)  # 7e-11 > tol / 2
    assert not ring.quadratic_stop(10, 0.0, tol)            # a backend that does not track the cosine never stops early
    assert not ring.quadratic_stop(1, 2e-6, 1.0)            # the cosine itself must be small


def test_row_partition_tiles_the_rows():
    """Row-sharded driver: contiguous slices that tile [0, m), whole 64-row tiles except the last, near even."""
    from importlib import import_module
    rows = import_module("svd-jacobi-mpi-cuda_b200.rows")
    for m, parts, mult in [(1048576, 8, 64), (1000, 3, 64), (512, 8, 1), (100, 8, 64), (65, 2, 64), (4096, 1, 64)]:
        p = rows.row_partition(m, parts, mult)
        assert len(p) == parts and p[0][0] == 0
        assert all(cnt > 0 for _, cnt in p)
        assert all(p[i][0] + p[i][1] == p[i + 1][0] for i in range(parts - 1))
        assert p[-1][0] + p[-1][1] == m
        if m >= parts * mult * 2:
            assert all(cnt % mult == 0 for _, cnt in p[:-1])
            assert max(c for _, c in p) - min(c for _, c in p) <= 2 * mult
    import pytest
    with pytest.raises(ValueError):
        rows.row_partition(3, 4)


class NumpyRowsBackend:
    """Stand-in for rows.LibBackend: the same protocol in numpy at column-pair granularity -- per step the partial
    (alpha, beta, gamma) of every pair of the reference's schedule are summed over the ranks with the solver's
    all-reduce, every rank derives the same rotations and applies them to its own rows of A and V."""

    def solve(self, solver, allow_not_converged):
        n, ml, vl = solver.n, solver.m_local, solver.v_local
        A = solver.A.numpy()[:, :ml]          # row j = column j, this rank's rows
        V = solver.V.numpy()[:, :vl]
        V[:] = 0.0
        V[solver.v_row0 + np.arange(vl), np.arange(vl)] = 1.0      # rows v_row0.. of the identity
        steps = H.oracle_lib().oracle_num_steps(n)
        sweeps, converged = 0, False
        for sweep in range(solver.max_sweeps):
            applied = 0
            for st in range(steps):
                pairs = H.oracle_schedule(n, st)
                pairs = pairs[(pairs[:, 0] < n) & (pairs[:, 1] < n)]
                trip = torch.from_numpy(np.stack([np.einsum("ij,ij->i", A[pairs[:, 0]], A[pairs[:, 1]]),
                                                  np.einsum("ij,ij->i", A[pairs[:, 0]], A[pairs[:, 0]]),
                                                  np.einsum("ij,ij->i", A[pairs[:, 1]], A[pairs[:, 1]])], axis=1))
                solver.allreduce_tensor(trip)
                for (p, q), (al, be, ga) in zip(pairs.tolist(), trip.numpy().tolist()):
                    if not abs(al) > solver.tol * np.sqrt(be * ga):
                        continue
                    tau = (ga - be) / (2.0 * al)
                    t = 1.0 / (tau + np.sqrt(1 + tau * tau)) if tau >= 0 else 1.0 / (tau - np.sqrt(1 + tau * tau))
                    c = 1.0 / np.sqrt(1 + t * t)
                    s = t * c
                    for X in (A, V):
                        xp, xq = X[p].copy(), X[q].copy()
                        X[p], X[q] = c * xp - s * xq, s * xp + c * xq
                    applied += 1
            sweeps += 1
            if applied == 0:
                converged = True
                break
        sq = torch.from_numpy(np.einsum("ij,ij->i", A, A))
        solver.allreduce_tensor(sq)
        sigma = np.sqrt(sq.numpy())
        order = np.argsort(-sigma, kind="stable")
        A[:] = A[order] / sigma[order][:, None]
        V[:] = V[order]
        solver.sigma.copy_(torch.from_numpy(sigma[order].copy()))
        return {"sweeps": sweeps, "converged": converged}


def _rows_worker(rank, world, port, m, n, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from importlib import import_module
        rows = import_module("svd-jacobi-mpi-cuda_b200.rows")
        solver = rows.RowShardSolver(m, n, torch.device("cpu"), backend=NumpyRowsBackend())
        parts = rows.row_partition(m, world)
        assert (solver.row0, solver.m_local) == parts[rank]
        assert (solver.v_row0, solver.v_local) == rows.row_partition(n, world, 1)[rank]
        A0 = H.normal_matrix(m, n, seed=m * 3 + n)
        solver.fill_local(torch.from_numpy(np.ascontiguousarray(A0[solver.row0:solver.row0 + solver.m_local].T)))
        st = solver.solve()
        assert st["converged"]
        # one all-reduce per step of every sweep, one for sigma
        steps = H.oracle_lib().oracle_num_steps(n)
        assert solver.allreduce_calls == st["sweeps"] * steps + 1
        np.savez(f"{out_path}.{rank}.npz", U=solver.U_local.numpy().T, V=solver.V_local.numpy().T,
                 s=solver.sigma.numpy(), sweeps=st["sweeps"])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,m,n", [(2, 300, 24), (3, 500, 20)])
def test_row_sharded_protocol_with_gloo(tmp_path, world, m, n):
    """The row-sharded driver's host side (partition, buffers, all-reduce plumbing, row slices of U and V) with a
    numpy stand-in for the per-rank compute: every rank ends with the same sigma, the stacked slices are an SVD."""
    out = str(tmp_path / "rows")
    mp.spawn(_rows_worker, args=(world, _free_port(), m, n, out), nprocs=world, join=True)
    parts = [np.load(f"{out}.{r}.npz") for r in range(world)]
    for p in parts[1:]:
        assert np.array_equal(p["s"], parts[0]["s"]) and int(p["sweeps"]) == int(parts[0]["sweeps"])
    U = np.concatenate([p["U"] for p in parts], axis=0)
    V = np.concatenate([p["V"] for p in parts], axis=0)
    A0 = H.normal_matrix(m, n, seed=m * 3 + n)
    s = parts[0]["s"]
    assert U.shape == (m, n) and V.shape == (n, n)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    Uo, so, Vo, ost, rc = H.oracle_svd(A0)                  # same sweeps as the single-process oracle
    assert int(parts[0]["sweeps"]) == ost.sweeps
