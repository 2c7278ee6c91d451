"""GPU tests of the row-sharded solve for tall inputs (svdj_dgesvd_rows_device, svd-jacobi-mpi-cuda_b200/rows.py).

Single rank: the entry point with a do-nothing all-reduce must equal the plain device solve (same kernels, the
Gram matrices take the detour through the reduce kernel and the callback).  Two ranks (needs 2 GPUs): each holds
half of the rows; sigma must match LAPACK, the stacked U and V rows must be orthogonal and reconstruct A to
10*max(m,n)*eps (BASELINE.json's bounds).
"""
import importlib
import os
import socket

import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _rows():
    return importlib.import_module("svd-jacobi-mpi-cuda_b200.rows")


def _check(A0, U, s, V):
    m, n = A0.shape
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.all(np.diff(s) <= 0)
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


@pytest.mark.parametrize("m,n", [(4096, 256), (3001, 192), (20000, 128)])
def test_rows_single_rank_equals_device_solve(svdj, m, n):
    rows = _rows()
    A0 = H.normal_matrix(m, n, seed=m + n)
    dev = torch.device("cuda", 0)
    solver = rows.RowShardSolver(m, n, dev)
    assert (solver.row0, solver.m_local, solver.v_row0, solver.v_local) == (0, m, 0, n)
    solver.fill_local(torch.from_numpy(np.ascontiguousarray(A0.T)).to(dev))
    st = solver.solve()
    assert st.converged == 1 and st.reserved == 32
    assert solver.allreduce_calls >= 2 + st.sweeps       # norms before and after, one per step in between
    U = solver.U_local.cpu().numpy().T
    V = solver.V_local.cpu().numpy().T
    s = solver.sigma.cpu().numpy()
    _check(A0, U, s, V)
    # the same input through the plain device entry point: same sweeps, sigma to rounding
    A = A0.copy(order="F"); s2 = np.zeros(n); V2 = np.zeros((n, n), order="F")
    st2 = svdj.dgesvd(svdj.AllVec, svdj.AllVec, m, n, A, m, s2, V2, n, svdj.default_opts(variant=svdj.VARIANT_BLOCK, block=32))
    assert abs(st2.sweeps - st.sweeps) <= 1
    assert H.sigma_error_normwise(s, s2) <= 1e-13


def test_rows_rejects_what_it_cannot_do(svdj):
    rows = _rows()
    dev = torch.device("cuda", 0)
    solver = rows.RowShardSolver(1000, 64, dev)          # n < 128: no 64-column block pairs to speak of
    solver.fill_local(torch.zeros(64, 1000, dtype=torch.float64, device=dev))
    with pytest.raises(RuntimeError):
        solver.solve()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, m, n, out_path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        rows = _rows()
        dev = torch.device("cuda", rank)
        solver = rows.RowShardSolver(m, n, dev)
        A0 = H.normal_matrix(m, n, seed=5)
        mine = A0[solver.row0:solver.row0 + solver.m_local]
        solver.fill_local(torch.from_numpy(np.ascontiguousarray(mine.T)).to(dev))
        st = solver.solve()
        np.savez(f"{out_path}.{rank}.npz", U=solver.U_local.cpu().numpy().T, V=solver.V_local.cpu().numpy().T,
                 s=solver.sigma.cpu().numpy(), sweeps=st.sweeps, converged=st.converged, row0=solver.row0,
                 v_row0=solver.v_row0)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_rows_two_gpus(tmp_path):
    import torch.multiprocessing as mp
    m, n = 30000, 256
    out = str(tmp_path / "rows")
    mp.spawn(_worker, args=(2, _free_port(), m, n, out), nprocs=2, join=True)
    parts = [np.load(f"{out}.{r}.npz") for r in range(2)]
    assert all(bool(p["converged"]) for p in parts)
    assert np.array_equal(parts[0]["s"], parts[1]["s"])          # every rank ends with the same, complete sigma
    U = np.concatenate([p["U"] for p in parts], axis=0)
    V = np.concatenate([p["V"] for p in parts], axis=0)
    assert U.shape == (m, n) and V.shape == (n, n)
    _check(H.normal_matrix(m, n, seed=5), U, parts[0]["s"], V)
