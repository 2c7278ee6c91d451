"""GPU tests of the multi-GPU path.  One GPU is enough for the per-rank compute (both phases, two buffers) and
for an in-process emulation of the block tournament; the NCCL exchange itself needs >= 2 GPUs and is
skipped otherwise (tests/test_ring_cpu.py covers the same host logic with gloo)."""
import importlib
import os
import socket

import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _ring():
    return importlib.import_module("svd-jacobi-mpi-cuda_b200.ring")


def _check(A0, U, s, V):
    m, n = A0.shape
    assert np.all(np.diff(s) <= 0)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


def test_single_rank_ring_solver_on_gpu():
    """P = 1: the whole sweep is one SVDJ_PHASE_ALL over two buffers of n/2 columns."""
    ring = _ring()
    m, n = 300, 256
    A0 = H.uniform_matrix(m, n, seed=12)
    dev = torch.device("cuda", 0)
    solver = ring.RingSolver(m, n, dev)
    solver.scatter_from_root(torch.from_numpy(np.ascontiguousarray(A0.T)).to(dev))
    st = solver.solve()
    assert st["converged"] and st["sweeps"] <= 20
    sigma, order = solver.finalize()
    U, V = solver.gather_to_root(order)
    _check(A0, U.cpu().numpy().T, sigma.cpu().numpy(), V.cpu().numpy().T)


def test_ring_solver_rank_deficient_input_converges():
    """The zero-column floor on the ring (svdj_ring_ws_set_floor): without it the null columns of a rank-deficient
    input keep every sweep rotating until max_sweeps."""
    ring = _ring()
    m, n = 256, 192
    rng = np.random.default_rng(3)
    A0 = np.asfortranarray(rng.standard_normal((m, 90)) @ rng.standard_normal((90, n)))
    dev = torch.device("cuda", 0)
    solver = ring.RingSolver(m, n, dev)
    solver.scatter_from_root(torch.from_numpy(np.ascontiguousarray(A0.T)).to(dev))
    st = solver.solve()
    assert st["converged"] and st["sweeps"] <= 25, st
    sigma, order = solver.finalize()
    s = sigma.cpu().numpy()
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    U, V = solver.gather_to_root(order)
    V = V.cpu().numpy().T
    assert np.abs(V.T @ V - np.eye(n)).max() <= 10 * n * H.EPS


@pytest.mark.parametrize("P", [2, 4])
def test_emulated_tournament_on_one_gpu(P):
    """All P ranks' blocks live on one GPU and the rounds run rank after rank: checks SVDJ_PHASE_CROSS (bipartite
    schedule, two buffers) and that the tournament orthogonalises every pair of columns."""
    ring = _ring()
    m, n = 64 * P + 40, 64 * P
    w = n // (2 * P)
    A0 = H.uniform_matrix(m, n, seed=P)
    dev = torch.device("cuda", 0)
    lda, ldv = (m + 15) // 16 * 16, (n + 15) // 16 * 16
    tol = m * 2.0 ** -53
    backend = ring.GpuBackend(m, n, w, lda, ldv, tol)
    blocks = []
    for b in range(2 * P):
        blk = ring.Block(w, m, n, lda, ldv, dev)
        blk.block_id = b
        blk.A[:, :m] = torch.from_numpy(np.ascontiguousarray(A0.T[b * w:(b + 1) * w])).to(dev)
        idx = torch.arange(w, device=dev)
        blk.V[idx, b * w + idx] = 1.0
        blocks.append(blk)
    counters = torch.zeros(2, dtype=torch.int64, device=dev)
    rounds = ring.tournament_rounds(P)
    sweeps = 0
    for sweep in range(40):
        counters.zero_()
        for rd, (top, bottom) in enumerate(rounds):
            for r in range(P):
                backend.phase(blocks[top[r]], blocks[bottom[r]], ring.PHASE_ALL if rd == 0 else ring.PHASE_CROSS, counters)
        sweeps += 1
        if int(counters[0].item()) == 0:
            break
    assert int(counters[0].item()) == 0 and sweeps <= 20
    sig = torch.cat([backend.finalize(b) for b in blocks]).cpu().numpy()
    U = np.concatenate([b.A[:, :m].cpu().numpy() for b in blocks]).T
    V = np.concatenate([b.V[:, :n].cpu().numpy() for b in blocks]).T
    order = np.argsort(-sig, kind="stable")
    _check(A0, U[:, order], sig[order], V[:, order])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, m, n, out_path, exchange):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        ring = _ring()
        dev = torch.device("cuda", rank)
        solver = ring.RingSolver(m, n, dev, exchange=exchange)
        A0 = H.uniform_matrix(m, n, seed=77)
        solver.scatter_from_root(torch.from_numpy(np.ascontiguousarray(A0.T)).to(dev) if rank == 0 else None)
        st = solver.solve()
        sigma, order = solver.finalize()
        U, V = solver.gather_to_root(order)
        if rank == 0:
            np.savez(out_path, U=U.cpu().numpy().T, s=sigma.cpu().numpy(), V=V.cpu().numpy().T, sweeps=st["sweeps"],
                     converged=st["converged"])
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("exchange", ["nccl", "fused"])
def test_ring_solver_two_gpus(tmp_path, exchange):
    """exchange='nccl': grouped send/recv between rounds; 'fused': the update kernel of the last step of a round
    stores the travelling blocks into the neighbour's IPC-mapped buffers (no send/recv)."""
    import torch.multiprocessing as mp
    m, n = 512, 512
    out = str(tmp_path / "r.npz")
    mp.spawn(_nccl_worker, args=(2, _free_port(), m, n, out, exchange), nprocs=2, join=True)
    r = np.load(out)
    assert bool(r["converged"])
    _check(H.uniform_matrix(m, n, seed=77), r["U"], r["s"], r["V"])
