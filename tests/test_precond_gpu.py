"""GPU tests of the QR front-end (svdj_opts.precondition; csrc/svdj_qr.cu), through the C ABI.

The reference has no factorisation step (it sweeps the raw columns, main.cu:478-1384), so the checker for the QR
itself is LAPACK (numpy.linalg.qr: same Householder sign convention, so R agrees entry by entry up to rounding) and
for the preconditioned solve the oracle / LAPACK singular values plus BASELINE.json's bounds:
sigma <= 1e-12 normwise, |U^T U - I|, |V^T V - I| and |A - U S V^T| / |A| <= 10 n eps.
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _qr_device(svdj, A0):
    """Factor A0 on the GPU; returns (Y, T) as torch tensors: Y (n, ld) holds the columns, T the panel factors."""
    m, n = A0.shape
    dY = torch.from_numpy(np.ascontiguousarray(A0.T)).cuda()       # row j = column j, ld = m
    dT = torch.zeros(int(svdj.lib.svdj_qr_tfactor_doubles(n)), dtype=torch.float64, device="cuda")
    rc = svdj.lib.svdj_qr_factor_device(m, n, dY.data_ptr(), m, dT.data_ptr(), None)
    assert rc == 0, svdj.last_error()
    torch.cuda.synchronize()
    return dY, dT


@pytest.mark.parametrize("m,n", [(64, 64), (300, 200), (777, 333), (1024, 1024), (4096, 96), (2500, 1000), (33, 33)])
def test_qr_factor_and_apply_q_match_lapack(svdj, m, n):
    A0 = H.normal_matrix(m, n, seed=3 * m + n)
    dY, dT = _qr_device(svdj, A0)
    Y = dY.cpu().numpy().T                                          # (m, n)
    R = np.triu(Y[:n, :n])
    Rl = np.linalg.qr(A0, mode="r")
    assert np.abs(R - Rl).max() <= 1e-12 * np.abs(Rl).max() * max(1.0, n / 64)
    # Q = Q [I; 0] through the apply kernel
    dC = torch.zeros(n, m, dtype=torch.float64, device="cuda")      # (m x n) column-major
    dC[torch.arange(n), torch.arange(n)] = 1.0
    rc = svdj.lib.svdj_qr_apply_q_device(m, n, dY.data_ptr(), m, dT.data_ptr(), dC.data_ptr(), m, n, None)
    assert rc == 0, svdj.last_error()
    torch.cuda.synchronize()
    Q = dC.cpu().numpy().T
    assert np.abs(Q.T @ Q - np.eye(n)).max() <= 10 * n * H.EPS
    assert np.linalg.norm(Q @ R - A0) <= 10 * n * H.EPS * np.linalg.norm(A0)


def test_qr_rank_deficient_and_zero_columns(svdj):
    """A zero column and a repeated column: tau = 0 reflectors, no NaN, Q still orthogonal, Q R = A."""
    m, n = 200, 96
    A0 = H.normal_matrix(m, n, seed=11)
    A0[:, 5] = 0.0
    A0[:, 40] = A0[:, 3]
    dY, dT = _qr_device(svdj, A0)
    Y = dY.cpu().numpy().T
    assert np.isfinite(Y).all()
    R = np.triu(Y[:n, :n])
    dC = torch.zeros(n, m, dtype=torch.float64, device="cuda")
    dC[torch.arange(n), torch.arange(n)] = 1.0
    assert svdj.lib.svdj_qr_apply_q_device(m, n, dY.data_ptr(), m, dT.data_ptr(), dC.data_ptr(), m, n, None) == 0
    Q = dC.cpu().numpy().T
    assert np.abs(Q.T @ Q - np.eye(n)).max() <= 10 * n * H.EPS
    assert np.linalg.norm(Q @ R - A0) <= 10 * n * H.EPS * np.linalg.norm(A0)


def _solve(svdj, A0, **kw):
    m, n = A0.shape
    A = A0.copy(order="F")
    s = np.zeros(min(m, n))
    V = np.zeros((n, n), order="F")
    st = svdj.dgesvd(svdj.AllVec, svdj.AllVec, m, n, A, m, s, V, n, svdj.default_opts(**kw))
    return A, s, V, st


@pytest.mark.parametrize("steps", [1, 2])
@pytest.mark.parametrize("m,n", [(512, 512), (700, 333), (1030, 257)])
def test_preconditioned_solve_matches_oracle(svdj, m, n, steps):
    A0 = H.uniform_matrix(m, n, seed=m + 5 * n)
    Uo, so, Vo, ost, orc = H.oracle_svd(A0)
    A, s, V, st = _solve(svdj, A0, precondition=steps)
    assert st.converged == 1 and st.precondition == steps
    assert H.sigma_error_normwise(s, so) <= 1e-12
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.all(np.diff(s) <= 0)
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    Ug, Vg = H.fix_signs(A, V)
    Uo2, Vo2 = H.fix_signs(Uo, Vo)
    gap = np.minimum(np.abs(np.diff(so, prepend=np.inf)), np.abs(np.diff(so, append=0.0))) / so[0]
    well = gap > 1e-3
    assert np.abs(Vg[:, well] - Vo2[:, well]).max() <= 1e-8
    assert np.abs(Ug[:, :n][:, well] - Uo2[:, :n][:, well]).max() <= 1e-8


@pytest.mark.parametrize("jobu,jobv", [(0, 2), (2, 0), (2, 2)])
def test_preconditioned_jobs(svdj, jobu, jobv):
    m, n = 400, 256
    A0 = H.normal_matrix(m, n, seed=77)
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.zeros((n, n), order="F")
    st = svdj.dgesvd(jobu, jobv, m, n, A, m, s, V if jobv != 2 else None, n, svdj.default_opts(precondition=1))
    assert st.converged == 1
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    if jobu != 2:
        assert np.abs(A.T @ A - np.eye(n)).max() <= 10 * n * H.EPS
    if jobv != 2:
        assert np.abs(V.T @ V - np.eye(n)).max() <= 10 * n * H.EPS


@pytest.mark.parametrize("cluster", [False, True])
def test_graded_spectrum_needs_few_sweeps_when_preconditioned(svdj, cluster):
    """cond 1e12, Haar U and V (BASELINE.json config 5's construction at n = 512): plain Jacobi needs ~40 sweeps,
    one QR step ~13, two ~9 (the oracle run on the same R^T agrees)."""
    n = 512
    A0, sigma = H.prescribed_spectrum(n, 1e12, seed=5, cluster=cluster)
    sl = np.linalg.svd(A0, compute_uv=False)
    sweeps = {}
    for steps in (0, 1, 2):
        A, s, V, st = _solve(svdj, A0, precondition=steps)
        assert st.converged == 1
        sweeps[steps] = st.sweeps
        assert H.sigma_error_normwise(s, sl) <= 1e-12
        assert H.sigma_error_normwise(s, sigma) <= 1e-12
        q = H.svd_quality(A0, A, s, V)
        bound = 10 * n * H.EPS
        assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, (steps, q)
    assert sweeps[1] <= 17 and sweeps[2] <= 14 and sweeps[0] >= 2 * sweeps[2], sweeps


def test_default_options_fall_back_to_the_qr_front_end(svdj):
    """opts.precondition < 0 (the default): a graded input on which plain Jacobi has not converged after 26 sweeps
    is solved again, preconditioned -- the caller gets an SVD, not status 1."""
    n = 768
    A0, sigma = H.prescribed_spectrum(n, 1e13, seed=9)
    A, s, V, st = _solve(svdj, A0)
    assert st.converged == 1
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert H.sigma_error_normwise(s, sigma) <= 1e-12
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    if st.precondition == 2:                     # the second attempt ran
        assert st.sweeps > 26


def _haar_graded_device(n, cond, seed, cluster):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    Q1, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device="cuda", generator=g))
    Q2, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device="cuda", generator=g))
    sigma = torch.logspace(0, -np.log10(cond), n, dtype=torch.float64, device="cuda")
    if cluster:
        sigma[n // 4: n // 4 + n // 8] = sigma[n // 4]
    A = ((Q1 * sigma) @ Q2.T).T.contiguous()       # (n, m): the column-major bytes of the m x n matrix
    return A, sigma


@pytest.mark.parametrize("cluster", [False, True])
def test_config5_full_size(svdj, cluster):
    """BASELINE.json config 5 at its full size: 8192 x 8192, prescribed spectrum with cond 1e12, two QR steps: the
    geometric spectrum converges in <= 15 sweeps (11 measured), the one with a cluster of 1024 EQUAL singular values
    in <= 20 (17 measured: inside a cluster every rotation has a large angle, convergence there is linear), both
    with the stated accuracy."""
    n = m = 8192
    A, sigma = _haar_graded_device(n, 1e12, 1234, cluster)
    A0 = A.clone()
    ds = torch.zeros(n, dtype=torch.float64, device="cuda")
    dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
    st = svdj.dgesvd_device(svdj.AllVec, svdj.AllVec, m, n, A.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n,
                            svdj.default_opts(precondition=2), allow_not_converged=True)
    assert st.converged == 1 and st.sweeps <= (20 if cluster else 15), (st.sweeps, st.last_offdiag)
    ref = torch.sort(sigma, descending=True).values
    assert float((ds - ref).abs().max() / ref[0]) <= 1e-12
    U, V = A.T, dV.T
    eye = torch.eye(n, dtype=torch.float64, device="cuda")
    bound = 10 * n * H.EPS
    assert float((U.T @ U - eye).abs().max()) <= bound
    assert float((V.T @ V - eye).abs().max()) <= bound
    assert float(torch.linalg.norm(A0.T - (U * ds) @ V.T) / torch.linalg.norm(A0)) <= bound


def test_config5_full_size_with_default_options(svdj):
    """The same 8192 x 8192 cond-1e12 input with DEFAULT options (precondition < 0): 26 plain sweeps do not converge,
    the library starts again from its copy of A with two QR steps, and the caller gets a converged SVD with the stated
    accuracy (round 1 returned status 1 and |U^T U - I| = 0.13 here)."""
    n = m = 8192
    A, sigma = _haar_graded_device(n, 1e12, 1234, False)
    A0 = A.clone()
    ds = torch.zeros(n, dtype=torch.float64, device="cuda")
    dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
    st = svdj.dgesvd_device(svdj.AllVec, svdj.AllVec, m, n, A.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n,
                            svdj.default_opts(), allow_not_converged=True)
    assert st.converged == 1 and st.precondition == 2 and 26 < st.sweeps <= 26 + 15, (st.sweeps, st.precondition)
    assert float((ds - sigma).abs().max() / sigma[0]) <= 1e-12
    U, V = A.T, dV.T
    eye = torch.eye(n, dtype=torch.float64, device="cuda")
    bound = 10 * n * H.EPS
    assert float((U.T @ U - eye).abs().max()) <= bound
    assert float((V.T @ V - eye).abs().max()) <= bound
    assert float(torch.linalg.norm(A0.T - (U * ds) @ V.T) / torch.linalg.norm(A0)) <= bound
