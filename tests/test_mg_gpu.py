"""GPU tests of the in-library multi-GPU driver (svdj_dgesvd_mg: C++ ranks as threads, fused block exchange, flags
in peer memory).  On a one-GPU box the ranks share device 0 (devices=[0]*P), which runs the whole driver --
tournament, peer stores, flags, mailboxes, gather -- with the "peers" on the same GPU; with >= 2 GPUs the same tests
run across devices as well.  Checker: LAPACK singular values and BASELINE.json's bounds (sigma <= 1e-12 normwise;
|U^T U - I|, |V^T V - I|, |A - U S V^T| / |A| <= 10 n eps)."""
import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _device_lists(P):
    out = [pytest.param([0] * P, id=f"P{P}-shared-gpu")]
    if torch.cuda.is_available() and torch.cuda.device_count() >= P:
        out.append(pytest.param(list(range(P)), id=f"P{P}-{P}gpus"))
    return out


def _check(A0, U, s, V):
    m, n = A0.shape
    assert np.all(np.diff(s) <= 0)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


@pytest.mark.parametrize("devices", _device_lists(2) + _device_lists(4))
@pytest.mark.parametrize("m,n", [(512, 512), (700, 384), (300, 200)])
def test_mg_solve_matches_lapack(svdj, m, n, devices):
    A0 = H.uniform_matrix(m, n, seed=m + n)
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.zeros((n, n), order="F")
    st = svdj.dgesvd_mg(svdj.AllVec, svdj.AllVec, m, n, A, m, s, V, n, len(devices), devices)
    assert st.converged == 1 and st.launches > 0
    _check(A0, A, s, V)
    # and once more on the cached ranks (stamps keep counting, buffers are reused)
    A2 = A0.copy(order="F")
    s2 = np.zeros(n)
    V2 = np.zeros((n, n), order="F")
    st2 = svdj.dgesvd_mg(svdj.AllVec, svdj.AllVec, m, n, A2, m, s2, V2, n, len(devices), devices)
    assert st2.sweeps == st.sweeps and np.array_equal(s, s2) and np.array_equal(A, A2) and np.array_equal(V, V2)


@pytest.mark.parametrize("devices", _device_lists(2))
def test_mg_matches_single_gpu_solver(svdj, devices):
    m = n = 1024
    A0 = H.normal_matrix(m, n, seed=4)
    A1 = A0.copy(order="F"); s1 = np.zeros(n); V1 = np.zeros((n, n), order="F")
    svdj.dgesvd(svdj.AllVec, svdj.AllVec, m, n, A1, m, s1, V1, n)
    A2 = A0.copy(order="F"); s2 = np.zeros(n); V2 = np.zeros((n, n), order="F")
    svdj.dgesvd_mg(svdj.AllVec, svdj.AllVec, m, n, A2, m, s2, V2, n, len(devices), devices)
    assert H.sigma_error_normwise(s2, s1) <= 1e-13
    U1, W1 = H.fix_signs(A1, V1)
    U2, W2 = H.fix_signs(A2, V2)
    assert np.abs(W1 - W2).max() <= 1e-8 and np.abs(U1 - U2).max() <= 1e-8
    _check(A0, A2, s2, V2)


@pytest.mark.parametrize("devices", _device_lists(2))
def test_mg_rank_deficient_and_novec(svdj, devices):
    m, n = 256, 192
    rng = np.random.default_rng(3)
    A0 = np.asfortranarray(rng.standard_normal((m, 100)) @ rng.standard_normal((100, n)))   # rank 100
    A = A0.copy(order="F"); s = np.zeros(n); V = np.zeros((n, n), order="F")
    st = svdj.dgesvd_mg(svdj.AllVec, svdj.AllVec, m, n, A, m, s, V, n, len(devices), devices)
    assert st.converged == 1 and st.sweeps < 30          # the null columns must not keep the sweeps alive
    sl = np.linalg.svd(A0, compute_uv=False)
    assert H.sigma_error_normwise(s, sl) <= 1e-12
    assert np.abs(V.T @ V - np.eye(n)).max() <= 10 * n * H.EPS
    assert np.linalg.norm(A0 - (A * s) @ V.T) <= 10 * n * H.EPS * np.linalg.norm(A0)
    # singular values only
    A = A0.copy(order="F"); s2 = np.zeros(n)
    svdj.dgesvd_mg(svdj.NoVec, svdj.NoVec, m, n, A, m, s2, None, n, len(devices), devices)
    assert H.sigma_error_normwise(s2, sl) <= 1e-12


@pytest.mark.parametrize("steps", [1, 2])
@pytest.mark.parametrize("devices", _device_lists(2))
def test_mg_graded_spectrum_with_qr_front_end(svdj, devices, steps):
    """cond 1e12 (config 5's construction at n = 512) on the ring: rank 0 factors, the ring iterates on R^T."""
    n = 512
    A0, sigma = H.prescribed_spectrum(n, 1e12, seed=5)
    A = A0.copy(order="F"); s = np.zeros(n); V = np.zeros((n, n), order="F")
    st = svdj.dgesvd_mg(svdj.AllVec, svdj.AllVec, n, n, A, n, s, V, n, len(devices), devices,
                        opts=svdj.default_opts(precondition=steps))
    assert st.converged == 1 and st.precondition == steps and st.sweeps <= 18
    assert H.sigma_error_normwise(s, sigma) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


def test_mg_argument_errors(svdj):
    A = np.zeros((64, 64), order="F"); s = np.zeros(64); V = np.zeros((64, 64), order="F")
    with pytest.raises(svdj.SvdjError):
        svdj.dgesvd_mg(0, 0, 64, 64, A, 64, s, V, 64, 2, [0, 99])
    with pytest.raises(svdj.SvdjError):
        svdj.dgesvd_mg(0, 0, 64, 64, A, 64, s, V, 64, 0)


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_solves_on_device_0_then_device_1(svdj):
    """The opt-in shared-memory sizes of the kernels and the SM count are per device: a process that solves on a
    second GPU through svdj_opts.device must find them set there too (round 1 set them once per process)."""
    A0 = H.uniform_matrix(640, 384, seed=5)
    for dev in (0, 1, 0):
        for variant in (svdj.VARIANT_BLOCK, svdj.VARIANT_LEVEL1):
            A = A0.copy(order="F")
            s = np.zeros(384)
            V = np.zeros((384, 384), order="F")
            st = svdj.dgesvd(svdj.AllVec, svdj.AllVec, 640, 384, A, 640, s, V, 384,
                             svdj.default_opts(device=dev, variant=variant))
            assert st.converged == 1
            _check(A0, A, s, V)
