"""GPU parity tests of the level-1 path, all through the C ABI (ctypes -> libsvdj.so).

Tolerances (BASELINE.json north_star): sorted sigma to <= 1e-12 (normwise: max|ds_i| / s_max, the only
meaningful reading once cond*eps approaches 1e-12), U and V up to column sign, ||U^T U - I||_max,
||V^T V - I||_max and ||A - U S V^T||_F / ||A||_F <= 10 * max(m,n) * eps with eps = 2^-52.
Single steps are compared with the oracle's fma-rotating mode to 64*m*2^-53 (relative to the
column scale): the device sums the Gram triple in a different order than the host loop.
"""
import glob
import os

import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev(a):
    """numpy (F order) -> torch cuda tensor holding the same column-major bytes."""
    t = torch.from_numpy(np.ascontiguousarray(a.T)).cuda()  # (n, m) row-major == (m, n) column-major
    return t


def _host(t, m, n):
    return np.asfortranarray(t.cpu().numpy().T[:m, :n])


STEP_CASES = [(64, 64), (96, 32), (130, 10), (65, 9), (512, 128), (1000, 37), (4096, 64), (8192, 16), (300, 2)]


@pytest.mark.parametrize("m,n", STEP_CASES)
@pytest.mark.parametrize("path", ["reg", "split"])
def test_single_steps_match_oracle(svdj, m, n, path):
    A0 = H.uniform_matrix(m, n, seed=m + n)
    V0 = np.asfortranarray(np.linalg.qr(H.normal_matrix(n, n, seed=n))[0])
    A, V = A0.copy(order="F"), V0.copy(order="F")
    dA, dV = _dev(A0), _dev(V0)
    ctr = torch.zeros(2, dtype=torch.int64, device="cuda")
    steps = svdj.num_steps(n)
    applied = 0
    tol = m * 2.0 ** -53
    for st in sorted(s_ for s_ in set([0, 1, steps // 2, steps - 1]) if 0 <= s_ < steps):
        ost = H.oracle_one_step(A, V, st, mode=H.ORACLE_REL, threshold=tol, fma=1)
        applied += ost.applied
        svdj.step_device(m, n, dA.data_ptr(), m, dV.data_ptr(), n, True, st, tol,
                         path=svdj.PATH_REG if path == "reg" else svdj.PATH_SPLIT, d_counters=ctr.data_ptr())
    torch.cuda.synchronize()
    gA, gV = _host(dA, m, n), _host(dV, n, n)
    # the Gram triple is summed in a different order on the device, so (c,s) differ by ~m*eps relative
    tol_cmp = 64 * m * 2.0 ** -53
    assert np.abs(gA - A).max() <= tol_cmp * np.abs(A0).max() * np.sqrt(m)
    assert np.abs(gV - V).max() <= tol_cmp
    assert int(ctr[0].item()) == applied
    off = np.array([ctr[1].item()], dtype=np.int64).view(np.float64)[0]
    assert 0.0 <= off <= 1.0 + 1e-12


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(H.GOLDEN, "*.npz"))), ids=os.path.basename)
def test_reference_mode_matches_reference_golden(svdj, path):
    """One (or k) sweeps with the reference's own stopping rule against fixtures made by the reference."""
    g = np.load(path)
    A0, sweeps = g["A0"], int(g["sweeps"])
    n = A0.shape[0]
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.zeros((n, n), order="F")
    st = svdj.omp_mpi_cuda_dgesvd_local_matrices(svdj.AllVec, svdj.AllVec, n, n, A, n, s, V, n,
                                                 reference_mode=True, max_sweeps=sweeps) if sweeps == 1 else \
        svdj.dgesvd(svdj.AllVec, svdj.AllVec, n, n, A, n, s, V, n,
                    svdj.default_opts(max_sweeps=sweeps, threshold_mode=svdj.THRESH_ABS, tol=1e-16, sort_desc=0),
                    allow_not_converged=True)
    assert st.sweeps == sweeps
    if sweeps <= 3:
        # same rotation sequence, different rounding: element-wise agreement
        assert np.abs(s - g["s"]).max() <= 1e-12 * g["s"].max()
        assert np.abs(A - g["U"]).max() <= 1e-10
        assert np.abs(V - g["V"]).max() <= 1e-10
    else:
        assert H.sigma_error_normwise(s, g["s"]) <= 1e-12
        q = H.svd_quality(A0, A, s, V)
        assert q["orth_u"] <= 10 * n * H.EPS and q["resid"] <= 10 * n * H.EPS


@pytest.mark.parametrize("m,n", [(64, 64), (200, 120), (257, 64), (512, 512), (129, 33)])
def test_solve_matches_oracle_and_lapack(svdj, m, n):
    A0 = H.uniform_matrix(m, n, seed=11)
    Uo, so, Vo, ost, orc = H.oracle_svd(A0)
    A = A0.copy(order="F")
    s = np.zeros(min(m, n))
    V = np.zeros((n, n), order="F")
    st = svdj.cuda_dgesvd_kernel(svdj.AllVec, svdj.AllVec, m, n, A, m, s, V, n, variant=svdj.VARIANT_LEVEL1)
    assert st.converged == 1
    assert abs(st.sweeps - ost.sweeps) <= 1, (st.sweeps, ost.sweeps)
    assert H.sigma_error_normwise(s, so) <= 1e-12
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.all(np.diff(s) <= 0)
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    # U and V up to column sign where sigma is well separated
    Ug, Vg = H.fix_signs(A, V)
    Uo2, Vo2 = H.fix_signs(Uo, Vo)
    gap = np.minimum(np.abs(np.diff(so, prepend=np.inf)), np.abs(np.diff(so, append=0.0))) / so[0]
    well = gap > 1e-3
    assert np.abs(Vg[:, well] - Vo2[:, well]).max() <= 1e-8
    assert np.abs(Ug[:, : len(so)][:, well] - Uo2[:, : len(so)][:, well]).max() <= 1e-8


def test_ill_conditioned_prescribed_spectrum(svdj):
    n = 256
    A0, sigma = H.prescribed_spectrum(n, 1e12, seed=5, cluster=True)
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.zeros((n, n), order="F")
    st = svdj.cuda_dgesvd_kernel(0, 0, n, n, A, n, s, V, n, variant=svdj.VARIANT_LEVEL1)
    assert st.converged == 1
    assert H.sigma_error_normwise(s, np.sort(sigma)[::-1]) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


def test_edge_cases(svdj):
    # zero matrix
    A = np.zeros((8, 6), order="F")
    s = np.ones(6)
    V = np.ones((6, 6), order="F")  # V need not be pre-zeroed
    st = svdj.cuda_dgesvd_kernel(0, 0, 8, 6, A, 8, s, V, 6)
    assert st.converged == 1 and np.all(s == 0) and np.array_equal(V, np.eye(6)) and np.all(A == 0)
    # rank deficient, duplicate columns
    B = H.uniform_matrix(20, 3, seed=1)
    A0 = np.asfortranarray(np.hstack([B, B[:, :2]]))
    A = A0.copy(order="F")
    s = np.zeros(5)
    V = np.zeros((5, 5), order="F")
    svdj.cuda_dgesvd_kernel(0, 0, 20, 5, A, 20, s, V, 5)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.linalg.norm(A0 - (A * s) @ V.T) <= 1e-13 * np.linalg.norm(A0)
    # single column, single row
    A = np.asfortranarray(np.array([[3.0], [4.0]]))
    s = np.zeros(1)
    V = np.zeros((1, 1), order="F")
    svdj.cuda_dgesvd_kernel(0, 0, 2, 1, A, 2, s, V, 1)
    assert abs(s[0] - 5.0) < 1e-15 and abs(V[0, 0]) == 1.0
    # leading dimensions larger than the matrix; padding must come back untouched
    m, n, lda, ldv = 30, 12, 37, 19
    A0 = H.uniform_matrix(m, n, seed=9)
    Abuf = np.full((lda, n), -7.0, order="F")
    Abuf[:m] = A0
    Vbuf = np.full((ldv, n), -7.0, order="F")
    s = np.zeros(n)
    svdj.dgesvd(0, 0, m, n, Abuf, lda, s, Vbuf, ldv)
    assert np.all(Abuf[m:] == -7.0) and np.all(Vbuf[n:] == -7.0)
    q = H.svd_quality(A0, Abuf[:m], s, Vbuf[:n])
    assert max(q.values()) <= 10 * n * H.EPS


def test_jobs_and_sort_options(svdj):
    m, n = 96, 40
    A0 = H.uniform_matrix(m, n, seed=4)
    ref = np.linalg.svd(A0, compute_uv=False)
    # NoVec / NoVec: A holds A*V (unnormalised), V untouched
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.full((n, n), 5.0, order="F")
    svdj.dgesvd(svdj.NoVec, svdj.NoVec, m, n, A, m, s, V, n)
    assert np.all(V == 5.0)
    assert H.sigma_error_normwise(s, ref) <= 1e-12
    assert np.allclose(np.linalg.norm(A, axis=0), s, rtol=1e-13)
    # unsorted (the reference's raw order) has the same multiset of sigma
    A = A0.copy(order="F")
    s2 = np.zeros(n)
    V = np.zeros((n, n), order="F")
    svdj.dgesvd(0, 0, m, n, A, m, s2, V, n, svdj.default_opts(sort_desc=0))
    assert H.sigma_error_normwise(s2, ref) <= 1e-12
    assert np.linalg.norm(A0 - (A * s2) @ V.T) <= 10 * n * H.EPS * np.linalg.norm(A0)
    # max_sweeps too small -> NOT_CONVERGED status, outputs still a valid factorisation of A
    A = A0.copy(order="F")
    with pytest.raises(svdj.SvdjError) as e:
        svdj.dgesvd(0, 0, m, n, A, m, s2, V, n, svdj.default_opts(max_sweeps=2))
    assert e.value.code == svdj.NOT_CONVERGED


def test_device_entry_point(svdj):
    m, n = 384, 256
    A0 = H.uniform_matrix(m, n, seed=21)
    dA = _dev(A0)
    ds = torch.zeros(n, dtype=torch.float64, device="cuda")
    dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        st = svdj.dgesvd_device(0, 0, m, n, dA.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n,
                                stream=stream.cuda_stream)
    torch.cuda.synchronize()
    assert st.converged == 1 and st.launches > 0 and st.seconds_sweeps > 0
    U, s, V = _host(dA, m, n), ds.cpu().numpy(), _host(dV, n, n)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    assert max(q.values()) <= 10 * n * H.EPS


def test_tall_skinny_uses_split_path(svdj):
    m, n = 20000, 24  # m > 8192 rows: pairs are split over CTAs through the scratch buffer
    A0 = H.normal_matrix(m, n, seed=2)
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.zeros((n, n), order="F")
    st = svdj.cuda_dgesvd_kernel(0, 0, m, n, A, m, s, V, n, variant=svdj.VARIANT_LEVEL1)
    assert st.converged == 1
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    # the stopping test is relative with tol = m * 2^-53, so U's bound scales with m (10 * max(m,n) * eps)
    assert q["orth_u"] <= 10 * n * H.EPS and q["orth_v"] <= 10 * n * H.EPS and q["resid"] <= 10 * n * H.EPS


def test_full_size_4096_properties(svdj):
    """BASELINE.json configs[1] at full size: size-independent properties + LAPACK singular values."""
    n = 4096
    A0 = H.uniform_matrix(n, n, seed=4096)
    A = A0.copy(order="F")
    s = np.zeros(n)
    V = np.zeros((n, n), order="F")
    st = svdj.cuda_dgesvd_kernel(0, 0, n, n, A, n, s, V, n, variant=svdj.VARIANT_LEVEL1)
    assert st.converged == 1 and st.sweeps <= 25
    ref = np.linalg.svd(A0, compute_uv=False)  # host LAPACK; cuSOLVER's svdvals are themselves 8.6e-13 off here
    assert H.sigma_error_normwise(s, ref) <= 1e-12  # measured 5.4e-13: ~78 000 rotations per column on this path
    dA0, dU, dV = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (A0, A, V))
    ds = torch.from_numpy(s).cuda()
    eye = torch.eye(n, dtype=torch.float64, device="cuda")
    bound = 10 * n * H.EPS
    assert float((dU.T @ dU - eye).abs().max()) <= bound
    assert float((dV.T @ dV - eye).abs().max()) <= bound
    assert float(torch.linalg.norm(dA0 - (dU * ds) @ dV.T) / torch.linalg.norm(dA0)) <= bound
