"""CPU tests of the oracle: pinned bit-for-bit to the reference's own solver (golden fixtures made
from it, and the freshly built oracle/_ref when present), then checked against LAPACK and the
accuracy bounds of BASELINE.json."""
import glob
import os

import numpy as np
import pytest

import helpers as H


@pytest.mark.parametrize("n", [2, 3, 8, 9, 33, 64, 101])
def test_schedule_is_a_complete_round_robin(n):
    steps = H.oracle_lib().oracle_num_steps(n)
    assert steps == (n - 1 if n % 2 == 0 else n)
    seen = set()
    for st in range(steps):
        pq = H.oracle_schedule(n, st)
        assert len(pq) == n // 2
        cols = pq.ravel()
        assert len(set(cols.tolist())) == len(cols), "pairs of a step must be disjoint"
        assert cols.min() >= 0 and cols.max() < n
        for p, q in pq:
            assert p != q
            key = (min(p, q), max(p, q))
            assert key not in seen
            seen.add(key)
    assert len(seen) == n * (n - 1) // 2, "every unordered pair exactly once per sweep"


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(H.GOLDEN, "*.npz"))), ids=os.path.basename)
def test_oracle_reproduces_reference_golden_bit_for_bit(path):
    g = np.load(path)
    A0, sweeps = g["A0"], int(g["sweeps"])
    # the reference's own stopping rule: fixed sweep count, absolute 1e-16 threshold, no sort
    U, s, V, st, rc = H.oracle_svd(A0, mode=H.ORACLE_ABS, threshold=1e-16, max_sweeps=sweeps, sort_desc=0, fma=0)
    assert rc == 0 and st.sweeps == sweeps
    assert np.array_equal(U, g["U"])
    assert np.array_equal(s, g["s"])
    assert np.array_equal(V, g["V"])


@pytest.mark.skipif(H.ref_cpu_lib() is None, reason="oracle/_ref not built (no /root/reference on this machine)")
@pytest.mark.parametrize("n,sweeps", [(16, 1), (47, 2), (128, 1)])
def test_oracle_matches_live_reference_build(n, sweeps):
    A0 = H.uniform_matrix(n, n, seed=100 + n)
    Ur, sr, Vr = H.ref_cpu_svd(A0, sweeps=sweeps)
    U, s, V, st, rc = H.oracle_svd(A0, mode=H.ORACLE_ABS, threshold=1e-16, max_sweeps=sweeps, sort_desc=0, fma=0)
    assert np.array_equal(U, Ur) and np.array_equal(s, sr) and np.array_equal(V, Vr)


def test_reference_input_generator_matches_golden():
    g = np.load(os.path.join(H.GOLDEN, "refinput_n48_s1.npz"))
    A = H.reference_input(48)
    assert np.array_equal(A, g["A0"])
    assert np.all(np.tril(A, -1) == 0.0) and A[0, 0] > 0


@pytest.mark.parametrize("m,n", [(64, 64), (200, 120), (257, 64)])
def test_oracle_converged_matches_lapack_and_bounds(m, n):
    A0 = H.uniform_matrix(m, n, seed=7)
    U, s, V, st, rc = H.oracle_svd(A0)
    assert rc == 0 and st.converged == 1 and st.sweeps < 30
    ref = np.linalg.svd(A0, compute_uv=False)
    assert H.sigma_error_normwise(s, ref) <= 1e-12
    assert np.all(np.diff(s) <= 0), "sorted descending"
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound


def test_oracle_one_sweep_is_not_an_svd_but_keeps_the_product():
    # what the reference ships: one sweep leaves U far from orthogonal, yet A = U S V^T holds
    A0 = H.uniform_matrix(64, 64, seed=3)
    U, s, V, st, rc = H.oracle_svd(A0, mode=H.ORACLE_ABS, threshold=1e-16, max_sweeps=1, sort_desc=0)
    q = H.svd_quality(A0, U, s, V)
    assert q["orth_u"] > 1e-2
    assert q["resid"] < 1e-13
    fro = H.oracle_lib().oracle_residual_fro(64, 64, A0, U, s, V)
    assert fro < 1e-12


def test_oracle_ill_conditioned_prescribed_spectrum():
    A0, sigma = H.prescribed_spectrum(96, 1e12, seed=5)
    U, s, V, st, rc = H.oracle_svd(A0)
    assert rc == 0
    assert H.sigma_error_normwise(s, sigma) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * 96 * H.EPS
    assert q["orth_v"] <= bound and q["resid"] <= bound and q["orth_u"] <= bound


def test_oracle_zero_and_rank_deficient():
    A0 = np.zeros((8, 6), order="F")
    U, s, V, st, rc = H.oracle_svd(A0)
    assert rc == 0 and np.all(s == 0) and np.array_equal(V, np.eye(6)) and np.all(U == 0)
    B = H.uniform_matrix(20, 3, seed=1)
    A1 = np.asfortranarray(np.hstack([B, B[:, :2]]))  # rank 3, 5 columns
    U, s, V, st, rc = H.oracle_svd(A1)
    assert rc == 0
    ref = np.linalg.svd(A1, compute_uv=False)
    assert H.sigma_error_normwise(s, ref) <= 1e-12
    assert s[3] <= 1e-13 * s[0] and s[4] <= 1e-13 * s[0]


def test_oracle_rejects_bad_arguments():
    A = np.zeros((4, 4), order="F")
    s = np.zeros(4)
    V = np.zeros((4, 4), order="F")
    import ctypes as C
    o = H.OracleOpts(H.ORACLE_REL, 1e-15, 5, 1, 1, 0)
    lib = H.oracle_lib()
    assert lib.oracle_dgesvd(7, 0, 4, 4, A, 4, s, V, 4, C.byref(o), None) == -1
    assert lib.oracle_dgesvd(0, 0, 4, 4, A, 3, s, V, 4, C.byref(o), None) == -6
    assert lib.oracle_dgesvd(0, 0, 4, 4, A, 4, s, V, 2, C.byref(o), None) == -9
