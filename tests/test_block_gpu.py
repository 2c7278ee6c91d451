"""GPU parity tests of the block-Jacobi variant (FP64 DMMA tiles), through the C ABI.

The block path rotates the same column pairs as the level-1 path but 32 columns at a time, from Gram
entries, so its rotation sequence differs from the reference's: parity is on the converged result, with the
tolerances of BASELINE.json (stated in tests/test_l1_gpu.py): sigma <= 1e-12 normwise against the oracle and
LAPACK, U and V up to sign where sigma is separated, orthogonality and residual <= 10*max(m,n)*eps.
"""
import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _solve(svdj, A0, variant, **kw):
    m, n = A0.shape
    A = A0.copy(order="F")
    s = np.zeros(min(m, n))
    V = np.zeros((n, n), order="F")
    st = svdj.dgesvd(svdj.AllVec, svdj.AllVec, m, n, A, m, s, V, n, svdj.default_opts(variant=variant, **kw))
    return A, s, V, st


# 120 and 100 columns: ragged last block; 37 columns: odd number of blocks (one idles per step)
@pytest.mark.parametrize("m,n", [(64, 64), (200, 120), (130, 37), (257, 64), (512, 512), (1000, 100), (96, 33)])
def test_block_solve_matches_oracle_and_lapack(svdj, m, n):
    A0 = H.uniform_matrix(m, n, seed=m * 7 + n)
    Uo, so, Vo, ost, orc = H.oracle_svd(A0)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st.variant == svdj.VARIANT_BLOCK and st.converged == 1
    assert st.sweeps <= ost.sweeps + 1, (st.sweeps, ost.sweeps)
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
    assert np.abs(Ug[:, : len(so)][:, well] - Uo2[:, : len(so)][:, well]).max() <= 1e-8


def test_block_and_level1_agree(svdj):
    A0 = H.normal_matrix(300, 160, seed=8)
    A1, s1, V1, st1 = _solve(svdj, A0, svdj.VARIANT_LEVEL1)
    A2, s2, V2, st2 = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st1.variant == svdj.VARIANT_LEVEL1 and st2.variant == svdj.VARIANT_BLOCK
    assert H.sigma_error_normwise(s2, s1) <= 1e-12
    U1, W1 = H.fix_signs(A1, V1)
    U2, W2 = H.fix_signs(A2, V2)
    assert np.abs(W1 - W2).max() <= 1e-9 and np.abs(U1 - U2).max() <= 1e-9


def test_block_ill_conditioned(svdj):
    n = 256
    A0, sigma = H.prescribed_spectrum(n, 1e12, seed=5, cluster=True)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st.converged == 1
    assert H.sigma_error_normwise(s, np.sort(sigma)[::-1]) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


def test_block_edge_cases(svdj):
    # zero matrix: nothing rotates, V = I
    A, s, V, st = _solve(svdj, np.zeros((80, 64), order="F"), svdj.VARIANT_BLOCK)
    assert st.converged == 1 and np.all(s == 0) and np.array_equal(V, np.eye(64))
    # rank deficient: duplicate columns spread over different blocks
    B = H.uniform_matrix(90, 40, seed=1)
    A0 = np.asfortranarray(np.hstack([B, B[:, :30]]))
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.linalg.norm(A0 - (A * s) @ V.T) <= 10 * 90 * H.EPS * np.linalg.norm(A0)
    assert np.all(s[40:] <= 1e-12 * s[0])
    # fewer than two blocks: the block variant hands over to the level-1 path
    A0 = H.uniform_matrix(40, 12, seed=2)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st.variant == svdj.VARIANT_LEVEL1
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    # NoVec / NoVec
    A0 = H.uniform_matrix(128, 96, seed=3)
    A = A0.copy(order="F")
    s = np.zeros(96)
    V = np.full((96, 96), 5.0, order="F")
    svdj.dgesvd(svdj.NoVec, svdj.NoVec, 128, 96, A, 128, s, V, 96, svdj.default_opts(variant=svdj.VARIANT_BLOCK))
    assert np.all(V == 5.0) and H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.allclose(np.linalg.norm(A, axis=0), s, rtol=1e-13)


def test_auto_picks_block_for_wide_problems_and_level1_for_reference_mode(svdj):
    A0 = H.uniform_matrix(128, 128, seed=5)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_AUTO)
    assert st.variant == svdj.VARIANT_BLOCK
    A = A0.copy(order="F")
    st = svdj.cuda_dgesvd_kernel(0, 0, 128, 128, A, 128, np.zeros(128), np.zeros((128, 128), order="F"), 128,
                                 reference_mode=True)
    assert st.variant == svdj.VARIANT_LEVEL1 and st.sweeps == 1


def test_block_tall_skinny(svdj):
    m, n = 50000, 64
    A0 = H.normal_matrix(m, n, seed=2)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st.converged == 1
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    assert q["orth_u"] <= 10 * n * H.EPS and q["orth_v"] <= 10 * n * H.EPS and q["resid"] <= 10 * n * H.EPS


def test_block_full_size_4096_properties(svdj):
    n = 4096
    A0 = H.uniform_matrix(n, n, seed=4096)
    dA = torch.from_numpy(np.ascontiguousarray(A0.T)).cuda()
    dA0 = dA.clone()
    ds = torch.zeros(n, dtype=torch.float64, device="cuda")
    dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
    st = svdj.dgesvd_device(0, 0, n, n, dA.data_ptr(), n, ds.data_ptr(), dV.data_ptr(), n,
                            svdj.default_opts(variant=svdj.VARIANT_BLOCK))
    assert st.converged == 1 and st.sweeps <= 25
    # host LAPACK (numpy, dgesdd) is the yardstick: cuSOLVER's own singular values are 8.6e-13 off at this size
    # (tools/sigma_check.py), ours agree with LAPACK to ~2e-15
    ref = np.linalg.svd(A0, compute_uv=False)
    assert H.sigma_error_normwise(ds.cpu().numpy(), ref) <= 1e-13
    U, V = dA.T, dV.T
    eye = torch.eye(n, dtype=torch.float64, device="cuda")
    bound = 10 * n * H.EPS
    assert float((U.T @ U - eye).abs().max()) <= bound
    assert float((V.T @ V - eye).abs().max()) <= bound
    assert float(torch.linalg.norm(dA0.T - (U * ds) @ V.T) / torch.linalg.norm(dA0)) <= bound


def test_unstaged_kernels_give_the_same_answer(svdj, monkeypatch):
    """SVDJ_BLOCK_PLAIN=1 forces the fragment-from-global kernels (the fallback for unaligned layouts)."""
    A0 = H.uniform_matrix(300, 200, seed=31)
    A1, s1, V1, st1 = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    monkeypatch.setenv("SVDJ_BLOCK_PLAIN", "1")
    A2, s2, V2, st2 = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    monkeypatch.delenv("SVDJ_BLOCK_PLAIN")
    assert st1.converged == 1 and st2.converged == 1
    assert H.sigma_error_normwise(s2, s1) <= 1e-13
    q = H.svd_quality(A0, A2, s2, V2)
    assert max(q.values()) <= 10 * 300 * H.EPS
    # an odd leading dimension on the device entry point takes the same fallback by itself
    torch_ = torch
    m, n, lda = 129, 96, 131
    B0 = H.uniform_matrix(m, n, seed=32)
    buf = torch_.zeros(n, lda, dtype=torch_.float64, device="cuda")
    buf[:, :m] = torch_.from_numpy(np.ascontiguousarray(B0.T)).cuda()
    ds = torch_.zeros(n, dtype=torch_.float64, device="cuda")
    dV = torch_.zeros(n, n + 1, dtype=torch_.float64, device="cuda")
    st = svdj.dgesvd_device(0, 0, m, n, buf.data_ptr(), lda, ds.data_ptr(), dV.data_ptr(), n + 1,
                            svdj.default_opts(variant=svdj.VARIANT_BLOCK))
    assert st.converged == 1
    assert H.sigma_error_normwise(ds.cpu().numpy(), np.linalg.svd(B0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(B0, buf[:, :m].cpu().numpy().T, ds.cpu().numpy(), dV[:, :n].cpu().numpy().T)
    assert max(q.values()) <= 10 * n * H.EPS


@pytest.mark.parametrize("qr_steps", [1, 2])
def test_qr_preconditioned_front_end(svdj, qr_steps):
    """Optional front-end (library QR + our Jacobi on R^T): graded spectrum, cond 1e12."""
    import importlib
    pre = importlib.import_module("svd-jacobi-mpi-cuda_b200.precond")
    n = 512
    A0, sigma = H.prescribed_spectrum(n, 1e12, seed=9)
    A = torch.from_numpy(np.ascontiguousarray(A0)).cuda()
    U, s, V, st = pre.dgesvd_preconditioned(A, svdj.default_opts(max_sweeps=60), qr_steps=qr_steps)
    assert st.converged == 1 and st.sweeps <= 16, st.sweeps   # 47 sweeps without the QR step
    U, s, V = U.cpu().numpy(), s.cpu().numpy(), V.cpu().numpy()
    assert H.sigma_error_normwise(s, np.sort(sigma)[::-1]) <= 1e-12
    q = H.svd_quality(A0, U, s, V)
    assert max(q.values()) <= 10 * n * H.EPS, q


@pytest.mark.parametrize("variant", ["level1", "block"])
@pytest.mark.parametrize("m,n", [(64, 100), (40, 72), (96, 96 + 33)])
def test_more_columns_than_rows(svdj, variant, m, n):
    """m < n: n - m columns are numerically zero at convergence; they must not keep the sweeps alive.
    (The reference assumes a square matrix; this is the contract of include/svdj.h for the general case:
    min(m,n) singular values, U in the first min(m,n) columns of A, V n x n orthogonal.)"""
    A0 = H.normal_matrix(m, n, seed=m + n)
    v = svdj.VARIANT_LEVEL1 if variant == "level1" else svdj.VARIANT_BLOCK
    A = A0.copy(order="F")
    s = np.zeros(m)
    V = np.zeros((n, n), order="F")
    st = svdj.dgesvd(svdj.AllVec, svdj.AllVec, m, n, A, m, s, V, n, svdj.default_opts(variant=v))
    assert st.converged == 1 and st.sweeps <= 30  # rank-deficient problems converge more slowly (23 sweeps at 64x100)
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    U = A[:, :m]
    bound = 10 * n * H.EPS
    assert np.abs(U.T @ U - np.eye(m)).max() <= bound
    assert np.abs(V.T @ V - np.eye(n)).max() <= bound
    assert np.linalg.norm(A0 - (U * s) @ V[:, :m].T) <= bound * np.linalg.norm(A0)
    # the trailing columns of A are the (numerically zero) images of the null space
    assert np.abs(A[:, m:]).max() <= 1e-12 * s[0]


@pytest.mark.parametrize("m,n", [(128, 128), (200, 120), (300, 72), (512, 512), (1030, 260)])
def test_block_width_32_option(svdj, m, n):
    """opts.block = 32: the same step with 64-column block pairs (svdj_block32.cu)."""
    A0 = H.uniform_matrix(m, n, seed=3 * m + n)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK, block=32)
    assert st.converged == 1 and st.reserved == 32
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    assert np.all(np.diff(s) <= 0)
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    # same answer as the default width
    A2, s2, V2, st2 = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st2.reserved == 16 and H.sigma_error_normwise(s, s2) <= 1e-12


@pytest.mark.parametrize("vranks", ["0", "2", "4", "8"])
@pytest.mark.parametrize("block", ["16", "32"])
def test_block_orderings_reach_the_same_svd(svdj, monkeypatch, vranks, block):
    """Flat round-robin over the column blocks (the reference's ordering one level up) and the two-level
    tournament with 2, 4 or 8 virtual ranks: different rotation sequences, the same converged SVD."""
    m = n = 1024
    A0 = H.uniform_matrix(m, n, seed=77)
    monkeypatch.setenv("SVDJ_VRANKS", vranks)
    monkeypatch.setenv("SVDJ_BLOCK", block)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st.converged == 1 and st.reserved == int(block)
    assert st.sweeps <= 18, st.sweeps
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


@pytest.mark.parametrize("m,n", [(512, 512), (700, 384), (2048, 256)])
def test_gram_cache_changes_nothing_but_the_time(svdj, monkeypatch, m, n):
    """64-column kernels with and without the cache of the blocks' Gram matrices: same sweeps (+-1), same SVD."""
    A0 = H.normal_matrix(m, n, seed=5 * m + n)
    monkeypatch.setenv("SVDJ_BLOCK", "32")
    out = {}
    for cache in ("1", "0"):
        monkeypatch.setenv("SVDJ_GCACHE", cache)
        out[cache] = _solve(svdj, A0, svdj.VARIANT_BLOCK)
        assert out[cache][3].converged == 1 and out[cache][3].reserved == 32
    (A1, s1, V1, st1), (A2, s2, V2, st2) = out["1"], out["0"]
    assert abs(st1.sweeps - st2.sweeps) <= 1
    assert H.sigma_error_normwise(s1, s2) <= 1e-13
    assert H.sigma_error_normwise(s1, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    bound = 10 * n * H.EPS
    for A, s, V in ((A1, s1, V1), (A2, s2, V2)):
        q = H.svd_quality(A0, A, s, V)
        assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


def test_tall_input_uses_wide_blocks_and_converges(svdj):
    """m >= 8 n: 64-column blocks and three inner sweeps (the eigen-solve does not grow with m)."""
    m, n = 20000, 256
    A0 = H.normal_matrix(m, n, seed=3)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_AUTO)
    assert st.variant == svdj.VARIANT_BLOCK and st.reserved == 32 and st.converged == 1
    assert st.sweeps <= 10, st.sweeps
    assert H.sigma_error_normwise(s, np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    q = H.svd_quality(A0, A, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q


def test_quadratic_stop_saves_the_confirming_sweep_only(svdj, monkeypatch):
    """With and without the quadratic-convergence stop: at most one sweep apart, the same SVD, and the state the
    early stop leaves passes the reference's test (a further solve on the result rotates nothing)."""
    m = n = 768
    A0 = H.uniform_matrix(m, n, seed=21)
    monkeypatch.setenv("SVDJ_EARLY_STOP", "1")   # off by default: its bound assumes separated singular values
    A1, s1, V1, st1 = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    monkeypatch.setenv("SVDJ_EARLY_STOP", "0")
    A2, s2, V2, st2 = _solve(svdj, A0, svdj.VARIANT_BLOCK)
    assert st1.converged == 1 and st2.converged == 1
    assert st2.sweeps - st1.sweeps in (0, 1)
    assert st2.last_offdiag <= 2 * m ** 0.5 * 2.0 ** -53              # the last sweep saw nothing above 2 tol
    assert H.sigma_error_normwise(s1, s2) <= 1e-14
    q = H.svd_quality(A0, A1, s1, V1)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    # U * diag(s) from the early stop is already orthogonal to working accuracy: solving it again applies nothing
    B = np.asfortranarray(A1 * s1)
    s3 = np.zeros(n); V3 = np.zeros((n, n), order="F")
    st3 = svdj.dgesvd(svdj.AllVec, svdj.AllVec, m, n, B, m, s3, V3, n, svdj.default_opts(variant=svdj.VARIANT_BLOCK))
    assert st3.sweeps <= 3 and st3.applied <= 64   # at most a few pairs sitting at the threshold


@pytest.mark.parametrize("m,n", [(2880, 336), (512, 2160), (2048, 1160)])
def test_shapes_whose_padded_block_count_needs_more_scratch(svdj, m, n):
    """Shapes for which the partial-Gram buffer of round 1 was too small (the number of row splits is not monotonic in
    the number of block pairs, and the buffer was sized before the block count was padded).  The buffer is now sized
    from the block count the kernels run with and the launch verifies its capacity: these solve, and solve right."""
    A0 = H.uniform_matrix(m, n, seed=m + 3 * n)
    A, s, V, st = _solve(svdj, A0, svdj.VARIANT_AUTO)
    assert st.converged == 1
    k = min(m, n)
    assert H.sigma_error_normwise(s[:k], np.linalg.svd(A0, compute_uv=False)) <= 1e-12
    bound = 10 * n * H.EPS
    U = A[:, :k]
    assert np.abs(U.T @ U - np.eye(k)).max() <= bound
    assert np.abs(V.T @ V - np.eye(n)).max() <= bound
    assert np.linalg.norm(A0 - (U * s[:k]) @ V[:, :k].T) <= bound * np.linalg.norm(A0)
