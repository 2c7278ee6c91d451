"""Test-side bindings: the oracle (oracle/liboracle.so), the reference built for checking
(oracle/_ref/*.so), input generators and the accuracy measures of BASELINE.json's north_star.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this file; the product
(svd-jacobi-mpi-cuda_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_CPU_SO = os.path.join(ROOT, "oracle", "_ref", "libref_cpu.so")
REF_GPU_SO = os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")

EPS = 2.0 ** -52  # machine epsilon of the accuracy bounds (10*n*eps)

_dp = np.ctypeslib.ndpointer(dtype=np.float64)


class OracleOpts(C.Structure):
    _fields_ = [("threshold_mode", C.c_int), ("threshold", C.c_double), ("max_sweeps", C.c_int),
                ("sort_desc", C.c_int), ("nthreads", C.c_int), ("fma_rotation", C.c_int)]


class OracleStats(C.Structure):
    _fields_ = [("sweeps", C.c_int), ("converged", C.c_int), ("tested", C.c_ulonglong),
                ("applied", C.c_ulonglong), ("last_offdiag", C.c_double), ("seconds_sweeps", C.c_double)]


ORACLE_ABS, ORACLE_REL = 0, 1
_oracle = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(ORACLE_SO)
        lib.oracle_num_steps.argtypes = [C.c_size_t]
        lib.oracle_num_steps.restype = C.c_size_t
        lib.oracle_schedule_step.argtypes = [C.c_size_t, C.c_size_t, C.POINTER(C.c_size_t)]
        lib.oracle_schedule_step.restype = C.c_size_t
        lib.oracle_dgesvd.argtypes = [C.c_int, C.c_int, C.c_size_t, C.c_size_t, _dp, C.c_size_t, _dp, _dp,
                                      C.c_size_t, C.POINTER(OracleOpts), C.POINTER(OracleStats)]
        lib.oracle_dgesvd.restype = C.c_int
        lib.oracle_step.argtypes = [C.c_size_t, C.c_size_t, _dp, C.c_size_t, _dp, C.c_size_t, C.c_int,
                                    C.c_size_t, C.POINTER(OracleOpts), C.POINTER(OracleStats)]
        lib.oracle_step.restype = None
        lib.oracle_fill_reference_input.argtypes = [C.c_size_t, C.c_uint, _dp]
        lib.oracle_fill_reference_input.restype = None
        lib.oracle_residual_fro.argtypes = [C.c_size_t, C.c_size_t, _dp, _dp, _dp, _dp]
        lib.oracle_residual_fro.restype = C.c_double
        _oracle = lib
    return _oracle


def oracle_schedule(n: int, step: int) -> np.ndarray:
    buf = (C.c_size_t * (2 * max(n // 2, 1)))()
    cnt = oracle_lib().oracle_schedule_step(n, step, buf)
    return np.array(buf[: 2 * cnt], dtype=np.int64).reshape(-1, 2)


def oracle_svd(A0: np.ndarray, mode=ORACLE_REL, threshold=None, max_sweeps=60, sort_desc=1, nthreads=0,
               fma=0, jobu=0, jobv=0):
    """Runs the oracle on a copy of A0; returns (U, s, V, stats, rc)."""
    m, n = A0.shape
    if threshold is None:
        threshold = m ** 0.5 * 2.0 ** -53 if mode == ORACLE_REL else 1e-16  # dgesvj CTOL with vectors
    A = np.array(A0, dtype=np.float64, order="F", copy=True)
    s = np.zeros(min(m, n))
    V = np.zeros((n, n), order="F")
    o = OracleOpts(mode, threshold, max_sweeps, sort_desc, nthreads, fma)
    st = OracleStats()
    rc = oracle_lib().oracle_dgesvd(jobu, jobv, m, n, A, m, s, V, n, C.byref(o), C.byref(st))
    return A, s, V, st, rc


def oracle_one_step(A: np.ndarray, V: np.ndarray, step: int, mode=ORACLE_REL, threshold=0.0, fma=1, nthreads=0):
    """In place: one round-robin step on F-ordered A (m x n) and V (n x n)."""
    m, n = A.shape
    o = OracleOpts(mode, threshold, 1, 0, nthreads, fma)
    st = OracleStats()
    oracle_lib().oracle_step(m, n, A, m, V, n, 1, step, C.byref(o), C.byref(st))
    return st


def reference_input(n: int, seed: int = 1000000) -> np.ndarray:
    A = np.zeros((n, n), order="F")
    oracle_lib().oracle_fill_reference_input(n, seed, A)
    return A


_ref_cpu = None


def ref_cpu_lib():
    """The reference's own solver built for host cores, or None when oracle/_ref is absent."""
    global _ref_cpu
    if _ref_cpu is None and os.path.exists(REF_CPU_SO):
        lib = C.CDLL(REF_CPU_SO)
        lib.ref_mpi_dgesvd.argtypes = [C.c_ulong, C.c_ulong, _dp, _dp, _dp, C.c_int]
        lib.ref_mpi_dgesvd.restype = C.c_int
        lib.ref_set_max_iterations.argtypes = [C.c_ulong]
        _ref_cpu = lib
    return _ref_cpu


def ref_cpu_svd(A0: np.ndarray, sweeps: int = 1, nthreads: int = 4):
    lib = ref_cpu_lib()
    m, n = A0.shape
    A = np.array(A0, dtype=np.float64, order="F", copy=True)
    s = np.zeros(min(m, n))
    V = np.zeros((n, n), order="F")
    lib.ref_set_max_iterations(sweeps)
    lib.ref_mpi_dgesvd(m, n, A, s, V, nthreads)
    lib.ref_set_max_iterations(1)
    return A, s, V


# ------------------------------------------------------------------------------------------------
# synthetic inputs.  Counter-based: element (i,j) depends only on (seed,i,j).
# ------------------------------------------------------------------------------------------------
def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def uniform_matrix(m: int, n: int, seed: int) -> np.ndarray:
    """Dense uniform(0,1) m x n matrix, column-major."""
    with np.errstate(over="ignore"):
        idx = np.arange(m * n, dtype=np.uint64) + np.uint64(seed) * np.uint64(0x100000001B3)
        bits = _splitmix64(idx)
    u = (bits >> np.uint64(11)).astype(np.float64) * 2.0 ** -53
    return np.asfortranarray(u.reshape((n, m)).T)


def normal_matrix(m: int, n: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return np.asfortranarray(rng.standard_normal((n, m)).T)


def prescribed_spectrum(n: int, cond: float, seed: int, cluster: bool = False):
    """A = U diag(sigma) V^T with Haar U, V and a geometric spectrum from 1 down to 1/cond."""
    rng = np.random.default_rng(seed)
    U, _ = np.linalg.qr(rng.standard_normal((n, n)))
    V, _ = np.linalg.qr(rng.standard_normal((n, n)))
    sigma = np.logspace(0.0, -np.log10(cond), n)
    if cluster:
        sigma[n // 4: n // 4 + max(n // 8, 2)] = sigma[n // 4]
    A = (U * sigma) @ V.T
    return np.asfortranarray(A), sigma


# ------------------------------------------------------------------------------------------------
# accuracy measures (BASELINE.json north_star)
# ------------------------------------------------------------------------------------------------
def svd_quality(A0, U, s, V):
    m, n = A0.shape
    k = min(m, n)
    I = np.eye(k)
    return {
        "orth_u": float(np.abs(U[:, :k].T @ U[:, :k] - I).max()),
        "orth_v": float(np.abs(V.T @ V - np.eye(n)).max()),
        "resid": float(np.linalg.norm(A0 - (U[:, :k] * s[:k]) @ V[:, :k].T) / np.linalg.norm(A0)),
    }


def sigma_error_normwise(s, s_ref):
    """max_i |s_i - s_ref_i| / s_ref_max on descending-sorted values."""
    a = np.sort(np.asarray(s))[::-1]
    b = np.sort(np.asarray(s_ref))[::-1]
    return float(np.abs(a - b).max() / b[0])


def fix_signs(U, V):
    """Column sign convention: largest-magnitude entry of each V column positive."""
    U = U.copy()
    V = V.copy()
    idx = np.abs(V).argmax(axis=0)
    sg = np.sign(V[idx, np.arange(V.shape[1])])
    sg[sg == 0] = 1.0
    V *= sg
    U[:, : V.shape[1]] *= sg[: U.shape[1]] if U.shape[1] <= V.shape[1] else np.pad(sg, (0, U.shape[1] - len(sg)), constant_values=1.0)
    return U, V


def uniform_columns(m: int, n: int, seed: int, c0: int, count: int) -> np.ndarray:
    """Columns [c0, c0+count) of uniform_matrix(m, n, seed) as a (count, m) array (row j = column c0+j),
    without building the whole matrix: every rank of a multi-GPU run generates its own columns."""
    with np.errstate(over="ignore"):
        idx = (np.arange(count * m, dtype=np.uint64) + np.uint64(c0 * m)) + np.uint64(seed) * np.uint64(0x100000001B3)
        bits = _splitmix64(idx)
    return ((bits >> np.uint64(11)).astype(np.float64) * 2.0 ** -53).reshape(count, m)
