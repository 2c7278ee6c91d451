"""ctypes binding of libsvdj.so, the B200-native one-sided Jacobi SVD.

The product is the CUDA library behind the C ABI in ``include/svdj.h``; this module only loads it
and mirrors the reference's solver interface (``/root/reference/lib/JacobiMethods.cuh:44-62``) for
the Python-side tests and ``bench.py``.  There is no CPU fallback and no import of ``oracle/``:
if the shared library has not been built the import fails loudly.

The directory name contains hyphens, so import it with::

    import importlib
    svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsvdj.so")

# lib/JacobiMethods.cuh:25-29
AllVec, SomeVec, NoVec = 0, 1, 2
VARIANT_AUTO, VARIANT_LEVEL1, VARIANT_BLOCK = 0, 1, 2
THRESH_REL, THRESH_ABS = 0, 1
PATH_AUTO, PATH_REG, PATH_SPLIT = 0, 1, 2
PHASE_ALL, PHASE_CROSS = 0, 1
OK, NOT_CONVERGED, ERR_CUDA, ERR_ALLOC, ERR_UNSUPPORTED = 0, 1, 2, 3, 4


class SvdjError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libsvdj status {code}: {msg}")
        self.code = code


# svdj_allreduce_fn (svdj.h): sum `count` doubles at device pointer `dptr` over all ranks, in place, on `stream`
AllreduceFn = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)


# svdj_allgather_fn (svdj.h): gather `bytes` bytes of host memory from every rank into recv, in rank order
AllgatherFn = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)


class MgComm(C.Structure):
    _fields_ = [("rank", C.c_int), ("size", C.c_int), ("allgather", AllgatherFn), ("ctx", C.c_void_p)]


class Opts(C.Structure):
    _fields_ = [
        ("tol", C.c_double),
        ("max_sweeps", C.c_int),
        ("sort_desc", C.c_int),
        ("variant", C.c_int),
        ("threshold_mode", C.c_int),
        ("block", C.c_int),
        ("device", C.c_int),
        ("precondition", C.c_int),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("sweeps", C.c_int),
        ("converged", C.c_int),
        ("variant", C.c_int),
        ("reserved", C.c_int),
        ("tested", C.c_ulonglong),
        ("applied", C.c_ulonglong),
        ("launches", C.c_ulonglong),
        ("last_offdiag", C.c_double),
        ("seconds_sweeps", C.c_double),
        ("seconds_finalize", C.c_double),
        ("seconds_h2d", C.c_double),
        ("seconds_d2h", C.c_double),
        ("seconds_total", C.c_double),
        ("step_kernel_ms", C.c_double),
        ("step_kernel_launches", C.c_ulonglong),
        ("probe_block_pairs", C.c_ulonglong),
        ("seconds_precond", C.c_double),
        ("precondition", C.c_int),
        ("reserved1", C.c_int),
    ]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


# every symbol include/svdj.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = (
    "svdj_default_opts",
    "svdj_dgesvd",
    "svdj_dgesvd_device",
    "svdj_step_device",
    "svdj_step_scratch_doubles",
    "svdj_step_device_noalloc",
    "svdj_ring_ws_create",
    "svdj_ring_ws_destroy",
    "svdj_ring_ws_set_floor",
    "svdj_qr_tfactor_doubles",
    "svdj_qr_factor_device",
    "svdj_qr_apply_q_device",
    "svdj_ring_phase_device",
    "svdj_ring_finalize_device",
    "svdj_ring_phase_send_device",
    "svdj_ipc_alloc",
    "svdj_ipc_free",
    "svdj_ipc_export",
    "svdj_ipc_open",
    "svdj_ipc_close",
    "svdj_dgesvd_mg",
    "svdj_mg_rank_create",
    "svdj_mg_rank_destroy",
    "svdj_mg_rank_dgesvd",
    "svdj_mg_rank_dgesvd_device",
    "svdj_mg_plan",
    "svdj_device_count",
    "svdj_num_steps",
    "svdj_schedule_step",
    "svdj_block_schedule_step",
    "svdj_dgesvd_rows_device",
    "svdj_last_error",
    "svdj_release_workspace",
    "svdj_version",
)


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)"
        )
    lib = C.CDLL(LIB_PATH)
    vp, sz, dbl, i = C.c_void_p, C.c_size_t, C.c_double, C.c_int
    lib.svdj_default_opts.argtypes = [C.POINTER(Opts)]
    lib.svdj_default_opts.restype = None
    lib.svdj_dgesvd.argtypes = [i, i, sz, sz, vp, sz, vp, vp, sz, C.POINTER(Opts), C.POINTER(Stats)]
    lib.svdj_dgesvd.restype = i
    lib.svdj_dgesvd_device.argtypes = [i, i, sz, sz, vp, sz, vp, vp, sz, C.POINTER(Opts), C.POINTER(Stats), vp]
    lib.svdj_dgesvd_device.restype = i
    lib.svdj_step_device.argtypes = [sz, sz, vp, sz, vp, sz, i, sz, dbl, i, i, vp, vp, vp]
    lib.svdj_step_device.restype = i
    lib.svdj_step_device_noalloc.argtypes = [sz, sz, vp, sz, vp, sz, i, sz, dbl, i, vp, vp]
    lib.svdj_step_device_noalloc.restype = i
    lib.svdj_ring_ws_create.argtypes = [sz, sz, C.POINTER(C.c_void_p)]
    lib.svdj_ring_ws_create.restype = i
    lib.svdj_ring_ws_destroy.argtypes = [vp]
    lib.svdj_ring_ws_destroy.restype = i
    lib.svdj_ring_ws_set_floor.argtypes = [vp, dbl]
    lib.svdj_ring_ws_set_floor.restype = i
    lib.svdj_qr_tfactor_doubles.argtypes = [sz]
    lib.svdj_qr_tfactor_doubles.restype = sz
    lib.svdj_qr_factor_device.argtypes = [sz, sz, vp, sz, vp, vp]
    lib.svdj_qr_factor_device.restype = i
    lib.svdj_qr_apply_q_device.argtypes = [sz, sz, vp, sz, vp, vp, sz, sz, vp]
    lib.svdj_qr_apply_q_device.restype = i
    lib.svdj_ring_phase_device.argtypes = [vp, sz, sz, sz, vp, vp, sz, vp, vp, sz, i, i, dbl, i, vp, vp]
    lib.svdj_ring_phase_device.restype = i
    lib.svdj_ring_finalize_device.argtypes = [vp, sz, sz, vp, sz, vp, i, vp]
    lib.svdj_ring_finalize_device.restype = i
    lib.svdj_ring_phase_send_device.argtypes = [vp, sz, sz, sz, vp, vp, sz, vp, vp, sz, i, dbl, i, vp, vp, vp, vp, vp, vp]
    lib.svdj_ring_phase_send_device.restype = i
    lib.svdj_ipc_alloc.argtypes = [sz, C.POINTER(C.c_void_p)]
    lib.svdj_ipc_alloc.restype = i
    lib.svdj_ipc_free.argtypes = [vp]
    lib.svdj_ipc_free.restype = i
    lib.svdj_ipc_export.argtypes = [vp, C.c_char_p]
    lib.svdj_ipc_export.restype = i
    lib.svdj_ipc_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.svdj_ipc_open.restype = i
    lib.svdj_ipc_close.argtypes = [vp]
    lib.svdj_ipc_close.restype = i
    lib.svdj_step_scratch_doubles.argtypes = [sz, sz]
    lib.svdj_step_scratch_doubles.restype = sz
    lib.svdj_num_steps.argtypes = [sz]
    lib.svdj_num_steps.restype = sz
    lib.svdj_schedule_step.argtypes = [sz, sz, C.POINTER(C.c_size_t)]
    lib.svdj_schedule_step.restype = sz
    lib.svdj_block_schedule_step.argtypes = [sz, C.c_int, sz, C.POINTER(C.c_size_t)]
    lib.svdj_block_schedule_step.restype = sz
    lib.svdj_dgesvd_rows_device.argtypes = [C.c_int, C.c_int, sz, sz, C.c_void_p, sz, C.c_void_p, C.c_void_p, sz, sz, sz,
                                            C.POINTER(Opts), C.POINTER(Stats), C.c_void_p, AllreduceFn, C.c_void_p]
    lib.svdj_dgesvd_rows_device.restype = C.c_int
    ip = C.POINTER(C.c_int)
    lib.svdj_dgesvd_mg.argtypes = [i, i, sz, sz, vp, sz, vp, vp, sz, C.POINTER(Opts), C.POINTER(Stats), i, ip]
    lib.svdj_dgesvd_mg.restype = i
    lib.svdj_mg_rank_create.argtypes = [C.POINTER(MgComm), sz, sz, C.POINTER(C.c_void_p)]
    lib.svdj_mg_rank_create.restype = i
    lib.svdj_mg_rank_destroy.argtypes = [vp]
    lib.svdj_mg_rank_destroy.restype = i
    lib.svdj_mg_rank_dgesvd.argtypes = [vp, i, i, sz, sz, vp, sz, vp, vp, sz, C.POINTER(Opts), C.POINTER(Stats)]
    lib.svdj_mg_rank_dgesvd.restype = i
    lib.svdj_mg_rank_dgesvd_device.argtypes = [vp, i, i, sz, sz, vp, sz, vp, vp, sz, C.POINTER(Opts), C.POINTER(Stats)]
    lib.svdj_mg_rank_dgesvd_device.restype = i
    lib.svdj_mg_plan.argtypes = [i, i, ip, ip, ip, ip, ip, ip]
    lib.svdj_mg_plan.restype = i
    lib.svdj_device_count.argtypes = []
    lib.svdj_device_count.restype = i
    lib.svdj_release_workspace.argtypes = []
    lib.svdj_release_workspace.restype = i
    lib.svdj_last_error.argtypes = []
    lib.svdj_last_error.restype = C.c_char_p
    lib.svdj_version.argtypes = []
    lib.svdj_version.restype = i
    return lib


lib = _load()


def default_opts(**kw) -> Opts:
    o = Opts()
    lib.svdj_default_opts(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError(f"unknown option {k!r}")
        setattr(o, k, v)
    return o


def last_error() -> str:
    return (lib.svdj_last_error() or b"").decode()


def _check(rc: int, allow_not_converged: bool = False) -> int:
    if rc == OK or (allow_not_converged and rc == NOT_CONVERGED):
        return rc
    if rc < 0:
        raise SvdjError(rc, f"argument {-rc} is invalid ({last_error()})")
    if rc == NOT_CONVERGED:
        raise SvdjError(rc, "not converged within max_sweeps")
    raise SvdjError(rc, last_error())


def num_steps(n: int) -> int:
    return int(lib.svdj_num_steps(n))


def schedule_step(n: int, step: int) -> np.ndarray:
    """Pairs (p, q), 0-based, of one round-robin step as an (n//2, 2) array."""
    buf = (C.c_size_t * (2 * max(n // 2, 1)))()
    cnt = lib.svdj_schedule_step(n, step, buf)
    return np.array(buf[: 2 * cnt], dtype=np.int64).reshape(-1, 2)


def block_schedule_step(nb: int, vranks: int, step: int) -> np.ndarray:
    """Block pairs (p, q), 0-based, of one step of the block path's ordering over nb column blocks as an
    (nb//2, 2) array; vranks >= 2 selects the two-level tournament (svdj.h: svdj_block_schedule_step)."""
    buf = (C.c_size_t * (2 * max(nb // 2, 1)))()
    cnt = lib.svdj_block_schedule_step(nb, vranks, step, buf)
    return np.array(buf[: 2 * cnt], dtype=np.int64).reshape(-1, 2)


def _host_ptr(a: np.ndarray, name: str) -> int:
    if a.dtype != np.float64:
        raise TypeError(f"{name} must be float64")
    if not a.flags["F_CONTIGUOUS"] and a.ndim == 2:
        raise ValueError(f"{name} must be column-major (Fortran order), as in the reference (lib/global.cuh:14)")
    if not a.flags["WRITEABLE"]:
        raise ValueError(f"{name} must be writable")
    return a.ctypes.data


def dgesvd(jobu, jobv, m, n, A, lda, s, V, ldv, opts: Opts | None = None, allow_not_converged=False) -> Stats:
    """Host-pointer solve with the reference's argument list (lib/JacobiMethods.cuh:44-62).

    A (m x n, column-major, overwritten with U), s (min(m,n)), V (n x n, column-major) are numpy arrays.
    """
    st = Stats()
    rc = lib.svdj_dgesvd(
        jobu, jobv, m, n, _host_ptr(A, "A"), lda, _host_ptr(s, "s"),
        None if V is None else _host_ptr(V, "V"), ldv,
        None if opts is None else C.byref(opts), C.byref(st),
    )
    _check(rc, allow_not_converged)
    st.rc = rc
    return st


# The reference's two entry points, same names and argument order.  In the reference they differ in
# MPI use and in the rotation threshold (1e-16 vs 1e-20, absolute); here both run the same device-
# resident solver to convergence.  `reference_mode=True` reproduces the reference's own stopping rule:
# exactly one sweep, absolute threshold, no sorting.
def cuda_dgesvd_kernel(jobu, jobv, m, n, A, lda, s, V, ldv, reference_mode=False, **optkw) -> Stats:
    if reference_mode:
        optkw = {**dict(max_sweeps=1, threshold_mode=THRESH_ABS, tol=1e-20, sort_desc=0), **optkw}
    return dgesvd(jobu, jobv, m, n, A, lda, s, V, ldv, default_opts(**optkw), allow_not_converged=reference_mode)


def omp_mpi_cuda_dgesvd_local_matrices(jobu, jobv, m, n, A, lda, s, V, ldv, reference_mode=False, **optkw) -> Stats:
    if reference_mode:
        optkw = {**dict(max_sweeps=1, threshold_mode=THRESH_ABS, tol=1e-16, sort_desc=0), **optkw}
    return dgesvd(jobu, jobv, m, n, A, lda, s, V, ldv, default_opts(**optkw), allow_not_converged=reference_mode)


def dgesvd_device(jobu, jobv, m, n, dA: int, lda, ds: int, dV: int, ldv, opts: Opts | None = None,
                  stream: int = 0, allow_not_converged=False) -> Stats:
    """Device-pointer solve; dA, ds, dV are raw device addresses (e.g. torch ``data_ptr()``)."""
    st = Stats()
    rc = lib.svdj_dgesvd_device(
        jobu, jobv, m, n, dA, lda, ds, dV, ldv, None if opts is None else C.byref(opts), C.byref(st),
        stream or None,
    )
    _check(rc, allow_not_converged)
    st.rc = rc
    return st


def step_device(m, n, dA: int, lda, dV: int, ldv, rotate_v, step, tol, threshold_mode=THRESH_REL,
                path=PATH_AUTO, d_scratch: int = 0, d_counters: int = 0, stream: int = 0) -> None:
    rc = lib.svdj_step_device(m, n, dA, lda, dV or None, ldv, int(rotate_v), step, tol, threshold_mode, path,
                              d_scratch or None, d_counters or None, stream or None)
    _check(rc)


def step_scratch_doubles(m: int, n: int) -> int:
    return int(lib.svdj_step_scratch_doubles(m, n))


NOALLOC_MAX_SCRATCH_DOUBLES = 98304  # SVDJ_NOALLOC_MAX_SCRATCH_DOUBLES in include/svdj.h


def step_device_noalloc(m, n, dA: int, lda, dV: int, ldv, rotate_v, step, tol, threshold_mode=THRESH_REL,
                        d_counters: int = 0, stream: int = 0) -> None:
    """Split-path step whose scratch is the array built into libsvdj.so: for callers that cannot
    allocate device memory (see include/svdj.h).  Calls that may overlap on one GPU race on that array."""
    rc = lib.svdj_step_device_noalloc(m, n, dA, lda, dV or None, ldv, int(rotate_v), step, tol, threshold_mode,
                                      d_counters or None, stream or None)
    _check(rc)


def dgesvd_mg(jobu, jobv, m, n, A, lda, s, V, ldv, ngpus: int, devices=None, opts: Opts | None = None,
              allow_not_converged=False) -> Stats:
    """The reference's distributed entry point as one call on `ngpus` GPUs of this process (svdj_dgesvd_mg): host
    arrays in and out like :func:`dgesvd`.  `devices`: CUDA ordinals of the ranks (default 0 .. ngpus-1)."""
    st = Stats()
    dev = None
    if devices is not None:
        if len(devices) != ngpus:
            raise ValueError("one device ordinal per rank")
        dev = (C.c_int * ngpus)(*devices)
    rc = lib.svdj_dgesvd_mg(jobu, jobv, m, n, _host_ptr(A, "A"), lda, _host_ptr(s, "s"),
                            None if V is None else _host_ptr(V, "V"), ldv, None if opts is None else C.byref(opts),
                            C.byref(st), ngpus, dev)
    _check(rc, allow_not_converged)
    st.rc = rc
    return st


def mg_plan(P: int, rnd: int):
    """(top_block, bottom_block, top_buf, bottom_buf, transfers) of the multi-GPU tournament after `rnd` moves."""
    arr = [(C.c_int * P)() for _ in range(4)]
    tr = (C.c_int * (8 * P))()
    cnt = C.c_int()
    rc = lib.svdj_mg_plan(P, rnd, arr[0], arr[1], arr[2], arr[3], tr, C.byref(cnt))
    _check(rc)
    return (list(arr[0]), list(arr[1]), list(arr[2]), list(arr[3]),
            [tuple(tr[4 * k: 4 * k + 4]) for k in range(cnt.value)])
