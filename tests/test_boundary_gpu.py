"""The drop-in boundary as a maintainer of the reference would use it, on the GPU:

* oracle/_ref/shim_check is a translation unit with the reference's OWN MatrixMPI / Matrix / SVD_OPTIONS definitions
  (extracted from /root/reference/main.cu at build time by oracle/Makefile) + SVDJ_HAVE_REFERENCE_TYPES +
  include/svdj_reference_api.hpp, linked against libsvdj.so.  It calls omp_mpi_cuda_dgesvd_local_matrices and
  cuda_dgesvd_kernel (global and Thesis::, main.cu:163, 440; lib/JacobiMethods.cuh:44-62); the results are checked
  against the oracle with BASELINE.json's tolerances.
* svd-jacobi-mpi-cuda_b200/svd_jacobi <n> is the reference driver's workflow (main.cu:1426-1676): its stdout lines,
  its report file (name pattern and contents) and the residual it prints.
"""
import glob
import os
import re
import subprocess

import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

SHIM = os.path.join(H.ROOT, "oracle", "_ref", "shim_check")
CLI = os.path.join(H.ROOT, "svd-jacobi-mpi-cuda_b200", "svd_jacobi")


@pytest.mark.parametrize("entry", ["mpi", "kernel", "thesis_mpi", "thesis_kernel"])
def test_reference_types_and_entry_points(tmp_path, entry):
    if not os.path.exists(SHIM):
        pytest.skip("oracle/_ref/shim_check not built (needs /root/reference at build time)")
    m, n = 200, 96
    A0 = H.uniform_matrix(m, n, seed=31)
    fin, fout = tmp_path / "a.bin", tmp_path / "out.bin"
    A0.T.copy().tofile(fin)                                   # column-major bytes
    r = subprocess.run([SHIM, entry, str(m), str(n), str(fin), str(fout)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "status 0" in r.stdout                             # svdj_shim::last_status() == SVDJ_OK
    raw = np.fromfile(fout)
    U = raw[: m * n].reshape(n, m).T
    s = raw[m * n: m * n + n]
    V = raw[m * n + n:].reshape(n, n).T
    Uo, so, Vo, ost, orc = H.oracle_svd(A0)
    assert H.sigma_error_normwise(s, so) <= 1e-12
    assert np.all(np.diff(s) <= 0)
    q = H.svd_quality(A0, U, s, V)
    bound = 10 * n * H.EPS
    assert q["orth_u"] <= bound and q["orth_v"] <= bound and q["resid"] <= bound, q
    Ug, Vg = H.fix_signs(U, V)
    Uo2, Vo2 = H.fix_signs(Uo, Vo)
    gap = np.minimum(np.abs(np.diff(so, prepend=np.inf)), np.abs(np.diff(so, append=0.0))) / so[0]
    well = gap > 1e-3
    assert np.abs(Vg[:, well] - Vo2[:, well]).max() <= 1e-8 and np.abs(Ug[:, well] - Uo2[:, well]).max() <= 1e-8


def _run_cli(tmp_path, *extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([CLI, "512", *extra], cwd=tmp_path, capture_output=True, text=True, timeout=600, env=e)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


def _check_cli_output(tmp_path, out):
    lines = out.strip().splitlines()
    assert lines[0] == "Dimensions, height: 512, width: 512"                       # main.cu:1583
    mt = re.fullmatch(r"SVD MPI\+OMP time with U,V calculation: ([0-9.e+-]+)", lines[1])   # main.cu:1638
    assert mt and float(mt.group(1)) > 0
    mr = re.fullmatch(r"\|\|A-USVt\|\|_F: ([0-9.e+-]+)", lines[2])                 # main.cu:1665
    assert mr
    # the input is the reference's own upper-triangular seed-1000000 matrix; |A|_F ~ 209
    assert float(mr.group(1)) <= 10 * 512 * H.EPS * 210.0
    reports = glob.glob(os.path.join(tmp_path, "reporte-dimension-512-time-*.txt"))        # main.cu:1667
    assert len(reports) == 1
    assert re.fullmatch(r"reporte-dimension-512-time-\d\d-\d\d-\d{4}-\d\d-\d\d-\d\d\.txt", os.path.basename(reports[0]))
    body = open(reports[0]).read().splitlines()
    assert body[0].startswith("Number of threads: ")                                # main.cu:1581
    assert body[1] == lines[0] and body[2] == lines[1] and body[3] == lines[2]
    assert '"converged": 1' in lines[-1]


def test_driver_cli_prints_the_reference_lines_and_writes_the_report(tmp_path):
    if not os.path.exists(CLI):
        pytest.skip("svd_jacobi not built")
    _check_cli_output(tmp_path, _run_cli(tmp_path))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_driver_cli_on_two_gpus(tmp_path):
    out = _run_cli(tmp_path, "--gpus", "2")
    _check_cli_output(tmp_path, out)
    assert '"gpus": 2' in out.strip().splitlines()[-1]
