"""Generates tests/golden/*.npz from the REFERENCE's own solver.

Run in the build container only (it needs /root/reference through oracle/_ref/libref_cpu.so, which
oracle/Makefile compiles from /root/reference/main.cu where it lies):

    make -C oracle && python tests/golden/make_golden.py

Each fixture holds an input A0 and the outputs (U in place of A, s, V) of
omp_mpi_cuda_dgesvd_local_matrices (main.cu:440) called as one rank with AllVec/AllVec, after
`sweeps` repetitions of its sweep loop (the shipped value is 1, main.cu:482).  The reference's
rotation kernel body runs on the host (see oracle/stubs/cuda_host_shim.h), so these are the results
of the reference's arithmetic without fused multiply-add.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import helpers as H  # noqa: E402

CASES = [
    # name, kind, n, sweeps
    ("dense_n8_s1", "uniform", 8, 1),
    ("dense_n9_s1", "uniform", 9, 1),      # odd n: one column idles per step
    ("dense_n33_s3", "uniform", 33, 3),
    ("dense_n64_s1", "uniform", 64, 1),
    ("dense_n64_s12", "uniform", 64, 12),   # converged
    ("refinput_n48_s1", "reference", 48, 1),  # the driver's own seed-1000000 upper-triangular input
]


def main():
    assert H.ref_cpu_lib() is not None, "oracle/_ref/libref_cpu.so missing: run make -C oracle"
    for name, kind, n, sweeps in CASES:
        A0 = H.uniform_matrix(n, n, seed=n) if kind == "uniform" else H.reference_input(n)
        U, s, V = H.ref_cpu_svd(A0, sweeps=sweeps, nthreads=2)
        np.savez_compressed(os.path.join(H.GOLDEN, name + ".npz"), A0=A0, U=U, s=s, V=V, sweeps=sweeps)
        print(name, "written; orth defect", float(np.abs(U.T @ U - np.eye(n)).max()))


if __name__ == "__main__":
    main()
