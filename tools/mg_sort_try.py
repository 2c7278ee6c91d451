# Sweep counts of the multi-GPU driver with and without ordering the columns of a block pair by size
# (SVDJ_MG_SORT) and with one or two inner sweeps, ranks sharing device 0 (the counts do not depend on the device
# placement; the times printed here do and are not bench numbers).
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import helpers as H
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")

sizes = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "1024,4096").split(",")]
for n in sizes:
    A0 = H.uniform_matrix(n, n, seed=n)
    ref = np.linalg.svd(A0, compute_uv=False) if n <= 2048 else None
    for P in (2, 4, 8):
        for sort in (0, 1):
            for inner in (1, 2):
                os.environ["SVDJ_MG_SORT"] = str(sort)
                os.environ["SVDJ_INNER_SWEEPS"] = str(inner)
                A = A0.copy(order="F"); s = np.zeros(n); V = np.zeros((n, n), order="F")
                t0 = time.time()
                try:
                    st = svdj.dgesvd_mg(0, 0, n, n, A, n, s, V, n, P, [0] * P)
                    err = H.sigma_error_normwise(s, ref) if ref is not None else float("nan")
                    q = H.svd_quality(A0, A, s, V) if n <= 2048 else None
                    print(f"n={n} P={P} sort={sort} inner={inner} sweeps={st.sweeps} conv={st.converged} "
                          f"t={time.time()-t0:.2f} sigma_err={err:.1e} q={q}", flush=True)
                except Exception as e:
                    print(f"n={n} P={P} sort={sort} inner={inner} FAIL {e}", flush=True)
