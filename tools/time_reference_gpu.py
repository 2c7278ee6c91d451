"""Times the reference's GENUINE CUDA path (oracle/_ref/libref_gpu.so = /root/reference/main.cu built by nvcc as it
is: host Gram triples, 8 cudaMemcpy and 2 launches of its jacobi_rotation kernel per pair, main.cu:727-757) for ONE
sweep -- all the reference ever runs (main.cu:482) -- next to our solver TO CONVERGENCE on the same input, and prints
one JSON record per size.  BASELINE.md section 5 / SURVEY.md section 8(d) ask for this once per round.

    python tools/time_reference_gpu.py [--sizes 512 1024] [--threads T]
"""
import argparse, ctypes as C, importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", type=int, nargs="*", default=[512, 1024])
ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
a = ap.parse_args()
dp = np.ctypeslib.ndpointer(dtype=np.float64)
ref = C.CDLL(H.REF_GPU_SO)
ref.ref_mpi_dgesvd.argtypes = [C.c_ulong, C.c_ulong, dp, dp, dp, C.c_int]
ref.ref_set_max_iterations.argtypes = [C.c_ulong]
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
for n in a.sizes:
    A0 = H.uniform_matrix(n, n, seed=n)
    ref.ref_set_max_iterations(1)
    best = None
    for rep in range(2):
        A = A0.copy(order="F"); s = np.zeros(n); V = np.zeros((n, n), order="F")
        t0 = time.perf_counter()
        ref.ref_mpi_dgesvd(n, n, A, s, V, a.threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    rotated = float(np.abs(V - np.eye(n)).max())
    A2 = A0.copy(order="F"); s2 = np.zeros(n); V2 = np.zeros((n, n), order="F")
    svdj.dgesvd(0, 0, n, n, A2, n, s2, V2, n)                     # warm-up
    A2 = A0.copy(order="F")
    t0 = time.perf_counter()
    st = svdj.dgesvd(0, 0, n, n, A2, n, s2, V2, n)
    ours = time.perf_counter() - t0
    pairs = n * (n - 1) // 2
    print(json.dumps({"n": n, "reference_gpu_one_sweep_s": best, "reference_pairs_per_s": pairs / best,
                      "reference_rotated_anything": rotated > 1e-3, "host_threads": a.threads,
                      "svdj_to_convergence_s": ours, "svdj_sweeps": st.sweeps,
                      "note": "the reference stops after this one sweep (its output is not an SVD); ours is the whole solve "
                              "from host buffers"}), flush=True)
