"""Times the solve with svdj_opts.precondition = 0 / 1 / 2 on dense uniform inputs and on the graded-spectrum
construction of BASELINE.json config 5; one JSON record per run.

    python tools/precond_sweep.py [--sizes 2048 4096] [--graded 4096 8192]
"""
import argparse, importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", type=int, nargs="*", default=[2048, 4096])
ap.add_argument("--graded", type=int, nargs="*", default=[4096, 8192])
ap.add_argument("--steps", type=int, nargs="*", default=[0, 1, 2])
ap.add_argument("--cluster", action="store_true", help="graded inputs: n/8 equal singular values from index n/4")
a = ap.parse_args()
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
dev = torch.device("cuda")


def run(name, A0, sigma_ref, steps):
    n, m = A0.shape
    eye = torch.eye(n, dtype=torch.float64, device=dev)
    for rep in range(2):
        A = A0.clone()
        ds = torch.zeros(n, dtype=torch.float64, device=dev); dV = torch.empty(n, n, dtype=torch.float64, device=dev)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        st = svdj.dgesvd_device(0, 0, m, n, A.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n,
                                svdj.default_opts(precondition=steps), allow_not_converged=True)
        torch.cuda.synchronize(); wall = time.perf_counter() - t0
    U, V = A.T, dV.T
    rec = {"input": name, "precondition": steps, "seconds": wall, "seconds_sweeps": st.seconds_sweeps,
           "seconds_precond": st.seconds_precond, "sweeps": st.sweeps, "converged": st.converged,
           "orth_u": float((U.T @ U - eye).abs().max()), "orth_v": float((V.T @ V - eye).abs().max()),
           "resid": float(torch.linalg.norm(A0.T - (U * ds) @ V.T) / torch.linalg.norm(A0)),
           "sigma_err": float((ds - sigma_ref).abs().max() / sigma_ref[0]), "bound_10_n_eps": 10 * n * 2.0 ** -52}
    print(json.dumps(rec), flush=True)


for n in a.sizes:
    A0 = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(n, n, seed=n).T)).to(dev)
    ref = torch.linalg.svdvals(A0)
    for s in a.steps:
        run(f"{n}x{n} uniform(0,1)", A0, ref, s)
for n in a.graded:
    g = torch.Generator(device=dev); g.manual_seed(1234)
    Q1, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
    Q2, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
    sig = torch.logspace(0, -12, n, dtype=torch.float64, device=dev)
    if a.cluster:
        sig[n // 4: n // 4 + n // 8] = sig[n // 4]
    A0 = ((Q1 * sig) @ Q2.T).T.contiguous(); del Q1, Q2
    for s in a.steps:
        if s == 0 and n > 4096:
            continue
        run(f"{n}x{n} graded cond 1e12" + (" clustered" if a.cluster else ""), A0, torch.sort(sig, descending=True).values, s)
