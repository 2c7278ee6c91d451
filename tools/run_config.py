"""Runs one of BASELINE.json's configs once on the visible GPU and prints a JSON record with time and accuracy.

    python tools/run_config.py --config 2|3|4|5 [--variant 0|1|2] [--cusolver]

config 2: 4096^2 uniform; 3: 16384^2 uniform; 4: 1,048,576 x 512 normal; 5: 8192^2 prescribed spectrum cond 1e12.
Accuracy is computed on the GPU with torch (orthogonality, residual; sigma against the prescribed spectrum for
config 5, against torch.linalg.svdvals otherwise).  --cusolver adds cuSOLVER's own Jacobi (gesvdj through
torch.linalg.svd) as a yardstick on the same matrix.
"""
import argparse, importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=2)
ap.add_argument("--variant", type=int, default=0)
ap.add_argument("--cusolver", action="store_true")
ap.add_argument("--no-svdvals", action="store_true")
a = ap.parse_args()
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
dev = torch.device("cuda")
g = torch.Generator(device=dev); g.manual_seed(1234)
sigma_true = None
if a.config == 2:
    m = n = 4096; A = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(m, n, seed=n).T)).to(dev); name = "4096x4096 uniform(0,1)"
elif a.config == 3:
    m = n = 16384; A = torch.rand(n, m, dtype=torch.float64, device=dev, generator=g); name = "16384x16384 uniform(0,1)"
elif a.config == 4:
    m, n = 1048576, 512; A = torch.randn(n, m, dtype=torch.float64, device=dev, generator=g); name = "1048576x512 normal"
elif a.config == 5:
    m = n = 8192
    Q1, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
    Q2, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
    sigma_true = torch.logspace(0, -12, n, dtype=torch.float64, device=dev)
    A = ((Q1 * sigma_true) @ Q2.T).T.contiguous(); del Q1, Q2   # stored (n, m): column-major m x n
    name = "8192x8192 prescribed geometric spectrum, cond 1e12"
A0 = A.clone()
ds = torch.zeros(n, dtype=torch.float64, device=dev); dV = torch.empty(n, n, dtype=torch.float64, device=dev)
torch.cuda.synchronize(); t0 = time.perf_counter()
st = svdj.dgesvd_device(0, 0, m, n, A.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n, svdj.default_opts(variant=a.variant), allow_not_converged=True)
torch.cuda.synchronize(); wall = time.perf_counter() - t0
U, V = A.T, dV.T   # (m, n), (n, n)
eye = torch.eye(n, dtype=torch.float64, device=dev)
rec = {"config": a.config, "workload": name, "variant": st.variant, "rc": st.rc, "sweeps": st.sweeps, "converged": st.converged,
       "seconds": wall, "seconds_sweeps": st.seconds_sweeps, "launches": st.launches, "rotations": st.applied,
       "orth_u": float((U.T @ U - eye).abs().max()), "orth_v": float((V.T @ V - eye).abs().max()),
       "bound_10_max_mn_eps": 10 * max(m, n) * 2.0 ** -52, "last_offdiag": st.last_offdiag}
R = A0.T - (U * ds) @ V.T
rec["resid"] = float(torch.linalg.norm(R) / torch.linalg.norm(A0)); del R
if sigma_true is not None:
    rec["sigma_err_vs_prescribed"] = float((ds - sigma_true).abs().max() / sigma_true[0])
elif not a.no_svdvals:
    ref = torch.linalg.svdvals(A0.T); rec["sigma_err_vs_lapack"] = float((ds - ref).abs().max() / ref[0])
if a.cusolver:
    torch.cuda.synchronize(); t0 = time.perf_counter()
    Uc, Sc, Vc = torch.linalg.svd(A0.T, full_matrices=False, driver="gesvdj")
    torch.cuda.synchronize(); rec["cusolver_gesvdj_seconds"] = time.perf_counter() - t0
    rec["cusolver_gesvdj_orth_u"] = float((Uc.T @ Uc - eye).abs().max())
    rec["sigma_err_vs_cusolver"] = float((ds - Sc).abs().max() / Sc[0])
print(json.dumps(rec))
