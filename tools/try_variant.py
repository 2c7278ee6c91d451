"""Quick look at a solver variant: accuracy against LAPACK and device time.  python tools/try_variant.py --variant 2 --sizes 256x256,4096x4096"""
import argparse, importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
ap = argparse.ArgumentParser(); ap.add_argument("--variant", type=int, default=2); ap.add_argument("--sizes", default="64x64,200x120,512x512,1024x1024")
ap.add_argument("--reps", type=int, default=1)
a = ap.parse_args()
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
for sz in a.sizes.split(","):
    m, n = map(int, sz.split("x"))
    A0 = H.uniform_matrix(m, n, seed=n)
    dA0 = torch.from_numpy(np.ascontiguousarray(A0.T)).cuda()
    for rep in range(a.reps):
        dA = dA0.clone(); ds = torch.zeros(min(m, n), dtype=torch.float64, device="cuda"); dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
        st = svdj.dgesvd_device(0, 0, m, n, dA.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n, svdj.default_opts(variant=a.variant), allow_not_converged=True)
    U, V, S = dA.T, dV.T, ds
    ref = torch.linalg.svdvals(dA0.T)
    eye = torch.eye(n, dtype=torch.float64, device="cuda")
    print(f"{m}x{n} variant={st.variant} rc={st.rc} sweeps={st.sweeps} conv={st.converged} applied={st.applied} t_sweeps={st.seconds_sweeps*1e3:.1f} ms fin={st.seconds_finalize*1e3:.1f} ms total={st.seconds_total*1e3:.1f} ms "
          f"launches={st.launches} probe_us={st.step_kernel_ms*1e3:.1f} sig={float((S-ref).abs().max()/ref[0]):.2e} orthU={float((U.T@U-eye).abs().max()):.2e} "
          f"orthV={float((V.T@V-eye).abs().max()):.2e} resid={float(torch.linalg.norm(dA0.T-(U*S)@V.T)/torch.linalg.norm(dA0)):.2e} off={st.last_offdiag:.2e}")
