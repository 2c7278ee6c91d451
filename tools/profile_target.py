"""One short solve for ncu: python tools/profile_target.py [n] [max_sweeps] [precondition]"""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
sweeps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
pre = int(sys.argv[3]) if len(sys.argv) > 3 else 0
A = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(n, n, seed=n).T)).cuda()
ds = torch.zeros(n, dtype=torch.float64, device="cuda"); dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
st = svdj.dgesvd_device(0, 0, n, n, A.data_ptr(), n, ds.data_ptr(), dV.data_ptr(), n,
                        svdj.default_opts(max_sweeps=sweeps, precondition=pre), allow_not_converged=True)
print("sweeps", st.sweeps, "seconds", st.seconds_sweeps)
