"""Who is off at 4096^2: compares the singular values of libsvdj, LAPACK on the host (numpy, dgesdd), torch on
the GPU (cuSOLVER) and an extended-precision check of sigma_1 (power iteration in float128 is too slow; instead
the Rayleigh quotient of our u_1, v_1 evaluated with compensated summation)."""
import importlib, os, sys, math
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
A0 = H.uniform_matrix(n, n, seed=n)
A = A0.copy(order="F"); s = np.zeros(n); V = np.zeros((n, n), order="F")
svdj.dgesvd(0, 0, n, n, A, n, s, V, n)
s_np = np.linalg.svd(A0, compute_uv=False)
s_t = torch.linalg.svdvals(torch.from_numpy(A0).cuda()).cpu().numpy()
def nw(a, b): return float(np.abs(a - b).max() / b[0])
print("ours vs numpy(host LAPACK) %.2e | ours vs torch(cuSOLVER) %.2e | numpy vs torch %.2e" % (nw(s, s_np), nw(s, s_t), nw(s_np, s_t)))
i = int(np.abs(s - s_np).argmax()); print("largest |ours - numpy| at index", i, "sigma", s_np[i], "diff", s[i] - s_np[i])
# sigma_1 from our singular vectors with long-double accumulation: u1^T A v1
u1 = A[:, 0].astype(np.longdouble); v1 = V[:, 0].astype(np.longdouble)
Av = (A0.astype(np.longdouble) @ v1); ray = float(u1 @ Av) / float(np.sqrt(u1 @ u1) * np.sqrt(v1 @ v1))
print("sigma_1: ours %.15e numpy %.15e torch %.15e longdouble-rayleigh %.15e" % (s[0], s_np[0], s_t[0], ray))
