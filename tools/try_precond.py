import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
pre = importlib.import_module("svd-jacobi-mpi-cuda_b200.precond")
dev = torch.device("cuda")
cases = [("graded", 1024, 1), ("graded", 1024, 2), ("graded", 8192, 1), ("graded", 8192, 2), ("uniform", 4096, 1), ("uniform", 4096, 2)]
for kind, n, steps in cases:
    g = torch.Generator(device=dev); g.manual_seed(n)
    if kind == "graded":
        Q1, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
        Q2, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
        sig = torch.logspace(0, -12, n, dtype=torch.float64, device=dev); A = (Q1 * sig) @ Q2.T; del Q1, Q2
    else:
        A = torch.rand(n, n, dtype=torch.float64, device=dev, generator=g); sig = torch.linalg.svdvals(A)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    U, s, V, st = pre.dgesvd_preconditioned(A, svdj.default_opts(max_sweeps=80), qr_steps=steps)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    eye = torch.eye(n, dtype=torch.float64, device=dev)
    print(f"{kind} n={n} qr_steps={steps}: total {dt:.2f}s jacobi {st.seconds_sweeps:.2f}s sweeps={st.sweeps} conv={st.converged} sig_err={float((s-sig).abs().max()/sig[0]):.2e} "
          f"orthU={float((U.T@U-eye).abs().max()):.2e} orthV={float((V.T@V-eye).abs().max()):.2e} resid={float(torch.linalg.norm(A-(U*s)@V.T)/torch.linalg.norm(A)):.2e}")
