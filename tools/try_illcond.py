import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
dev = torch.device("cuda")
for n, cond, variant in [(512, 1e12, 2), (1024, 1e12, 2), (2048, 1e12, 2), (2048, 1e6, 2), (1024, 1e12, 1)]:
    g = torch.Generator(device=dev); g.manual_seed(n)
    Q1, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
    Q2, _ = torch.linalg.qr(torch.randn(n, n, dtype=torch.float64, device=dev, generator=g))
    sig = torch.logspace(0, -np.log10(cond), n, dtype=torch.float64, device=dev)
    A = ((Q1 * sig) @ Q2.T).T.contiguous(); A0 = A.clone()
    ds = torch.zeros(n, dtype=torch.float64, device=dev); dV = torch.empty(n, n, dtype=torch.float64, device=dev)
    os.environ["SVDJ_TRACE"] = "1" if n == 1024 and variant == 2 else ""
    if not os.environ["SVDJ_TRACE"]: del os.environ["SVDJ_TRACE"]
    st = svdj.dgesvd_device(0, 0, n, n, A.data_ptr(), n, ds.data_ptr(), dV.data_ptr(), n, svdj.default_opts(variant=variant, max_sweeps=40), allow_not_converged=True)
    U, V = A.T, dV.T; eye = torch.eye(n, dtype=torch.float64, device=dev)
    print(f"n={n} cond={cond:g} variant={variant}: rc={st.rc} sweeps={st.sweeps} off={st.last_offdiag:.2e} sig_err={float((ds-sig).abs().max()):.2e} "
          f"orthU={float((U.T@U-eye).abs().max()):.2e} orthV={float((V.T@V-eye).abs().max()):.2e} t={st.seconds_sweeps:.2f}s")
