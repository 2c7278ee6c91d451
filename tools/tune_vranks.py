# One GPU: time-to-solution and sweeps for several numbers of virtual ranks of the two-level ordering (SVDJ_VRANKS),
# device-resident, CUDA events, 1 warm-up + 2 timed solves each.  usage: python tools/tune_vranks.py 4096:4,2,8,16 8192:4,8
import importlib, os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import torch
import helpers as H
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")

for spec in sys.argv[1:]:
    n_s, ps = spec.split(":")
    n = int(n_s)
    A0 = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(n, n, seed=n).T)).cuda()  # column-major bytes
    dA = torch.empty_like(A0)
    ds = torch.zeros(n, dtype=torch.float64, device="cuda")
    dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
    opts = svdj.default_opts()
    stream = torch.cuda.current_stream().cuda_stream
    for P in ps.split(","):
        os.environ["SVDJ_VRANKS"] = P
        times = []
        try:
            for it in range(3):
                dA.copy_(A0)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                st = svdj.dgesvd_device(svdj.AllVec, svdj.AllVec, n, n, dA.data_ptr(), n, ds.data_ptr(), dV.data_ptr(), n,
                                        opts, stream=stream)
                e1.record()
                torch.cuda.synchronize()
                if it > 0:
                    times.append(e0.elapsed_time(e1) * 1e-3)
            U = dA.T
            orth = float((U.T @ U - torch.eye(n, dtype=torch.float64, device="cuda")).abs().max())
            print(f"n={n} vranks={P} t={np.mean(times):.4f} sweeps={st.sweeps} conv={st.converged} orth_u={orth:.1e}", flush=True)
        except Exception as e:
            print(f"n={n} vranks={P} FAIL {e}", flush=True)
