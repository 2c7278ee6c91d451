"""Records how many sweeps the reference's algorithm (its ordering, relative test, tol = sqrt(m) 2^-53) needs on
bench.py's uniform(0,1) input, for the extrapolation in bench.py's CPU baseline and reference arm:

    python tools/record_sweeps.py --oracle 512 1024 --out profiles/sweeps_to_converge.json      (host, here)
    python tools/record_sweeps.py --level1 2048 4096 8192 --out gpurun_out/sweeps_level1.json   (GPU box)

--oracle runs oracle/liboracle.so to convergence; --level1 runs the CUDA level-1 path (SVDJ_VARIANT_LEVEL1: the same
pairs in the same order).  Existing entries of the output file are kept; every entry says where it came from."""
import argparse, importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H

ap = argparse.ArgumentParser()
ap.add_argument("--oracle", type=int, nargs="*", default=[])
ap.add_argument("--level1", type=int, nargs="*", default=[])
ap.add_argument("--out", required=True)
a = ap.parse_args()
rec = {"sweeps": {}, "source": {}}
if os.path.exists(a.out):
    rec = json.load(open(a.out))
for n in a.oracle:
    A0 = H.uniform_matrix(n, n, seed=n)
    t0 = time.time()
    _, _, _, st, rc = H.oracle_svd(A0, nthreads=os.cpu_count() or 1)
    assert rc == 0 and st.converged == 1
    rec["sweeps"][str(n)] = int(st.sweeps)
    rec["source"][str(n)] = f"oracle/liboracle.so to convergence, {time.time() - t0:.0f} s on {os.cpu_count()} host threads"
    print(n, st.sweeps, flush=True)
if a.level1:
    import torch
    svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
    for n in a.level1:
        A = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(n, n, seed=n).T)).cuda()
        ds = torch.zeros(n, dtype=torch.float64, device="cuda"); dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
        st = svdj.dgesvd_device(0, 0, n, n, A.data_ptr(), n, ds.data_ptr(), dV.data_ptr(), n,
                                svdj.default_opts(variant=svdj.VARIANT_LEVEL1, precondition=0))
        assert st.converged == 1 and st.variant == svdj.VARIANT_LEVEL1
        rec["sweeps"][str(n)] = int(st.sweeps)
        rec["source"][str(n)] = f"CUDA level-1 path (reference ordering) on B200, {st.seconds_sweeps:.1f} s"
        print(n, st.sweeps, flush=True)
os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
json.dump(rec, open(a.out, "w"), indent=1, sort_keys=True)
