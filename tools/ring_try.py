"""torchrun tools/ring_try.py --n N [--sweeps S]: ring solver on the visible GPUs, prints timing per rank 0."""
import argparse, importlib, os, sys, time
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
ap = argparse.ArgumentParser(); ap.add_argument("--size", "--n", dest="n", type=int, default=4096); ap.add_argument("--sweeps", type=int, default=60); ap.add_argument("--exchange", default="nccl")
a = ap.parse_args()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
ring = importlib.import_module("svd-jacobi-mpi-cuda_b200.ring")
n = m = a.n
solver = ring.RingSolver(m, n, dev, max_sweeps=a.sweeps, exchange=a.exchange)
w = solver.w
init = {b.block_id * w: torch.from_numpy(H.uniform_columns(m, n, n, b.block_id * w, w)).to(dev) for b in (solver.top, solver.bottom)}
solver.fill_local(lambda c0, cnt: init[c0])
dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
st = solver.solve()
sigma, order = solver.finalize()
torch.cuda.synchronize(); dist.barrier(); dt = time.perf_counter() - t0
if rank == 0:
    print(f"n={n} gpus={world} exchange={a.exchange}: {dt:.2f} s, {st}, sigma[0]={float(sigma[0]):.6f} launches/rank={solver.backend.launches}")
dist.destroy_process_group()
