"""Runs a fixed number of level-1 round-robin steps on a resident matrix: the command ncu profiles.

    python tools/profile_steps.py [--n 4096] [--m 0] [--steps 40] [--path auto|reg|split]
"""
import argparse
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4096)
ap.add_argument("--m", type=int, default=0)
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--path", default="auto")
a = ap.parse_args()
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
n, m = a.n, a.m or a.n
A = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(m, n, seed=n).T)).cuda()
V = torch.eye(n, dtype=torch.float64, device="cuda")
ctr = torch.zeros(2, dtype=torch.int64, device="cuda")
path = {"auto": svdj.PATH_AUTO, "reg": svdj.PATH_REG, "split": svdj.PATH_SPLIT}[a.path]
scratch = torch.empty(max(svdj.step_scratch_doubles(m, n), 1), dtype=torch.float64, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize()
for it in range(a.steps):
    if it == 5:
        ev[0].record()
    svdj.step_device(m, n, A.data_ptr(), m, V.data_ptr(), n, True, it % svdj.num_steps(n), m * 2.0 ** -53,
                     path=path, d_scratch=scratch.data_ptr(), d_counters=ctr.data_ptr())
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / max(a.steps - 5, 1)
by = (n // 2) * 32.0 * (m + n)
print(f"{m}x{n} path={a.path}: {ms * 1e3:.1f} us/step, {by / ms / 1e6:.0f} GB/s algorithmic, rotated {int(ctr[0])}")
