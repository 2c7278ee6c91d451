import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
ring = importlib.import_module("svd-jacobi-mpi-cuda_b200.ring")
m, n, w = 200, 128, 32
dev = torch.device("cuda", 0)
lda, ldv = 208, 128
backend = ring.GpuBackend(m, n, w, lda, ldv, m * 2.0 ** -53)
A0 = H.uniform_matrix(m, n, seed=2)
def mk(b):
    blk = ring.Block(w, m, n, lda, ldv, dev); blk.block_id = b
    blk.A[:, :m] = torch.from_numpy(np.ascontiguousarray(A0.T[b * w:(b + 1) * w])).to(dev)
    idx = torch.arange(w, device=dev); blk.V[idx, b * w + idx] = 1.0; return blk
def maxcos(blks):
    A = torch.cat([b.A[:, :m] for b in blks]).T
    G = A.T @ A; d = torch.sqrt(torch.diag(G)); C = (G / torch.outer(d, d)).abs(); C.fill_diagonal_(0); return C.max().item()
for order in ["xy", "yx"]:
    X, Y = mk(0), mk(1)
    top, bot = (X, Y) if order == "xy" else (Y, X)
    print(order, "ptr order top<bot:", top.A.data_ptr() < bot.A.data_ptr())
    ctr = torch.zeros(2, dtype=torch.int64, device=dev)
    for it in range(8):
        ctr.zero_(); backend.phase(top, bot, ring.PHASE_ALL, ctr)
        print("  it", it, "rot", int(ctr[0]), "maxcos %.2e" % maxcos([X, Y]), "pad max %.1e" % max(X.A[:, m:].abs().max().item(), Y.A[:, m:].abs().max().item()),
              "V colnorm err %.1e" % (torch.cat([X.V, Y.V]).norm(dim=1) - 1).abs().max().item())
