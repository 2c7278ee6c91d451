import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
ring = importlib.import_module("svd-jacobi-mpi-cuda_b200.ring")
P = int(sys.argv[1]); only_all = len(sys.argv) > 2 and sys.argv[2] == "all"
m, n = 200, 64 * P; w = n // (2 * P)
A0 = H.uniform_matrix(m, n, seed=P); dev = torch.device("cuda", 0)
lda, ldv = (m + 15) // 16 * 16, (n + 15) // 16 * 16
backend = ring.GpuBackend(m, n, w, lda, ldv, m * 2.0 ** -53)
blocks = []
for b in range(2 * P):
    blk = ring.Block(w, m, n, lda, ldv, dev); blk.block_id = b
    blk.A[:, :m] = torch.from_numpy(np.ascontiguousarray(A0.T[b * w:(b + 1) * w])).to(dev)
    idx = torch.arange(w, device=dev); blk.V[idx, b * w + idx] = 1.0; blocks.append(blk)
counters = torch.zeros(2, dtype=torch.int64, device=dev)
rounds = ring.tournament_rounds(P)
for sweep in range(16):
    counters.zero_()
    for rd, (top, bottom) in enumerate(rounds):
        for r in range(P):
            ph = ring.PHASE_ALL if (rd == 0 or only_all) else ring.PHASE_CROSS
            backend.phase(blocks[top[r]], blocks[bottom[r]], ph, counters)
    Afull = torch.cat([b.A[:, :m] for b in blocks]).T
    G = Afull.T @ Afull; d = torch.sqrt(torch.diag(G)); Cn = (G / torch.outer(d, d)).abs(); Cn.fill_diagonal_(0)
    print(sweep + 1, "rot", int(counters[0]), "reported off %.2e" % counters[1:2].view(torch.float64).item(), "true max cos %.2e" % Cn.max().item())
