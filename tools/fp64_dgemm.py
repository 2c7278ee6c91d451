"""cuBLAS DGEMM throughput on this GPU: the FP64 denominator BASELINE.md asks for."""
import json, time, torch
n = 8192
a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda")
for _ in range(2): c = a @ b
torch.cuda.synchronize()
best = 1e9
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(json.dumps({"dgemm_n": n, "ms": best, "fp64_tflops_cublas": 2 * n ** 3 / best / 1e9}))
# skinny shapes of the block path: (m x 32)^T (m x 32) and (m x 32)(32 x 32), batched 128
m, d, nb = 4096, 32, 128
x = torch.randn(nb, m, d, dtype=torch.float64, device="cuda"); w = torch.randn(nb, d, d, dtype=torch.float64, device="cuda")
for name, f, fl in [("gram", lambda: x.transpose(1, 2) @ x, 2 * nb * m * d * d), ("update", lambda: x @ w, 2 * nb * m * d * d)]:
    f(); torch.cuda.synchronize(); best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    print(json.dumps({"shape": name, "ms": best, "tflops": fl / best / 1e9, "GBps_in": nb * m * d * 8 / best / 1e6}))
