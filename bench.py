#!/usr/bin/env python
"""bench.py -- time-to-solution of the FP64 one-sided Jacobi SVD (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl svdj|reference] [--n 4096]

A "step" is one full solve (sweeps to convergence, sigma/U/V out) of the synthetic workload.
At N=1 the workload is BASELINE.json configs[1]: 4096 x 4096 dense uniform(0,1) FP64.

Keys of the JSON line (one line, rank 0):
  value      device-resident time-to-solution in seconds (inputs already in HBM; svdj_dgesvd_device)
  e2e        the same solve through svdj_dgesvd with HOST buffers: H2D of A and D2H of U, s, V inside
  roofline   HBM roofline of the dominant kernel (the fused level-1 step): algorithmic bytes
             (32(m+n) per rotated pair, 16m per pair only tested) / CUDA-event time of the sweep loop
  cpu_baseline  the oracle (kind "port") timed on this box's host cores on a bounded sample
--impl reference times the reference's own solver (oracle/_ref/libref_cpu.so: main.cu built for the
host, rotation kernel body run on CPU cores) -- one reference sweep per step, extrapolated to the
sweeps a converged solve needs.
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "svd_time_to_solution"
UNIT = "s"
# sweeps the reference's algorithm (its ordering, relative test tol = m*2^-53) needs on the bench input;
# 512/1024: the oracle's own count; 4096: the CUDA level-1 path (same ordering, same counts as the oracle wherever
# both were run); 2048/8192/16384 extrapolated at +2 sweeps per doubling of n.
SWEEPS_TO_CONVERGE = {512: 14, 1024: 15, 2048: 17, 4096: 19, 8192: 21, 16384: 23}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smmax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smmax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smmax) if smmax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def algorithmic_bytes(m, n, tested, applied):
    return applied * 32.0 * (m + n) + (tested - applied) * 16.0 * m


def cpu_baseline_port(A0, sweeps_needed, budget_s=15.0):
    """Oracle on host cores: time whole round-robin steps of sweep 1 until ~budget_s, extrapolate."""
    import helpers as H

    m, n = A0.shape
    cores = os.cpu_count() or 1
    A = A0.copy(order="F")
    V = np.asfortranarray(np.eye(n))
    steps = H.oracle_lib().oracle_num_steps(n)
    o = H.OracleOpts(H.ORACLE_REL, m * 2.0 ** -53, 1, 0, cores, 0)
    st = H.OracleStats()
    t0 = time.perf_counter()
    done = 0
    while done < steps and (time.perf_counter() - t0) < budget_s:
        H.oracle_lib().oracle_step(m, n, A, m, V, n, 1, done, C.byref(o), C.byref(st))
        done += 1
    dt = time.perf_counter() - t0
    per_step = dt / max(done, 1)
    value = per_step * steps * sweeps_needed
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{done} of {steps} round-robin steps of sweep 1 on {m}x{n} in {dt:.1f} s, "
                      f"extrapolated x{steps}/{done} steps x{sweeps_needed} sweeps",
            "rotations_per_s": (st.applied / dt) if dt > 0 else None}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.

    n <= 1024: oracle/_ref/libref_cpu.so, i.e. the reference's own solver (main.cu:440-1423) built for the
    host; one step = one call = one sweep (its maxIterations = 1, main.cu:482).  Larger n: a whole reference
    sweep does not fit a bounded run (it cannot be stopped part-way), so the step is a bounded number of
    round-robin steps of the oracle port (same algorithm, same OpenMP-over-pairs structure, without the
    reference's per-pair staging copies -- i.e. a faster CPU baseline than the reference itself).
    Either way the value is extrapolated to the sweeps a converged solve needs and says so."""
    import helpers as H

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = args.n
    cores = os.cpu_count() or 1
    sweeps_needed = SWEEPS_TO_CONVERGE.get(n, 15)
    A0 = H.uniform_matrix(n, n, seed=n)
    lib = H.ref_cpu_lib()
    pairs = n * (n - 1) // 2
    steps_per_sweep = H.oracle_lib().oracle_num_steps(n)
    if lib is not None and n <= 1024:
        kind = "reference"

        def one():
            t0 = time.perf_counter()
            H.ref_cpu_svd(A0, sweeps=1, nthreads=cores)
            return time.perf_counter() - t0, pairs

        what = f"one whole reference sweep of {n}x{n} per step"
    else:
        kind = "port"
        budget = 8.0
        state = {"A": A0.copy(order="F"), "V": np.asfortranarray(np.eye(n)), "next": 0}
        o = H.OracleOpts(H.ORACLE_ABS, 1e-16, 1, 0, cores, 0)

        def one():
            st = H.OracleStats()
            t0 = time.perf_counter()
            done = 0
            while (time.perf_counter() - t0) < budget and done < steps_per_sweep:
                H.oracle_lib().oracle_step(n, n, state["A"], n, state["V"], n, 1, state["next"] % steps_per_sweep,
                                           C.byref(o), C.byref(st))
                state["next"] += 1
                done += 1
            return (time.perf_counter() - t0) * steps_per_sweep / done, pairs

        what = (f"~{budget:.0f} s of consecutive round-robin steps of {n}x{n} per step (oracle port of "
                f"main.cu:684-760), scaled to the {steps_per_sweep} steps of a sweep")
    for _ in range(args.warmup):
        one()
    times = [one()[0] for _ in range(args.steps)]
    per_sweep = float(np.mean(times))
    value = per_sweep * sweeps_needed
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": value * 1e3, "higher_is_better": False,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{n}x{n} dense uniform(0,1) FP64, full SVD (U, S, V) to convergence",
                   "note": f"value = seconds per sweep x {sweeps_needed} sweeps needed to converge; the reference "
                           "itself stops after 1 sweep (main.cu:482)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{what}; {per_sweep:.2f} s per sweep x{sweeps_needed} sweeps"},
        "seconds_per_sweep": per_sweep,
        "rotations_per_s": pairs / per_sweep,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def run_ring(args, svdj, world, rank, local_rank):
    """N > 1: the same n x n solve spread over N GPUs (strong scaling): 2N column blocks on a ring, NCCL send/recv
    between rounds, one all-reduce per sweep (svd-jacobi-mpi-cuda_b200/ring.py)."""
    import torch
    import torch.distributed as dist

    import helpers as H

    ring = importlib.import_module("svd-jacobi-mpi-cuda_b200.ring")
    n = m = args.n
    dev = torch.device("cuda", local_rank)
    solver = ring.RingSolver(m, n, dev, exchange=args.exchange)
    w = solver.w

    # every rank generates its own columns (counter-based generator: same matrix as the 1-GPU run)
    init = {}
    for blk in (solver.top, solver.bottom):
        init[blk.block_id * w] = torch.from_numpy(H.uniform_columns(m, n, n, blk.block_id * w, w)).to(dev)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def one_solve():
        solver.fill_local(lambda c0, cnt: init[c0])
        st = solver.solve()
        sigma, order = solver.finalize()
        return st, sigma, order

    for _ in range(args.warmup):
        one_solve()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    solver.exchanged_bytes = 0
    solver.backend.launches = 0
    ev0.record()
    for _ in range(args.steps):
        st, sigma, order = one_solve()
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ev0.elapsed_time(ev1) * 1e-3], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    per_step_s = float(t.item()) / args.steps
    sent_per_solve = solver.exchanged_bytes / args.steps
    launches = torch.tensor([solver.backend.launches], dtype=torch.int64, device=dev)
    dist.all_reduce(launches, op=dist.ReduceOp.SUM)

    # e2e: the matrix starts in rank 0's pinned host memory and U, s, V end there (the reference's contract: only
    # the root holds A, s, V -- main.cu:1455, 1548-1556)
    hA0 = torch.from_numpy(np.ascontiguousarray(H.uniform_matrix(m, n, seed=n).T)).pin_memory() if rank == 0 else None
    # results land in pinned host buffers the caller owns (as the single-GPU entry point's do); pageable
    # destinations allocated inside the timed region cost ~0.1 s of page faults and staging per solve
    if rank == 0:
        hU = torch.empty(n, m, dtype=torch.float64).pin_memory()
        hV = torch.empty(n, n, dtype=torch.float64).pin_memory()
        hs = torch.empty(n, dtype=torch.float64).pin_memory()
    e2e = []
    for it in range(1 + args.steps):
        barrier()
        t0 = time.perf_counter()
        solver.scatter_from_root(hA0.to(dev, non_blocking=True) if rank == 0 else None)
        st, sigma, order = solver.solve(), None, None
        sigma, order = solver.finalize()
        U, V = solver.gather_to_root(order)
        if rank == 0:
            hU.copy_(U, non_blocking=True)
            hV.copy_(V, non_blocking=True)
            hs.copy_(sigma, non_blocking=True)
            torch.cuda.synchronize()
        barrier()
        if it > 0:
            e2e.append(time.perf_counter() - t0)
    e2e_s = float(np.mean(e2e))

    if rank == 0:
        Ud, Vd, Sd = hU.to(dev).T, hV.to(dev).T, hs.to(dev)
        A0d = hA0.to(dev).T
        eye = torch.eye(n, dtype=torch.float64, device=dev)
        acc = {"orth_u": float((Ud.T @ Ud - eye).abs().max()), "orth_v": float((Vd.T @ Vd - eye).abs().max()),
               "resid": float(torch.linalg.norm(A0d - (Ud * Sd) @ Vd.T) / torch.linalg.norm(A0d)),
               "bound_10_n_eps": 10 * n * 2.0 ** -52}
        line = {
            "metric": METRIC, "value": per_step_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step_s * 1e3, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{m}x{n} dense uniform(0,1) FP64, full SVD (U, S, V) to convergence",
                       "variant": "block", "tol": "m*2^-53 relative",
                       "l2": "inputs_larger_than_l2 (A+V = %d MiB)" % (2 * m * n * 8 >> 20),
                       "parallelism": f"{2 * world} column blocks of {w} columns on a ring of {world} GPUs "
                                      + ("(blocks stored into the neighbours' buffers by the update kernel over NVLink peer mappings, 1 all-reduce per sweep)" if args.exchange == "fused" else "(NCCL send/recv between rounds, 1 all-reduce per sweep)")},
            "sweeps": st["sweeps"], "rotations_applied": st["applied"],
            "ring": {"bytes_sent_per_rank_per_solve": sent_per_solve, "rounds_per_sweep": 2 * world - 1, "exchange": args.exchange},
            "roofline": None, "cpu_baseline": None,
            "e2e": {"value": e2e_s, "unit": UNIT, "h2d_bytes_per_step": m * n * 8,
                    "d2h_bytes_per_step": (m * n + n * n + n) * 8},
            "gpu_launches": int(launches.item()), "clocks": clocks, "accuracy": acc,
        }
        print(json.dumps(line))
    dist.destroy_process_group()
    return 0


def run_rows(args, svdj, world, rank, local_rank):
    """--rows M with M >= 8 n: the tall-skinny workload (BASELINE.json config 4 is 1048576 x 512), sharded by rows
    over the N GPUs (svd-jacobi-mpi-cuda_b200/rows.py): one all-reduce of the Gram matrices per step, no column
    traffic.  Strong scaling: the matrix is the same for every N (generated in 65536-row chunks, seed per chunk)."""
    import torch
    import torch.distributed as dist

    rows = importlib.import_module("svd-jacobi-mpi-cuda_b200.rows")
    m, n = args.rows, args.n
    dev = torch.device("cuda", local_rank)
    solver = rows.RowShardSolver(m, n, dev)
    r0, ml = solver.row0, solver.m_local

    def my_rows():  # (n, m_local), standard normal, the same global matrix whatever the partition
        CH = 65536
        out = torch.empty(n, ml, dtype=torch.float64, device=dev)
        for c in range(r0 // CH, (r0 + ml - 1) // CH + 1):
            g = torch.Generator(device=dev)
            g.manual_seed(1234 + c)
            chunk = torch.randn(n, min(CH, m - c * CH), dtype=torch.float64, device=dev, generator=g)
            lo, hi = max(r0, c * CH), min(r0 + ml, (c + 1) * CH)
            out[:, lo - r0:hi - r0] = chunk[:, lo - c * CH:hi - c * CH]
        return out

    A_rows = my_rows()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_solve():
        solver.fill_local(A_rows)
        return solver.solve()

    for _ in range(args.warmup):
        one_solve()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    solver.allreduce_calls = solver.allreduce_bytes = 0
    stats = []
    ev0.record()
    for _ in range(args.steps):
        stats.append(one_solve())
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ev0.elapsed_time(ev1) * 1e-3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    per_step_s = float(t.item()) / args.steps
    st = stats[-1]
    ar_calls, ar_bytes = solver.allreduce_calls / args.steps, solver.allreduce_bytes / args.steps

    # e2e: every rank's rows start in its own pinned host memory and its rows of U and V (and sigma) end there
    hA = A_rows.cpu().pin_memory()
    hU = torch.empty(n, ml, dtype=torch.float64).pin_memory()
    hV = torch.empty(n, solver.v_local, dtype=torch.float64).pin_memory()
    hs = torch.empty(n, dtype=torch.float64).pin_memory()
    e2e = []
    for it in range(1 + args.steps):
        barrier()
        t0 = time.perf_counter()
        solver.fill_local(hA.to(dev, non_blocking=True))
        solver.solve()
        hU.copy_(solver.U_local, non_blocking=True)
        hV.copy_(solver.V_local, non_blocking=True)
        hs.copy_(solver.sigma, non_blocking=True)
        barrier()
        if it > 0:
            e2e.append(time.perf_counter() - t0)
    e2e_s = float(np.mean(e2e))

    # accuracy from the row slices: U^T U and V^T V are sums over the ranks; the residual needs the whole (small) V
    U, V, sg = solver.U_local.T, solver.V_local.T, solver.sigma        # (m_local, n), (v_local, n)
    utu, vtv = U.T @ U, V.T @ V
    Vfull = torch.zeros(n, n, dtype=torch.float64, device=dev)
    Vfull[solver.v_row0:solver.v_row0 + solver.v_local] = V
    if world > 1:
        dist.all_reduce(utu)
        dist.all_reduce(vtv)
        dist.all_reduce(Vfull)
    res2 = torch.stack([torch.linalg.norm(A_rows.T - (U * sg) @ Vfull.T) ** 2, torch.linalg.norm(A_rows) ** 2])
    if world > 1:
        dist.all_reduce(res2)
    if rank == 0:
        eye = torch.eye(n, dtype=torch.float64, device=dev)
        acc = {"orth_u": float((utu - eye).abs().max()), "orth_v": float((vtv - eye).abs().max()),
               "resid": float(torch.sqrt(res2[0] / res2[1])), "bound_10_m_eps": 10 * m * 2.0 ** -52}
        launches = sum(s_.step_kernel_launches for s_ in stats)
        pairs_upd = sum(s_.probe_block_pairs for s_ in stats)
        k_ms = float(np.mean([s_.step_kernel_ms for s_ in stats]))
        flops_per_launch = pairs_upd * 2.0 * (ml + solver.v_local) * 64 * 64 / max(launches, 1)
        tfs = flops_per_launch / (k_ms * 1e-3) / 1e12 if k_ms > 0 else 0.0
        line = {
            "metric": METRIC, "value": per_step_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step_s * 1e3, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{m}x{n} dense N(0,1) FP64, full SVD (U, S, V) to convergence",
                       "variant": "block", "block_width": 32, "tol": "m*2^-53 relative",
                       "l2": "inputs_larger_than_l2 (A = %d MiB)" % (m * n * 8 >> 20),
                       "parallelism": f"rows sharded over {world} GPU(s): {ml} rows of A and {solver.v_local} rows of V "
                                      "per rank, one all-reduce of the block pairs' Gram matrices per step"},
            "sweeps": st.sweeps, "rotations_applied": int(st.applied),
            "rows": {"allreduce_calls_per_solve": ar_calls, "allreduce_bytes_per_solve": ar_bytes},
            "roofline": {"bound": "tensor", "achieved": tfs, "peak": 36.9, "unit": "TFLOP/s", "frac": tfs / 36.9,
                         "traffic": None, "kernel": "blk32_update_kernel on rank 0's rows",
                         "peak_source": "measured DMMA issue peak, profiles/r1_fp64_peaks.md",
                         "flops_per_launch": flops_per_launch, "avg_launch_us": k_ms * 1e3,
                         "launches_probed": int(launches)},
            "cpu_baseline": None,
            "e2e": {"value": e2e_s, "unit": UNIT, "h2d_bytes_per_step": m * n * 8,
                    "d2h_bytes_per_step": (m * n + n * n + n) * 8},
            "gpu_launches": int(sum(s_.launches for s_ in stats)) * world, "clocks": clocks, "accuracy": acc,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_svdj(args):
    import torch
    import torch.distributed as dist

    import helpers as H

    svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if args.rows:
        if args.rows < 8 * args.n:
            raise SystemExit("--rows M is the tall-skinny workload: M >= 8 * --size")
        return run_rows(args, svdj, world, rank, local_rank)
    if world > 1:
        return run_ring(args, svdj, world, rank, local_rank)

    n = m = args.n
    A0 = H.uniform_matrix(m, n, seed=n)
    dA0 = torch.from_numpy(np.ascontiguousarray(A0.T)).cuda()  # column-major bytes
    dA = torch.empty_like(dA0)
    ds = torch.zeros(n, dtype=torch.float64, device="cuda")
    dV = torch.empty(n, n, dtype=torch.float64, device="cuda")
    opts = svdj.default_opts(variant=args.variant)
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def solve_device():
        dA.copy_(dA0)
        return svdj.dgesvd_device(svdj.AllVec, svdj.AllVec, m, n, dA.data_ptr(), m, ds.data_ptr(), dV.data_ptr(), n,
                                  opts, stream=stream)

    for _ in range(args.warmup):
        st = solve_device()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats = []
    ev0.record()
    for _ in range(args.steps):
        stats.append(solve_device())
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_s = ev0.elapsed_time(ev1) * 1e-3
    if world > 1:
        t = torch.tensor([total_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_s = float(t.item())
    per_step_s = total_s / args.steps

    # roofline of the dominant kernel, live from this run's CUDA events (recorded inside the library on the
    # stream the kernels are launched on)
    last = stats[-1]
    sweep_s = float(np.mean([s_.seconds_sweeps for s_ in stats]))
    peak, peak_src = peaks()
    if last.variant == svdj.VARIANT_BLOCK:
        # block path: the update kernel X <- X W, Y <- Y W is the dominant kernel (~55-60 % of a step).  Every 8th
        # launch is bracketed by CUDA events inside the library; the probes also count the block pairs updated.
        width = last.reserved or 16          # columns per block: 16 or 32
        d = 2 * width                        # columns of a block pair
        launches = sum(s_.step_kernel_launches for s_ in stats)
        pairs_upd = sum(s_.probe_block_pairs for s_ in stats)
        k_ms = float(np.mean([s_.step_kernel_ms for s_ in stats]))
        # per updated block pair: d columns of A and V read once and written once, 2 (m+n) d^2 flops
        bytes_per_launch = pairs_upd * 16.0 * d * (m + n) / max(launches, 1)
        flops_per_launch = pairs_upd * 2.0 * (m + n) * d * d / max(launches, 1)
        gbs = bytes_per_launch / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
        tfs = flops_per_launch / (k_ms * 1e-3) / 1e12 if k_ms > 0 else 0.0
        dmma_peak = 36.9  # TFLOP/s, mma.sync.m8n8k4.f64 issue peak measured on this pool (profiles/r1_fp64_peaks.md)
        full = (m, n) == (4096, 4096)
        if width == 32:
            # 0.31 * 32 = 10 flop/B against a ridge of 36.9 TF / 6.45 TB/s = 5.7: the FP64 tensor pipe bounds it
            roofline = {"bound": "tensor", "achieved": tfs, "peak": dmma_peak, "unit": "TFLOP/s", "frac": tfs / dmma_peak,
                        # ncu --set full, one launch with every pair active: dram read 272.0 MB + write 212.5 MB
                        # (profiles/r1_final.md); the algorithmic figure of such a launch is 268 + 268 MB
                        "traffic": 484.5e6 if full else None,
                        "peak_source": "measured DMMA (mma.sync.m8n8k4.f64) issue peak on this pool, tools/fp64_peak.cu; "
                                       "MEASURED_PEAKS.json has no FP64 figure (cuBLAS DGEMM here: 35.5)",
                        "kernel": "blk32_update_kernel (X<-XW, Y<-YW for 64-column block pairs; FP64 DMMA tiles, "
                                  "W in shared memory in fragment order, cp.async staged, per-pair column pointer tables)",
                        "flops_per_launch": flops_per_launch, "avg_launch_us": k_ms * 1e3,
                        "launches_probed": int(launches),
                        "hbm": {"alg_bytes_per_launch": bytes_per_launch, "achieved_gbs": gbs, "peak_gbs": peak,
                                "frac": gbs / peak, "peak_source": peak_src}}
        else:
            roofline = {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                        "traffic": 484.7e6 if full else None, "peak_source": peak_src,
                        "kernel": "blk_update_staged_kernel (X<-XW, Y<-YW; FP64 DMMA tiles, cp.async staged)",
                        "frac_of_8TBps_nominal": gbs / 8000.0, "alg_bytes_per_launch": bytes_per_launch,
                        "avg_launch_us": k_ms * 1e3, "launches_probed": int(launches),
                        "fp64_tensor": {"flops_per_launch": flops_per_launch, "achieved_tflops": tfs,
                                        "peak_tflops": dmma_peak,
                                        "peak_source": "measured DMMA issue peak, profiles/r1_fp64_peaks.md"}}
    else:
        # level-1 path: the fused Gram + rotation step is the only kernel of the sweep loop
        alg_bytes = float(np.mean([algorithmic_bytes(m, n, s_.tested, s_.applied) for s_ in stats]))
        achieved = alg_bytes / sweep_s / 1e9
        launches_per_solve = last.step_kernel_launches
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": 478.6e6 if (m, n) == (4096, 4096) else None, "peak_source": peak_src,
                    "kernel": "l1_pair_reg_kernel (fused Gram+rotation step)",
                    "frac_of_8TBps_nominal": achieved / 8000.0,
                    "alg_bytes_per_launch": alg_bytes / max(launches_per_solve, 1),
                    "avg_launch_us": sweep_s / max(launches_per_solve, 1) * 1e6}

    # e2e: host buffers (pinned), H2D + solve + D2H inside the timed call
    hA0 = torch.from_numpy(np.ascontiguousarray(A0.T)).pin_memory()
    hA = torch.empty_like(hA0).pin_memory()
    hs = torch.zeros(n, dtype=torch.float64).pin_memory()
    hV = torch.empty(n, n, dtype=torch.float64).pin_memory()
    e2e_times = []
    for it in range(1 + args.steps):
        hA.copy_(hA0)
        barrier()
        t0 = time.perf_counter()
        st_h = svdj.Stats()
        rc = svdj.lib.svdj_dgesvd(svdj.AllVec, svdj.AllVec, m, n, hA.data_ptr(), m, hs.data_ptr(), hV.data_ptr(), n,
                                  C.byref(opts), C.byref(st_h))
        assert rc == 0, svdj.last_error()
        dt = time.perf_counter() - t0
        if it > 0:
            e2e_times.append(dt)
    e2e_s = float(np.mean(e2e_times))
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())

    if rank == 0:
        # accuracy of the last e2e result (cheap, on the GPU) so that the line carries its own parity evidence
        U = hA.cuda().T
        V = hV.cuda().T
        S = hs.cuda()
        eye = torch.eye(n, dtype=torch.float64, device="cuda")
        A0d = dA0.T
        acc = {"orth_u": float((U.T @ U - eye).abs().max()), "orth_v": float((V.T @ V - eye).abs().max()),
               "resid": float(torch.linalg.norm(A0d - (U * S) @ V.T) / torch.linalg.norm(A0d)),
               "bound_10_n_eps": 10 * n * 2.0 ** -52}
        del U, V, eye
        cpu = cpu_baseline_port(A0, SWEEPS_TO_CONVERGE.get(n, last.sweeps)) if not args.no_cpu_baseline else None
        pairs = n * (n - 1) // 2
        line = {
            "metric": METRIC, "value": per_step_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step_s * 1e3, "higher_is_better": False,
            "scaling": "strong" if world == 1 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{m}x{n} dense uniform(0,1) FP64, full SVD (U, S, V) to convergence",
                       "variant": {1: "level1", 2: "block"}.get(last.variant, str(last.variant)),
                       "block_width": int(last.reserved) if last.variant == svdj.VARIANT_BLOCK else None,
                       "tol": "m*2^-53 relative", "l2": "inputs_larger_than_l2 (A+V = %d MiB)" % (2 * m * n * 8 >> 20)
                       if 2 * m * n * 8 > 126e6 else "inputs fit L2; restored from a master copy each step",
                       "parallelism": "1 GPU" if world == 1 else f"{world} independent replicas"},
            "sweeps": last.sweeps, "rotations_applied": last.applied, "pairs_tested": last.tested,
            "rotations_per_s": last.tested / last.seconds_sweeps,
            "seconds_sweeps": sweep_s, "seconds_finalize": last.seconds_finalize,
            "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": e2e_s, "unit": UNIT, "h2d_bytes_per_step": m * n * 8,
                    "d2h_bytes_per_step": (m * n + n * n + min(m, n)) * 8,
                    "h2d_s": st_h.seconds_h2d, "d2h_s": st_h.seconds_d2h},
            "gpu_launches": int(sum(s_.launches for s_ in stats)),
            "clocks": clocks, "accuracy": acc,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="svdj", choices=["svdj", "reference"])
    ap.add_argument("--size", "--n", dest="n", type=int, default=4096)
    ap.add_argument("--rows", type=int, default=0, help="M >= 8 * size: the tall M x size workload, sharded by rows over the GPUs")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"], help="N > 1: how column blocks move")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_svdj(args)


if __name__ == "__main__":
    sys.exit(main())
