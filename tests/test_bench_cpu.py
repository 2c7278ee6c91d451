"""CPU tests of bench.py's measurement helpers: the algorithmic-byte figure, the clock
sampler's parsing of nvidia-smi lines, the CPU baseline's sample/extrapolation, and the command line."""
import importlib.util
import os
import sys

import numpy as np

import helpers as H

ROOT = H.ROOT


def _bench():
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_algorithmic_bytes_match_the_survey_figures():
    b = _bench()
    # one full sweep, every pair rotated: n (n-1) / 2 pairs x 32 (m + n) bytes
    for n, tb in [(4096, 2.1985e12), (512, 4.29e9), (8192, 17.59e12)]:
        pairs = n * (n - 1) // 2
        assert abs(b.algorithmic_bytes(n, n, pairs, pairs) / tb - 1) < 2e-3
    m, n = 1048576, 512
    pairs = n * (n - 1) // 2
    assert abs(b.algorithmic_bytes(m, n, pairs, pairs) / 4.3916e12 - 1) < 2e-3
    # a pair that is only tested costs its Gram read: 16 m bytes
    assert b.algorithmic_bytes(100, 50, 10, 4) == 4 * 32.0 * 150 + 6 * 16.0 * 100


def test_clock_sampler_parses_nvidia_smi_lines():
    b = _bench()
    s = b.ClockSampler(0)
    s.proc = type("P", (), {"terminate": lambda self: None, "wait": lambda self, timeout=None: 0, "kill": lambda self: None})()
    s.lines = ["0, 1965, 1965, 640.1, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
               "0, 1950, 1965, 980.3, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
               "garbage", "0, [N/A], 1965, 1.0, 0x0, Not Active, Not Active, Not Active, Not Active"]
    out = s.stop()
    assert out["samples"] == 2 and out["sm_mhz"] == 1957.5 and out["sm_max_mhz"] == 1965.0
    assert out["reasons"] == ["sw_power_cap"]
    assert b.ClockSampler(0).stop()["reasons"] == ["nvidia-smi unavailable"]   # never started


def test_cpu_baseline_times_whole_steps_and_extrapolates():
    b = _bench()
    A0 = H.uniform_matrix(96, 64, seed=1)
    r = b.cpu_baseline_port(A0, sweeps_needed=11, budget_s=0.2)
    assert r["kind"] == "port" and r["unit"] == "s" and r["cores"] == (os.cpu_count() or 1)
    assert r["value"] > 0 and "round-robin steps of sweep 1 on 96x64" in r["sample"]
    assert r["rotations_per_s"] is None or r["rotations_per_s"] > 0


def test_peaks_reads_the_driver_file_or_says_fallback():
    b = _bench()
    v, src = b.peaks()
    assert v > 1000 and ("measured" in src or "fallback" in src)


def test_command_line_defaults(monkeypatch):
    b = _bench()
    seen = {}
    monkeypatch.setattr(b, "run_svdj", lambda args: seen.update(vars(args)) or 0)
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    assert b.main() == 0
    assert seen["gpus"] == 1 and seen["n"] == 4096 and seen["warmup"] >= 3 and seen["impl"] == "svdj" and seen["rows"] == 0
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "8", "--n", "16384", "--steps", "1", "--warmup", "1"])
    b.main()
    assert seen["gpus"] == 8 and seen["n"] == 16384
