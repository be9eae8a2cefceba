"""CPU-side checks of bench.py's host logic (no GPU, no timing)."""
import importlib.util
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_module", ROOT / "bench.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_sweep_placements_are_aligned_disjoint_and_different():
    b = load_bench()
    bU, bG = 1_000_000 * 27 * 4, 1_000_000 * 27 * 4 * 6      # fp32 P = 2: field sizes that are not multiples of 256
    seen = set()
    for i in range(len(b.PLACEMENTS)):
        off, end = b.place(i, bU, bG)
        spans = sorted((off[f], off[f] + (bG if f == "G" else bU)) for f in "UGW")
        assert all(o % 256 == 0 for o, _ in spans)
        assert all(a1 <= b0 for (_, a1), (b0, _) in zip(spans, spans[1:]))
        assert spans[-1][1] <= end
        seen.add(tuple(off[f] for f in "UGW"))
    assert len(seen) == len(b.PLACEMENTS) >= 5


def test_both_arms_print_the_same_config(monkeypatch):
    b = load_bench()
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "4"])
    a = b.parse_args()
    ours, ref = b.config_dict(a, 4), b.config_dict(a, a.gpus)
    assert ours == ref
    assert ours["dofs_per_step"] == 4 * 1_000_000 * 64 and ours["cells_per_gpu"] == 1_000_000
    assert "degree-3" in ours["workload"] and "fp64" in ours["workload"]


def test_clock_sampler_windows():
    b = load_bench()
    c = b.ClockSampler(0)
    row = "1500, 1965, 900.0, Not Active, Not Active, Not Active, Active"
    c.rows = [(10.0, row), (10.05, row.replace("1500", "1900")), (11.0, row.replace("1500", "1200"))]
    whole = c.summary()
    assert whole["samples"] == 3 and whole["reasons"] == ["sw_power_cap"] and "window" in whole
    part = c.summary(9.9, 10.1)
    assert part["samples"] == 2 and part["sm_mhz"] == 1700.0
    assert c.summary(20.0, 21.0)["samples"] == 0


def test_sweep_front_end_prints_the_reference_report_line():
    """tools/sweep.py reproduces the columns of bench_mass.cpp:140-146 with the reference's own flop / byte models
    (bench_mass.cpp:122,125, bench_stiffness.cpp:122,125)."""
    spec = importlib.util.spec_from_file_location("sweep_tool", ROOT / "tools" / "sweep.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    e = {"op": "stiffness", "degree": 3, "dtype": "f64", "cells": 1000, "ms": {"median": 2.0}}
    f = mod.reference_line(e).split()
    assert f[:4] == ["3", "1", "double", "64000"] and f[5:8] == ["gpu", "0", "0"]
    m, n, k = 4, 16, 4
    flops = 3 * 3 * 2 * (2.0 * m * n * k + n * k) * 1000
    assert abs(float(f[8]) - flops / 2e-3 / 1e9) < 1e-3 * float(f[8])
    assert abs(float(f[9]) - 9 * 64000 * 8 / 2e-3 / 1e9) < 1e-3 * float(f[9])
    assert abs(float(f[10]) - 64000 / 2e-3 / 1e9) < 1e-6
    e = {"op": "mass", "degree": 2, "dtype": "f32", "cells": 10, "ms": {"median": 1.0}}
    f = mod.reference_line(e).split()
    assert f[2] == "float" and abs(float(f[9]) - 4 * 270 * 4 / 1e-3 / 1e9) < 1e-9
