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


def test_the_printed_line_stays_small_and_keeps_every_contract_key():
    """Round 2's N = 1 line was 52 KB (48 sweep dicts inside it) and the driver's capture held only its tail.  The
    printed form of that very record (profiles/r02_bench_n1.json) must fit in 6 KB with value / roofline / e2e /
    cpu_baseline / clocks / strong and a compact sweep table still in it, and fit_line must bound any record."""
    import json
    b = load_bench()
    full = json.loads((ROOT / "profiles" / "r02_bench_n1.json").read_text())
    assert len(json.dumps(full)) > 40_000 and len(full["sweep"]) == 48
    full["leg_errors"] = {"stream_peaks": "RuntimeError: " + "x" * 120}
    line = b.slim_line(full)
    line["detail_file"] = "gpurun_out/bench_detail_n1.json"
    printed = b.fit_line(line)
    text = json.dumps(printed)
    assert len(text) < 6144, len(text)
    assert "dropped_for_size" not in printed
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "roofline", "e2e", "gpu_launches", "clocks", "cpu_baseline",
                "strong", "sweep", "sweep_summary"):
        assert key in printed, key
    assert printed["config"] == full["config"]                       # both arms print the same config
    for key in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert key in printed["roofline"]
    for key in ("value", "unit", "cores", "kind", "sample"):
        assert key in printed["cpu_baseline"]
    for key in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert key in printed["e2e"]
    assert abs(printed["value"] / full["value"] - 1) < 1e-5 and printed["checksum"] == full["checksum"]
    sw = printed["sweep"]
    assert sw["P"] == list(range(1, 9)) and len(sw["groups"]) == 6 and sw["cols"][0] == "frac"
    e = next(x for x in full["sweep"] if x["name"] == "stiffness_P3_f64")
    assert abs(sw["groups"]["stiffness_f64"][0][2] / e["frac"] - 1) < 5e-3
    # the guard: a record that is still too large loses optional keys, never contract keys
    line["sweep_summary"]["junk"] = "y" * 5000
    small = b.fit_line(line)
    assert len(json.dumps(small)) < 6144 and "dropped_for_size" in small
    for key in ("value", "roofline", "e2e", "cpu_baseline", "clocks", "config"):
        assert key in small
