#!/usr/bin/env python
"""Degree sweep front-end: the GPU counterpart of the reference's run.sh:13-23
(degree x loop-order sweep of ./build/gemm_order) and of the one-line report of bench_mass.cpp:118-146.

    python tools/sweep.py [--filter stiffness_P] [--cells 1000000] [--csv out.csv] [--json out.json]
    python tools/sweep.py --from-json RECORD.json      # re-print a recorded full bench.py record, no GPU needed
    python tools/sweep.py --from-json RECORD.json --markdown   # the DESIGN.md section 5 table, generated, never typed

RECORD.json is the full record bench.py writes next to its compact stdout line (gpurun_out/bench_detail_n1.json, committed
copies under profiles/), or a round-2 line that still carried the 48 entries itself.

Runs bench.py's `sweep` leg (48 entries: mass / stiffness REF_EXACT / stiffness FULL x P = 1..8 x fp64 / fp32; per entry
five placements, clocks, oracle spot check) and prints, per entry,
  * the reference's report line:  degree size type Ndofs elapsed order Mb Nb GFLOP/s GB/s GDoF/s
    (size = 1 rank; order/Mb/Nb are CPU loop-nest knobs the GPU ignores: printed as "gpu 0 0"; GFLOP/s and GB/s use the
    reference's own models, bench_mass.cpp:122,125 / bench_stiffness.cpp:122,125, so the columns compare like for like),
  * and, with --csv, a machine-readable row with the roofline fraction, kernel family, clocks and parity columns.
"""
import argparse
import csv
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def reference_line(e):
    """bench_mass.cpp:140-146 / bench_stiffness.cpp:140-146 for one sweep entry (median placement)."""
    P, cells = e["degree"], e["cells"]
    n1 = P + 1
    nd = n1 ** 3
    elapsed = e["ms"]["median"] * 1e-3
    size_t = 8 if e["dtype"] == "f64" else 4
    gemm = 2.0 * n1 * n1 * n1 * n1 + n1 * n1 * n1          # 2 m n k + n k  with m = k = P+1, n = (P+1)^2
    if e["op"] == "mass":
        flops = 3 * 2 * gemm * cells                         # bench_mass.cpp:122
        mem = (1 + 2 + 1) * cells * nd * size_t               # U + 2 W + detJ, bench_mass.cpp:125
    else:
        flops = 3 * 3 * 2 * gemm * cells                     # bench_stiffness.cpp:122
        mem = (1 + 2 + 6) * cells * nd * size_t               # U + 2 W + G, bench_stiffness.cpp:125
    tname = "double" if e["dtype"] == "f64" else "float"
    return (f"{P} 1 {tname} {cells * nd} {elapsed:.6g} gpu 0 0 "
            f"{flops / elapsed / 1e9:.6g} {mem / elapsed / 1e9:.6g} {cells * nd / elapsed / 1e9:.6g}")


CSV_COLS = ("name", "op", "mode", "degree", "dtype", "cells", "family", "gdofs_min", "gdofs_median", "gdofs_max",
            "frac_min", "frac_median", "frac_max", "frac_of_nominal_8TBs", "ms_median", "sm_mhz", "clock_samples",
            "throttle_reasons", "max_rel_err", "tol", "parity_ok", "regs", "smem_bytes", "blocks_per_sm")


def csv_row(e):
    return {"name": e["name"], "op": e["op"], "mode": e["mode"], "degree": e["degree"], "dtype": e["dtype"],
            "cells": e["cells"], "family": e["family"], "gdofs_min": e["gdofs"]["min"], "gdofs_median": e["gdofs"]["median"],
            "gdofs_max": e["gdofs"]["max"], "frac_min": e["frac_min"], "frac_median": e["frac"], "frac_max": e["frac_max"],
            "frac_of_nominal_8TBs": e["frac_of_nominal_8TBs"], "ms_median": e["ms"]["median"],
            "sm_mhz": e["clocks"].get("sm_mhz"), "clock_samples": e["clocks"].get("samples"),
            "throttle_reasons": "+".join(e["clocks"].get("reasons", [])), "max_rel_err": e["max_rel_err"], "tol": e["tol"],
            "parity_ok": e["parity_ok"], "regs": e["kernel"]["regs"], "smem_bytes": e["kernel"]["smem_bytes"],
            "blocks_per_sm": e["kernel"]["blocks_per_sm"]}


FAMILY_TAG = {"slab": "s", "cell": "c", "column": "col", "tensor core": "tc", "tensor-core": "tc", "generic": "g"}


def markdown_table(sweep, peak):
    """per degree and operator, median fraction of the measured HBM peak / of the nominal 8 TB/s,
    and the kernel family that served the entry."""
    cols = (("mass", "ref_exact", "f64", "mass fp64"), ("mass", "ref_exact", "f32", "mass fp32"),
            ("stiffness", "ref_exact", "f64", "stiffness fp64"), ("stiffness", "ref_exact", "f32", "stiffness fp32"),
            ("stiffness", "full", "f64", "FULL fp64"), ("stiffness", "full", "f32", "FULL fp32"))
    by = {(e["op"], e["mode"], e["dtype"], e["degree"]): e for e in sweep}
    out = [f"Median over five placements: fraction of the measured peak ({peak:.1f} GB/s) / of nominal 8 TB/s, kernel family "
           "(s slab, c cell, col column, tc tensor core).", "",
           "| P | " + " | ".join(c[3] for c in cols) + " |", "|---|" + "---|" * len(cols)]
    for P in sorted({e["degree"] for e in sweep}):
        cells = []
        for op, mode, dt, _ in cols:
            e = by.get((op, mode, dt, P))
            cells.append("–" if e is None else f"{e['frac']:.2f} / {e['frac_of_nominal_8TBs']:.2f} {FAMILY_TAG.get(e['family'], e['family'])}")
        out.append(f"| {P} | " + " | ".join(cells) + " |")
    fr = sorted(e["frac"] for e in sweep)
    low = [e["frac"] for e in sweep if e["degree"] <= 4 and e["mode"] == "ref_exact"]
    out += ["", f"Entries {len(sweep)}, parity ok {sum(1 for e in sweep if e['parity_ok'])}; median of all {fr[len(fr) // 2 - 1] / 2 + fr[len(fr) // 2] / 2:.3f}"
            if len(fr) % 2 == 0 else f"Entries {len(sweep)}; median {fr[len(fr) // 2]:.3f}",
            f"minimum over mass / REF_EXACT at degree <= 4: {min(low):.3f} of measured, {min(low) * peak / 8000:.3f} of 8 TB/s." if low else ""]
    return "\n".join(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--markdown", action="store_true", help="print the DESIGN.md section 5 table instead of the report lines")
    ap.add_argument("--filter", default="")
    ap.add_argument("--cells", type=int, default=1_000_000)
    ap.add_argument("--csv")
    ap.add_argument("--json")
    ap.add_argument("--from-json", help="print from a recorded bench.py JSON line instead of running")
    a = ap.parse_args()
    if a.from_json:
        text = Path(a.from_json).read_text().strip()
        try:
            line = json.loads(text)                          # a full record (indented or one line)
        except json.JSONDecodeError:
            line = json.loads(text.splitlines()[-1])         # a log whose last line is the record
        line = line.get("parsed", line)
    else:
        # bench.py prints a compact line; the 48 full entries are in the record it writes beside it
        detail = ROOT / "gpurun_out" / "sweep_tool_detail.json"
        cmd = [sys.executable, str(ROOT / "bench.py"), "--steps", "5", "--warmup", "3", "--no-e2e", "--no-cpu-baseline",
               "--no-strong", "--sweep-cells", str(a.cells), "--sweep-filter", a.filter, "--detail-file", str(detail)]
        subprocess.run(cmd, check=True, capture_output=True, text=True)
        line = json.loads(detail.read_text())
    sweep = [e for e in line["sweep"] if a.filter in e["name"]]
    if a.markdown:
        print(markdown_table(sweep, line["roofline"]["peak"]))
        return 0
    print("# degree size type Ndofs elapsed order Mb Nb GFLOP/s GB/s GDoF/s   (bench_mass.cpp:140-146)")
    for e in sweep:
        print(reference_line(e), f"  # {e['name']}: {e['family']}, {e['frac']:.3f} of measured HBM "
              f"[{e['frac_min']:.3f}..{e['frac_max']:.3f}], {e['clocks'].get('sm_mhz')} MHz, "
              f"err {e['max_rel_err']:.1e} {'ok' if e['parity_ok'] else 'FAIL'}")
    if a.csv:
        with open(a.csv, "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=CSV_COLS)
            w.writeheader()
            for e in sweep:
                w.writerow(csv_row(e))
    if a.json:
        Path(a.json).write_text(json.dumps(sweep, indent=1))
    return 0 if all(e["parity_ok"] for e in sweep) else 1


if __name__ == "__main__":
    raise SystemExit(main())
