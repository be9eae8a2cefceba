"""Summarise an ncu report (read here, without a GPU) into profiles/<name>.md and update
profiles/traffic.json.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep <name> [--traffic-key KEY]
"""
import csv
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "smsp__inst_executed.sum",
]


def to_bytes(v, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
    return float(v) * scale


def main():
    rep, name = sys.argv[1], sys.argv[2]
    tkey = sys.argv[sys.argv.index("--traffic-key") + 1] if "--traffic-key" in sys.argv else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    lines = [f"# ncu summary: {name}\n", f"source report: `{Path(rep).name}` (ncu --set full --clock-control none --import-source on)\n"]
    traffic = []
    for r in rows[2:]:
        lines.append(f"\n## {r[hdr.index('Kernel Name')]}\n\n| metric | value | unit |\n|---|---:|---|\n")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                lines.append(f"| {k} | {r[i]} | {units[i]} |\n")
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        t = to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])
        traffic.append(t)
        lines.append(f"| **dram traffic (read+write)** | {t / 1e6:.2f} | MB |\n")
    (ROOT / "profiles").mkdir(exist_ok=True)
    (ROOT / "profiles" / f"{name}.md").write_text("".join(lines))
    if tkey and traffic:
        p = ROOT / "profiles" / "traffic.json"
        d = json.loads(p.read_text()) if p.exists() else {}
        d[tkey] = sum(traffic) / len(traffic)
        p.write_text(json.dumps(d, indent=1, sort_keys=True) + "\n")
    print("".join(lines))


if __name__ == "__main__":
    main()
