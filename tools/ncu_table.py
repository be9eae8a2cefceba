"""Compact one-line-per-kernel table from an ncu report (read without a GPU).
    python tools/ncu_table.py gpurun_out/x.ncu-rep
"""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
cols = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
        ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue%"),
        ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64%"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wf"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "conflicts"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "blk"),
        ("launch__shared_mem_per_block_dynamic", "smemKB"),
        ("smsp__inst_executed.sum", "inst"),
        ("smsp__average_warp_latency_issue_stalled_long_scoreboard.pct", "st_long"),
        ]
extra = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
print(" | ".join(["kernel"] + [c[1] for c in cols]))
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    import re
    m = re.search(r"(\w+_kernel)<([^>]*)>", name)
    short = (m.group(1) + "<" + m.group(2) + ">") if m else name[:40]
    vals = []
    for k, _ in cols:
        if k in hdr:
            i = hdr.index(k)
            v = r[i]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.4g}" + (units[i][0] if units[i] in ("Gbyte", "Mbyte", "Kbyte") else "")
            except ValueError:
                pass
            vals.append(v)
        else:
            vals.append("-")
    print(" | ".join([short] + vals))
    if "--stalls" in sys.argv:
        st = sorted(((float(r[hdr.index(h)] or 0), h.split("issue_stalled_")[1].split("_per_issue")[0]) for h in extra), reverse=True)[:6]
        print("      stalls: " + ", ".join(f"{n}={v:.2f}" for v, n in st))
