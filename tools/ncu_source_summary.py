"""Per-opcode summary of an ncu report's source page (read here, no GPU): share of executed instructions and of stall
samples, shared-memory wavefronts (total / ideal / excessive) per opcode, and the most-sampled instructions.

    python tools/ncu_source_summary.py gpurun_out/x.ncu-rep [--top 12]
"""
import csv
import subprocess
import sys
from collections import Counter


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 10
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True, check=True).stdout
    kern, cur = [], None
    for r in csv.reader(out.splitlines()):
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            kern.append(cur)
        elif r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and len(r) > 10:
            cur["rows"].append(r)
    seen = set()
    for k in kern:
        if k["name"] in seen:
            continue
        seen.add(k["name"])
        h = k["hdr"]
        iS, iX, iN = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
        iW, iI, iE = h.index("L1 Wavefronts Shared"), h.index("L1 Wavefronts Shared Ideal"), h.index("L1 Wavefronts Shared Excessive")
        stall = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
        tot = sum(int(r[iX]) for r in k["rows"])
        ts = sum(int(r[iN]) for r in k["rows"])
        wf = sum(int(r[iW]) for r in k["rows"])
        ex = sum(int(r[iE]) for r in k["rows"])
        print(f"== {k['name'][:110]}")
        print(f"   executed warp instructions {tot}, stall samples {ts}, shared-memory wavefronts {wf} (excessive {ex} = {100 * ex / max(1, wf):.0f} %)")
        ops, smp, w, wi, we = Counter(), Counter(), Counter(), Counter(), Counter()
        for r in k["rows"]:
            src = r[iS].split()
            op = (src[1] if src and src[0].startswith("@") else (src[0] if src else "?"))
            op = ".".join(op.split(".")[:2]) if op.startswith(("LDS", "STS", "LDGSTS")) else op.split(".")[0]
            ops[op] += int(r[iX]); smp[op] += int(r[iN]); w[op] += int(r[iW]); wi[op] += int(r[iI]); we[op] += int(r[iE])
        print("   opcode        instr %  samples %   smem wavefronts      ideal  excessive  wf/instr")
        for op, n in ops.most_common(top):
            print(f"   {op:12s} {100 * n / tot:7.1f} {100 * smp[op] / max(1, ts):9.1f} {w[op]:16d} {wi[op]:10d} {we[op]:10d} {w[op] / max(1, n):9.2f}")
        st = Counter()
        for r in k["rows"]:
            for i in stall:
                st[h[i]] += int(r[i])
        print("   stall samples: " + ", ".join(f"{a[6:]} {100 * b / max(1, ts):.0f} %" for a, b in st.most_common(6)))


if __name__ == "__main__":
    main()
