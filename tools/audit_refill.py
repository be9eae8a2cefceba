"""Static audit of the refill hazard (tma.cuh, staging_reads_done): in the SASS of every kernel that refills a staging
buffer by bulk copy (UBLKCP.S.G) inside its main loop, report which instruction class precedes each refill -- a store
(STS: the staged values were consumed), arithmetic (the loads' scoreboards were waited on), or a bare load (LDS: the
window the fence has to cover).

usage: python tools/audit_refill.py femtc_b200/csrc/_obj/tiled_kernels.o femtc_b200/csrc/_obj/column_kernels.o
"""
import re, subprocess, sys

CLASSES = (("LDS", "LDS"), ("STS", "STS"), ("UBLKCP.S.G", "REFILL"), ("UBLKCP.G.S", "STORE"), ("MEMBAR", "MEMBAR"),
           ("SYNCS.PHASECHK", "WAIT"), ("SYNCS.ARRIVE", "ARRIVE"), ("BAR.", "BAR"))
FP = re.compile(r"^(D|F|H)(FMA|MUL|ADD)")

for obj in sys.argv[1:]:
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    name, seq = None, []

    def report():
        if name is None:
            return
        refills = [i for i, c in enumerate(seq) if c == "REFILL"]
        if not refills:
            return
        after = {"STS": 0, "FP": 0, "LDS": 0, "WAIT": 0, "BAR": 0, "other": 0}
        fenced_lds = 0
        for i in refills:
            j = i - 1
            fence = False
            while j >= 0 and seq[j] in ("ARRIVE", "MEMBAR", "REFILL"):
                fence |= seq[j] == "MEMBAR"
                j -= 1
            prev = seq[j] if j >= 0 else "other"
            after[prev if prev in after else "other"] += 1
            fenced_lds += prev == "LDS" and fence
        short = re.sub(r"_GLOBAL__N__\w+?_cu_\w{8}", "", name)
        print(f"{short[:110]:110s} refills {len(refills):3d}  preceded by " +
              " ".join(f"{k}:{v}" for k, v in after.items() if v) + (f"  (LDS cases fenced: {fenced_lds})" if after["LDS"] else ""))

    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            report()
            name, seq = m.group(1), []
            continue
        m = re.search(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        for pre, cls in CLASSES:
            if op.startswith(pre):
                seq.append(cls)
                break
        else:
            if FP.match(op):
                seq.append("FP")
    report()
