"""Side-by-side table of a tools/ab.sh output: % of measured HBM per configuration and library."""
import re, sys
sec, order, cur = {}, [], None
for l in open(sys.argv[1]):
    if l.startswith("=="):
        cur = l.strip("= \n")
        if cur not in order:
            order.append(cur)
        continue
    m = re.match(r"(\w+)\s+(f\d+) P=(\d) mode=(\d) variant=(\d+).*?([\d.]+)% of", l)
    if m:
        sec.setdefault(m.groups()[:5], {})[cur] = float(m.group(6))
print("config".ljust(28) + "".join(o[-14:].rjust(15) for o in order))
for k, v in sec.items():
    print(" ".join(k).ljust(28) + "".join((f"{v[o]:.1f}" if o in v else "-").rjust(15) for o in order))
