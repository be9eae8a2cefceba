"""Device-resident timing sweep (development tool; bench.py is the contract benchmark).

    python tools/quick_bench.py [--ops mass,stiffness] [--degrees 1-8] [--dtypes f64,f32]
                                [--modes 0,1] [--cells 1000000] [--reps 20] [--variants 0] [--kernel 0]
Prints one line per configuration: GDoF/s, algorithmic GB/s and fraction of the measured HBM peak.
"""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import femtc_b200 as femtc  # noqa: E402
from femtc_b200.distributed import fill_uniform  # noqa: E402


def peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    return json.loads(p.read_text())["hbm_gbs"] if p.exists() else 6650.0


def parse_range(s):
    out = []
    for part in s.split(","):
        if "-" in part:
            a, b = part.split("-")
            out += list(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ops", default="mass,stiffness")
    ap.add_argument("--degrees", default="1-8")
    ap.add_argument("--dtypes", default="f64,f32")
    ap.add_argument("--modes", default="0,1")
    ap.add_argument("--cells", type=int, default=1_000_000)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--variants", default="0")
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--max-gb", type=float, default=60.0)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--fixed", action="store_true", help="exactly --warmup + --reps launches per configuration (for ncu)")
    ap.add_argument("--tc", type=int, default=-1, help="tensor-core family: 0 off, 1 on, -1 library default")
    ap.add_argument("--configs", default="", help="explicit list op:dtype:P:mode[,...] instead of the cross product")
    a = ap.parse_args()
    if a.configs:
        for item in a.configs.split(","):
            op, dt, P, mode = item.split(":")
            sweep(a, [op], [dt], [int(P)], [int(mode)])
    else:
        sweep(a, a.ops.split(","), a.dtypes.split(","), parse_range(a.degrees), parse_range(a.modes))


def sweep(a, ops, dtypes, degrees, modes):
    peak = peak_gbs()
    femtc.set_option("kernel", a.kernel)
    femtc.set_option("tc", a.tc)
    rng = np.random.default_rng(1)
    for op in ops:
        for dt in dtypes:
            tdt = torch.float64 if dt == "f64" else torch.float32
            sz = 8 if dt == "f64" else 4
            for P in degrees:
                nd = (P + 1) ** 3
                gw = 6 if op == "stiffness" else 1
                cells = a.cells
                while cells * nd * (2 + gw) * sz > a.max_gb * 1e9:
                    cells //= 2
                phi = rng.random(2 * (P + 1) ** 2).astype(np.float64 if dt == "f64" else np.float32)
                U = fill_uniform(torch.empty(cells * nd, dtype=tdt, device="cuda"), 1, 1)
                G = fill_uniform(torch.empty(cells * nd * gw, dtype=tdt, device="cuda"), 1, 2)
                W = torch.empty_like(U)
                for mode in ([0] if op == "mass" else modes):
                    for variant in parse_range(a.variants):
                        femtc.set_option("variant", variant)

                        def run():
                            if op == "mass":
                                femtc.mass_operator(phi, U, W, G, cells, P)
                            else:
                                femtc.stiffness_operator(phi, U, W, G, cells, P, mode=mode)
                        # warm-up: at least `warmup` launches AND at least ~30 ms of them (a 20 us kernel launched three
                        # times leaves the clocks where the previous configuration's idle gap put them); timed: at least
                        # `reps` launches and ~20 ms
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        for _ in range(a.warmup):
                            run()
                        e1.record()
                        torch.cuda.synchronize()
                        per = max(e0.elapsed_time(e1) / a.warmup, 1e-3)
                        for _ in range(0 if a.fixed else min(2000, int(30.0 / per))):
                            run()
                        reps = a.reps if a.fixed else max(a.reps, min(2000, int(20.0 / per)))
                        torch.cuda.synchronize()
                        e0.record()
                        for _ in range(reps):
                            run()
                        e1.record()
                        torch.cuda.synchronize()
                        ms = e0.elapsed_time(e1) / reps
                        gdofs = cells * nd / ms / 1e6
                        gbs = gdofs * (2 + gw) * sz
                        print(f"{op:9s} {dt} P={P} mode={mode} variant={variant} cells={cells:8d} "
                              f"{ms:8.3f} ms  {gdofs:8.2f} GDoF/s  {gbs:8.1f} GB/s  {100 * gbs / peak:5.1f}% of measured HBM",
                              flush=True)
                del U, G, W
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
