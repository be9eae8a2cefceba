"""Device-resident throughput of the batched rectangular contraction (tmm.hpp apply_contractions).

    python tools/tmm_bench.py [--batch 1000000] [--reps 20] [--dtypes f64,f32] [--kernel 0]
One line per (Na, Nb): tensors/s, algorithmic GB/s = (Nb^3 + Na^3) sizeof(T) per tensor, fraction of the measured HBM peak.
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import femtc_b200 as femtc  # noqa: E402
from femtc_b200.distributed import fill_uniform  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1_000_000)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--dtypes", default="f64,f32")
    ap.add_argument("--kernel", type=int, default=0, help="1 = force the generic kernel")
    ap.add_argument("--shapes", default="3:2,4:3,5:4,6:4,4:4,6:5,7:6,8:6,6:6,8:8,10:8")
    a = ap.parse_args()
    p = ROOT / "MEASURED_PEAKS.json"
    peak = json.loads(p.read_text())["hbm_gbs"] if p.exists() else 6650.0
    femtc.set_option("kernel", a.kernel)
    rng = np.random.default_rng(5)
    for dt in a.dtypes.split(","):
        tdt, npdt, sz = (torch.float64, np.float64, 8) if dt == "f64" else (torch.float32, np.float32, 4)
        for shp in a.shapes.split(","):
            Na, Nb = (int(v) for v in shp.split(":"))
            batch = max(a.batch, int(1.5e9 / ((Nb ** 3 + Na ** 3) * sz)))   # small shapes: time the device, not the launch
            I = [rng.random(Na * Nb).astype(npdt) for _ in range(3)]
            B = fill_uniform(torch.empty(batch * Nb ** 3, dtype=tdt, device="cuda"), 1, 1)
            A = torch.empty(batch * Na ** 3, dtype=tdt, device="cuda")
            t0 = time.perf_counter()                       # at least 50 ms of warm-up: the SM clock has to ramp
            while time.perf_counter() - t0 < 0.05:
                for _ in range(5):
                    femtc.apply_contractions(*I, B, A, Na, Nb, True, batch=batch)
                torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.reps):
                femtc.apply_contractions(*I, B, A, Na, Nb, True, batch=batch)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.reps
            gbs = batch * (Nb ** 3 + Na ** 3) * sz / ms / 1e6
            print(f"tmm {dt} Nb={Nb} -> Na={Na} batch={batch:8d} {ms:8.3f} ms  {batch / ms / 1e6:7.2f} Gtensor/s  "
                  f"{gbs:8.1f} GB/s  {100 * gbs / peak:5.1f}% of measured HBM", flush=True)
            del A, B
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
