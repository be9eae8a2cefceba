"""Does the relative placement of U, detJ/G and W in HBM matter?  Carves the three fields out of one
allocation with a varying gap between them and times the operator (development probe)."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import femtc_b200 as femtc  # noqa: E402
from femtc_b200.distributed import fill_uniform  # noqa: E402

def run(op, P, dt, cells, gaps, reps=20):
    tdt = torch.float64 if dt == "f64" else torch.float32
    sz = 8 if dt == "f64" else 4
    nd = (P + 1) ** 3
    gw = 6 if op == "stiffness" else 1
    rng = np.random.default_rng(1)
    phi = rng.random(2 * (P + 1) ** 2).astype(np.float64 if dt == "f64" else np.float32)
    nU, nG = cells * nd, cells * nd * gw
    big = torch.empty((nU * 2 + nG) + 3 * (max(gaps) // sz + 64), dtype=tdt, device="cuda")
    fill_uniform(big, 1, 1)
    for gap in gaps:
        g = gap // sz
        U = big[0:nU]
        G = big[nU + g: nU + g + nG]
        W = big[nU + g + nG + g: nU + g + nG + g + nU]
        def go():
            if op == "mass":
                femtc.mass_operator(phi, U, W, G, cells, P)
            else:
                femtc.stiffness_operator(phi, U, W, G, cells, P, mode=0)
        for _ in range(3):
            go()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            go()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        gbs = cells * nd * (2 + gw) * sz / ms / 1e6
        print(f"{op} {dt} P={P} gap={gap:>9d} B  U%2MB={U.data_ptr() % (1<<21):>8d} G%2MB={G.data_ptr() % (1<<21):>8d} "
              f"W%2MB={W.data_ptr() % (1<<21):>8d}  {ms:.3f} ms {gbs:7.1f} GB/s", flush=True)

if __name__ == "__main__":
    gaps = [0, 16, 256, 1024, 4096, 16384, 65536, 1 << 20, (1 << 21), (1 << 21) + 4096, 3 << 20, 5 << 20, (1 << 24)]
    for op, P, dt in (("mass", 4, "f64"), ("mass", 3, "f64"), ("mass", 4, "f32"), ("stiffness", 3, "f64")):
        run(op, P, dt, 1_000_000, gaps)
