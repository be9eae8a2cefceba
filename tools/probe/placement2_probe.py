"""Placement probe 2 (development): U, detJ/G and W all on 2 MB boundaries (what separate large allocations get),
with the distance between them varied in 2 MB steps; and the same with a sub-2 MB skew added to detJ/G and W."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import femtc_b200 as femtc  # noqa: E402

MB2 = 2 << 20


def run(op, P, dt, cells, reps=200):
    tdt = torch.float64 if dt == "f64" else torch.float32
    sz = 8 if dt == "f64" else 4
    nd = (P + 1) ** 3
    gw = 6 if op == "stiffness" else 1
    rng = np.random.default_rng(1)
    phi = rng.random(2 * (P + 1) ** 2).astype(np.float64 if dt == "f64" else np.float32)
    nU, nG = cells * nd, cells * nd * gw
    up = lambda b: (b + MB2 - 1) // MB2 * MB2
    arena = torch.rand((up(nU * sz) * 2 + up(nG * sz) + 40 * MB2) // sz, dtype=tdt, device="cuda")
    base = (-arena.data_ptr()) % MB2 // sz
    for kj, kw, skew in [(0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 1, 0), (2, 3, 0), (3, 5, 0), (4, 4, 0), (7, 2, 0),
                         (0, 0, 4096), (0, 0, 65536), (0, 0, 256 * 1024), (0, 0, 1 << 20), (0, 0, 1536 * 1024)]:
        oU = base
        oG = oU + (up(nU * sz) + kj * MB2 + skew) // sz
        oW = oG + (up(nG * sz) + kw * MB2 + skew) // sz
        U, G, W = arena[oU:oU + nU], arena[oG:oG + nG], arena[oW:oW + nU]

        def go():
            if op == "mass":
                femtc.mass_operator(phi, U, W, G, cells, P)
            else:
                femtc.stiffness_operator(phi, U, W, G, cells, P, mode=0)
        for _ in range(100):
            go()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            go()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(f"{op} {dt} P={P} dJ=+{kj}x2MB dW=+{kw}x2MB skew={skew:8d} B  (G-U)%32MB={((oG - oU) * sz) % (32 << 20) >> 20:3d} MB "
              f"(W-U)%32MB={((oW - oU) * sz) % (32 << 20) >> 20:3d} MB  {ms:7.4f} ms  {cells * nd * (2 + gw) * sz / ms / 1e6:7.1f} GB/s", flush=True)


if __name__ == "__main__":
    run("mass", 4, "f32", 1_000_000)
    run("stiffness", 5, "f64", 1_000_000, reps=50)
    run("mass", 5, "f64", 1_000_000, reps=100)
