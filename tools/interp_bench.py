"""Forward-interpolation micro-benchmark (restates the reference's dead main.cpp:64-94):
one W = (A x A x A) U per cell, n = P + 1 points per direction, two ways on the same device-resident fields:

  factorised  three axis contractions, 3 * 2 n^4 flops per cell: femtc_apply_contractions_device_* with Na = Nb = n
              (the batched TMA kernel of tmm_kernels.cu, tmm.hpp:91-126)
  naive       the dense n^3 x n^3 Kronecker operator applied to every cell, 2 n^6 flops per cell: one library GEMM
              (torch.matmul; TF32 off), as main.cpp's "naive" loop nest does on the CPU

    python tools/interp_bench.py [--cells 1000000] [--degrees 1-7] [--dtypes f64,f32] [--reps 10]
One line per (degree, dtype): GDoF/s of both, algorithmic GB/s (2 n^3 sizeof(T) per cell) and fraction of the measured
HBM peak for the factorised kernel, and the largest relative difference between the two results.  Parity of the
factorised operator against the CPU oracle is a GPU test (tests/test_gpu_parity.py::
test_forward_interpolation_operator_matches_oracle, degrees 1..8, both dtypes); the GEMM here is only the "naive" arm.
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


def interpolate(A1, U, W, n, cells, Tr=True):
    """The benchmark's operator: W[e] = (A x A x A) U[e] for `cells` cell tensors of n^3 values resident in HBM,
    through femtc_apply_contractions_device_* with I0 = I1 = I2 = A (n x n, flat; A(a,b) = A1[a*n+b] if Tr else A1[b*n+a])."""
    flat = np.ascontiguousarray(A1).ravel()
    femtc.apply_contractions(flat, flat, flat, U, W, n, n, bool(Tr), batch=cells)


def timed(fn, reps):
    # warm up for at least 50 ms of device time: a handful of sub-millisecond launches leaves the SM clock ramping
    import time
    t0 = time.perf_counter()
    while True:
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        if time.perf_counter() - t0 > 0.05:
            break
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=int, default=1_000_000)
    ap.add_argument("--degrees", default="1-7")
    ap.add_argument("--dtypes", default="f64,f32")
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    lo, hi = (int(v) for v in a.degrees.split("-")) if "-" in a.degrees else (int(a.degrees),) * 2
    p = ROOT / "MEASURED_PEAKS.json"
    peak = json.loads(p.read_text())["hbm_gbs"] if p.exists() else 6650.0
    torch.backends.cuda.matmul.allow_tf32 = False
    rng = np.random.default_rng(9)
    for dt in a.dtypes.split(","):
        tdt, npdt, sz = (torch.float64, np.float64, 8) if dt == "f64" else (torch.float32, np.float32, 4)
        for P in range(lo, hi + 1):
            n = P + 1
            nd = n ** 3
            cells = max(a.cells, int(1.5e9 / (2 * nd * sz)))   # small degrees: enough cells that the device, not the launch, is timed
            A1 = (rng.random((n, n)) - 0.5).astype(npdt)
            U = fill_uniform(torch.empty(cells * nd, dtype=tdt, device="cuda"), 1, 1)
            W = torch.empty_like(U)
            run = lambda: interpolate(A1, U, W, n, cells, True)  # noqa: E731
            ms_f = timed(run, max(a.reps, 20))
            # naive: W_cell = K U_cell with K = A (x) A (x) A; the storage order of A under Tr is settled by the data
            Wn = torch.empty_like(U)
            best = None
            for M in (A1, A1.T):
                K = torch.from_numpy(np.kron(np.kron(M, M), M).astype(npdt)).cuda()
                naive = lambda K=K: torch.matmul(U.view(cells, nd), K.T, out=Wn.view(cells, nd))  # noqa: E731
                naive()
                diff = ((Wn - W).abs().max() / W.abs().max()).item()
                if best is None or diff < best[0]:
                    best = (diff, naive)
            ms_n = timed(best[1], max(2, a.reps // 2))
            gbs = cells * 2 * nd * sz / ms_f / 1e6
            print(f"interp {dt} P={P} cells={cells:8d}  factorised {ms_f:8.3f} ms {cells * nd / ms_f / 1e6:8.2f} GDoF/s "
                  f"{gbs:7.1f} GB/s {100 * gbs / peak:5.1f}% of measured HBM |  naive GEMM {ms_n:8.3f} ms "
                  f"{cells * nd / ms_n / 1e6:8.2f} GDoF/s  ({ms_n / ms_f:5.1f}x slower)  max rel diff {best[0]:.2e}", flush=True)
            del U, W, Wn, K
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
