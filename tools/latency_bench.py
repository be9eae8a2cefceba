"""Small-batch latency: the reference's default problem (run.sh:9, Ndofs = 1e6) is 15 625 cells at degree 3, a ~64 MB
launch that is over in ~10 us, so host-side launch cost decides the rate.  Times one operator application
  (a) through the Python mirror, (b) through the C ABI with prepared arguments, (c) replayed from a CUDA graph that
  captured 20 calls of the C ABI (tables travel as kernel parameters, the ragged tail is one more launch: nothing in the
  call allocates or copies, so stream capture works),
and checks that the replayed graph reproduces the eager result bit for bit.

    python tools/latency_bench.py [--ndofs 1000000] [--degrees 1-8] [--dtype f64]
"""
import argparse
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import femtc_b200 as femtc  # noqa: E402
from femtc_b200 import _lib  # noqa: E402
from femtc_b200.distributed import fill_uniform  # noqa: E402


def time_us(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ndofs", type=int, default=1_000_000)
    ap.add_argument("--degrees", default="1,2,3,4,6,8")
    ap.add_argument("--dtype", default="f64")
    a = ap.parse_args()
    tdt = torch.float64 if a.dtype == "f64" else torch.float32
    size = 8 if a.dtype == "f64" else 4
    peak = 6451.8
    print(f"# Ndofs = {a.ndofs}, {a.dtype}; us per application and % of {peak} GB/s (algorithmic bytes)")
    print("# op P cells | python wrapper | C ABI | CUDA graph replay (20 calls per graph) | graph == eager")
    for op in ("mass", "stiffness"):
        gw = 6 if op == "stiffness" else 1
        for P in [int(x) for x in a.degrees.split(",")]:
            nd = (P + 1) ** 3
            cells = a.ndofs // nd
            phi = np.random.default_rng(0).random(2 * (P + 1) ** 2).astype(np.float64 if a.dtype == "f64" else np.float32)
            U = fill_uniform(torch.empty(cells * nd, dtype=tdt, device="cuda"), 1, 1)
            G = fill_uniform(torch.empty(cells * nd * gw, dtype=tdt, device="cuda"), 1, 2)
            W = torch.empty_like(U)
            fn = getattr(_lib.lib, f"femtc_{op}_apply_device_{a.dtype}")

            def wrapper():
                if op == "mass":
                    femtc.mass_operator(phi, U, W, G, cells, P)
                else:
                    femtc.stiffness_operator(phi, U, W, G, cells, P)

            def make_abi(stream):
                args = [phi.ctypes.data, phi.size, U.data_ptr(), W.data_ptr(), G.data_ptr(), cells, P]
                args += [0, stream] if op == "stiffness" else [stream]
                return lambda: _lib.check(fn(*args))
            abi = make_abi(torch.cuda.current_stream().cuda_stream)
            t_wrap = time_us(wrapper, 200)
            t_abi = time_us(abi, 200)
            eager = W.clone()
            side = torch.cuda.Stream()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                cap = make_abi(side.cuda_stream)
                cap()
                torch.cuda.synchronize()
                with torch.cuda.graph(graph, stream=side):
                    for _ in range(20):
                        cap()
            W.zero_()
            graph.replay()
            torch.cuda.synchronize()
            same = bool(torch.equal(W, eager))
            t_graph = time_us(graph.replay, 20) / 20
            gb = cells * nd * (2 + gw) * size / 1e9
            pct = lambda us: 100 * gb / (us * 1e-6) / peak
            print(f"{op:9s} P={P} {cells:7d} | {t_wrap:7.2f} us {pct(t_wrap):5.1f}% | {t_abi:7.2f} us {pct(t_abi):5.1f}% | "
                  f"{t_graph:7.2f} us {pct(t_graph):5.1f}% | {same}", flush=True)
            assert same


if __name__ == "__main__":
    main()
