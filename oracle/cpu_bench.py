"""TEST / MEASUREMENT INFRASTRUCTURE ONLY -- times the reference's own CPU kernels.

N worker processes, each with private arrays for its own contiguous cell range, a barrier per
repetition, throughput = total DoFs / slowest worker -- the semantics of `mpirun -n N ./bench_*`
in the reference (bench_mass.cpp:64,92,101,135-137), except that here the N ranks share one
problem of `cells` cells instead of each running all of it.  Processes rather than threads:
short ctypes calls from Python threads serialise on GIL hand-offs.

Used only by bench.py (cpu_baseline leg and --impl reference).
"""
from __future__ import annotations

import multiprocessing as mp
import time

import numpy as np


def _worker(idx, nworkers, op, degree, dtype, cells, reps, warm, barrier, out_q, want_output):
    from oracle import Oracle, Reference, have_reference, splitmix

    o = Oracle()
    nd = (degree + 1) ** 3
    c0, c1 = (idx * cells) // nworkers, ((idx + 1) * cells) // nworkers
    n = c1 - c0
    phi, U, geom = splitmix(o, op, degree, n, dtype, seed=1, cell0=c0)
    W = np.zeros(n * nd, dtype=dtype)
    if have_reference():
        ref = Reference(fast=True)
        kind, isa = "reference", ref.isa
        if op == "mass":
            def run(): return ref.mass(phi, U, geom, degree, 1, W)[0]
        else:
            def run(): return ref.stiffness(phi, U, geom, degree, 1, W)[0]
    else:  # the compiled reference did not travel to this host: time the C restatement
        kind, isa = "port", "oracle.c -O2"
        if op == "mass":
            def run():
                rc, out = o.mass(phi, U, geom, degree, 1)
                W[:] = out
                return rc
        else:
            def run():
                rc, out = o.stiffness(phi, U, geom, degree, 1)
                W[:] = out
                return rc
    for _ in range(warm):
        assert run() == 0
    times = []
    for _ in range(reps):
        barrier.wait()
        t0 = time.perf_counter()
        assert run() == 0
        times.append(time.perf_counter() - t0)
    head = W[: min(n, want_output) * nd].copy() if (idx == 0 and want_output) else None
    out_q.put((idx, times, kind, isa, head))


def run(op: str, degree: int, dtype, cells: int, workers: int, reps: int, warm: int = 1, want_output: int = 0):
    """Returns dict(rate GDoF/s, sec per pass, kind, desc, head=W of the first `want_output` cells)."""
    ctx = mp.get_context("spawn")
    barrier = ctx.Barrier(workers)
    q = ctx.Queue()
    dtype = np.dtype(dtype).name
    procs = [ctx.Process(target=_worker, args=(i, workers, op, degree, dtype, cells, reps, warm, barrier, q, want_output))
             for i in range(workers)]
    for p in procs:
        p.start()
    results = [q.get(timeout=1800) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    results.sort(key=lambda r: r[0])
    per_rep = np.max(np.array([r[1] for r in results]), axis=0)  # slowest worker per repetition
    sec = float(per_rep.mean())
    nd = (degree + 1) ** 3
    gw = 6 if op == "stiffness" else 1
    kind, isa = results[0][2], results[0][3]
    nbytes = (2 + gw) * cells * nd * np.dtype(dtype).itemsize
    desc = (f"{cells} cells x {reps} passes split over {workers} processes (private arrays, barrier per pass), "
            f"loop order ikj Mb=Nb=0, {isa} -Ofast, {nbytes / 1e9:.2f} GB working set")
    return {"rate": cells * nd / sec / 1e9, "sec": sec, "kind": kind, "desc": desc, "head": results[0][4],
            "cores": workers}
