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


ORDERS = ("ijk", "ikj", "jik", "jki", "kij", "kji")   # linalg.hpp:11-19
TILES = ((0, 0), (4, 8), (8, 16))                        # (Mb, Nb) pairs instantiated by ref_shim_tune.cpp
TUNED_DEGREES = (3, 4)


def _tune_lib():
    """oracle/_ref/libfemtc_ref_tune_*.so (the reference's six loop orders x three tile pairs at the headline degrees),
    or None where it did not travel."""
    import ctypes as C
    from pathlib import Path

    from oracle.pyoracle import _cpu_flags
    flags = _cpu_flags()
    v4 = {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= flags
    path = Path(__file__).resolve().parent / "_ref" / f"libfemtc_ref_tune_{'v4' if v4 else 'v3'}.so"
    if not path.exists():
        return None
    lib = C.CDLL(str(path))
    for op in ("mass", "stiffness"):
        for s in ("f64", "f32"):
            getattr(lib, f"femtc_ref_tune_{op}_{s}").argtypes = [C.c_void_p] * 4 + [C.c_size_t, C.c_int, C.c_int, C.c_int]
    return lib


def _tuned_call(lib, op, dtype, phi, U, W, geom, n, degree, order, tile):
    suf = "f64" if np.dtype(dtype) == np.float64 else "f32"
    return getattr(lib, f"femtc_ref_tune_{op}_{suf}")(phi.ctypes.data, U.ctypes.data, W.ctypes.data, geom.ctypes.data,
                                                      n, degree, order, tile)


def tune(op: str, degree: int, dtype, ndofs: int = 1_000_000, reps: int = 10, keep: int = 5):
    """The reference's own tuning protocol on ONE core (run.sh:9-23, bench_mass.cpp:95-106: 10 repetitions, the last 5
    timed) over six loop orders x three tile pairs at Ndofs = 10^6.  Returns dict(best=(order, tile), rate GDoF/s,
    table={"ikj 4x8": rate, ...}) or None when the tuning library or the degree is not available."""
    lib = _tune_lib()
    if lib is None or degree not in TUNED_DEGREES:
        return None
    from oracle import Oracle, splitmix
    o = Oracle()
    nd = (degree + 1) ** 3
    n = max(1, ndofs // nd)
    phi, U, geom = splitmix(o, op, degree, n, np.dtype(dtype).name, seed=1)
    W = np.zeros(n * nd, dtype=dtype)
    table, best = {}, None
    for oi, oname in enumerate(ORDERS):
        for ti, (mb, nb) in enumerate(TILES):
            ts = []
            for r in range(reps):
                t0 = time.perf_counter()
                rc = _tuned_call(lib, op, dtype, phi, U, W, geom, n, degree, oi, ti)
                ts.append(time.perf_counter() - t0)
                assert rc == 0
                if r == 1 and ts[-1] * reps > 3.0:   # a slow triple (kij: 6x slower) is not worth ten repetitions
                    break
            rate = n * nd / (sum(ts[-keep:]) / len(ts[-keep:])) / 1e9
            table[f"{oname} {mb}x{nb}"] = rate
            if best is None or rate > best[2]:
                best = (oi, ti, rate)
    return {"best": (best[0], best[1]), "rate": best[2], "table": table,
            "best_name": f"{ORDERS[best[0]]} Mb={TILES[best[1]][0]} Nb={TILES[best[1]][1]}"}


def _worker(idx, nworkers, op, degree, dtype, cells, reps, warm, barrier, out_q, want_output, triple=None):
    from oracle import Oracle, Reference, have_reference, splitmix

    o = Oracle()
    nd = (degree + 1) ** 3
    c0, c1 = (idx * cells) // nworkers, ((idx + 1) * cells) // nworkers
    n = c1 - c0
    phi, U, geom = splitmix(o, op, degree, n, dtype, seed=1, cell0=c0)
    W = np.zeros(n * nd, dtype=dtype)
    tlib = _tune_lib() if (triple is not None and degree in TUNED_DEGREES) else None
    if have_reference() and tlib is not None:
        ref = Reference(fast=True)
        kind, isa = "reference", ref.isa
        def run(): return _tuned_call(tlib, op, dtype, phi, U, W, geom, n, degree, triple[0], triple[1])
    elif have_reference():
        triple = None
        ref = Reference(fast=True)
        kind, isa = "reference", ref.isa
        if op == "mass":
            def run(): return ref.mass(phi, U, geom, degree, 1, W)[0]
        else:
            def run(): return ref.stiffness(phi, U, geom, degree, 1, W)[0]
    else:  # the compiled reference did not travel to this host: time the C restatement
        triple = None
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
    out_q.put((idx, times, kind, isa, head, triple))


def run(op: str, degree: int, dtype, cells: int, workers: int, reps: int, warm: int = 1, want_output: int = 0,
        triple=None):
    """Returns dict(rate GDoF/s, sec per pass, kind, desc, head=W of the first `want_output` cells).
    `triple` = (order index, tile index) from tune(): the reference's knobs; None = ikj, Mb = Nb = 0."""
    ctx = mp.get_context("spawn")
    barrier = ctx.Barrier(workers)
    q = ctx.Queue()
    dtype = np.dtype(dtype).name
    procs = [ctx.Process(target=_worker, args=(i, workers, op, degree, dtype, cells, reps, warm, barrier, q, want_output,
                                               triple))
             for i in range(workers)]
    for p in procs:
        p.start()
    import queue
    results, deadline = [], time.time() + 1800
    while len(results) < len(procs):
        try:
            results.append(q.get(timeout=2))
        except queue.Empty:
            dead = [p.exitcode for p in procs if p.exitcode not in (None, 0)]
            if dead or time.time() > deadline:
                for p in procs:
                    if p.is_alive():
                        p.terminate()   # the exact processes started above
                raise RuntimeError(f"cpu_bench worker failed (exit codes {dead})" if dead else "cpu_bench timed out")
    for p in procs:
        p.join(timeout=60)
    results.sort(key=lambda r: r[0])
    per_rep = np.max(np.array([r[1] for r in results]), axis=0)  # slowest worker per repetition
    sec = float(per_rep.mean())
    nd = (degree + 1) ** 3
    gw = 6 if op == "stiffness" else 1
    kind, isa, used = results[0][2], results[0][3], results[0][5]
    nbytes = (2 + gw) * cells * nd * np.dtype(dtype).itemsize
    knobs = f"loop order {ORDERS[used[0]]} Mb={TILES[used[1]][0]} Nb={TILES[used[1]][1]}" if used else "loop order ikj Mb=Nb=0"
    desc = (f"{cells} cells x {reps} passes split over {workers} processes (private arrays, barrier per pass), "
            f"{knobs}, {isa} -Ofast, {nbytes / 1e9:.2f} GB working set")
    return {"rate": cells * nd / sec / 1e9, "sec": sec, "kind": kind, "desc": desc, "head": results[0][4],
            "cores": workers}
