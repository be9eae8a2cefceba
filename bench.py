#!/usr/bin/env python
"""Benchmark of the hot path: cell-batched stiffness / mass operator action on hexahedra.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

Default workload = BASELINE.json configs[1]: Laplace stiffness action, degree 3, fp64,
1 000 000 cells per GPU (64 M DoFs, 4.1 GB of fields), reference semantics (REF_EXACT).
A step is one operator application over the whole batch.  One JSON line is printed by rank 0:

  value     GDoF/s, inputs resident in HBM, CUDA events around the K steps, max over ranks,
            whole job (all GPUs).  With N > 1 the cells are partitioned (weak scaling: the
            per-GPU batch is fixed) and the only collective, the 4-double checksum reduction over
            NCCL, runs once at the end of the timed region.
  e2e       same metric through the host-pointer C ABI (femtc_stiffness_apply_f64): pinned HOST
            buffers in, HOST buffer out, H2D/D2H copies inside the timed region.
  roofline  algorithmic HBM bytes (8 values per DoF for stiffness, 3 for mass) per launch over the
            measured kernel time, against the measured copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline   the UNMODIFIED reference (oracle/_ref, -Ofast) on the host cores, rank 0, N = 1.

  sweep     (N = 1) BASELINE config 3 under this run's clock: mass / stiffness REF_EXACT / stiffness FULL x
            P = 1..8 x fp32 / fp64 at 1 M cells = 48 entries, each with GDoF/s (min / median / max over five
            placements of U, geometry, W inside one arena), fraction of the measured HBM peak, the kernel family
            that served it, a clocks record for its own window and a spot check of three regenerated 40-cell
            ranges (first / middle / last) against the CPU oracle.
  strong    BASELINE config 5: stiffness P = 4 fp64, 64 M cells in total partitioned over the N GPUs (resident
            chunks), spot checks at chunk boundaries, clocks over the whole leg.

--impl reference times only that CPU reference, with every host thread, on the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "GDoF/s"
OPS = {"stiffness": 6, "mass": 1}  # geometry values per point


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--op", default="stiffness", choices=tuple(OPS))
    ap.add_argument("--degree", type=int, default=3)
    ap.add_argument("--dtype", default="f64", choices=("f64", "f32"))
    ap.add_argument("--mode", default="ref_exact", choices=("ref_exact", "full"))
    ap.add_argument("--cells", type=int, default=1_000_000, help="cells per GPU")
    ap.add_argument("--total-cells", type=int, default=0,
                    help="strong scaling (BASELINE config 5): this many cells in total, partitioned over the GPUs; "
                         "a rank whose share exceeds --resident-gb sweeps it in resident chunks")
    ap.add_argument("--resident-gb", type=float, default=72.0, help="per-GPU field bytes held at a time with --total-cells")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true", help="skip the 48-entry degree sweep (N = 1 only)")
    ap.add_argument("--detail-file", default="", help="where the full record behind the compact stdout line is written "
                                                      "(default gpurun_out/bench_detail_n<N>.json)")
    ap.add_argument("--sweep-cells", type=int, default=1_000_000)
    ap.add_argument("--sweep-filter", default="", help="development: only entries whose name contains this")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (BASELINE config 5)")
    ap.add_argument("--strong-cells", type=int, default=64_000_000)
    ap.add_argument("--strong-degree", type=int, default=4)
    ap.add_argument("--strong-steps", type=int, default=3)
    return ap.parse_args()


def workload_name(a):
    geom = "G (6 per point)" if a.op == "stiffness" else "detJ"
    return (f"{'bench_stiffness: Laplace stiffness' if a.op == 'stiffness' else 'bench_mass: mass'} action, "
            f"degree-{a.degree} hexahedra, {'fp64' if a.dtype == 'f64' else 'fp32'}, {a.cells} cells per GPU, "
            f"U + {geom} -> W, synthetic uniform [0,1) inputs")


def config_dict(a, world: int):
    """The workload both arms are quoted on (the reference arm times a bounded sample of it and says so in
    cpu_baseline.sample); identical in both arms' lines."""
    nd, gw, size = (a.degree + 1) ** 3, OPS[a.op], 8 if a.dtype == "f64" else 4
    return {"workload": workload_name(a), "op": a.op, "degree": a.degree, "mode": a.mode, "cells_per_gpu": a.cells,
            "dofs_per_step": world * a.cells * nd,
            "parallelism": f"cells partitioned over {world} GPU(s), no halo; one 4-double checksum all-gather (NCCL) "
                           "at the end of the timed region",
            "placement": "U | geometry | W back to back in one device allocation",
            "l2": f"no flush needed: each step streams {a.cells * nd * (2 + gw) * size / 1e9:.2f} GB per GPU, 126 MB L2"}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, device copy)"
    return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def recorded_traffic(a):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture."""
    p = ROOT / "profiles" / "traffic.json"
    if not p.exists():
        return None
    key = f"{a.op}_P{a.degree}_{a.dtype}_{a.mode}_{a.cells}"
    return json.loads(p.read_text()).get(key)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self, t0: float | None = None, t1: float | None = None):
        """all samples, or those received in the wall-clock window [t0, t1] (+ one sampling period)"""
        sm, mx, pw, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for t, row in list(self.rows):
            if t0 is not None and not (t0 <= t <= t1 + 0.06):
                continue
            f = [x.strip() for x in row.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        busy = [s for s, p in zip(sm, pw) if p >= 0.5 * max(pw)] or sm
        out = {"sm_mhz": float(np.median(busy)), "sm_max_mhz": max(mx), "power_w_max": max(pw),
               "reasons": sorted(reasons), "samples": len(sm)}
        if t0 is None:
            out["window"] = "timed region plus ~1 s of the same step repeated right after it"
        return out


# ------------------------------------------------------------------------------------------------------
# CPU reference (oracle/_ref = the unmodified reference headers, -Ofast, widest ISA the host supports)
# ------------------------------------------------------------------------------------------------------
def cpu_reference_rate(a, cells: int, workers: int, reps: int, warm: int = 1, want_output: int = 0, triple=None):
    """GDoF/s of the reference's own CPU kernel (oracle/_ref) with `workers` processes on disjoint
    cell ranges of the same synthetic inputs; see oracle/cpu_bench.py."""
    from oracle import cpu_bench
    return cpu_bench.run(a.op, a.degree, np.float64 if a.dtype == "f64" else np.float32, cells, workers, reps,
                         warm, want_output, triple)


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_tuning(a):
    """the reference's knob sweep on ONE core at Ndofs = 10^6 (run.sh:9-23), six loop
    orders x three (Mb, Nb) pairs, best kept.  None where the tuning library does not cover this degree."""
    from oracle import cpu_bench
    t = cpu_bench.tune(a.op, a.degree, np.float64 if a.dtype == "f64" else np.float32)
    if t is None:
        return None, None
    return t["best"], {"one_core_best": t["best_name"], "one_core_best_gdofs": t["rate"],
                       "one_core_table_gdofs": {k: round(v, 5) for k, v in t["table"].items()},
                       "protocol": "1 core, Ndofs = 1e6, 10 repetitions, mean of the last 5 (bench_mass.cpp:95-106); "
                                   "orders ijk..kji x (Mb,Nb) in {(0,0),(4,8),(8,16)} (run.sh:13-23)"}


def run_reference(a, rank: int):
    if rank != 0:
        return
    workers = os.cpu_count() or 1
    triple, tuning = cpu_tuning(a)
    # bounded sample: shrink the batch until one pass takes about a second on this host
    cells = a.cells
    probe = cpu_reference_rate(a, min(cells, 16384 * workers), workers, reps=1, warm=0, triple=triple)
    est = cells / min(cells, 16384 * workers) * probe["sec"]
    while est > 1.5 and cells > 16384 * workers:
        cells //= 2
        est /= 2
    res = cpu_reference_rate(a, cells, workers, reps=a.steps, warm=max(1, min(a.warmup, 3)), triple=triple)
    rate, sec, kind, desc = res["rate"], res["sec"], res["kind"], res["desc"]
    threads = workers
    nd = (a.degree + 1) ** 3
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "GDoF/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic",
        "config": config_dict(a, a.gpus),
        "sample_cells_per_step": cells, "sample_dofs_per_step": cells * nd,
        "note": "a rate: the CPU reference is timed on a bounded sample of ONE GPU's batch whatever --gpus says "
                "(the repo arm at N GPUs processes N batches per step)",
        "cpu_baseline": {"value": rate, "unit": "GDoF/s", "cores": threads, "kind": kind, "sample": desc,
                         "cpu": cpu_model(), "tuning": tuning},
        "e2e": {"value": rate, "unit": "GDoF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------
def spot_check(femtc_mod, oracle_mod, W_dev, op, P, ndt, mode, phi, ranges, nd):
    """max |w - ref| / max |ref| over regenerated cell ranges: [(first global cell, cells, offset of that cell in W_dev)].
    The oracle (CPU restatement, pinned to the compiled reference) is the checker here, after the timing."""
    o = oracle_mod["o"]
    worst = 0.0
    for c0, n, off in ranges:
        _, u, g = oracle_mod["splitmix"](o, op, P, n, ndt, 1, cell0=c0)
        rc, ref = (o.mass(phi, u, g, P) if op == "mass" else o.stiffness(phi, u, g, P, mode=mode))
        assert rc == 0
        got = W_dev[off * nd:(off + n) * nd].cpu().numpy().astype(np.float64)
        ref = ref.astype(np.float64)
        worst = max(worst, float(np.abs(got - ref).max() / np.abs(ref).max()))
    return worst


def load_oracle():
    from oracle import Oracle, splitmix
    return {"o": Oracle(), "splitmix": splitmix}


def strong_leg(a, rank, local_rank, world, dev, *, op, degree, dtype, mode_name, total_cells, steps, warm, resident_gb):
    """BASELINE config 5: a fixed global batch partitioned over the GPUs, no halo.  A rank holds at most
    `resident_gb` of fields; a larger share is swept in chunks (inputs regenerated on the device between chunks,
    outside the timed spans).  Timed: the operator launches only (CUDA events per chunk, summed; max over ranks).
    The one collective is the 4-double checksum all-gather at the end.  Every chunk's first and last 40 cells are
    regenerated on the host and compared with the oracle.  Returns the result dict on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist

    import femtc_b200 as femtc
    from femtc_b200.distributed import combine_checksums, device_checksum, fill_uniform, partition

    P, nd, gw = degree, (degree + 1) ** 3, OPS[op]
    tdt = torch.float64 if dtype == "f64" else torch.float32
    ndt = np.float64 if dtype == "f64" else np.float32
    size = 8 if dtype == "f64" else 4
    mode = femtc.FULL if mode_name == "full" else femtc.REF_EXACT
    lo, hi = partition(total_cells, rank, world)
    bytes_per_cell = nd * (2 + gw) * size
    chunk = max(32, int(resident_gb * 1e9 / bytes_per_cell) // 32 * 32)
    chunk = min(chunk, hi - lo)
    n2 = (P + 1) ** 2
    phi = fill_uniform(torch.empty(2 * n2, dtype=tdt, device=dev), 1, 0, 0).cpu().numpy().astype(ndt)

    def up(n):
        return (n + 31) // 32 * 32
    nU, nG = chunk * nd, chunk * nd * gw
    arena = torch.empty(up(nU) + up(nG) + up(nU), dtype=tdt, device=dev)
    U, G, W = arena[:nU], arena[up(nU):up(nU) + nG], arena[up(nU) + up(nG):up(nU) + up(nG) + nU]
    orc = load_oracle()

    def apply(n):
        rc = (femtc.mass_operator(phi, U, W, G, n, P) if op == "mass"
              else femtc.stiffness_operator(phi, U, W, G, n, P, mode=mode))
        assert rc == 0

    def sweep(reps_per_chunk, timed):
        """one pass over the rank's cells; returns (kernel ms, checksum partials, worst spot-check error)"""
        ms, parts, worst = 0.0, [], 0.0
        for c0 in range(lo, hi, chunk):
            n = min(chunk, hi - c0)
            fill_uniform(U[:n * nd], 1, 1, c0 * nd)
            fill_uniform(G[:n * nd * gw], 1, 2, c0 * nd * gw)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps_per_chunk):
                apply(n)
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1) / reps_per_chunk
            if timed:
                parts.append(device_checksum(W[:n * nd]).cpu().numpy())
                k = min(40, n)
                worst = max(worst, spot_check(femtc, orc, W, op, P, ndt, mode, phi,
                                              [(c0, k, 0), (c0 + n - k, k, n - k)], nd))
        return ms, parts, worst

    sweep(warm, False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = femtc.get_option("launches")
    with ClockSampler(local_rank) as clocks:
        ms, parts, worst = sweep(steps, True)
        # keep the same load up for a moment so that a short leg still collects clock samples
        t_end = time.perf_counter() + 1.0
        while time.perf_counter() < t_end:
            apply(min(chunk, hi - lo))
            torch.cuda.synchronize()
    launches = femtc.get_option("launches") - launches0
    loc = np.array([sum(p[0] for p in parts), sum(p[1] for p in parts), max(p[2] for p in parts), sum(p[3] for p in parts)])
    loc_t = torch.tensor(loc, dtype=torch.float64, device=dev)
    glob = combine_checksums(loc_t) if world > 1 else loc_t
    t = torch.tensor([ms, worst], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step, worst = t[0].item(), t[1].item()
    del arena, U, G, W
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = peaks()
    dofs = total_cells * nd
    achieved = (hi - lo) * nd * (2 + gw) * size / (ms_step * 1e-3) / 1e9
    cs = [float(x) for x in glob.cpu().numpy()]
    tol = 1e-12 if dtype == "f64" else 1e-5
    return {
        "value": dofs / ms_step / 1e6, "unit": "GDoF/s", "scaling": "strong", "n_gpus": world, "steps": steps,
        "warmup": warm, "ms_per_step": ms_step, "dtype": dtype,
        "config": {"workload": f"{op} action degree {P}, {dtype}, {total_cells} cells in total partitioned over "
                               f"{world} B200" + (" with NCCL checksum reduction" if world > 1 else ""),
                   "op": op, "degree": P, "mode": mode_name, "total_cells": total_cells, "cells_this_rank": hi - lo,
                   "resident_chunk_cells": chunk, "chunks_per_rank": -(-(hi - lo) // chunk), "dofs_per_step": dofs,
                   "timed": "operator launches only, CUDA events per resident chunk summed, max over ranks; inputs are "
                            "regenerated on the device between chunks outside the timed spans",
                   "l2": f"no flush needed: every chunk streams {chunk * bytes_per_cell / 1e9:.1f} GB"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_dof": (2 + gw) * size},
        "gpu_launches": int(launches), "clocks": clocks.summary(),
        "parity": {"max_rel_err": worst, "tol": tol, "ok": bool(worst <= tol),
                   "checked": "first and last 40 cells of every resident chunk on every rank vs the CPU oracle"},
        "checksum": {"sum": cs[0], "sum_sq": cs[1], "max_abs": cs[2], "n": cs[3]},
    }


def run_strong(a, rank: int, local_rank: int, world: int):
    """`--total-cells N`: only the strong-scaling leg, printed as the line."""
    import torch
    import torch.distributed as dist

    from femtc_b200 import _lib

    torch.cuda.set_device(local_rank)
    _lib.check(_lib.lib.femtc_set_device(local_rank))
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    res = strong_leg(a, rank, local_rank, world, dev, op=a.op, degree=a.degree, dtype=a.dtype, mode_name=a.mode,
                     total_cells=a.total_cells, steps=a.steps, warm=max(1, min(a.warmup, 3)), resident_gb=a.resident_gb)
    if rank == 0:
        line = {"metric": METRIC, "value": res["value"], "unit": "GDoF/s", "n_gpus": world, "steps": a.steps,
                "warmup": res["warmup"], "ms_per_step": res["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": a.dtype, "data": "synthetic", "e2e": None}
        line.update({k: res[k] for k in ("config", "roofline", "gpu_launches", "clocks", "parity", "checksum")})
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------
# BASELINE config 3: the degree sweep, under this run's clock
# ------------------------------------------------------------------------------------------------------
FAMILY = {0: "generic", 1: "slab", 2: "cell", 3: "column", 4: "tensor-core"}
MIB = 1 << 20
# five placements of the three fields inside one arena: (first byte, gap between fields, order)
PLACEMENTS = ((0, 0, "UGW"), (34 * MIB + 768, 6 * MIB + 1280, "UGW"), (10 * MIB, 0, "WUG"),
              (0, 14 * MIB + 256, "GWU"), (82 * MIB, 13 * 65536, "UGW"))


def place(i: int, bU: int, bG: int):
    base, gap, order = PLACEMENTS[i]
    size = {"U": bU, "G": bG, "W": bU}
    off, cur = {}, base
    for f in order:
        off[f] = cur
        cur = (cur + size[f] + 255) // 256 * 256 + gap
    return off, cur


def run_sweep(a, dev, local_rank):
    """48 entries: (mass, stiffness REF_EXACT, stiffness FULL) x P = 1..8 x (fp64, fp32), a.sweep_cells cells each."""
    import torch

    import femtc_b200 as femtc
    from femtc_b200 import _lib
    from femtc_b200.distributed import fill_uniform

    cells = a.sweep_cells
    peak, _ = peaks()
    orc = load_oracle()
    stream = torch.cuda.current_stream().cuda_stream
    todo = [(op, mode_name, dtype, P, f"{op}{'_full' if mode_name == 'full' else ''}_P{P}_{dtype}")
            for op, mode_name in (("mass", "ref_exact"), ("stiffness", "ref_exact"), ("stiffness", "full"))
            for dtype in ("f64", "f32") for P in range(1, 9)]
    todo = [t for t in todo if a.sweep_filter in t[4]]
    need = 0
    for op, _, dtype, P, _ in todo:
        b = cells * (P + 1) ** 3 * (8 if dtype == "f64" else 4)
        need = max(need, max(place(i, b, b * OPS[op])[1] for i in range(len(PLACEMENTS))))
    arena = torch.empty(need, dtype=torch.uint8, device=dev)
    entries = []
    with ClockSampler(local_rank) as clocks:
        for op, mode_name, dtype, P, name in todo:
            nd, gw, n2 = (P + 1) ** 3, OPS[op], (P + 1) ** 2
            tdt = torch.float64 if dtype == "f64" else torch.float32
            ndt = np.float64 if dtype == "f64" else np.float32
            size = 8 if dtype == "f64" else 4
            mode = femtc.FULL if mode_name == "full" else femtc.REF_EXACT
            phi = fill_uniform(torch.empty(2 * n2, dtype=tdt, device=dev), 1, 0, 0).cpu().numpy().astype(ndt)
            bU, bG = cells * nd * size, cells * nd * gw * size
            fn = getattr(_lib.lib, f"femtc_{op}_apply_device_{dtype}")
            t_entry0 = time.time()
            ms_all = []
            for i in range(len(PLACEMENTS)):
                off, _ = place(i, bU, bG)
                U = arena[off["U"]:off["U"] + bU].view(tdt)
                G = arena[off["G"]:off["G"] + bG].view(tdt)
                W = arena[off["W"]:off["W"] + bU].view(tdt)
                fill_uniform(U, 1, 1, 0)
                fill_uniform(G, 1, 2, 0)
                # the C ABI entry itself, arguments prepared once: at P = 1 the kernel lasts ~20 us and the
                # Python wrapper's checks would be part of the measurement
                args = [phi.ctypes.data, phi.size, U.data_ptr(), W.data_ptr(), G.data_ptr(), cells, P]
                args += [mode, stream] if op == "stiffness" else [stream]

                def launch(reps):
                    for _ in range(reps):
                        rc = fn(*args)
                        if rc:
                            _lib.check(rc)
                launch(2)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); launch(1); e1.record()
                torch.cuda.synchronize()
                est = max(e0.elapsed_time(e1), 1e-3)
                launch(max(3, int(20.0 / est)))  # >= 20 ms of warm-up: the clocks have ramped
                reps = min(4000, max(5, int(40.0 / est)))
                torch.cuda.synchronize()
                e0.record(); launch(reps); e1.record()
                torch.cuda.synchronize()
                ms_all.append(e0.elapsed_time(e1) / reps)
            t_entry1 = time.time()
            err = spot_check(femtc, orc, W, op, P, ndt, mode, phi,
                             [(0, 40, 0), (cells // 2 + 3, 40, cells // 2 + 3), (cells - 40, 40, cells - 40)], nd)
            tol = 1e-12 if dtype == "f64" else 1e-5
            info = femtc.kernel_info(op, size, P, mode)
            rate = lambda ms: cells * nd / ms / 1e6
            gbs = lambda ms: cells * nd * (2 + gw) * size / (ms * 1e-3) / 1e9
            med = float(np.median(ms_all))
            entries.append({
                "name": name, "op": op, "mode": mode_name, "degree": P, "dtype": dtype, "cells": cells,
                "gdofs": {"min": rate(max(ms_all)), "median": rate(med), "max": rate(min(ms_all))},
                "ms": {"min": min(ms_all), "median": med, "max": max(ms_all), "per_placement": ms_all},
                "wall_s": t_entry1 - t_entry0,
                "achieved_gbs": gbs(med), "frac": gbs(med) / peak, "frac_min": gbs(max(ms_all)) / peak,
                "frac_max": gbs(min(ms_all)) / peak, "frac_of_nominal_8TBs": gbs(med) / 8000.0,
                "placements": len(ms_all), "family": FAMILY.get(info["family"], "?"),
                "kernel": {k: info[k] for k in ("cells_per_tile", "threads", "smem_bytes", "blocks_per_sm", "regs", "detail")},
                "clocks": clocks.summary(t_entry0, t_entry1),
                "max_rel_err": err, "tol": tol, "parity_ok": bool(err <= tol)})
    del arena
    torch.cuda.empty_cache()
    return entries


def stream_peaks(dev):
    """Read-dominated streaming peaks measured now with the library's probe kernel (reads : 1 write of 16-byte words
    with no arithmetic to speak of): the denominators that match the operators' traffic mix.  GB/s, best of 5."""
    import torch

    from femtc_b200 import _lib

    per = 512 * MIB
    src = torch.empty(7 * per, dtype=torch.uint8, device=dev).zero_()
    dst = torch.empty(per, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    out = {}
    for reads in (1, 2, 7):
        best = 0.0
        for rep in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(4):
                _lib.check(_lib.lib.femtc_stream_probe_device(src.data_ptr(), dst.data_ptr(), per, reads, stream))
            e1.record()
            torch.cuda.synchronize()
            if rep:
                best = max(best, 4 * (reads + 1) * per / (e0.elapsed_time(e1) * 1e-3) / 1e9)
        out[f"{reads}:1"] = best
    del src, dst
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------------
# The ONE stdout line stays small at every N (round 2 lost its N = 1 line to a 52 KB print): the line carries a compact
# form, the full record goes to a side file.
# ------------------------------------------------------------------------------------------------------
LINE_LIMIT = 6000  # bytes of the printed line, enforced by fit_line
SWEEP_COLS = ["frac", "frac_min", "gdofs", "err_over_tol", "sm_mhz"]


def sig(x, n=6):
    """Floats rounded to n significant digits, recursively (ints, strings, None untouched)."""
    if isinstance(x, float):
        return float(f"{x:.{n}g}") if np.isfinite(x) else None
    if isinstance(x, dict):
        return {k: sig(v, n) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [sig(v, n) for v in x]
    return x


def compact_sweep(entries):
    """48 dicts (~46 KB) -> one table (~1.8 KB).  groups["mass_f64"][c][i] is column SWEEP_COLS[c] at degree P[i]:
    median fraction of the measured HBM peak, worst-placement fraction, median GDoF/s, spot-check error / tolerance,
    median SM MHz; 3 significant digits.  Group names: mass | stiffness (REF_EXACT) | stiffness_full, then the dtype."""
    groups = {}
    for e in sorted(entries, key=lambda e: e["degree"]):
        g = groups.setdefault(e["name"].replace(f"_P{e['degree']}_", "_"), {"P": [], "v": [[] for _ in SWEEP_COLS]})
        mhz = (e.get("clocks") or {}).get("sm_mhz")
        row = sig([e["frac"], e["frac_min"], e["gdofs"]["median"], e["max_rel_err"] / e["tol"],
                   float(mhz) if mhz is not None else None], 3)
        g["P"].append(e["degree"])
        row[-1] = int(row[-1]) if row[-1] is not None else None
        for col, v in zip(g["v"], row):
            col.append(v)
    degrees = [g["P"] for g in groups.values()]
    if degrees and all(d == degrees[0] for d in degrees):
        return {"cols": SWEEP_COLS, "P": degrees[0], "groups": {k: g["v"] for k, g in groups.items()}}
    return {"cols": SWEEP_COLS, "groups": groups}


def _pick(d, keys):
    return {k: d[k] for k in keys if isinstance(d, dict) and k in d}


def slim_line(line):
    """The printed form of a full bench record: every contract key kept (value, roofline, e2e, cpu_baseline, clocks,
    strong, ...), prose and per-entry detail dropped, floats at 6 significant digits (checksums keep all digits)."""
    out = dict(line)
    if isinstance(out.get("roofline"), dict):
        out["roofline"] = {k: v for k, v in out["roofline"].items() if k != "stream_peak_note"}
    if isinstance(out.get("e2e"), dict) and isinstance(out["e2e"].get("copy_bound"), dict):
        out["e2e"] = dict(out["e2e"])
        out["e2e"]["copy_bound"] = _pick(out["e2e"]["copy_bound"], ("ms_per_step", "h2d_gbs", "value_at_bound",
                                                                     "pipeline_efficiency"))
    if isinstance(out.get("strong"), dict):
        st = out["strong"]
        slim = _pick(st, ("value", "unit", "scaling", "n_gpus", "steps", "warmup", "ms_per_step", "dtype", "gpu_launches"))
        slim["config"] = _pick(st.get("config", {}), ("workload", "total_cells", "chunks_per_rank"))
        slim["roofline"] = _pick(st.get("roofline", {}), ("achieved", "peak", "frac"))
        slim["clocks"] = _pick(st.get("clocks", {}), ("sm_mhz", "sm_max_mhz", "reasons"))
        slim["parity"] = _pick(st.get("parity", {}), ("max_rel_err", "tol", "ok"))
        out["strong"] = slim
    if isinstance(out.get("sweep"), list):
        out["sweep"] = compact_sweep(out["sweep"])
    if isinstance(out.get("sweep_summary"), dict):
        out["sweep_summary"] = {k: v for k, v in out["sweep_summary"].items() if k not in ("timing", "peak_source")}
    cb = out.get("cpu_baseline")
    if isinstance(cb, dict):
        cb = dict(cb)
        if isinstance(cb.get("tuning"), dict):
            cb["tuning"] = _pick(cb["tuning"], ("one_core_best", "one_core_best_gdofs"))
        if isinstance(cb.get("cache_resident"), dict):
            cb["cache_resident"] = _pick(cb["cache_resident"], ("value", "unit", "cores"))
        out["cpu_baseline"] = cb
    checksum = out.get("checksum")
    out = sig(out)
    if checksum is not None:
        out["checksum"] = checksum
    return out


# optional keys given up (in this order) if a line would still exceed LINE_LIMIT; contract keys are never touched
DROP_ORDER = ("other_mode", "checksum", "e2e_cached_geometry", "sweep", "strong", "e2e_resident_geometry", "sweep_summary", "leg_errors")


def fit_line(line, limit=LINE_LIMIT):
    line = dict(line)
    dropped = []
    for key in DROP_ORDER:
        if len(json.dumps(line)) <= limit:
            break
        if key in line:
            del line[key]
            dropped.append(key)
            line["dropped_for_size"] = dropped
    return line


def write_detail(a, world, line):
    """Full record (per-entry sweep detail, CPU tuning table, notes) next to the run; never fatal."""
    path = Path(a.detail_file) if a.detail_file else ROOT / "gpurun_out" / f"bench_detail_n{world}.json"
    try:
        path.parent.mkdir(parents=True, exist_ok=True)
        path.write_text(json.dumps(line, indent=1) + "\n")
        return str(path.relative_to(ROOT)) if path.is_relative_to(ROOT) else str(path)
    except OSError:
        return None


def run_ours(a, rank: int, local_rank: int, world: int):
    import torch
    import torch.distributed as dist

    import femtc_b200 as femtc
    from femtc_b200 import _lib
    from femtc_b200.distributed import combine_checksums, device_checksum, fill_uniform

    torch.cuda.set_device(local_rank)
    _lib.check(_lib.lib.femtc_set_device(local_rank))
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    P, nd, gw = a.degree, (a.degree + 1) ** 3, OPS[a.op]
    tdt = torch.float64 if a.dtype == "f64" else torch.float32
    ndt = np.float64 if a.dtype == "f64" else np.float32
    size = 8 if a.dtype == "f64" else 4
    mode = femtc.FULL if a.mode == "full" else femtc.REF_EXACT
    cells = a.cells
    cell0 = rank * cells  # rank r owns global cells [r*cells, (r+1)*cells): contiguous, no halo
    n2 = (P + 1) ** 2

    # inputs generated on the owning device with the counter-based generator (seed 1; streams
    # phi = 0, U = 1, detJ|G = 2), offset by the rank's first cell
    phi = np.empty(2 * n2, dtype=ndt)
    phi_t = fill_uniform(torch.empty(2 * n2, dtype=tdt, device=dev), 1, 0, 0)
    phi[:] = phi_t.cpu().numpy()
    # The three fields live back to back in one allocation (256-byte aligned starts), the way an
    # application's field arena would hold them.  Their relative placement in HBM moves the kernel
    # time by several percent (profiles/r01_placement_probe.txt); back to back is a favourable one.
    def up(n):
        return (n + 31) // 32 * 32
    nU, nG = cells * nd, cells * nd * gw
    arena = torch.empty(up(nU) + up(nG) + up(nU), dtype=tdt, device=dev)
    U = fill_uniform(arena[:nU], 1, 1, cell0 * nd)
    G = fill_uniform(arena[up(nU):up(nU) + nG], 1, 2, cell0 * nd * gw)
    W = arena[up(nU) + up(nG):up(nU) + up(nG) + nU].zero_()

    def step(checksum=None):
        if a.op == "mass":
            rc = femtc.mass_operator(phi, U, W, G, cells, P, checksum=checksum)
        else:
            rc = femtc.stiffness_operator(phi, U, W, G, cells, P, mode=mode, checksum=checksum)
        assert rc == 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # The checksum the job reports at the end (bench_mass.cpp:135-137) comes out of the LAST step's own epilogue
    # (femtc_*_apply_device_ex: per-warp partials from the W values in registers, no second pass over W).
    local = torch.zeros(4, dtype=torch.float64, device=dev)
    for _ in range(max(a.warmup, 3)):
        step()
    step(local)  # one warm-up of the checksum variant of the kernel and of the collective path as well
    glob = combine_checksums(local) if world > 1 else local
    assert torch.equal(local.cpu()[2:], device_checksum(W).cpu()[2:]), "fused and stand-alone checksum disagree"
    barrier()

    launches0 = femtc.get_option("launches")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        e0.record()
        k0.record()
        for i in range(a.steps):
            step(local if i == a.steps - 1 else None)
        k1.record()
        glob = combine_checksums(local) if world > 1 else local.clone()
        e1.record()
        barrier()
        launches = femtc.get_option("launches") - launches0
        # The timed region lasts milliseconds, shorter than one nvidia-smi sampling period: keep the
        # same step running (untimed) for about a second so the sampler sees the clocks under this load.
        t_end = time.perf_counter() + 1.0
        while time.perf_counter() < t_end:
            for _ in range(20):
                step()
            torch.cuda.synchronize()
    ms_total = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    ms_kern = torch.tensor([k0.elapsed_time(k1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms_kern, op=dist.ReduceOp.MAX)
    ms_step = ms_total.item() / a.steps
    ms_kernel = ms_kern.item() / a.steps
    dofs_step = world * cells * nd
    value = dofs_step / ms_step / 1e6
    checksum = [float(x) for x in glob.cpu().numpy()]

    W_main_head = None  # filled below for the comparison with the CPU baseline's output

    # ---- end to end through the host-pointer C ABI, pinned host buffers --------------------------------------
    e2e = e2e_resident = e2e_cached = None
    if not a.no_e2e:
        hU = torch.empty(cells * nd, dtype=tdt).pin_memory()
        hG = torch.empty(cells * nd * gw, dtype=tdt).pin_memory()
        hW = torch.empty(cells * nd, dtype=tdt).pin_memory()
        hU.copy_(U)
        hG.copy_(G)
        nU, nG, nW = hU.numpy(), hG.numpy(), hW.numpy()

        def host_step():
            if a.op == "mass":
                rc = femtc.mass_operator(phi, nU, nW, nG, cells, P, femtc.Order.ikj, 0, 0)
            else:
                rc = femtc.stiffness_operator(phi, nU, nW, nG, cells, P, femtc.Order.ikj, 0, 0, mode)
            assert rc == 0

        host_step()  # warm-up: allocates the library's device staging slots
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            host_step()  # synchronous: W is in host memory on return
        sec = torch.tensor([(time.perf_counter() - t0) / a.e2e_steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(sec, op=dist.ReduceOp.MAX)
        assert np.array_equal(nW[:4096 * nd], W[:4096 * nd].cpu().numpy()), "host and device paths differ"
        e2e = {"value": dofs_step / sec.item() / 1e9, "unit": "GDoF/s",
               "h2d_bytes_per_step": int(cells * nd * (1 + gw) * size), "d2h_bytes_per_step": int(cells * nd * size),
               "ms_per_step": sec.item() * 1e3, "steps": a.e2e_steps,
               "api": f"femtc_{a.op}_apply_{a.dtype} (host pointers, pinned memory)"}
        # The bound of that figure: the same pinned buffers moved by bare cudaMemcpyAsync, H2D (U, geometry) on one
        # stream and D2H (W) on another, no kernel.  e2e.ms_per_step / this = how close the pipeline is to the link.
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        def copies():
            with torch.cuda.stream(s_in):
                U.copy_(hU, non_blocking=True)
                G.copy_(hG, non_blocking=True)
            with torch.cuda.stream(s_out):
                hW.copy_(W, non_blocking=True)
            s_in.synchronize()
            s_out.synchronize()
        torch.cuda.synchronize()
        copies()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            copies()
        csec = torch.tensor([(time.perf_counter() - t0) / a.e2e_steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(csec, op=dist.ReduceOp.MAX)
        e2e["copy_bound"] = {"ms_per_step": csec.item() * 1e3,
                             "h2d_gbs": e2e["h2d_bytes_per_step"] / csec.item() / 1e9,
                             "value_at_bound": dofs_step / csec.item() / 1e9, "unit": "GDoF/s",
                             "pipeline_efficiency": csec.item() / sec.item(),
                             "what": "same pinned buffers, cudaMemcpyAsync H2D (U, geometry) || D2H (W), no kernel; "
                                     "max over ranks"}
        # Same metric with the geometry held by an operator handle (femtc_operator_*: detJ / G uploaded
        # once at set-up, as a solver would; the reference's own drivers re-apply the operator to the same
        # arrays, bench_stiffness.cpp:95-105).  Per step only U travels in and W out.  Reported beside the
        # strict figure above, never instead of it.
        K = femtc.ResidentOperator(a.op, phi, nG, cells, P, mode=mode)
        K.apply(nU, nW)
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            K.apply(nU, nW)
        sec = torch.tensor([(time.perf_counter() - t0) / a.e2e_steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(sec, op=dist.ReduceOp.MAX)
        assert np.array_equal(nW[:4096 * nd], W[:4096 * nd].cpu().numpy()), "operator handle and device paths differ"
        e2e_resident = {"value": dofs_step / sec.item() / 1e9, "unit": "GDoF/s",
                        "h2d_bytes_per_step": int(cells * nd * size), "d2h_bytes_per_step": int(cells * nd * size),
                        "setup_h2d_bytes_once": int(cells * nd * gw * size),
                        "ms_per_step": sec.item() * 1e3, "steps": a.e2e_steps,
                        "api": "femtc_operator_apply (geometry resident in HBM; U, W host pointers, pinned memory)"}
        K.destroy()
        # The same figure from the UNCHANGED drop-in call with the opt-in geometry cache (option geom_cache: G stays in HBM
        # keyed by address + sampled hash, section 8f-1): what bench_stiffness.cpp:95-105 gets without touching its loop.
        femtc.set_option("geom_cache", 1)
        try:
            host_step()  # uploads the geometry once
            barrier()
            t0 = time.perf_counter()
            for _ in range(a.e2e_steps):
                host_step()
            sec = torch.tensor([(time.perf_counter() - t0) / a.e2e_steps], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(sec, op=dist.ReduceOp.MAX)
            assert np.array_equal(nW[:4096 * nd], W[:4096 * nd].cpu().numpy()), "cached-geometry and device paths differ"
            e2e_cached = {"value": dofs_step / sec.item() / 1e9, "unit": "GDoF/s",
                          "h2d_bytes_per_step": int(cells * nd * size), "d2h_bytes_per_step": int(cells * nd * size),
                          "ms_per_step": sec.item() * 1e3, "steps": a.e2e_steps,
                          "api": f"femtc_{a.op}_apply_{a.dtype} with option geom_cache = 1"}
        finally:
            femtc.set_option("geom_cache", 0)
            _lib.lib.femtc_release_cache()
        del hU, hG, hW

    if rank == 0 and not a.no_cpu_baseline:
        W_main_head = W[:min(cells, 262144) * nd].cpu().numpy()

    # ---- secondary figure: the other stiffness mode, same inputs ---------------------------------------------
    other = None
    if a.op == "stiffness":
        om = femtc.REF_EXACT if mode == femtc.FULL else femtc.FULL
        for _ in range(3):
            femtc.stiffness_operator(phi, U, W, G, cells, P, mode=om)
        torch.cuda.synchronize()
        k0.record()
        for _ in range(10):
            femtc.stiffness_operator(phi, U, W, G, cells, P, mode=om)
        k1.record()
        torch.cuda.synchronize()
        other = {"mode": "full" if om == femtc.FULL else "ref_exact", "value": cells * nd / (k0.elapsed_time(k1) / 10) / 1e6,
                 "unit": "GDoF/s per GPU"}

    # ---- free the headline fields, then the other BASELINE configurations under the same clock ----------------------
    del arena, U, G, W
    torch.cuda.empty_cache()
    femtc_release = _lib.lib.femtc_release_cache
    femtc_release()
    # The extra legs must never cost the headline line: on one rank a failure is recorded in the line instead of raised
    # (with several ranks every rank takes part in the leg's collectives, so an exception there is fatal anyway).
    leg_errors = {}

    def guarded(name, fn):
        if world > 1:
            return fn()
        try:
            return fn()
        except Exception as exc:  # noqa: BLE001
            leg_errors[name] = f"{type(exc).__name__}: {exc}"
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
            return None

    strong = None
    if not a.no_strong:
        strong = guarded("strong", lambda: strong_leg(
            a, rank, local_rank, world, dev, op="stiffness", degree=a.strong_degree, dtype="f64", mode_name="ref_exact",
            total_cells=a.strong_cells, steps=a.strong_steps, warm=1, resident_gb=a.resident_gb))
    sweep = speaks = None
    if world == 1 and not a.no_sweep:
        speaks = guarded("stream_peaks", lambda: stream_peaks(dev))
        sweep = guarded("sweep", lambda: run_sweep(a, dev, local_rank))

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    bytes_per_dof = (2 + gw) * size
    achieved = cells * nd * bytes_per_dof / (ms_kernel * 1e-3) / 1e9  # per launch = per rank
    line = {
        "metric": METRIC, "value": value, "unit": "GDoF/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic",
        "config": config_dict(a, world),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": recorded_traffic(a), "peak_source": peak_src,
                     "algorithmic_bytes_per_dof": bytes_per_dof, "kernel_ms": ms_kernel,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "e2e": e2e,
        "e2e_resident_geometry": e2e_resident,
        "e2e_cached_geometry": e2e_cached,
        "gpu_launches": int(launches),
        "clocks": clocks.summary(),
        "checksum": {"sum": checksum[0], "sum_sq": checksum[1], "max_abs": checksum[2], "n": checksum[3]},
    }
    if other:
        line["other_mode"] = other
    if speaks:
        mix = "7:1" if a.op == "stiffness" else "2:1"
        line["roofline"]["stream_peaks_gbs"] = speaks
        line["roofline"]["frac_of_stream_peak"] = achieved / speaks[mix]
        line["roofline"]["stream_peak_note"] = (f"read-dominated streaming kernel of the library, measured in this run; this "
                                                f"operator's mix is {mix} reads:writes, the MEASURED_PEAKS copy is 1:1")
    if strong:
        line["strong"] = strong
    if sweep is not None:
        line["sweep"] = sweep
        ok = [e for e in sweep if e["parity_ok"]]
        low = [e for e in sweep if e["degree"] <= 4 and e["mode"] == "ref_exact"]
        line["sweep_summary"] = {
            "entries": len(sweep), "parity_ok": len(ok), "cells": a.sweep_cells,
            "min_median_frac_degree_le4_mass_and_ref_exact": min((e["frac"] for e in low), default=None),
            "median_frac_all": float(np.median([e["frac"] for e in sweep])) if sweep else None,
            "peak": peak, "peak_source": peak_src,
            "timing": "per entry and placement: >= 20 ms warm-up, then >= 40 ms (>= 5 launches) between two CUDA "
                      "events on the launching stream; every entry streams more than the 126 MB L2 per launch except "
                      "mass P=1 fp32 (96 MB)"}
    if leg_errors:
        line["leg_errors"] = leg_errors
    if not a.no_cpu_baseline:  # rank 0, at every N (the other ranks wait at the final barrier)
        try:
            threads = os.cpu_count() or 1
            triple, tuning = cpu_tuning(a)
            sample_cells = min(cells, 262144)
            res = cpu_reference_rate(a, sample_cells, threads, reps=20, want_output=4096, triple=triple)
            W_cpu = res["head"]
            W_main_head = W_main_head[:W_cpu.size]
            line["cpu_baseline"] = {"value": res["rate"], "unit": "GDoF/s", "cores": threads, "kind": res["kind"],
                                    "sample": res["desc"], "cpu": cpu_model(), "tuning": tuning}
            # the reference's own problem size (run.sh:9: Ndofs = 1e6, cache resident) per worker, all cores, same knobs
            small = cpu_reference_rate(a, threads * max(1, 1_000_000 // nd), threads, reps=20, triple=triple)
            line["cpu_baseline"]["cache_resident"] = {"value": small["rate"], "unit": "GDoF/s", "cores": threads,
                                                      "sample": small["desc"]}
            if a.mode == "ref_exact" or a.op == "mass":
                # the CPU baseline ran the same cells of the same synthetic inputs: compare its output
                err = float(np.abs(W_main_head.astype(np.float64) - W_cpu.astype(np.float64)).max()
                            / np.abs(W_cpu.astype(np.float64)).max())
                line["cpu_baseline"]["gpu_vs_cpu_max_rel_err"] = err
        except Exception as exc:  # noqa: BLE001  (the CPU leg must not cost the GPU line)
            line["cpu_baseline"] = {"value": None, "unit": "GDoF/s", "cores": os.cpu_count() or 1, "kind": "unavailable",
                                    "sample": f"{type(exc).__name__}: {exc}"}
    detail = write_detail(a, world, line)
    printed = slim_line(line)
    if detail:
        printed["detail_file"] = detail
    print(json.dumps(fit_line(printed)), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _keep_stdout_for_the_json_line():
    """The contract is ONE JSON line on stdout.  Libraries underneath (NCCL prints its version banner on
    stdout when NCCL_DEBUG=VERSION is set on the box) must not add to it: from here on file descriptor 1
    points at stderr, and sys.stdout keeps a private copy of the original stdout for our own print."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real


def main():
    a = parse_args()
    _keep_stdout_for_the_json_line()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "reference":
        run_reference(a, rank)
        return
    if world != a.gpus and world == 1 and a.gpus > 1:
        # convenience: `python bench.py --gpus N` without torchrun re-launches itself under it
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", str(Path(__file__).resolve())] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd, stdout=sys.stdout))  # the child inherits the real stdout
    if a.total_cells:
        run_strong(a, rank, local_rank, world)
    else:
        run_ours(a, rank, local_rank, world)


if __name__ == "__main__":
    main()
