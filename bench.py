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
    return ap.parse_args()


def workload_name(a):
    geom = "G (6 per point)" if a.op == "stiffness" else "detJ"
    return (f"{'bench_stiffness: Laplace stiffness' if a.op == 'stiffness' else 'bench_mass: mass'} action, "
            f"degree-{a.degree} hexahedra, {'fp64' if a.dtype == 'f64' else 'fp32'}, {a.cells} cells per GPU, "
            f"U + {geom} -> W, synthetic uniform [0,1) inputs")


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
            self.rows.append(line.strip())

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, pw, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for row in self.rows:
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
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "reasons": sorted(reasons), "samples": len(sm),
                "window": "timed region plus ~1 s of the same step repeated right after it"}


# ------------------------------------------------------------------------------------------------------
# CPU reference (oracle/_ref = the unmodified reference headers, -Ofast, widest ISA the host supports)
# ------------------------------------------------------------------------------------------------------
def cpu_reference_rate(a, cells: int, workers: int, reps: int, warm: int = 1, want_output: int = 0):
    """GDoF/s of the reference's own CPU kernel (oracle/_ref) with `workers` processes on disjoint
    cell ranges of the same synthetic inputs; see oracle/cpu_bench.py."""
    from oracle import cpu_bench
    return cpu_bench.run(a.op, a.degree, np.float64 if a.dtype == "f64" else np.float32, cells, workers, reps,
                         warm, want_output)


def run_reference(a, rank: int):
    if rank != 0:
        return
    workers = os.cpu_count() or 1
    # bounded sample: shrink the batch until one pass takes about a second on this host
    cells = a.cells
    probe = cpu_reference_rate(a, min(cells, 16384 * workers), workers, reps=1, warm=0)
    est = cells / min(cells, 16384 * workers) * probe["sec"]
    while est > 1.5 and cells > 16384 * workers:
        cells //= 2
        est /= 2
    res = cpu_reference_rate(a, cells, workers, reps=a.steps, warm=max(1, min(a.warmup, 3)))
    rate, sec, kind, desc = res["rate"], res["sec"], res["kind"], res["desc"]
    threads = workers
    nd = (a.degree + 1) ** 3
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "GDoF/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic",
        "config": {"workload": workload_name(a), "op": a.op, "degree": a.degree, "mode": a.mode,
                   "cells_per_step": cells, "dofs_per_step": cells * nd},
        "cpu_baseline": {"value": rate, "unit": "GDoF/s", "cores": threads, "kind": kind, "sample": desc},
        "e2e": {"value": rate, "unit": "GDoF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------
def run_strong(a, rank: int, local_rank: int, world: int):
    """BASELINE config 5: a fixed global batch (--total-cells) partitioned over the GPUs, no halo.  A rank holds
    at most --resident-gb of fields; a larger share is swept in chunks (inputs regenerated on the device between
    chunks, outside the timed spans).  Timed: the operator launches only (CUDA events per chunk, summed; max over
    ranks).  The one collective is the 4-double checksum all-gather at the end."""
    import torch
    import torch.distributed as dist

    import femtc_b200 as femtc
    from femtc_b200 import _lib
    from femtc_b200.distributed import combine_checksums, device_checksum, fill_uniform, partition

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
    lo, hi = partition(a.total_cells, rank, world)
    bytes_per_cell = nd * (2 + gw) * size
    chunk = max(32, int(a.resident_gb * 1e9 / bytes_per_cell) // 32 * 32)
    chunk = min(chunk, hi - lo)
    n2 = (P + 1) ** 2
    phi = fill_uniform(torch.empty(2 * n2, dtype=tdt, device=dev), 1, 0, 0).cpu().numpy().astype(ndt)

    def up(n):
        return (n + 31) // 32 * 32
    nU, nG = chunk * nd, chunk * nd * gw
    arena = torch.empty(up(nU) + up(nG) + up(nU), dtype=tdt, device=dev)
    U, G, W = arena[:nU], arena[up(nU):up(nU) + nG], arena[up(nU) + up(nG):up(nU) + up(nG) + nU]

    def apply(n):
        rc = (femtc.mass_operator(phi, U, W, G, n, P) if a.op == "mass"
              else femtc.stiffness_operator(phi, U, W, G, n, P, mode=mode))
        assert rc == 0

    def sweep(reps_per_chunk, timed):
        """one pass over the rank's cells; returns (kernel ms, checksum partials)"""
        ms, parts = 0.0, []
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
        return ms, parts

    sweep(max(1, min(a.warmup, 3)), False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = femtc.get_option("launches")
    with ClockSampler(local_rank) as clocks:
        ms, parts = sweep(a.steps, True)
    launches = femtc.get_option("launches") - launches0
    loc = np.array([sum(p[0] for p in parts), sum(p[1] for p in parts), max(p[2] for p in parts), sum(p[3] for p in parts)])
    loc_t = torch.tensor(loc, dtype=torch.float64, device=dev)
    glob = combine_checksums(loc_t) if world > 1 else loc_t
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item()
    if rank == 0:
        peak, peak_src = peaks()
        dofs = a.total_cells * nd
        achieved = (hi - lo) * nd * (2 + gw) * size / (ms_step * 1e-3) / 1e9
        cs = [float(x) for x in glob.cpu().numpy()]
        print(json.dumps({
            "metric": METRIC, "value": dofs / ms_step / 1e6, "unit": "GDoF/s", "n_gpus": world, "steps": a.steps,
            "warmup": max(1, min(a.warmup, 3)), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
            "config": {"workload": f"stiffness action degree {P}, {a.dtype}, {a.total_cells} cells in total partitioned over "
                                   f"{world} B200 with NCCL checksum reduction" if a.op == "stiffness" else
                                   f"mass action degree {P}, {a.dtype}, {a.total_cells} cells in total over {world} B200",
                       "op": a.op, "degree": P, "mode": a.mode, "total_cells": a.total_cells, "cells_this_rank": hi - lo,
                       "resident_chunk_cells": chunk, "chunks_per_rank": -(-(hi - lo) // chunk), "dofs_per_step": dofs,
                       "timed": "operator launches only, CUDA events per resident chunk summed, max over ranks; inputs are "
                                "regenerated on the device between chunks outside the timed spans",
                       "l2": f"no flush needed: every chunk streams {chunk * bytes_per_cell / 1e9:.1f} GB"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_dof": (2 + gw) * size},
            "e2e": None, "gpu_launches": int(launches), "clocks": clocks.summary(),
            "checksum": {"sum": cs[0], "sum_sq": cs[1], "max_abs": cs[2], "n": cs[3]},
        }), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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

    def step():
        if a.op == "mass":
            rc = femtc.mass_operator(phi, U, W, G, cells, P)
        else:
            rc = femtc.stiffness_operator(phi, U, W, G, cells, P, mode=mode)
        assert rc == 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        step()
    # one warm-up of the collective path as well
    glob = combine_checksums(device_checksum(W)) if world > 1 else device_checksum(W)
    barrier()

    launches0 = femtc.get_option("launches")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        e0.record()
        k0.record()
        for _ in range(a.steps):
            step()
        k1.record()
        local = device_checksum(W)
        glob = combine_checksums(local) if world > 1 else local
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
    e2e = e2e_resident = None
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
        del hU, hG, hW

    if rank == 0 and world == 1 and not a.no_cpu_baseline:
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
        "config": {"workload": workload_name(a), "op": a.op, "degree": P, "mode": a.mode, "cells_per_gpu": cells,
                   "dofs_per_step": dofs_step, "parallelism": f"cells partitioned over {world} GPU(s), no halo; "
                   "one 4-double checksum all-gather (NCCL) at the end of the timed region",
                   "placement": "U | geometry | W back to back in one device allocation",
                   "l2": f"no flush needed: each step streams {cells * nd * bytes_per_dof / 1e9:.2f} GB per GPU, 126 MB L2"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": recorded_traffic(a), "peak_source": peak_src,
                     "algorithmic_bytes_per_dof": bytes_per_dof, "kernel_ms": ms_kernel,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "e2e": e2e,
        "e2e_resident_geometry": e2e_resident,
        "gpu_launches": int(launches),
        "clocks": clocks.summary(),
        "checksum": {"sum": checksum[0], "sum_sq": checksum[1], "max_abs": checksum[2], "n": checksum[3]},
    }
    if other:
        line["other_mode"] = other
    if world == 1 and not a.no_cpu_baseline:
        threads = os.cpu_count() or 1
        sample_cells = min(cells, 262144)
        res = cpu_reference_rate(a, sample_cells, threads, reps=20, want_output=4096)
        W_cpu = res["head"]
        W_main_head = W_main_head[:W_cpu.size]
        line["cpu_baseline"] = {"value": res["rate"], "unit": "GDoF/s", "cores": threads, "kind": res["kind"],
                                "sample": res["desc"]}
        if a.mode == "ref_exact" or a.op == "mass":
            # the CPU baseline ran the same cells of the same synthetic inputs: compare its output
            err = float(np.abs(W_main_head.astype(np.float64) - W_cpu.astype(np.float64)).max()
                        / np.abs(W_cpu.astype(np.float64)).max())
            line["cpu_baseline"]["gpu_vs_cpu_max_rel_err"] = err
    print(json.dumps(line), flush=True)
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
