"""Where does the host-pointer path's time go at N GPUs?  (the strict e2e step takes 68 ms on one
GPU and 185 ms per rank when eight ranks run together.)

Separates the three candidates with bare copies, no kernels:
  1. each GPU alone: H2D of a pinned buffer               -> the link
  2. all GPUs together, every buffer allocated by a thread running on NUMA node 0  -> host DRAM / socket interconnect
  3. all GPUs together, every buffer allocated by a thread bound to the CPUs of the GPU's own NUMA node
and the same for D2H and for both directions at once.  One process, one worker thread per GPU (the shape of the
library's host_devices fan-out).  Prints a table and one JSON line.

    python tools/probe/pcie_probe.py [--mb 2048] [--reps 3]
"""
import argparse
import json
import os
import re
import subprocess
import threading
import time

import torch


def numa_nodes():
    nodes = {}
    base = "/sys/devices/system/node"
    if not os.path.isdir(base):
        return nodes
    for d in sorted(os.listdir(base)):
        m = re.fullmatch(r"node(\d+)", d)
        if not m:
            continue
        cpus = set()
        for part in open(f"{base}/{d}/cpulist").read().strip().split(","):
            if not part:
                continue
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        nodes[int(m.group(1))] = cpus
    return nodes


def gpu_numa(n):
    """NUMA node of each GPU from the PCI sysfs entry (falls back to 0)."""
    out = []
    for i in range(n):
        bus = torch.cuda.get_device_properties(i).pci_bus_id if hasattr(torch.cuda.get_device_properties(i), "pci_bus_id") else None
        node = 0
        try:
            q = subprocess.run(["nvidia-smi", "-i", str(i), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                               capture_output=True, text=True).stdout.strip().lower()
            q = q[4:] if q.startswith("0000") and len(q) > 12 else q
            p = f"/sys/bus/pci/devices/{q}/numa_node"
            if os.path.exists(p):
                node = max(0, int(open(p).read().strip()))
        except Exception:
            pass
        out.append(node)
    return out


def run_case(devs, alloc_cpus, mb, reps, direction):
    """every device in `devs` copies `mb` MB `reps` times concurrently; returns (aggregate GB/s, per-device GB/s)."""
    n = mb << 20
    res = {}
    start = threading.Barrier(len(devs))

    def work(d):
        allowed = os.sched_getaffinity(0)
        cpus = alloc_cpus(d) & allowed if alloc_cpus(d) else allowed
        if cpus:
            os.sched_setaffinity(0, cpus)          # 0 = the calling thread
        torch.cuda.set_device(d)
        h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
        h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
        h_in.fill_(1)                               # first touch on this thread's node
        h_out.fill_(2)
        g_in = torch.empty(n, dtype=torch.uint8, device=f"cuda:{d}")
        g_out = torch.ones(n, dtype=torch.uint8, device=f"cuda:{d}")
        s1, s2 = torch.cuda.Stream(d), torch.cuda.Stream(d)

        def once():
            if direction in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    g_in.copy_(h_in, non_blocking=True)
            if direction in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_out.copy_(g_out, non_blocking=True)
            s1.synchronize()
            s2.synchronize()
        once()
        start.wait()
        t0 = time.perf_counter()
        for _ in range(reps):
            once()
        res[d] = (time.perf_counter() - t0) / reps
        os.sched_setaffinity(0, allowed)

    ts = [threading.Thread(target=work, args=(d,)) for d in devs]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    mult = 2 if direction == "both" else 1
    per = {d: mult * n / s / 1e9 for d, s in res.items()}
    return mult * n * len(devs) / max(res.values()) / 1e9, per


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    ndev = torch.cuda.device_count()
    nodes = numa_nodes()
    gnode = gpu_numa(ndev)
    print(f"GPUs: {ndev}; NUMA nodes: { {k: len(v) for k, v in nodes.items()} } CPUs; GPU -> node: {gnode}; "
          f"this process may run on {len(os.sched_getaffinity(0))} CPUs")
    topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout
    print(topo)
    node0 = nodes.get(0, set())
    out = {"gpus": ndev, "numa_nodes": {k: len(v) for k, v in nodes.items()}, "gpu_node": gnode, "mb": a.mb, "cases": {}}
    for direction in ("h2d", "d2h", "both"):
        alone = [run_case([d], lambda d: nodes.get(gnode[d], set()), a.mb, a.reps, direction)[0] for d in range(ndev)]
        print(f"{direction:5s} each GPU alone (buffer on its own node):   " + " ".join(f"{x:6.1f}" for x in alone) + " GB/s")
        out["cases"][f"{direction}_alone"] = alone
        if ndev > 1:
            agg0, per0 = run_case(list(range(ndev)), lambda d: node0, a.mb, a.reps, direction)
            aggl, perl = run_case(list(range(ndev)), lambda d: nodes.get(gnode[d], set()), a.mb, a.reps, direction)
            print(f"{direction:5s} all together, buffers on node 0:            " + " ".join(f"{per0[d]:6.1f}" for d in range(ndev)) + f"  = {agg0:7.1f} GB/s aggregate")
            print(f"{direction:5s} all together, buffers on each GPU's node:   " + " ".join(f"{perl[d]:6.1f}" for d in range(ndev)) + f"  = {aggl:7.1f} GB/s aggregate")
            out["cases"][f"{direction}_all_node0"] = {"aggregate": agg0, "per_gpu": [per0[d] for d in range(ndev)]}
            out["cases"][f"{direction}_all_local"] = {"aggregate": aggl, "per_gpu": [perl[d] for d in range(ndev)]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
