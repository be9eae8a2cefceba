"""CPU tests (gloo, world_size 2) of the multi-GPU host logic: contiguous cell partition with no
halo and the single 4-double reduction.  The local operator here is the oracle -- the CUDA
kernels are exercised by the -m gpu tests; what is under test is the partition/reduction code."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from femtc_b200.distributed import aligned_partition, partition  # noqa: E402


def test_partition_covers_everything_once():
    for n in (0, 1, 7, 15625, 10 ** 6 + 3):
        for world in (1, 2, 3, 4, 8):
            for part in (partition, aligned_partition):
                cuts = [part(n, r, world) for r in range(world)]
                assert cuts[0][0] == 0 and cuts[-1][1] == n
                for (a0, a1), (b0, b1) in zip(cuts, cuts[1:]):
                    assert a1 == b0 and a0 <= a1
    for r in range(8):
        s, e = aligned_partition(10 ** 6 + 3, r, 8)
        assert s % 16 == 0
    with pytest.raises(ValueError):
        partition(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, op, P, ncells, out_q):
    import torch
    import torch.distributed as dist
    from oracle import Oracle, splitmix
    from femtc_b200.distributed import aligned_partition, combine_checksums

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    o = Oracle()
    c0, c1 = aligned_partition(ncells, rank, world)
    phi, U, geom = splitmix(o, op, P, c1 - c0, "float64", seed=5, cell0=c0)
    rc, W = (o.mass(phi, U, geom, P) if op == "mass" else o.stiffness(phi, U, geom, P))
    assert rc == 0
    local = torch.from_numpy(o.checksum(W))
    glob = combine_checksums(local)
    out_q.put((rank, c0, c1, glob.numpy().copy()))
    dist.destroy_process_group()


@pytest.mark.parametrize("op,P", [("mass", 3), ("stiffness", 2)])
def test_two_rank_partition_and_reduction(op, P):
    import torch.multiprocessing as mp
    from oracle import Oracle, splitmix

    ncells, world = 101, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, op, P, ncells, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    o = Oracle()
    phi, U, geom = splitmix(o, op, P, ncells, "float64", seed=5)
    rc, W = (o.mass(phi, U, geom, P) if op == "mass" else o.stiffness(phi, U, geom, P))
    want = o.checksum(W)
    for rank, c0, c1, got in results:
        assert got[3] == want[3] == W.size
        assert got[2] == want[2]
        assert abs(got[0] - want[0]) <= 1e-12 * abs(want[0])
        assert abs(got[1] - want[1]) <= 1e-12 * abs(want[1])
