"""Cell partitioning across the GPUs of one box and the one collective the path needs.

The operator action is cell-local (every array is cell-major, there is no global numbering and
no halo: reference src/mass.hpp:25-29, src/stiffness.hpp:55-58), so ranks never exchange field
data.  The only exchange is the final reduction of four doubles per rank -- the analogue of
the three MPI_Reduce calls of bench_mass.cpp:135-137 -- carried by NCCL (or gloo in CPU tests).
"""
from __future__ import annotations

from typing import Tuple


def partition(num_cells: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous cell range [start, stop) of `rank`: floor(r*N/R) .. floor((r+1)*N/R)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    return (rank * num_cells) // world, ((rank + 1) * num_cells) // world


def aligned_partition(num_cells: int, rank: int, world: int, align: int = 16) -> Tuple[int, int]:
    """Like partition() but interior boundaries are multiples of `align` cells, so every shard
    starts on a 16-byte boundary of U, W, detJ and G for any degree and dtype (TMA path)."""
    def cut(r):
        if r == world:
            return num_cells
        return min(num_cells, ((r * num_cells) // world) // align * align)
    return cut(rank), cut(rank + 1)


def combine_checksums(local4, group=None):
    """All ranks contribute {sum, sum of squares, max|x|, n}; everyone gets the global four.

    `local4` is a float64 tensor of 4 values on the backend's device (CUDA for nccl, CPU for
    gloo).  One all_gather of 32 bytes per rank; the combine is done redundantly on every rank.
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    gathered = torch.empty(world * 4, dtype=torch.float64, device=local4.device)
    dist.all_gather_into_tensor(gathered, local4.contiguous(), group=group)
    g = gathered.view(world, 4)
    out = torch.empty(4, dtype=torch.float64, device=local4.device)
    out[0] = g[:, 0].sum()
    out[1] = g[:, 1].sum()
    out[2] = g[:, 2].max()
    out[3] = g[:, 3].sum()
    return out


def device_checksum(x):
    """{sum, sum of squares, max|x|, n} of a CUDA tensor, computed by the library's reduction
    kernel on torch's current stream; returns a float64 CUDA tensor of 4 values."""
    import torch
    from . import _lib

    out = torch.empty(4, dtype=torch.float64, device=x.device)
    s = {torch.float64: "f64", torch.float32: "f32"}[x.dtype]
    _lib.check(getattr(_lib.lib, f"femtc_checksum_device_{s}")(
        x.data_ptr(), x.numel(), out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out


def fill_uniform(x, seed: int = 1, stream_id: int = 0, offset: int = 0, centred: bool = False):
    """Fill a CUDA tensor in place with the counter-based uniform generator (see femtc_b200.h)."""
    import torch
    from . import _lib

    s = {torch.float64: "f64", torch.float32: "f32"}[x.dtype]
    _lib.check(getattr(_lib.lib, f"femtc_fill_uniform_device_{s}")(
        x.data_ptr(), x.numel(), seed, stream_id, offset, int(centred),
        torch.cuda.current_stream().cuda_stream))
    return x
