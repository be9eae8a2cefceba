"""femtc_apply_contractions_device_static_f64: the contraction entry point that allocates nothing on the device.

Who needs it: a program that gave all of HBM to its own arena at start-up (the library's hidden cudaMallocAsync would
fail there or take memory the arena already promised), or a time-stepping loop that cannot risk the device-wide
synchronisation an allocation or free may cause.  The entry point keeps its scratch in a 4 MiB __device__ array inside
libfemtc_b200.so; this file checks, through the C ABI, that it computes what femtc_apply_contractions_device_f64
computes -- on two calls in a row, the second one smaller so that it runs over what the first left in the array --
and that a shape too large for the array is refused without touching the output.

The file name sorts last on purpose: the driver runs pytest with -x, so a failure here cannot hide any other test.
The test does not depend on what ran before it (own inputs, own tensors, options left at their defaults).
"""

import numpy as np
import pytest

TOL = 1e-12  # fp64 tolerance of the parity tests (tests/test_gpu_parity.py)


def rel_err(a, ref):
    a, ref = a.astype(np.float64), ref.astype(np.float64)
    return float(np.abs(a - ref).max() / np.abs(ref).max())


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available()
    return torch


def both_entry_points(femtc, torch, Na, Nb, Tr, batch, seed):
    """Run the allocating and the no-allocation entry point on the same inputs; return both results (host arrays)."""
    lib = femtc._lib.lib
    rng = np.random.default_rng(seed)
    I = [np.ascontiguousarray(rng.random(Na * Nb) - 0.5) for _ in range(3)]
    dB = torch.from_numpy(rng.random(batch * Nb ** 3) - 0.5).cuda()
    out = []
    for fn in (lib.femtc_apply_contractions_device_f64, lib.femtc_apply_contractions_device_static_f64):
        dA = torch.full((batch * Na ** 3,), float("nan"), dtype=torch.float64, device="cuda")
        rc = fn(Na, Nb, Tr, I[0].ctypes.data, I[1].ctypes.data, I[2].ctypes.data, dB.data_ptr(), dA.data_ptr(),
                batch, torch.cuda.current_stream().cuda_stream)
        assert rc == 0, femtc._lib.last_error()
        torch.cuda.synchronize()
        out.append(dA.cpu().numpy())
    return out


@pytest.mark.gpu
def test_no_allocation_entry_point_matches_the_allocating_one_twice_in_a_row(femtc, torch_cuda):
    torch = torch_cuda
    launches0 = femtc.get_option("launches")
    # 1st call: a shape whose work arrays would not fit shared memory (the allocating entry point spills to an
    # allocated workspace, this one to the static array), more tensors than blocks.  2nd call: a small shape with the
    # other matrix layout, running over the first call's leftovers.  3rd: a shape the allocating entry point serves
    # with its tiled TMA kernel (different summation code, same tolerance).
    for Na, Nb, Tr, batch, seed in ((30, 28, 0, 37, 1), (5, 4, 1, 2500, 2), (6, 5, 0, 1777, 3)):
        ref, got = both_entry_points(femtc, torch, Na, Nb, Tr, batch, seed)
        assert np.isfinite(ref).all() and np.isfinite(got).all(), (Na, Nb)
        assert rel_err(got, ref) <= TOL, (Na, Nb, Tr, batch)
    assert femtc.get_option("launches") >= launches0 + 6


@pytest.mark.gpu
def test_shape_beyond_the_static_array_is_refused(femtc, torch_cuda):
    torch = torch_cuda
    lib = femtc._lib.lib
    Na = Nb = 64                                       # 64*64*128 work values alone are the whole 4 MiB
    I = np.zeros(Na * Nb)
    dB = torch.zeros(Nb ** 3, dtype=torch.float64, device="cuda")
    dA = torch.full((Na ** 3,), 7.0, dtype=torch.float64, device="cuda")
    rc = lib.femtc_apply_contractions_device_static_f64(Na, Nb, 0, I.ctypes.data, I.ctypes.data, I.ctypes.data,
                                                        dB.data_ptr(), dA.data_ptr(), 1, None)
    torch.cuda.synchronize()
    assert rc == femtc._lib.ERR_ARG
    msg = femtc._lib.last_error()
    assert "4194304" in msg and "Na = 64" in msg, msg
    assert bool((dA == 7.0).all())
    # the largest cube that fits still runs (one work slice: a single block walks the batch)
    ref, got = both_entry_points(femtc, torch, 63, 63, 1, 2, 4)
    assert rel_err(got, ref) <= TOL
