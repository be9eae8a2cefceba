"""Mirror of reference src/tmm.hpp (the alternative, rectangular formulation) over the C ABI."""
from __future__ import annotations

import numpy as np

from . import _lib
from .linalg import _suffix, _host


def apply_contractions(I0, I1, I2, B, A, Na: int, Nb: int, Tr: bool, batch: int = 1) -> None:
    """apply_contractions<T,Na,Nb,Tr>(I0, I1, I2, B, A, buffer) (reference src/tmm.hpp:91-126):
    A[a0,a1,a2] = sum I0(a0,b0) I1(a1,b1) I2(a2,b2) B[b0,b1,b2]; A is overwritten.  The scratch
    `Buffer` of the reference (tmm.hpp:46-60) has no counterpart: scratch lives in shared memory."""
    dev = [hasattr(x, "is_cuda") and x.is_cuda for x in (B, A)]
    if all(dev):
        # tensors already in HBM: femtc_apply_contractions_device_* on the current stream; matrices stay on the host
        import torch
        s = {torch.float64: "f64", torch.float32: "f32"}[B.dtype]
        npdt = np.float64 if s == "f64" else np.float32
        if B.numel() < batch * Nb ** 3 or A.numel() < batch * Na ** 3 or A.dtype != B.dtype:
            raise ValueError("B needs batch*Nb^3 and A batch*Na^3 values of one dtype")
        if not (B.is_contiguous() and A.is_contiguous()):
            raise TypeError("contiguous tensors required")
        mats = [np.ascontiguousarray(I, dtype=npdt) for I in (I0, I1, I2)]
        if any(I.size < Na * Nb for I in mats):
            raise ValueError("interpolation matrices need Na*Nb values")
        _lib.check(getattr(_lib.lib, f"femtc_apply_contractions_device_{s}")(
            Na, Nb, int(bool(Tr)), mats[0].ctypes.data, mats[1].ctypes.data, mats[2].ctypes.data,
            B.data_ptr(), A.data_ptr(), batch, torch.cuda.current_stream().cuda_stream))
        return
    if any(dev):
        raise TypeError("B and A must both be host arrays or both be CUDA tensors")
    s = _suffix(B)
    if B.size < batch * Nb ** 3 or A.size < batch * Na ** 3:
        raise ValueError("B needs batch*Nb^3 and A batch*Na^3 values")
    for I in (I0, I1, I2):
        if I.size < Na * Nb or I.dtype != B.dtype:
            raise ValueError("interpolation matrices need Na*Nb values of the field dtype")
    _lib.check(getattr(_lib.lib, f"femtc_apply_contractions_{s}")(
        Na, Nb, int(bool(Tr)), _host(I0, "I0"), _host(I1, "I1"), _host(I2, "I2"), _host(B, "B"),
        _host(A, "A", True), batch))
