"""Mirror of reference src/tmm.hpp (the alternative, rectangular formulation) over the C ABI."""
from __future__ import annotations

import numpy as np

from . import _lib
from .linalg import _suffix, _host


def apply_contractions(I0, I1, I2, B, A, Na: int, Nb: int, Tr: bool, batch: int = 1) -> None:
    """apply_contractions<T,Na,Nb,Tr>(I0, I1, I2, B, A, buffer) (reference src/tmm.hpp:91-126):
    A[a0,a1,a2] = sum I0(a0,b0) I1(a1,b1) I2(a2,b2) B[b0,b1,b2]; A is overwritten.  The scratch
    `Buffer` of the reference (tmm.hpp:46-60) has no counterpart: scratch lives in shared memory."""
    s = _suffix(B)
    if B.size < batch * Nb ** 3 or A.size < batch * Na ** 3:
        raise ValueError("B needs batch*Nb^3 and A batch*Na^3 values")
    for I in (I0, I1, I2):
        if I.size < Na * Nb or I.dtype != B.dtype:
            raise ValueError("interpolation matrices need Na*Nb values of the field dtype")
    _lib.check(getattr(_lib.lib, f"femtc_apply_contractions_{s}")(
        Na, Nb, int(bool(Tr)), _host(I0, "I0"), _host(I1, "I1"), _host(I2, "I2"), _host(B, "B"),
        _host(A, "A", True), batch))
