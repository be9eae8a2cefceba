"""Mirror of the reference's namespace `operators` (reference src/mass.hpp, src/stiffness.hpp).

Fields may be numpy arrays (host-pointer ABI: synchronous, staged through the GPU in pipelined
chunks) or torch CUDA tensors (device-pointer ABI: enqueued on torch's current stream).  phi is
always host data (numpy array or CPU tensor).  There is no CPU implementation behind these calls.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .linalg import Order

REF_EXACT = _lib.STIFFNESS_REF_EXACT
FULL = _lib.STIFFNESS_FULL

# what the reference prints to std::cout before returning with W untouched
_MESSAGES = {
    _lib.ERR_DEGREE: "degree not supported",      # mass.hpp:135, stiffness.hpp:185
    _lib.ERR_ORDER: "order not supported",        # mass.hpp:166, stiffness.hpp:215
    _lib.ERR_BLOCK: "Block size not supported",   # mass.hpp:191,216, stiffness.hpp:240,265
}


def _is_device(x) -> bool:
    return hasattr(x, "data_ptr") and getattr(x, "is_cuda", False)


def _phi_host(phi):
    if hasattr(phi, "data_ptr"):
        if phi.is_cuda:
            raise TypeError("phi must live on the host (it is passed to the kernels as parameters)")
        phi = phi.numpy()
    phi = np.ascontiguousarray(phi)
    return phi


def _suffix_np(dt) -> str:
    dt = np.dtype(dt)
    if dt == np.float64:
        return "f64"
    if dt == np.float32:
        return "f32"
    raise TypeError(f"unsupported dtype {dt}")


def _suffix_any(x) -> str:
    if _is_device(x):
        import torch
        return {torch.float64: "f64", torch.float32: "f32"}[x.dtype]
    return _suffix_np(x.dtype)


def _ptr(x, name, suffix):
    if _is_device(x):
        if not x.is_contiguous() or _suffix_any(x) != suffix:
            raise TypeError(f"{name}: contiguous tensor of matching dtype required")
        return x.data_ptr()
    if not isinstance(x, np.ndarray) or not x.flags.c_contiguous or _suffix_np(x.dtype) != suffix:
        raise TypeError(f"{name}: C-contiguous numpy array of matching dtype required")
    return x.ctypes.data


def _finish(rc: int) -> int:
    if rc in _MESSAGES:
        print(_MESSAGES[rc])  # the reference's behaviour: message on stdout, W untouched, no throw
        return rc
    _lib.check(rc)
    return rc


def _current_stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def mass_operator(phi, U, W, detJ, num_cells: int, degree: int, order=Order.ijk, Mb: int = 0, Nb: int = 0,
                  checksum=None) -> int:
    """operators::mass_operator<T>(phi, U, W, detJ, num_cells, degree, order, Mb, Nb)
    (reference src/mass.hpp:196-199).  W = (AxAxA) diag(detJ) (AxAxA) U per cell.
    `checksum` (device path only): a float64 CUDA tensor of 4 values that receives {sum, sum of squares, max |w|, n}
    of W from the operator's own epilogue (femtc_mass_apply_device_ex_*)."""
    phi = _phi_host(phi)
    s = _suffix_np(phi.dtype)
    _check_sizes(num_cells, degree, (U, "U", 1), (W, "W", 1), (detJ, "detJ", 1))
    dev = [_is_device(x) for x in (U, W, detJ)]
    if all(dev):
        # knobs are validated like the reference does even though the device ABI does not take them
        rc = _knob_code(degree, order, Mb, Nb)
        if rc:
            return _finish(rc)
        if checksum is not None:
            rc = getattr(_lib.lib, f"femtc_mass_apply_device_ex_{s}")(
                phi.ctypes.data, phi.size, _ptr(U, "U", s), _ptr(W, "W", s), _ptr(detJ, "detJ", s),
                int(num_cells), int(degree), _checksum_ptr(checksum), _current_stream())
        else:
            rc = getattr(_lib.lib, f"femtc_mass_apply_device_{s}")(
                phi.ctypes.data, phi.size, _ptr(U, "U", s), _ptr(W, "W", s), _ptr(detJ, "detJ", s),
                int(num_cells), int(degree), _current_stream())
    elif not any(dev):
        rc = getattr(_lib.lib, f"femtc_mass_apply_{s}")(
            phi.ctypes.data, phi.size, _ptr(U, "U", s), _ptr(W, "W", s), _ptr(detJ, "detJ", s),
            int(num_cells), int(degree), int(order), int(Mb), int(Nb))
    else:
        raise TypeError("U, W, detJ must all be host arrays or all be CUDA tensors")
    return _finish(rc)


def _checksum_ptr(x):
    import torch
    if not (_is_device(x) and x.dtype == torch.float64 and x.numel() >= 4 and x.is_contiguous()):
        raise TypeError("checksum: contiguous float64 CUDA tensor of 4 values required")
    return x.data_ptr()


def stiffness_operator(phi, U, W, G, num_cells: int, degree: int, order=Order.ijk, Mb: int = 0, Nb: int = 0,
                       mode: int = REF_EXACT, checksum=None) -> int:
    """operators::stiffness_operator<T>(phi, U, W, G, num_cells, degree, order, Mb, Nb)
    (reference src/stiffness.hpp:245-248).  mode=REF_EXACT reproduces the reference as coded,
    mode=FULL the intended x+y+z operator (see include/femtc_b200.h)."""
    phi = _phi_host(phi)
    s = _suffix_np(phi.dtype)
    _check_sizes(num_cells, degree, (U, "U", 1), (W, "W", 1), (G, "G", 6))
    dev = [_is_device(x) for x in (U, W, G)]
    if all(dev):
        rc = _knob_code(degree, order, Mb, Nb)
        if rc:
            return _finish(rc)
        if checksum is not None:
            rc = getattr(_lib.lib, f"femtc_stiffness_apply_device_ex_{s}")(
                phi.ctypes.data, phi.size, _ptr(U, "U", s), _ptr(W, "W", s), _ptr(G, "G", s),
                int(num_cells), int(degree), int(mode), _checksum_ptr(checksum), _current_stream())
        else:
            rc = getattr(_lib.lib, f"femtc_stiffness_apply_device_{s}")(
                phi.ctypes.data, phi.size, _ptr(U, "U", s), _ptr(W, "W", s), _ptr(G, "G", s),
                int(num_cells), int(degree), int(mode), _current_stream())
    elif not any(dev):
        rc = getattr(_lib.lib, f"femtc_stiffness_apply_{s}")(
            phi.ctypes.data, phi.size, _ptr(U, "U", s), _ptr(W, "W", s), _ptr(G, "G", s),
            int(num_cells), int(degree), int(order), int(Mb), int(Nb), int(mode))
    else:
        raise TypeError("U, W, G must all be host arrays or all be CUDA tensors")
    return _finish(rc)


def _check_sizes(num_cells, degree, *fields) -> None:
    """The reference trusts its spans; here a field shorter than num_cells*(P+1)^3 is an error."""
    if not 1 <= degree <= 25:
        return  # reported through the reference's "degree not supported" path
    need = int(num_cells) * (degree + 1) ** 3
    for x, name, mult in fields:
        have = x.numel() if _is_device(x) else x.size
        if have < need * mult:
            raise ValueError(f"{name} holds {have} values, {need * mult} needed")


def _knob_code(degree, order, Mb, Nb) -> int:
    """The reference's dispatch order: Nb, Mb (mass.hpp:201,176), order (:145), degree (:57)."""
    if Nb not in (0, 4, 8, 16) or Mb not in (0, 4, 8, 16):
        return _lib.ERR_BLOCK
    if int(order) not in range(6):
        return _lib.ERR_ORDER
    if not 1 <= degree <= 25:
        return _lib.ERR_DEGREE
    return _lib.OK


class ResidentOperator:
    """Operator handle with the geometry resident in HBM (femtc_operator_* in include/femtc_b200.h).

    The reference's drivers re-apply the operator to the same detJ / G (bench_mass.cpp:95-105,
    bench_stiffness.cpp:95-105); a handle uploads them once, `apply` then moves only U and W.
    kind: "mass" (geom = detJ) or "stiffness" (geom = G).  `geom` is a host numpy array."""

    def __init__(self, kind: str, phi, geom, num_cells: int, degree: int, mode: int = REF_EXACT):
        import ctypes as C
        if kind not in ("mass", "stiffness"):
            raise ValueError("kind must be 'mass' or 'stiffness'")
        phi = _phi_host(phi)
        s = _suffix_np(phi.dtype)
        self.kind, self.degree, self.num_cells, self.suffix = kind, int(degree), int(num_cells), s
        if 1 <= degree <= 25:
            need = int(num_cells) * (degree + 1) ** 3 * (6 if kind == "stiffness" else 1)
            if geom.size < need:
                raise ValueError(f"geometry holds {geom.size} values, {need} needed")
        h = C.c_void_p()
        rc = getattr(_lib.lib, f"femtc_operator_create_{s}")(
            C.byref(h), 1 if kind == "stiffness" else 0, phi.ctypes.data, phi.size,
            _ptr(geom, "geom", s), int(num_cells), int(degree), int(mode))
        self._h = h if rc == 0 else None
        if rc:
            _finish(rc)
            raise _lib.FemtcError(rc, _lib.last_error())

    @property
    def device_bytes(self) -> int:
        return int(_lib.lib.femtc_operator_device_bytes(self._h))

    def apply(self, U, W) -> int:
        """W = op(U).  U, W both host arrays (synchronous, pipelined copies) or both CUDA tensors
        (asynchronous on the current stream)."""
        if self._h is None:
            raise RuntimeError("operator handle was destroyed")
        _check_sizes(self.num_cells, self.degree, (U, "U", 1), (W, "W", 1))
        dev = [_is_device(x) for x in (U, W)]
        if all(dev):
            rc = _lib.lib.femtc_operator_apply_device(self._h, _ptr(U, "U", self.suffix), _ptr(W, "W", self.suffix),
                                                      _current_stream())
        elif not any(dev):
            rc = _lib.lib.femtc_operator_apply(self._h, _ptr(U, "U", self.suffix), _ptr(W, "W", self.suffix))
        else:
            raise TypeError("U and W must both be host arrays or both be CUDA tensors")
        return _finish(rc)

    def destroy(self) -> None:
        if getattr(self, "_h", None) is not None:
            lib = getattr(_lib, "lib", None) if _lib is not None else None  # module globals may be gone at exit
            if lib is not None:
                lib.femtc_operator_destroy(self._h)
            self._h = None

    __del__ = destroy

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.destroy()
