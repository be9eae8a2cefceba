"""Mirror of the reference's namespace `linalg` (reference src/linalg.hpp) over the C ABI."""
from __future__ import annotations

import enum

import numpy as np

from . import _lib


class Order(enum.IntEnum):
    """linalg::Order (reference src/linalg.hpp:11-19); the integer values are ABI."""
    ijk = 0
    ikj = 1
    jik = 2
    jki = 3
    kij = 4
    kji = 5


def string2order(order_str: str) -> Order:
    """linalg::string2order (reference src/linalg.hpp:21-37): raises RuntimeError("Invalid loop
    order") where the reference throws std::runtime_error with that text."""
    code = _lib.lib.femtc_string2order(str(order_str).encode())
    if code < 0:
        raise RuntimeError("Invalid loop order")
    return Order(code)


def _suffix(a: np.ndarray) -> str:
    if a.dtype == np.float64:
        return "f64"
    if a.dtype == np.float32:
        return "f32"
    raise TypeError(f"unsupported dtype {a.dtype}: T must be float or double (bench_mass.cpp:192-195)")


def _host(a, name, writable=False):
    if not isinstance(a, np.ndarray) or not a.flags.c_contiguous or (writable and not a.flags.writeable):
        raise TypeError(f"{name} must be a C-contiguous numpy array")
    return a.ctypes.data


def gemm_blocked(k: int, m: int, n: int, order, MB: int, NB: int, a, b, c) -> None:
    """linalg::gemm_blocked<T,k,m,n,order,MB,NB>(a,b,c) (reference src/linalg.hpp:101-182).
    C (m x n, column-major) = A (m x k, column-major) . B (k x n, row-major); C is overwritten."""
    s = _suffix(a)
    _lib.check(getattr(_lib.lib, f"femtc_gemm_blocked_{s}")(
        k, m, n, int(order), MB, NB, _host(a, "a"), _host(b, "b"), _host(c, "c", True)))


def gemm_micro(k: int, m: int, nc: int, lda: int, ldb: int, ldc: int, a, b, c, order=Order.ijk) -> None:
    """linalg::gemm_micro<T,k,m,nc,lda,ldb,ldc,order>(a,b,c) (reference src/linalg.hpp:45-94).
    c (m x nc, row-major) += a (m x k, column-major) . b (k x nc, row-major)."""
    s = _suffix(a)
    _lib.check(getattr(_lib.lib, f"femtc_gemm_micro_{s}")(
        k, m, nc, lda, ldb, ldc, int(order), _host(a, "a"), _host(b, "b"), _host(c, "c", True)))
