"""TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ORDER_NAMES = ("ijk", "ikj", "jik", "jki", "kij", "kji")  # linalg.hpp:11-19
# (order, Mb, Nb) triples instantiated by oracle/ref_shim_*.cpp
REF_CFGS = ((0, 0, 0), (1, 0, 0), (4, 8, 16), (3, 4, 8))

_DT = {np.dtype("float64"): ("f64", C.c_double), np.dtype("float32"): ("f32", C.c_float)}


def _suf(dtype):
    return _DT[np.dtype(dtype)][0]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def build(verbose: bool = False) -> None:
    """Compile liboracle.so and, where /root/reference exists, oracle/_ref/*.so."""
    r = subprocess.run(["make", "-C", str(HERE), "all"], capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout, r.stderr)
    if r.returncode:
        raise RuntimeError("oracle build failed")


def have_reference() -> bool:
    return (HERE / "_ref" / "libfemtc_ref.so").exists()


class Oracle:
    """The plain-C restatement (oracle.c)."""

    def __init__(self):
        path = HERE / "liboracle.so"
        if not path.exists():
            build()
        self.lib = C.CDLL(str(path))
        self.lib.oracle_string2order.argtypes = [C.c_char_p]
        for s in ("f64", "f32"):
            getattr(self.lib, f"oracle_fill_splitmix_{s}").argtypes = [
                C.c_void_p, C.c_size_t, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int]
            getattr(self.lib, f"oracle_fill_splitmix_{s}").restype = None
            getattr(self.lib, f"oracle_fill_rand_{s}").argtypes = [C.c_void_p, C.c_size_t]
            getattr(self.lib, f"oracle_fill_rand_{s}").restype = None
            getattr(self.lib, f"oracle_checksum_{s}").argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
            getattr(self.lib, f"oracle_checksum_{s}").restype = None
            getattr(self.lib, f"oracle_mass_{s}").argtypes = [C.c_void_p] * 4 + [C.c_size_t] + [C.c_int] * 4
            getattr(self.lib, f"oracle_stiffness_{s}").argtypes = [C.c_void_p] * 4 + [C.c_size_t] + [C.c_int] * 5
            getattr(self.lib, f"oracle_gemm_blocked_{s}").argtypes = [C.c_int] * 6 + [C.c_void_p] * 3
            getattr(self.lib, f"oracle_gemm_micro_{s}").argtypes = [C.c_int] * 7 + [C.c_void_p] * 3
            getattr(self.lib, f"oracle_apply_contractions_{s}").argtypes = [C.c_int] * 3 + [C.c_void_p] * 5

    def string2order(self, s: str) -> int:
        return self.lib.oracle_string2order(s.encode())

    def mass(self, phi, U, detJ, degree, order=0, Mb=0, Nb=0):
        nd = (degree + 1) ** 3
        nc = U.size // nd
        W = np.zeros_like(U)
        rc = getattr(self.lib, f"oracle_mass_{_suf(U.dtype)}")(
            _p(phi), _p(U), _p(W), _p(detJ), nc, degree, order, Mb, Nb)
        return rc, W

    def stiffness(self, phi, U, G, degree, order=0, Mb=0, Nb=0, mode=0):
        nd = (degree + 1) ** 3
        nc = U.size // nd
        W = np.zeros_like(U)
        rc = getattr(self.lib, f"oracle_stiffness_{_suf(U.dtype)}")(
            _p(phi), _p(U), _p(W), _p(G), nc, degree, order, Mb, Nb, mode)
        return rc, W

    def gemm_blocked(self, k, m, n, order, MB, NB, a, b, c):
        return getattr(self.lib, f"oracle_gemm_blocked_{_suf(a.dtype)}")(
            k, m, n, order, MB, NB, _p(a), _p(b), _p(c))

    def gemm_micro(self, k, m, nc, lda, ldb, ldc, order, a, b, c):
        return getattr(self.lib, f"oracle_gemm_micro_{_suf(a.dtype)}")(
            k, m, nc, lda, ldb, ldc, order, _p(a), _p(b), _p(c))

    def apply_contractions(self, Na, Nb, Tr, I0, I1, I2, B):
        A = np.zeros(Na ** 3, dtype=B.dtype)
        rc = getattr(self.lib, f"oracle_apply_contractions_{_suf(B.dtype)}")(
            Na, Nb, int(Tr), _p(I0), _p(I1), _p(I2), _p(B), _p(A))
        return rc, A

    def srand(self, seed: int):
        self.lib.oracle_srand(C.c_uint(seed))

    def fill_rand(self, n, dtype):
        x = np.empty(n, dtype=dtype)
        getattr(self.lib, f"oracle_fill_rand_{_suf(dtype)}")(_p(x), n)
        return x

    def fill_splitmix(self, n, dtype, seed=1, stream=0, offset=0, centred=False):
        x = np.empty(n, dtype=dtype)
        getattr(self.lib, f"oracle_fill_splitmix_{_suf(dtype)}")(_p(x), n, seed, stream, offset, int(centred))
        return x

    def checksum(self, x):
        out = np.zeros(4)
        getattr(self.lib, f"oracle_checksum_{_suf(x.dtype)}")(_p(x), x.size, _p(out))
        return out


def _cpu_flags() -> set:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


class Reference:
    """The unmodified reference headers compiled behind a C ABI (oracle/_ref).

    fast=False -> -O2 IEEE build (parity anchor); fast=True -> -Ofast build at the widest
    ISA level this host supports (timed CPU baseline).
    """

    def __init__(self, fast: bool = False):
        if fast:
            flags = _cpu_flags()
            v4 = {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= flags
            name = "libfemtc_ref_fast_v4.so" if v4 else "libfemtc_ref_fast_v3.so"
            self.isa = "x86-64-v4" if v4 else "x86-64-v3"
        else:
            name = "libfemtc_ref.so"
            self.isa = "x86-64 -O2"
        path = HERE / "_ref" / name
        if not path.exists():
            raise FileNotFoundError(f"{path} missing: run `make -C oracle` where /root/reference exists")
        self.path = path
        self.lib = C.CDLL(str(path))
        for s in ("f64", "f32"):
            getattr(self.lib, f"femtc_ref_mass_{s}").argtypes = [C.c_void_p] * 4 + [C.c_size_t, C.c_int, C.c_int]
            getattr(self.lib, f"femtc_ref_stiffness_{s}").argtypes = [C.c_void_p] * 4 + [C.c_size_t, C.c_int, C.c_int]
            getattr(self.lib, f"femtc_ref_gemm_blocked_{s}").argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 3
            getattr(self.lib, f"femtc_ref_gemm_micro_{s}").argtypes = [C.c_int] + [C.c_void_p] * 3
            getattr(self.lib, f"femtc_ref_apply_contractions_{s}").argtypes = [C.c_int] * 3 + [C.c_void_p] * 5
        self.lib.femtc_ref_string2order.argtypes = [C.c_char_p]

    def string2order(self, s: str) -> int:
        return self.lib.femtc_ref_string2order(s.encode())

    def mass(self, phi, U, detJ, degree, cfg=0, W=None):
        nd = (degree + 1) ** 3
        if W is None:
            W = np.zeros_like(U)
        rc = getattr(self.lib, f"femtc_ref_mass_{_suf(U.dtype)}")(
            _p(phi), _p(U), _p(W), _p(detJ), U.size // nd, degree, cfg)
        return rc, W

    def stiffness(self, phi, U, G, degree, cfg=0, W=None):
        nd = (degree + 1) ** 3
        if W is None:
            W = np.zeros_like(U)
        rc = getattr(self.lib, f"femtc_ref_stiffness_{_suf(U.dtype)}")(
            _p(phi), _p(U), _p(W), _p(G), U.size // nd, degree, cfg)
        return rc, W

    def gemm_blocked(self, P, cfg, a, b, c):
        return getattr(self.lib, f"femtc_ref_gemm_blocked_{_suf(a.dtype)}")(P, cfg, _p(a), _p(b), _p(c))

    def gemm_micro(self, order, a, b, c):
        return getattr(self.lib, f"femtc_ref_gemm_micro_{_suf(a.dtype)}")(order, _p(a), _p(b), _p(c))

    def apply_contractions(self, Na, Nb, Tr, I0, I1, I2, B):
        A = np.zeros(Na ** 3, dtype=B.dtype)
        Bc = B.copy()
        rc = getattr(self.lib, f"femtc_ref_apply_contractions_{_suf(B.dtype)}")(
            Na, Nb, int(Tr), _p(I0), _p(I1), _p(I2), _p(Bc), _p(A))
        return rc, A


def bench_inputs(oracle: Oracle, op: str, degree: int, ndofs: int, dtype, seed: int = 1,
                 full_phi: bool = True):
    """Inputs exactly as bench_mass.cpp:74-88 / bench_stiffness.cpp:74-88 make them:
    glibc rand() in the order phi -> U -> detJ|G.  For stiffness phi gets 2*(P+1)^2 values
    (the operator's real contract) unless full_phi=False."""
    n1 = degree + 1
    nd = n1 ** 3
    nc = ndofs // nd
    oracle.srand(seed)
    nphi = n1 * n1 * (2 if (op == "stiffness" and full_phi) else 1)
    phi = oracle.fill_rand(nphi, dtype)
    U = oracle.fill_rand(nc * nd, dtype)
    geom = oracle.fill_rand(nc * nd * (6 if op == "stiffness" else 1), dtype)
    return phi, U, geom, nc


def splitmix(oracle: Oracle, op: str, degree: int, num_cells: int, dtype, seed: int = 1,
             centred: bool = False, cell0: int = 0):
    """Counter-based inputs (streams: phi=0, U=1, detJ/G=2) for cells [cell0, cell0+num_cells)."""
    n1 = degree + 1
    nd = n1 ** 3
    phi = oracle.fill_splitmix(2 * n1 * n1, dtype, seed, 0, 0, centred)
    U = oracle.fill_splitmix(num_cells * nd, dtype, seed, 1, cell0 * nd, centred)
    w = 6 if op == "stiffness" else 1
    geom = oracle.fill_splitmix(num_cells * nd * w, dtype, seed, 2, cell0 * nd * w, centred)
    return phi, U, geom
