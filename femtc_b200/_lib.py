"""ctypes binding of libfemtc_b200.so (the C ABI declared in include/femtc_b200.h).

There is no fallback of any kind: if the shared library has not been built the import fails.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG = Path(__file__).resolve().parent
# FEMTC_B200_LIB: another build of the same library (development A/B runs); never a fallback
LIB_PATH = Path(os.environ.get("FEMTC_B200_LIB") or PKG / "libfemtc_b200.so")

OK, ERR_DEGREE, ERR_ORDER, ERR_BLOCK, ERR_MODE, ERR_LENGTH, ERR_CUDA, ERR_ARG = range(8)
STIFFNESS_REF_EXACT, STIFFNESS_FULL = 0, 1


class FemtcError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"femtc_b200 error {code}: {msg}")
        self.code = code


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing. Build it with `python femtc_b200/build.py` "
            "(nvcc, sm_100a). femtc_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(str(LIB_PATH))
    vp, sz, i, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64
    for s in ("f64", "f32"):
        getattr(lib, f"femtc_mass_apply_{s}").argtypes = [vp, sz, vp, vp, vp, sz, i, i, i, i]
        getattr(lib, f"femtc_stiffness_apply_{s}").argtypes = [vp, sz, vp, vp, vp, sz, i, i, i, i, i]
        getattr(lib, f"femtc_mass_apply_device_{s}").argtypes = [vp, sz, vp, vp, vp, sz, i, vp]
        getattr(lib, f"femtc_stiffness_apply_device_{s}").argtypes = [vp, sz, vp, vp, vp, sz, i, i, vp]
        getattr(lib, f"femtc_mass_apply_device_ex_{s}").argtypes = [vp, sz, vp, vp, vp, sz, i, vp, vp]
        getattr(lib, f"femtc_stiffness_apply_device_ex_{s}").argtypes = [vp, sz, vp, vp, vp, sz, i, i, vp, vp]
        getattr(lib, f"femtc_gemm_blocked_{s}").argtypes = [i] * 6 + [vp] * 3
        getattr(lib, f"femtc_gemm_micro_{s}").argtypes = [i] * 7 + [vp] * 3
        getattr(lib, f"femtc_gemm_strided_acc_{s}").argtypes = [i] * 8 + [vp] * 3
        getattr(lib, f"femtc_apply_contractions_{s}").argtypes = [i] * 3 + [vp] * 5 + [sz]
        getattr(lib, f"femtc_apply_contractions_device_{s}").argtypes = [i] * 3 + [vp] * 5 + [sz, vp]
        getattr(lib, f"femtc_checksum_device_{s}").argtypes = [vp, sz, vp, vp]
        getattr(lib, f"femtc_fill_uniform_device_{s}").argtypes = [vp, sz, u64, u64, u64, i, vp]
        getattr(lib, f"femtc_operator_create_{s}").argtypes = [C.POINTER(vp), i, vp, sz, vp, sz, i, i]
    lib.femtc_apply_contractions_device_static_f64.argtypes = [i] * 3 + [vp] * 5 + [sz, vp]
    lib.femtc_stream_probe_device.argtypes = [vp, vp, sz, i, vp]
    lib.femtc_operator_apply.argtypes = [vp, vp, vp]
    lib.femtc_operator_apply_device.argtypes = [vp, vp, vp, vp]
    lib.femtc_operator_device_bytes.argtypes = [vp]
    lib.femtc_operator_device_bytes.restype = sz
    lib.femtc_operator_destroy.argtypes = [vp]
    lib.femtc_operator_destroy.restype = None
    lib.femtc_string2order.argtypes = [C.c_char_p]
    lib.femtc_set_device.argtypes = [i]
    lib.femtc_last_error.restype = C.c_char_p
    lib.femtc_version.restype = C.c_char_p
    lib.femtc_host_alloc.argtypes = [sz]
    lib.femtc_host_alloc.restype = vp
    lib.femtc_host_free.argtypes = [vp]
    lib.femtc_host_free.restype = None
    lib.femtc_release_cache.restype = None
    lib.femtc_set_option.argtypes = [C.c_char_p, C.c_longlong]
    lib.femtc_get_option.argtypes = [C.c_char_p]
    lib.femtc_get_option.restype = C.c_longlong
    lib.femtc_kernel_info.argtypes = [i, i, i, i, C.POINTER(C.c_longlong)]
    return lib


lib = _load()


def last_error() -> str:
    return lib.femtc_last_error().decode()


def check(rc: int) -> None:
    if rc != OK:
        raise FemtcError(rc, last_error())


def set_option(name: str, value: int) -> None:
    check(lib.femtc_set_option(name.encode(), int(value)))


def get_option(name: str) -> int:
    return int(lib.femtc_get_option(name.encode()))


def kernel_info(op: str, dtype_bytes: int, degree: int, mode: int = 0) -> dict:
    info = (C.c_longlong * 8)()
    check(lib.femtc_kernel_info(0 if op == "mass" else 1, dtype_bytes, degree, mode, info))
    keys = ("family", "cells_per_tile", "threads", "smem_bytes", "blocks_per_sm", "regs", "grid", "detail")
    return dict(zip(keys, [int(v) for v in info]))
