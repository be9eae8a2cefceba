"""femtc_b200 -- B200 (sm_100a) mass / stiffness operator action, drop-in for the hot path of
IgorBaratta/femtc.  The product is libfemtc_b200.so (C ABI, include/femtc_b200.h) plus the
replacement C++ headers in include/; this package is the thin Python mirror of the same
interface used by the tests and by bench.py."""
from . import _lib
from ._lib import FemtcError, get_option, kernel_info, last_error, set_option
from .linalg import Order, gemm_blocked, gemm_micro, string2order
from .operators import FULL, REF_EXACT, ResidentOperator, mass_operator, stiffness_operator
from .tmm import apply_contractions

__all__ = [
    "FemtcError", "Order", "string2order", "gemm_blocked", "gemm_micro", "mass_operator",
    "stiffness_operator", "ResidentOperator", "apply_contractions", "REF_EXACT", "FULL", "set_option", "get_option",
    "kernel_info", "last_error", "device",
]


def version() -> str:
    return _lib.lib.femtc_version().decode()


def device_count() -> int:
    return int(_lib.lib.femtc_device_count())
