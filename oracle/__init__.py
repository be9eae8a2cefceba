"""TEST INFRASTRUCTURE ONLY.

ctypes bindings for the CPU oracle (oracle/liboracle.so, the plain-C restatement) and,
when present, for the unmodified reference compiled behind a C ABI (oracle/_ref/*.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (femtc_b200/, include/) never does.
"""
from .pyoracle import (  # noqa: F401
    Oracle,
    Reference,
    build,
    have_reference,
    ORDER_NAMES,
    REF_CFGS,
    bench_inputs,
    splitmix,
)
