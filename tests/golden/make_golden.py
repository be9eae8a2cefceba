"""Generates tests/golden/reference_outputs.npz from the UNMODIFIED reference
(oracle/_ref/libfemtc_ref.so = /root/reference/src/*.hpp compiled -O2, IEEE, no contraction).

Run in the build container (where /root/reference exists):
    python tests/golden/make_golden.py
Inputs are not stored: they are regenerated from the counter-based generator
(oracle.splitmix, seed/stream/centred as recorded in the key), only the reference's outputs are.
Key format:  <op>_P<degree>_<dtype>_c<cells>_s<seed>_<centred>   value: W (or C / A for the
small-GEMM and contraction cases).
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import Oracle, Reference, splitmix  # noqa: E402

CELLS = {1: 37, 2: 21, 3: 13, 4: 9, 5: 7, 6: 5, 7: 4, 8: 3}
SEED = 7


def main():
    o, r = Oracle(), Reference()
    out = {}
    for dt in ("float64", "float32"):
        for P in range(1, 9):
            nc = CELLS[P]
            for centred in (0, 1):
                phi, U, J = splitmix(o, "mass", P, nc, dt, SEED, bool(centred))
                rc, W = r.mass(phi, U, J, P, cfg=0)
                assert rc == 0
                out[f"mass_P{P}_{dt}_c{nc}_s{SEED}_{centred}"] = W
                phi, U, G = splitmix(o, "stiffness", P, nc, dt, SEED, bool(centred))
                rc, W = r.stiffness(phi, U, G, P, cfg=0)
                assert rc == 0
                out[f"stiffness_P{P}_{dt}_c{nc}_s{SEED}_{centred}"] = W
        # gemm_blocked at the operator shape, tiling (8,16) = cfg 2 (tests/test_gemm.cpp:41)
        for P in (1, 2, 3, 4, 7, 8, 12, 25):
            n1 = P + 1
            a = o.fill_splitmix(n1 * n1, dt, SEED, 3, 0, True)
            b = o.fill_splitmix(n1 ** 3, dt, SEED, 4, 0, True)
            c = np.zeros(n1 ** 3, dtype=dt)
            assert r.gemm_blocked(P, 2, a, b, c) == 0
            out[f"gemmblocked_P{P}_{dt}"] = c
        # gemm_micro at the shape of tests/test_micro.cpp:44, accumulating into a non-zero c
        a = o.fill_splitmix(4 * 26, dt, SEED, 5, 0, True)
        b = o.fill_splitmix(26 * 16, dt, SEED, 6, 0, True)
        c = o.fill_splitmix(4 * 16, dt, SEED, 7, 0, True)
        assert r.gemm_micro(0, a, b, c) == 0
        out[f"gemmmicro_{dt}"] = c
        for (Na, Nb) in ((4, 4), (6, 4), (3, 5), (9, 9)):
            for Tr in (0, 1):
                I = [o.fill_splitmix(Na * Nb, dt, SEED, 8 + i, 0, True) for i in range(3)]
                B = o.fill_splitmix(Nb ** 3, dt, SEED, 11, 0, True)
                rc, A = r.apply_contractions(Na, Nb, Tr, *I, B)
                assert rc == 0
                out[f"tmm_{Na}_{Nb}_{Tr}_{dt}"] = A
    path = Path(__file__).with_name("reference_outputs.npz")
    np.savez_compressed(path, **out)
    print(path, path.stat().st_size, "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
