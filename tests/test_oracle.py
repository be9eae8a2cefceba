"""CPU tests: pin the oracle (oracle/oracle.c) to the reference.

Anchors, strongest first:
  1. bit-for-bit equality with the UNMODIFIED reference headers compiled into
     oracle/_ref/libfemtc_ref.so (skipped only when that file is absent);
  2. bit-for-bit equality with the committed outputs of that library (tests/golden);
  3. the known-answer table of SURVEY.md section 8c (glibc rand, srand(1), Ndofs = 1e6);
  4. the closed forms of SURVEY.md section 8c, via numpy einsum.
"""
from pathlib import Path

import numpy as np
import pytest

from oracle import REF_CFGS, bench_inputs, splitmix

GOLDEN = np.load(Path(__file__).parent / "golden" / "reference_outputs.npz")
DTYPES = ("float64", "float32")


def _parse(key):
    op, P, dt, c, s, centred = key.split("_")
    return op, int(P[1:]), dt, int(c[1:]), int(s[1:]), bool(int(centred))


# ---- 2. golden fixtures ----------------------------------------------------------------------
@pytest.mark.parametrize("key", [k for k in GOLDEN.files if k.startswith(("mass_", "stiffness_"))])
def test_operator_matches_golden(oracle, key):
    op, P, dt, nc, seed, centred = _parse(key)
    phi, U, geom = splitmix(oracle, op, P, nc, dt, seed, centred)
    if op == "mass":
        rc, W = oracle.mass(phi, U, geom, P)
    else:
        rc, W = oracle.stiffness(phi, U, geom, P, mode=0)
    assert rc == 0
    assert np.array_equal(W, GOLDEN[key]), key


@pytest.mark.parametrize("dt", DTYPES)
def test_small_gemm_and_tmm_match_golden(oracle, dt):
    o = oracle
    for P in (1, 2, 3, 4, 7, 8, 12, 25):
        n1 = P + 1
        a = o.fill_splitmix(n1 * n1, dt, 7, 3, 0, True)
        b = o.fill_splitmix(n1 ** 3, dt, 7, 4, 0, True)
        for order, MB, NB in ((4, 8, 16), (0, 0, 0), (3, 4, 8), (5, 16, 4)):
            c = np.full(n1 ** 3, 123.0, dtype=dt)  # must be overwritten
            assert o.gemm_blocked(n1, n1, n1 * n1, order, MB, NB, a, b, c) == 0
            assert np.array_equal(c, GOLDEN[f"gemmblocked_P{P}_{dt}"])
    a = o.fill_splitmix(4 * 26, dt, 7, 5, 0, True)
    b = o.fill_splitmix(26 * 16, dt, 7, 6, 0, True)
    for order in range(6):
        c = o.fill_splitmix(4 * 16, dt, 7, 7, 0, True)
        assert o.gemm_micro(26, 4, 16, 4, 16, 16, order, a, b, c) == 0
        assert np.array_equal(c, GOLDEN[f"gemmmicro_{dt}"])
    for (Na, Nb) in ((4, 4), (6, 4), (3, 5), (9, 9)):
        for Tr in (0, 1):
            I = [o.fill_splitmix(Na * Nb, dt, 7, 8 + i, 0, True) for i in range(3)]
            B = o.fill_splitmix(Nb ** 3, dt, 7, 11, 0, True)
            rc, A = o.apply_contractions(Na, Nb, Tr, *I, B)
            assert rc == 0
            assert np.array_equal(A, GOLDEN[f"tmm_{Na}_{Nb}_{Tr}_{dt}"])


# ---- 1. the compiled reference itself -----------------------------------------------------------
@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("P", range(1, 11))
def test_operators_bitexact_vs_reference(oracle, reference, P, dt):
    nc = 5
    for centred in (False, True):
        phi, U, J = splitmix(oracle, "mass", P, nc, dt, 3, centred)
        _, _, G = splitmix(oracle, "stiffness", P, nc, dt, 3, centred)
        for cfg, (order, Mb, Nb) in enumerate(REF_CFGS):
            rc, Wr = reference.mass(phi, U, J, P, cfg)
            rc2, Wo = oracle.mass(phi, U, J, P, order, Mb, Nb)
            assert rc == 0 and rc2 == 0
            assert np.array_equal(Wr, Wo)
            rc, Wr = reference.stiffness(phi, U, G, P, cfg)
            rc2, Wo = oracle.stiffness(phi, U, G, P, order, Mb, Nb, mode=0)
            assert rc == 0 and rc2 == 0
            assert np.array_equal(Wr, Wo)


def test_gemm_layer_bitexact_vs_reference(oracle, reference):
    rng = np.random.default_rng(0)
    for dt in DTYPES:
        for P in range(1, 26):
            n1 = P + 1
            a = rng.random(n1 * n1).astype(dt) - 0.5
            b = rng.random(n1 ** 3).astype(dt) - 0.5
            for cfg, (order, MB, NB) in enumerate(REF_CFGS):
                cr = np.zeros(n1 ** 3, dtype=dt)
                co = np.ones(n1 ** 3, dtype=dt)
                assert reference.gemm_blocked(P, cfg, a, b, cr) == 0
                assert oracle.gemm_blocked(n1, n1, n1 * n1, order, MB, NB, a, b, co) == 0
                assert np.array_equal(cr, co)
        a = rng.random(4 * 26).astype(dt)
        b = rng.random(26 * 16).astype(dt)
        c0 = rng.random(4 * 16).astype(dt)
        for order in range(6):
            cr, co = c0.copy(), c0.copy()
            assert reference.gemm_micro(order, a, b, cr) == 0
            assert oracle.gemm_micro(26, 4, 16, 4, 16, 16, order, a, b, co) == 0
            assert np.array_equal(cr, co)
        for (Na, Nb) in ((2, 2), (3, 3), (5, 5), (8, 6), (5, 4), (4, 7)):
            for Tr in (0, 1):
                I = [rng.random(Na * Nb).astype(dt) - 0.5 for _ in range(3)]
                B = rng.random(Nb ** 3).astype(dt) - 0.5
                _, Ar = reference.apply_contractions(Na, Nb, Tr, *I, B)
                _, Ao = oracle.apply_contractions(Na, Nb, Tr, *I, B)
                assert np.array_equal(Ar, Ao)


def test_string2order_vs_reference(oracle, reference):
    for s in ("ijk", "ikj", "jik", "jki", "kij", "kji", "", "xyz", "IJK", "ijkk"):
        assert oracle.string2order(s) == reference.string2order(s)


# ---- 3. known-answer table ---------------------------------------------------------
KAT = [
    # op, dtype, P, cells, U[0], W[0], W[last], sum W
    ("mass", "float64", 1, 125000, 0.911647357936784, 3.13506034250607, 0.505407201594529, 1839852.15173637),
    ("mass", "float64", 3, 15625, 0.635711727959901, 34.2241917029131, 90.1750055494434, 53300189.7580895),
    ("mass", "float64", 4, 8000, 0.400944394246183, 52.3660109789621, 142.499209158142, 78591281.9853732),
    ("mass", "float64", 8, 1371, 0.165974166321556, 2796.96318901602, 1767.55356411748, 2925805575.85453),
    ("mass", "float32", 3, 15625, 0.63571173, 34.224194, 90.175003, 53300194),
    ("stiffness", "float64", 1, 125000, 0.277774710803188, 7.61411278163771, 2.16075018187585, 3800150.45247559),
    ("stiffness", "float64", 3, 15625, 0.612639832595661, 99.8039269774196, 188.837293386733, 91770408.7513072),
    ("stiffness", "float64", 4, 8000, 0.525995350222101, 189.310940684047, 424.694454500219, 261744389.133521),
    ("stiffness", "float64", 8, 1371, 0.0519388187918527, 8690.92286397276, 6582.81502646409, 8330517081.3165),
    ("stiffness", "float32", 3, 15625, 0.61263984, 99.80394, 188.83731, 91770413),
]


@pytest.mark.parametrize("op,dt,P,cells,u0,w0,wl,sw", KAT)
def test_known_answers(oracle, op, dt, P, cells, u0, w0, wl, sw):
    phi, U, geom, nc = bench_inputs(oracle, op, P, 10 ** 6, dt)
    assert nc == cells
    tol = 2e-14 if dt == "float64" else 2e-7
    assert abs(phi[0] - 0.84018771715471) < 1e-7
    assert abs(U[0] - u0) <= tol * abs(u0) + 1e-15
    rc, W = (oracle.mass(phi, U, geom, P) if op == "mass" else oracle.stiffness(phi, U, geom, P))
    assert rc == 0
    assert abs(W[0] - w0) <= tol * abs(w0)
    assert abs(W[-1] - wl) <= tol * abs(wl)
    # the table's fp32 sums were accumulated in fp32 by the survey probe: compare loosely there
    assert abs(W.astype(np.float64).sum() - sw) <= (1e-13 if dt == "float64" else 1e-6) * abs(sw)


# ---- 4. closed forms -------------------------------------------------------------------------------
def _mats(phi, n1):
    A = phi[: n1 * n1].reshape(n1, n1).T.astype(np.float64)      # A[i,p] = phi[i + p*n1]
    D = phi[n1 * n1: 2 * n1 * n1].reshape(n1, n1).T.astype(np.float64)
    return A, D


def _kron3(M0, M1, M2, X):
    return np.einsum("ia,jb,kc,nabc->nijk", M0, M1, M2, X, optimize=True)


@pytest.mark.parametrize("P", range(1, 9))
def test_closed_forms(oracle, P):
    n1, nc = P + 1, 4
    phi, U, J = splitmix(oracle, "mass", P, nc, "float64", 11, True)
    _, _, G = splitmix(oracle, "stiffness", P, nc, "float64", 11, True)
    A, D = _mats(phi, n1)
    Ut = U.reshape(nc, n1, n1, n1)
    # mass: W = (AxAxA) diag(detJ) (AxAxA) U, same A both ways
    Wm = _kron3(A, A, A, _kron3(A, A, A, Ut) * J.reshape(Ut.shape)).ravel()
    rc, W = oracle.mass(phi, U, J, P)
    assert np.abs(W - Wm).max() <= 1e-13 * np.abs(Wm).max()
    g = G.reshape(nc, n1, n1, n1, 6)
    wx, wy = _kron3(D, A, A, Ut), _kron3(A, D, A, Ut)
    # REF_EXACT
    wz = _kron3(A, A, A, Ut)
    f2 = g[..., 2] * wx + g[..., 1] * wy + g[..., 0] * wz
    Wr = _kron3(A, A, D, f2).ravel()
    rc, W = oracle.stiffness(phi, U, G, P, mode=0)
    assert np.abs(W - Wr).max() <= 1e-13 * np.abs(Wr).max()
    # FULL
    wz = _kron3(A, A, D, Ut)
    f0 = g[..., 5] * wx + g[..., 4] * wy + g[..., 2] * wz
    f1 = g[..., 4] * wx + g[..., 3] * wy + g[..., 1] * wz
    f2 = g[..., 2] * wx + g[..., 1] * wy + g[..., 0] * wz
    Wf = (_kron3(D, A, A, f0) + _kron3(A, D, A, f1) + _kron3(A, A, D, f2)).ravel()
    rc, W = oracle.stiffness(phi, U, G, P, mode=1)
    assert np.abs(W - Wf).max() <= 1e-13 * np.abs(Wf).max()


def test_oracle_error_codes(oracle):
    phi = np.ones(32)
    U = np.ones(64)
    assert oracle.mass(phi, U, U, 3, 0, 5, 0)[0] == 3      # Block size not supported
    assert oracle.mass(phi, U, U, 3, 7, 0, 0)[0] == 2      # order not supported
    assert oracle.mass(phi, U, U, 26, 0, 0, 0)[0] == 1     # degree not supported
    assert oracle.mass(phi, U, U, 26, 9, 3, 0)[0] == 3     # Nb/Mb are dispatched first (mass.hpp:201,176)
    assert oracle.stiffness(phi, U, np.ones(384), 3, mode=2)[0] == 4


def test_splitmix_ranges_and_offsets(oracle):
    a = oracle.fill_splitmix(1000, "float64", 1, 2, 0)
    b = oracle.fill_splitmix(400, "float64", 1, 2, 600)
    assert np.array_equal(a[600:], b)
    assert 0 <= a.min() and a.max() < 1
    c = oracle.fill_splitmix(1000, "float32", 1, 2, 0, True)
    assert -0.5 <= c.min() and c.max() < 0.5
