"""Choose the shared-memory layouts of the column kernels (degrees 5..8 and the shapes the
slab kernels do not cover) with the bank-conflict model of tools/bank_sim.py, and write
femtc_b200/csrc/column_layouts.cuh.

Column kernels: a block owns C cells; thread t = c*n1^2 + hi*n1 + lo holds, in every phase, one
*column* of its cell tensor X[i0][i1][i2] (all values of the phase's register axis R for fixed
indices on the two thread axes (H, L) <-> (hi, lo)).

    phase C: R = 0, (H, L) = (1, 2)   -- natural order: staged U / W / G are read coalesced
    phase B: R = 1, thread axes {0, 2} in an order chosen here
    phase A: R = 2, thread axes {0, 1} in an order chosen here (also reads a natural-layout
             staging buffer with lane stride n1 or n1^2)

Exchanges X_BC (between B and C) and X_AB (between A and B) go through padded scratch
    addr = c*Sc + i0*s0 + i1*s1 + i2*s2       (elements)
The search minimises the modelled wavefronts of all writes and reads of both exchanges (both
directions cost the same, the access sets are identical) plus the phase-A staging access.

    python tools/gen_column_layouts.py            # writes the header, prints the table
"""
from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tools"))
from bank_sim import wavefronts  # noqa: E402

CELLS = 4  # cells in the modelled block: enough to cover warps that straddle cell boundaries


def phase_addrs(n1, tsize, R, H, L, strides, Sc, rval, cells=CELLS):
    """Per-thread byte address of element (R = rval) for all threads of a block."""
    out = []
    for c in range(cells):
        for hi in range(n1):
            for lo in range(n1):
                idx = [0, 0, 0]
                idx[R] = rval
                idx[H] = hi
                idx[L] = lo
                e = c * Sc + idx[0] * strides[0] + idx[1] * strides[1] + idx[2] * strides[2]
                out.append(e * tsize)
    return out


def access_cost(n1, tsize, R, H, L, strides, Sc):
    """wavefronts and ideal wavefronts for every thread touching every element of its column.
    The pattern of one register value repeats for the others up to a uniform shift when the
    register-axis stride is a multiple of 32 banks, which it is not in general, so all are modelled."""
    tot = 0
    ideal = 0.0
    for r in range(n1):
        a = phase_addrs(n1, tsize, R, H, L, strides, Sc, r)
        for w0 in range(0, len(a), 32):
            lanes = a[w0:w0 + 32]
            lanes = lanes + [None] * (32 - len(lanes))
            tot += wavefronts(lanes, tsize)
            ideal += sum(x is not None for x in lanes) * tsize / 128
    return tot, ideal


def best_exchange(n1, tsize, ph_w, ph_r, families):
    """families: iterable of (strides, minimal cell stride); the cell stride is padded by 0..M."""
    M = 32 if tsize == 4 else 16
    best = None
    for s, cell in families:
        for p3 in range(0, M + 1):
            Sc = cell + p3
            cw, iw = access_cost(n1, tsize, *ph_w, s, Sc)
            cr, ir = access_cost(n1, tsize, *ph_r, s, Sc)
            key = (cw + cr, Sc)
            if best is None or key < best[0]:
                best = (key, s, Sc, cw, cr, iw + ir)
            if cw + cr <= iw + ir + 1e-9:
                break
    return best


def fam_bc(n1, tsize):
    # addr = i2 + i1*n1 + i0*S0 with S0 = n1^2 + pad: writer (phase B, lanes (i0, i2)) and reader
    # (phase C, lanes (i1, i2)) both see bank = thread index when S0 == n1 (mod banks)
    M = 32 if tsize == 4 else 16
    for pad in range(0, M + 1):
        S0 = n1 * n1 + pad
        yield (S0, n1, 1), n1 * S0


def fam_ab(n1, tsize):
    # addr = i2 + i0*s0 + i1*s1, s0 = n1 + p1, s1 = n1*s0 + p2
    for p1 in range(0, 3):
        s0 = n1 + p1
        for p2 in range(0, 9):
            s1 = n1 * s0 + p2
            yield (s0, s1, 1), n1 * s1
    # and the transposed nesting addr = i2 + i1*s1 + i0*s0
    for p1 in range(0, 3):
        s1 = n1 + p1
        for p2 in range(0, 9):
            s0 = n1 * s1 + p2
            yield (s0, s1, 1), n1 * s0


def fam_ca(n1, tsize):
    # every nesting order, inner stride 1, middle = n1 + p1, outer = n1*middle + p2
    import itertools
    for order in itertools.permutations(range(3)):
        for p1 in range(0, 3):
            mid = n1 + p1
            for p2 in range(0, 9):
                outer = n1 * mid + p2
                s = [0, 0, 0]
                s[order[0]], s[order[1]], s[order[2]] = 1, mid, outer
                yield tuple(s), n1 * outer


def search(n1, tsize):
    nat = (n1 * n1, n1, 1)
    phC = (0, 1, 2)
    phB = (1, 0, 2)
    xbc = best_exchange(n1, tsize, phB, phC, fam_bc(n1, tsize))
    results = []
    for ordA in ((0, 1), (1, 0)):
        phA = (2, ordA[0], ordA[1])
        xab = best_exchange(n1, tsize, phA, phB, fam_ab(n1, tsize))
        stA, idA = access_cost(n1, tsize, *phA, nat, n1 ** 3)
        total = xbc[0][0] + xab[0][0] + stA
        results.append([total, xbc[2] + xab[2], (0, 2), ordA, xbc, xab, stA, idA])
    results.sort(key=lambda r: (r[0], r[1]))
    best = results[0]
    # direct exchange between phase C and phase A: used where the natural-layout staging access of
    # phase A conflicts badly (even n1) -- the staged tensor is read in phase C order and handed over
    phA = (2, best[3][0], best[3][1])
    xca = best_exchange(n1, tsize, phC, phA, fam_ca(n1, tsize)) if n1 % 2 == 0 else None
    best.append(xca)
    return best


def emit(rows):
    out = ["// GENERATED by tools/gen_column_layouts.py -- do not edit by hand.",
           "// Shared-memory layouts of the column kernels (tools/bank_sim.py conflict model).",
           "// Strides in elements.  hB/lB, hA/lA: the tensor axes the (hi, lo) thread indices stand for",
           "// in phase B (register axis 1) and phase A (register axis 2).",
           "#pragma once", "", "namespace femtc", "{", "template <typename T, int N1>", "struct ColumnLayout;", ""]
    for (tname, n1), r in rows.items():
        _, _, ordB, ordA, xbc, xab, stA, idA, xca = r
        if xca is None:
            xca = xab  # unused for odd n1 (phase A reads the staging buffer conflict-free)
        out += [f"template <>", f"struct ColumnLayout<{tname}, {n1}>", "{",
                f"  static constexpr int hB = {ordB[0]}, lB = {ordB[1]}, hA = {ordA[0]}, lA = {ordA[1]};",
                f"  struct XBC {{ static constexpr int Sc = {xbc[2]}, s0 = {xbc[1][0]}, s1 = {xbc[1][1]}, s2 = {xbc[1][2]}; }};"
                f" // wavefronts {xbc[3]}+{xbc[4]} (ideal {xbc[5]:.0f})",
                f"  struct XAB {{ static constexpr int Sc = {xab[2]}, s0 = {xab[1][0]}, s1 = {xab[1][1]}, s2 = {xab[1][2]}; }};"
                f" // wavefronts {xab[3]}+{xab[4]} (ideal {xab[5]:.0f})",
                f"  struct XCA {{ static constexpr int Sc = {xca[2]}, s0 = {xca[1][0]}, s1 = {xca[1][1]}, s2 = {xca[1][2]}; }};"
                f" // wavefronts {xca[3]}+{xca[4]} (ideal {xca[5]:.0f})",
                f"  static constexpr int stageA_wavefronts = {stA}, stageA_ideal = {int(round(idA))};"
                f" // phase-A access to a natural-layout staging buffer",
                "};", ""]
    out += ["} // namespace femtc", ""]
    return "\n".join(out)


def main():
    rows = {}
    for tname, tsize in (("double", 8), ("float", 4)):
        for n1 in range(2, 10):
            r = search(n1, tsize)
            rows[(tname, n1)] = r
            print(f"{tname} n1={n1}: total {r[0]} ordB={r[2]} ordA={r[3]} XBC={r[4][1]} Sc={r[4][2]} "
                  f"({r[4][3]}+{r[4][4]} / {r[4][5]:.0f})  XAB={r[5][1]} Sc={r[5][2]} ({r[5][3]}+{r[5][4]} / {r[5][5]:.0f}) "
                  f"stageA {r[6]} / {r[7]:.0f} XCA={r[8] and (r[8][1], r[8][2], r[8][3], r[8][4])}", flush=True)
    (ROOT / "femtc_b200" / "csrc" / "column_layouts.cuh").write_text(emit(rows))


if __name__ == "__main__":
    main()
