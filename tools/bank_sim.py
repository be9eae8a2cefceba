"""Shared-memory bank-conflict model used to choose the staging/exchange layouts of the tiled
kernels (design tool, not shipped code).

Model (sm_70+): 32 banks x 4 B.  A warp-wide request of width w bytes per lane is served in
w/4 phases of 32*4/w lanes (8-byte: two half-warps, 16-byte: four quarter-warps).  Within a phase
the cost is the largest number of DISTINCT 4-byte words that fall into one bank (equal words are
broadcast).  wavefronts = sum over phases; ideal = active_bytes/128.
"""
from __future__ import annotations

import itertools


def wavefronts(addrs, width):
    """addrs: per-lane byte address (None = inactive lane); width: 4, 8 or 16."""
    lanes_per_phase = 32 * 4 // width
    total = 0
    for ph in range(0, 32, lanes_per_phase):
        banks = {}
        for a in addrs[ph:ph + lanes_per_phase]:
            if a is None:
                continue
            assert a % width == 0, "misaligned vector access"
            for k in range(width // 4):
                w = a // 4 + k
                banks.setdefault(w % 32, set()).add(w)
        total += max((len(s) for s in banks.values()), default=0)
    return total


def ownership_cost(n1, tsize, tau, cell_stride, width_elems=1, elem_stride=1, rec=1, base=0):
    """Average wavefronts per request, and the ideal, for a warp whose lane (c, t) reads every
    element of its slab: cells c = lane // n1, thread-axis index t = lane % n1 on axis `tau`.
    Element (c, i0, i1, i2) of record width `rec` lives at
        base + (c*cell_stride + ((i0*n1+i1)*n1+i2)*rec) * tsize        [cell_stride in elements]
    width_elems: elements per vector access along the innermost register axis (or along the
    record for rec > 1)."""
    cpw = 32 // n1
    reg_axes = [a for a in range(3) if a != tau]
    tot = ideal = nreq = 0
    inner = range(0, n1, width_elems) if rec == 1 else range(1)
    for r_outer in range(n1):
        for r_inner in inner:
            for k in (range(0, rec, width_elems) if rec > 1 else range(1)):
                addrs = []
                for lane in range(32):
                    c, t = divmod(lane, n1)
                    if c >= cpw:
                        addrs.append(None)
                        continue
                    idx = [0, 0, 0]
                    idx[tau] = t
                    if rec == 1:
                        idx[reg_axes[0]] = r_outer
                        idx[reg_axes[1]] = r_inner
                    else:
                        idx[reg_axes[0]] = r_outer
                        idx[reg_axes[1]] = 0
                    e = c * cell_stride + ((idx[0] * n1 + idx[1]) * n1 + idx[2]) * rec + k
                    addrs.append(base + e * tsize)
                if rec > 1:
                    # one request per (r_outer, second reg axis, k): loop the second axis here
                    for r2 in range(n1):
                        a2 = []
                        for lane in range(32):
                            c, t = divmod(lane, n1)
                            if c >= cpw:
                                a2.append(None)
                                continue
                            idx = [0, 0, 0]
                            idx[tau] = t
                            idx[reg_axes[0]] = r_outer
                            idx[reg_axes[1]] = r2
                            e = c * cell_stride + ((idx[0] * n1 + idx[1]) * n1 + idx[2]) * rec + k
                            a2.append(base + e * tsize)
                        w = wavefronts(a2, width_elems * tsize)
                        tot += w
                        ideal += sum(x is not None for x in a2) * width_elems * tsize / 128
                        nreq += 1
                else:
                    w = wavefronts(addrs, width_elems * tsize)
                    tot += w
                    ideal += sum(x is not None for x in addrs) * width_elems * tsize / 128
                    nreq += 1
    return tot, ideal, nreq


def report():
    for tsize, name in ((8, "f64"), (4, "f32")):
        for n1 in range(2, 10):
            nd = n1 ** 3
            print(f"--- {name} n1={n1} nd={nd} cells/warp={32 // n1}")
            for tau in (0, 1, 2):
                # candidate cell strides: natural, and padded ones that keep 16-byte alignment
                cands = [nd]
                if (nd * tsize) % 16 == 0:
                    step = 16 // tsize
                    cands += [nd + k * step for k in range(1, 9)]
                best = []
                for S in cands:
                    for wv in (1, 2, 4):
                        if wv * tsize > 16:
                            continue
                        if wv > 1:
                            # vector along the innermost register axis: must be contiguous & aligned
                            reg_axes = [a for a in range(3) if a != tau]
                            if reg_axes[1] != 2 or n1 % wv or (S * tsize) % (wv * tsize) or (n1 * tsize) % (wv * tsize):
                                continue
                        tot, ideal, nreq = ownership_cost(n1, tsize, tau, S, wv)
                        best.append((tot, nreq, S, wv, ideal))
                best.sort()
                s = "; ".join(f"S={S} v{wv}: {tot}wf/{nreq}req (ideal {ideal:.1f})" for tot, nreq, S, wv, ideal in best[:4])
                print(f"   tau={tau}: {s}")


if __name__ == "__main__":
    report()


def exchange_search(n1, tsize, ta, tb, max_pad=6, top=5):
    """Scratch layouts for the register exchange from thread axis `ta` (writers) to thread axis
    `tb` (readers); g = the remaining axis (a register axis on both sides).
    Layout: element (c, x0, x1, x2) at c*Sc + sum_k x_k * stride_k.  Searched: axis order
    (slow, mid, fast) and paddings.  Vector width > 1 only along the fast axis when it is a register
    axis of that side.  Returns the best (wavefronts, instructions) candidates."""
    g = 3 - ta - tb
    cpw = 32 // n1
    results = []
    maxv = 16 // tsize
    for order in itertools.permutations((0, 1, 2)):
        slow, mid, fast = order
        for p_mid in range(0, max_pad):
            s_fast = 1
            s_mid = n1 + p_mid
            for p_slow in range(0, max_pad):
                s_slow = n1 * s_mid + p_slow
                base_cell = (n1 - 1) * (s_slow + s_mid + s_fast) + 1
                for p_cell in range(0, 2 * max_pad):
                    Sc = base_cell + p_cell
                    stride = {slow: s_slow, mid: s_mid, fast: s_fast}
                    tot_w = tot_i = 0
                    ok = True
                    for (tau, is_write) in ((ta, True), (tb, False)):
                        regs = [a for a in range(3) if a != tau]
                        # vector width along `fast` if it is a register axis here
                        wv = 1
                        if fast in regs:
                            for cand in (4, 2):
                                if cand <= maxv and n1 % cand == 0 and all(
                                        (s * tsize) % (cand * tsize) == 0 for s in (s_mid, s_slow, Sc)):
                                    wv = cand
                                    break
                        other = [a for a in regs if a != fast] if fast in regs else regs
                        # enumerate requests
                        if fast in regs:
                            o = other[0]
                            reqs = [(ro, rf) for ro in range(n1) for rf in range(0, n1, wv)]
                            def idx_of(t, r, tau=tau, o=o):
                                d = {tau: t, o: r[0], fast: r[1]}
                                return d
                        else:
                            reqs = [(r0, r1) for r0 in range(n1) for r1 in range(n1)]
                            def idx_of(t, r, tau=tau, regs=regs):
                                return {tau: t, regs[0]: r[0], regs[1]: r[1]}
                        for r in reqs:
                            addrs = []
                            for lane in range(32):
                                c, t = divmod(lane, n1)
                                if c >= cpw:
                                    addrs.append(None)
                                    continue
                                d = idx_of(t, r)
                                e = c * Sc + sum(d[k] * stride[k] for k in range(3))
                                addrs.append(e * tsize)
                            tot_w += wavefronts(addrs, wv * tsize)
                            tot_i += 1
                    ideal = 2 * cpw * n1 ** 3 * tsize / 128
                    results.append((tot_w, tot_i, order, s_mid, s_slow, Sc, ideal))
    results.sort(key=lambda r: (r[0], r[1], r[5]))
    return results[:top]


def report_exchange():
    for tsize, name in ((8, "f64"), (4, "f32")):
        for n1 in range(2, 10):
            print(f"--- exchange {name} n1={n1}")
            for ta, tb in ((0, 1), (0, 2), (1, 0), (1, 2), (2, 0), (2, 1)):
                best = exchange_search(n1, tsize, ta, tb, top=2)
                s = "; ".join(f"order={o} s_mid={sm} s_slow={ss} Sc={sc}: {w}wf/{i}ins (ideal {idl:.0f})"
                              for w, i, o, sm, ss, sc, idl in best)
                print(f"   {ta}->{tb}: {s}")


def g_plane_cost(n1, tsize, tau2, Sg, ref_exact):
    """Wavefronts for reading the geometry records of one i0-plane: element k of point (i1,i2)
    of cell c at (c*Sg + (i1*n1+i2)*6 + k) elements.  Lanes (c,t) with t on axis tau2 (1 or 2)."""
    cpw = 32 // n1
    # accesses per point: list of (k, width_elems)
    if tsize == 8:
        acc = [(0, 2), (2, 1)] if ref_exact else [(0, 2), (2, 2), (4, 2)]
    else:
        acc = [(0, 2), (2, 1)] if ref_exact else [(0, 2), (2, 2), (4, 2)]
    tot = ideal = ins = 0
    for r in range(n1):
        for (k, wv) in acc:
            addrs = []
            for lane in range(32):
                c, t = divmod(lane, n1)
                if c >= cpw:
                    addrs.append(None)
                    continue
                i1, i2 = (t, r) if tau2 == 1 else (r, t)
                addrs.append((c * Sg + (i1 * n1 + i2) * 6 + k) * tsize)
            tot += wavefronts(addrs, wv * tsize)
            ideal += sum(a is not None for a in addrs) * wv * tsize / 128
            ins += 1
    return tot, ideal, ins


def report_g():
    for tsize, name in ((8, "f64"), (4, "f32")):
        for n1 in range(2, 7):
            plane = n1 * n1 * 6
            step = 16 // tsize
            for ref_exact in (True, False):
                for tau2 in (1, 2):
                    best = []
                    for pad in range(0, 16 * step, step):
                        Sg = plane + pad
                        if (Sg * tsize) % 16:
                            continue
                        tot, ideal, ins = g_plane_cost(n1, tsize, tau2, Sg, ref_exact)
                        best.append((tot, pad, ideal, ins))
                    best.sort()
                    print(f"G {name} n1={n1} {'REF ' if ref_exact else 'FULL'} tau2={tau2}: " +
                          "; ".join(f"pad={p}: {t}wf (ideal {i:.1f}, {n} ins)" for t, p, i, n in best[:3]))
