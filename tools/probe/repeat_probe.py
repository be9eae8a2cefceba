"""Development check: launch one operator configuration many times on L2-resident inputs and count launches
whose output differs bit-for-bit from the first (a data race shows up as a non-zero count).

usage: python tools/probe/repeat_probe.py [launches]        (FEMTC_B200_LIB selects another build of the library)
"""
import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import femtc_b200 as femtc
from femtc_b200.distributed import fill_uniform

REPS = int(sys.argv[1]) if len(sys.argv) > 1 else 300
CASES = [  # op, mode, P, dtype, variant, cells
    ("mass", 0, 7, torch.float32, 20, 8887),
    ("mass", 0, 6, torch.float32, 20, 8887),
    ("mass", 0, 5, torch.float32, 0, 8887),
    ("mass", 0, 5, torch.float64, 20, 8887),
    ("mass", 0, 3, torch.float64, 0, 40003),
    ("mass", 0, 3, torch.float32, 0, 40003),
    ("mass", 0, 4, torch.float64, 0, 20011),
    ("mass", 0, 2, torch.float64, 0, 80021),
    ("stiffness", 0, 3, torch.float64, 0, 20011),
    ("stiffness", 0, 3, torch.float32, 0, 20011),
    ("stiffness", 1, 3, torch.float64, 0, 20011),
    ("stiffness", 0, 4, torch.float64, 0, 10007),
    ("stiffness", 0, 4, torch.float32, 0, 10007),
    ("stiffness", 0, 2, torch.float64, 0, 40003),
    ("stiffness", 0, 5, torch.float64, 0, 8887),
    ("stiffness", 0, 6, torch.float32, 0, 8887),
    ("stiffness", 0, 7, torch.float32, 0, 8887),
    ("stiffness", 0, 8, torch.float32, 0, 4099),
    ("stiffness", 0, 8, torch.float64, 0, 4099),
    ("mass", 0, 8, torch.float64, 0, 4099),
    ("mass", 0, 7, torch.float32, 0, 8887),
]
rng = np.random.default_rng(5)
bad_total = 0
for op, mode, P, tdt, variant, nc in CASES:
    n1 = P + 1
    nd = n1 ** 3
    npdt = np.float64 if tdt == torch.float64 else np.float32
    phi = rng.random(2 * n1 * n1).astype(npdt)
    U = fill_uniform(torch.empty(nc * nd, dtype=tdt, device="cuda"), 1, 1)
    G = fill_uniform(torch.empty(nc * nd * (1 if op == "mass" else 6), dtype=tdt, device="cuda"), 1, 2)
    femtc.set_option("variant", variant)
    first, bad = None, 0
    for i in range(REPS):
        W = torch.full_like(U, float("nan"))
        if op == "mass":
            rc = femtc.mass_operator(phi, U, W, G, nc, P)
        else:
            rc = femtc.stiffness_operator(phi, U, W, G, nc, P, mode=mode)
        assert rc == 0
        if first is None:
            first = W.clone()
        elif not torch.equal(W.view(torch.uint8), first.view(torch.uint8)):
            bad += 1
    torch.cuda.synchronize()
    femtc.set_option("variant", 0)
    bad_total += bad
    print(f"{op:9s} mode {mode} P {P} {str(tdt)[6:]:8s} variant {variant:2d} cells {nc:6d}: "
          f"{bad} of {REPS - 1} repeats differ from the first launch", flush=True)
print("TOTAL differing launches:", bad_total)
