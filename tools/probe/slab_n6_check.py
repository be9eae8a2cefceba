"""Development check: the slab kernels forced on at n1 = 6..8 (variant 20) against the default kernels."""
import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import femtc_b200 as femtc
from femtc_b200.distributed import fill_uniform
rng = np.random.default_rng(3)
for P in (5, 6, 7):
    nd = (P + 1) ** 3; nc = 20011
    for dt, tdt in (("f64", torch.float64), ("f32", torch.float32)):
        phi = rng.random(2 * (P + 1) ** 2).astype(np.float64 if dt == "f64" else np.float32)
        U = fill_uniform(torch.empty(nc * nd, dtype=tdt, device="cuda"), 1, 1)
        for op, gw in (("mass", 1),):
            G = fill_uniform(torch.empty(nc * nd * gw, dtype=tdt, device="cuda"), 1, 2)
            out = {}
            for v in (0, 20):
                femtc.set_option("variant", v)
                W = torch.full_like(U, float("nan"))
                femtc.mass_operator(phi, U, W, G, nc, P)
                torch.cuda.synchronize()
                out[v] = W.cpu().numpy().astype(np.float64)
            femtc.set_option("variant", 0)
            err = np.abs(out[20] - out[0]).max() / np.abs(out[0]).max()
            print(op, dt, "P", P, "variant 20 vs 0: max rel diff", err, "finite", np.isfinite(out[20]).all(), flush=True)
