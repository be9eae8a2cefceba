import sys, time, numpy as np, torch
sys.path.insert(0, "/root/repo")
import femtc_b200 as femtc
P, cells = 3, 1_000_000
nd = (P + 1) ** 3
phi = np.random.default_rng(0).random(2 * (P + 1) ** 2)
hU = torch.rand(cells * nd, dtype=torch.float64).pin_memory()
hG = torch.rand(cells * nd * 6, dtype=torch.float64).pin_memory()
hW = torch.empty(cells * nd, dtype=torch.float64).pin_memory()
nU, nG, nW = hU.numpy(), hG.numpy(), hW.numpy()
K = femtc.ResidentOperator("stiffness", phi, nG, cells, P)
for mb in (4, 8, 16, 32, 64, 128):
    femtc.set_option("host_chunk_mb", mb)
    K.apply(nU, nW)
    t0 = time.perf_counter()
    for _ in range(5):
        K.apply(nU, nW)
    dt = (time.perf_counter() - t0) / 5
    print(f"resident chunk {mb:4d} MB: {dt*1e3:7.2f} ms  {cells*nd/dt/1e9:6.2f} GDoF/s", flush=True)
for mb in (16, 32, 64, 128, 256):
    femtc.set_option("host_chunk_mb", mb)
    femtc.stiffness_operator(phi, nU, nW, nG, cells, P)
    t0 = time.perf_counter()
    for _ in range(3):
        femtc.stiffness_operator(phi, nU, nW, nG, cells, P)
    dt = (time.perf_counter() - t0) / 3
    print(f"strict   chunk {mb:4d} MB: {dt*1e3:7.2f} ms  {cells*nd/dt/1e9:6.2f} GDoF/s", flush=True)
