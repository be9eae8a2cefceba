"""Development probe for the tcgen05 kind::tf32 arm: identity tables make every stage a permutation."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import femtc_b200 as femtc  # noqa: E402

P, nc = 7, 4
n1, nd = 8, 512
phi = np.zeros(2 * 64, dtype=np.float32)
mode = sys.argv[1] if len(sys.argv) > 1 else "identity"
if mode == "identity":
    for i in range(8):
        phi[i + i * 8] = 1.0
else:
    rng = np.random.default_rng(0)
    phi[:64] = rng.random(64, dtype=np.float32)
U = np.arange(nc * nd, dtype=np.float32) + 1
J = np.ones(nc * nd, dtype=np.float32)
dU, dJ = torch.from_numpy(U).cuda(), torch.from_numpy(J).cuda()
out = {}
for tc in (0, 1):
    femtc.set_option("tc", tc)
    dW = torch.full_like(dU, float("nan"))
    assert femtc.mass_operator(phi, dU, dW, dJ, nc, P) == 0
    torch.cuda.synchronize()
    out[tc] = dW.cpu().numpy()
    print("tc", tc, "family", femtc.kernel_info("mass", 4, P)["family"], out[tc][:12], "...", out[tc][-4:])
d = np.abs(out[1] - out[0]) / np.maximum(np.abs(out[0]), 1e-30)
print("max rel diff", d.max(), "zeros in tc=1:", int((out[1] == 0).sum()), "nan:", int(np.isnan(out[1]).sum()))
bad = np.nonzero(d > 1e-5)[0]
print("bad count", bad.size, "first bad", bad[:16])
