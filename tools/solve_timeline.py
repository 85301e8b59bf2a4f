"""Timeline of one batched solve from the kernels' own %globaltimer stamps (no profiler): which launches overlap.

    PYTHONPATH=. python tools/solve_timeline.py [B] [n_obs] [host|dev] [chunks]
"""
import ctypes as C
import sys

import numpy as np
import torch

import cilqr_b200  # noqa: E402
from cilqr_b200 import scenarios  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
n_obs = int(sys.argv[2]) if len(sys.argv) > 2 else 20
host = len(sys.argv) > 3 and sys.argv[3] == "host"
chunks = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dev = torch.device("cuda:0")
lib = cilqr_b200.load_library()
p = cilqr_b200.Parameters()
lib.cilqr_params_default(C.byref(p))
p = p.replace(horizon=50)
N = 50
sc = scenarios.make_scenarios(p, 0, B, n_obs, dense=False)
keys = ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")
s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, options={cilqr_b200.OPT_TRACE: 1, cilqr_b200.OPT_CHUNKS: chunks})
if host:
    d = {k: torch.from_numpy(sc[k]).pin_memory().numpy() for k in keys}
    U = torch.empty((B, 2 * N), dtype=torch.float64).pin_memory().numpy()
    res = {"X": torch.empty((B, 4 * (N + 1)), dtype=torch.float64).pin_memory().numpy(), "J": torch.empty(B, dtype=torch.float64).pin_memory().numpy(),
           "iters": torch.empty(B, dtype=torch.int32).pin_memory().numpy(), "status": torch.empty(B, dtype=torch.int32).pin_memory().numpy()}
else:
    d = {k: torch.from_numpy(sc[k]).to(dev) for k in keys}
    U = torch.empty_like(d["U0"])
    res = {"X": torch.empty((B, 4 * (N + 1)), dtype=torch.float64, device=dev), "J": torch.empty(B, dtype=torch.float64, device=dev),
           "iters": torch.empty(B, dtype=torch.int32, device=dev), "status": torch.empty(B, dtype=torch.int32, device=dev)}
for _ in range(4):
    s.set_obstacles_cv(d["obs_state0"], d["obs_LW"], d["obs_sigma"], sc["kappa"])
    if host:
        U[...] = d["U0"]
    else:
        U.copy_(d["U0"])
    s.get_optimal_control_seq(d["x0"], U, d["poly"], d["xlp"], outputs=res)
    torch.cuda.synchronize()
tl = s.solve_trace()
end = max(e[4] for e in tl)
print("launches %d, span %.1f us" % (len(tl), end / 1e3))
busy = np.zeros(end // 100 + 2, dtype=np.int32)
for c, r, name, a, b in tl:
    busy[a // 100:b // 100 + 1] += 1
print("time with 0 / 1 / 2 / 3+ launches in flight: %s us" % ", ".join("%.0f" % (np.sum(busy[:end // 100 + 1] == k if k < 3 else busy[:end // 100 + 1] >= 3) / 10) for k in range(4)))
for c, r, name, a, b in tl:
    print("chunk %d round %2d %-6s %8.1f -> %8.1f us  (%6.1f)" % (c, r, name, a / 1e3, b / 1e3, (b - a) / 1e3))
