"""Where the host-buffer (CILQR_MEM_HOST) step spends its time: PCIe rates of this box, then the two C-ABI calls timed apart.
Usage: python tools/e2e_breakdown.py [B] [n_obs]   (on the GPU box)"""
import ctypes as C
import sys
import time

import numpy as np
import torch

import cilqr_b200
from cilqr_b200 import scenarios

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
n_obs = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dev = torch.device("cuda:0")


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


n = 160 * 1024 * 1024
hbuf = torch.empty(n, dtype=torch.uint8).pin_memory()
hbuf2 = torch.empty(n, dtype=torch.uint8).pin_memory()
dbuf = torch.empty(n, dtype=torch.uint8, device=dev)
dbuf2 = torch.empty(n, dtype=torch.uint8, device=dev)
ms = timed(lambda: dbuf.copy_(hbuf, non_blocking=True))
print(f"H2D pinned 160 MiB: {ms:.2f} ms = {n / ms / 1e6:.1f} GB/s")
ms = timed(lambda: hbuf.copy_(dbuf, non_blocking=True))
print(f"D2H pinned 160 MiB: {ms:.2f} ms = {n / ms / 1e6:.1f} GB/s")
s2 = torch.cuda.Stream()


def both():
    dbuf.copy_(hbuf, non_blocking=True)
    with torch.cuda.stream(s2):
        hbuf2.copy_(dbuf2, non_blocking=True)


ms = timed(both)
print(f"H2D + D2H on two streams, 160 MiB each: {ms:.2f} ms = {2 * n / ms / 1e6:.1f} GB/s aggregate")

lib = cilqr_b200.load_library()
p = cilqr_b200.Parameters()
lib.cilqr_params_default(C.byref(p))
p = p.replace(horizon=50)
N = 50
sc = scenarios.make_scenarios(p, 0, B, n_obs, dense=False)
keys = ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")
pin = {k: torch.from_numpy(sc[k]).pin_memory() for k in keys}
Uh = torch.empty((B, 2 * N), dtype=torch.float64).pin_memory()
outs = {"X": torch.empty((B, 4 * (N + 1)), dtype=torch.float64).pin_memory().numpy(), "J": torch.empty(B, dtype=torch.float64).pin_memory().numpy(),
        "iters": torch.empty(B, dtype=torch.int32).pin_memory().numpy(), "status": torch.empty(B, dtype=torch.int32).pin_memory().numpy()}
hs = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)


def obstacles():
    hs.set_obstacles_cv(pin["obs_state0"].numpy(), pin["obs_LW"].numpy(), pin["obs_sigma"].numpy(), sc["kappa"])


def solve():
    hs.get_optimal_control_seq(pin["x0"].numpy(), Uh.numpy(), pin["poly"].numpy(), pin["xlp"].numpy(), outputs=outs)


def reset():
    Uh.copy_(pin["U0"])


obstacles()
for _ in range(3):
    reset()
    solve()
print(f"host memcpy of U0 (pinned -> pinned, {Uh.numel() * 8 / 1e6:.0f} MB): {timed(reset):.2f} ms")
print(f"cilqr_set_obstacles_cv (host arrays): {timed(obstacles):.2f} ms")
print(f"cilqr_solve_batch (host arrays, includes the U0 reset): {timed(lambda: (reset(), solve())):.2f} ms")
print(f"both: {timed(lambda: (obstacles(), reset(), solve())):.2f} ms")
d = {k: torch.from_numpy(sc[k]).to(dev) for k in keys}
U = torch.empty_like(d["U0"])
res = {"X": torch.empty((B, 4 * (N + 1)), dtype=torch.float64, device=dev), "J": torch.empty(B, dtype=torch.float64, device=dev),
       "iters": torch.empty(B, dtype=torch.int32, device=dev), "status": torch.empty(B, dtype=torch.int32, device=dev)}
ds = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
ds.set_obstacles_cv(d["obs_state0"], d["obs_LW"], d["obs_sigma"], sc["kappa"])


def dsolve():
    U.copy_(d["U0"])
    ds.get_optimal_control_seq(d["x0"], U, d["poly"], d["xlp"], outputs=res)


for _ in range(3):
    dsolve()
print(f"cilqr_solve_batch (device arrays): {timed(dsolve):.2f} ms")
print(f"cilqr_set_obstacles_cv (device arrays): {timed(lambda: ds.set_obstacles_cv(d['obs_state0'], d['obs_LW'], d['obs_sigma'], sc['kappa'])):.2f} ms")
