"""cilqr_solve_cv_static — the solve entry point that allocates nothing (scratch in a static device array,
no reference counterpart) — against the existing entry points cilqr_set_obstacles_cv + cilqr_solve_batch,
through the C ABI, on two calls in a row.  Collected last; passes alone in a fresh process and
after the other tests.

Tolerance: the tests' 1e-9 relative (BASELINE.json north_star); the two paths run the same device functions, so the results are
in fact identical and the iteration counts / status words are compared exactly."""
import ctypes as C

import numpy as np
import pytest

import cilqr_b200
from cilqr_b200 import scenarios

pytestmark = pytest.mark.gpu
TOL = 1e-9


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def params(horizon=50):
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(C.byref(p))
    return p.replace(horizon=horizon)


def outputs(B, N, dev):
    import torch
    return {"X": torch.full((B, 4 * (N + 1)), float("nan"), dtype=torch.float64, device=dev), "J": torch.full((B,), float("nan"), dtype=torch.float64, device=dev),
            "iters": torch.full((B,), -7, dtype=torch.int32, device=dev), "status": torch.full((B,), -7, dtype=torch.int32, device=dev)}


def test_static_scratch_solve_matches_the_allocating_entry_points_twice_in_a_row():
    import torch
    dev = torch.device("cuda:0")
    p = params()
    N = p.horizon
    lib = cilqr_b200.load_library()
    keys = ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")
    # two calls in a row on ONE handle, different shapes: the second (fewer scenarios, other obstacle count) must not see the first's leftovers
    cases = [(300, 6, 4242, True), (77, 11, 999, False)]
    s = cilqr_b200.iLQR(p, max_batch=max(c[0] for c in cases), max_obs=max(c[1] for c in cases))
    got = []
    for B, n_obs, first, curved in cases:
        sc = scenarios.make_scenarios(p, first, B, n_obs, curved=curved, dense=False)
        t = {k: torch.from_numpy(sc[k]).to(dev) for k in keys}
        need = C.c_int64(0)
        assert lib.cilqr_static_scratch_bytes(C.byref(need), C.byref(p), C.c_int(B), C.c_int(n_obs)) == 0
        assert need.value == 256 + B * n_obs * (8 + N) * 8 <= 16 << 20
        U = t["U0"].clone()
        out = outputs(B, N, dev)
        s.solve_cv_static(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"], t["x0"], U, t["poly"], t["xlp"], out)
        got.append((sc, t, U, out))
    torch.cuda.synchronize()
    for (B, n_obs, first, curved), (sc, t, U, out) in zip(cases, got):
        r = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)  # the existing entry points, on a handle of their own
        r.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
        Ur = t["U0"].clone()
        r.get_optimal_control_seq(t["x0"], Ur, t["poly"], t["xlp"])
        torch.cuda.synchronize()
        assert torch.equal(out["iters"], r.iterations) and torch.equal(out["status"], r.status)
        Uh, Xh, Jh = U.cpu().numpy(), out["X"].cpu().numpy(), out["J"].cpu().numpy()
        Uw, Xw, Jw = Ur.cpu().numpy(), r.X_result.cpu().numpy(), r.J_result.cpu().numpy()
        for b in range(B):
            assert rel(Uh[b], Uw[b]) < TOL and rel(Xh[b], Xw[b]) < TOL and abs(Jh[b] - Jw[b]) <= TOL * abs(Jw[b]), b
        r.close()
    # the same call again on the first case's buffers: same answer (the control block is cleared by the call itself)
    sc, t, U, out = got[0]
    U2 = t["U0"].clone()
    out2 = outputs(cases[0][0], N, dev)
    s.solve_cv_static(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"], t["x0"], U2, t["poly"], t["xlp"], out2)
    torch.cuda.synchronize()
    assert torch.equal(U2, U) and torch.equal(out2["X"], out["X"]) and torch.equal(out2["J"], out["J"]) and torch.equal(out2["iters"], out["iters"])
    # no obstacles at all, X / J / iters / status not wanted
    U3 = t["U0"].clone()
    s.solve_cv_static(None, None, None, sc["kappa"], t["x0"], U3, t["poly"], t["xlp"], {})
    r = cilqr_b200.iLQR(p, max_batch=cases[0][0], max_obs=1)
    U3r = t["U0"].clone()
    r.get_optimal_control_seq(t["x0"], U3r, t["poly"], t["xlp"])
    torch.cuda.synchronize()
    assert rel(U3.cpu().numpy(), U3r.cpu().numpy()) < TOL
    r.close()
    s.close()


def test_static_scratch_solve_rejects_what_does_not_fit():
    import torch
    dev = torch.device("cuda:0")
    p = params()
    N = p.horizon
    B, n_obs = 2048, 20  # 256 + 2048 * 20 * 58 * 8 = 19.0 MB > 16 MiB
    s = cilqr_b200.iLQR(p, max_batch=8, max_obs=1)
    z = lambda *shape: torch.zeros(shape, dtype=torch.float64, device=dev)  # noqa: E731
    with pytest.raises(cilqr_b200.CilqrError, match="CILQR_STATIC_SCRATCH_BYTES"):
        s.solve_cv_static(z(B, n_obs, 4), z(B, n_obs, 2), z(B, n_obs, 3), 2.4477, z(B, 4), z(B, 2 * N), z(B, 6), z(B, 2), {})
    with pytest.raises(ValueError, match="CUDA tensors"):
        s.solve_cv_static(None, None, None, 2.4477, np.zeros((1, 4)), np.zeros((1, 2 * N)), np.zeros((1, 6)), np.zeros((1, 2)), {})
    s.close()
