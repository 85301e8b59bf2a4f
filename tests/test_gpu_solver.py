"""GPU parity tests of the CILQR solver (run on the B200 box: pytest -m gpu).

Every call goes through the C ABI of include/cilqr.h (via the ctypes host mirror).  The checker is the compiled
reference itself (oracle/_ref, built from the unmodified sources; the prebuilt .so travels to the GPU box) where it
exists, else the plain-C port; committed golden vectors generated from the reference are checked as well.

Tolerance: BASELINE.json north_star — optimal control sequence and cost within 1e-9 relative in FP64."""
import os

import numpy as np
import pytest

import cilqr_b200
from cilqr_b200 import scenarios
from cilqr_b200.params import ST_TOLERANCE, ST_MAX_ITER, ST_LAMB_MAX

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-9
CASES = ["n50_obs3", "n50_obs0_curved", "n40_obs2_curved", "n50_obs10"]


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "solver_golden.npz"))


def make_solver(p, sc, B, n_obs):
    s = cilqr_b200.iLQR(p, max_batch=max(B, 1), max_obs=max(n_obs, 1))
    if n_obs:
        s.set_obstacles_dense(sc["dim"], sc["pos"])
    return s


def case_inputs(checker, gold, name):
    N, n_obs, curved, B = (int(gold[name + "/" + k]) for k in ("horizon", "n_obs", "curved", "B"))
    p = checker.params_default(horizon=N)
    return p, scenarios.make_scenarios(p, 0, B, n_obs, curved=bool(curved)), n_obs, B


# ------------------------------------------------------------------------------------------------ golden vectors
@pytest.mark.parametrize("name", CASES)
def test_solve_matches_reference_golden(checker, gold, name):
    p, sc, n_obs, B = case_inputs(checker, gold, name)
    s = make_solver(p, sc, B, n_obs)
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    assert np.array_equal(s.iterations, gold[name + "/iters"])
    for b in range(B):
        assert rel(U[b], gold[name + "/U"][b]) < TOL
        assert rel(s.X_result[b], gold[name + "/X"][b]) < TOL
        assert abs(s.J_result[b] - gold[name + "/J"][b]) <= TOL * abs(gold[name + "/J"][b])
    s.close()


@pytest.mark.parametrize("name", CASES)
def test_stages_match_reference_golden(checker, gold, name):
    p, sc, n_obs, B = case_inputs(checker, gold, name)
    s = make_solver(p, sc, B, n_obs)
    g = lambda k: np.tile(gold[name + "/stage/" + k], (B, 1))  # noqa: E731  (scenario 0's stage vectors in every slot)
    sc0 = {k: np.tile(sc[k][0], (B,) + (1,) * (sc[k].ndim - 1)) for k in ("x0", "U0", "poly", "xlp")}
    if n_obs:
        s.set_obstacles_dense(np.tile(sc["dim"][0], (B, 1, 1)), np.tile(sc["pos"][0], (B, 1, 1)))
    X0 = s.get_nominal_trajectory(sc0["x0"], sc0["U0"])
    assert rel(X0, g("X0")) < 1e-12
    lx, lxx = s.get_state_cost(g("X0"), sc0["poly"], sc0["xlp"])
    assert rel(lx, g("l_x")) < 1e-11 and rel(lxx, g("l_xx")) < 1e-11
    lu, luu = s.get_control_cost(g("X0"), sc0["U0"])
    assert rel(lu, g("l_u")) < 1e-12 and rel(luu, g("l_uu")) < 1e-12
    J = s.get_J(g("X0"), sc0["U0"], sc0["poly"], sc0["xlp"])
    assert rel(J, np.full(B, float(gold[name + "/stage/J0"]))) < 1e-12
    ok = s.backward_pass(g("X0"), sc0["U0"], sc0["poly"], sc0["xlp"], 1.0)
    assert np.all(ok == 1) and rel(s.k, g("k")) < 1e-10 and rel(s.K, g("K")) < 1e-10
    Xn, Un = s.forward_pass(g("X0"), sc0["U0"], g("k"), g("K"))
    assert rel(Xn, g("X_new")) < 1e-12 and rel(Un, g("U_new")) < 1e-12
    Xh, Uh = s.forward_pass(g("X0"), sc0["U0"], g("k"), g("K"), alpha=0.25)
    assert rel(Xh, g("X_new_a025")) < 1e-12 and rel(Uh, g("U_new_a025")) < 1e-12
    s.close()


def test_alpha_extension_matches_reference_pinned_wrapper(checker, gold):
    """n_alphas = 8 with the tracking merit, pinned on a wrapper that only calls reference functions."""
    p = checker.params_default(horizon=50)
    sc = scenarios.make_scenarios(p, 100, 3, 3)
    s = make_solver(p, sc, 3, 3)
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"], n_alphas=8, merit=cilqr_b200.MERIT_TRACKING)
    for b in range(3):
        assert int(s.iterations[b]) == int(gold["alpha8/%d/iters" % b])
        assert rel(U[b], gold["alpha8/%d/U" % b]) < TOL
        assert abs(s.J_result[b] - float(gold["alpha8/%d/J" % b])) < TOL * abs(float(gold["alpha8/%d/J" % b]))
    s.close()


# ------------------------------------------------------------------------------------------------ fresh seeded inputs vs the checker
@pytest.mark.parametrize("n_obs,curved,first", [(0, False, 5000), (3, True, 6000), (20, False, 7000), (7, True, 8000)])
def test_batched_solve_matches_checker(checker, n_obs, curved, first):
    p = checker.params_default(horizon=50)
    B = 48
    sc = scenarios.make_scenarios(p, first, B, n_obs, curved=curved)
    s = make_solver(p, sc, B, n_obs)
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    flips = []
    for b in range(B):
        dim = sc["dim"][b] if n_obs else None
        pos = sc["pos"][b] if n_obs else None
        r = checker.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        if int(s.iterations[b]) != r["iters"]:
            flips.append(b)  # a discrete decision (accept test / arg-min) flipped by a last-ulp difference
            continue
        assert rel(U[b], r["U"]) < TOL, b
        assert rel(s.X_result[b], r["X"]) < TOL, b
        assert abs(s.J_result[b] - r["J"]) <= TOL * abs(r["J"]), b
    assert flips == [], "branch flips in scenarios %s" % flips
    st = np.asarray(s.status)
    assert np.all((st == ST_TOLERANCE) | (st == ST_MAX_ITER) | (st == ST_LAMB_MAX))
    s.close()


@pytest.mark.parametrize("fused", [1, 0])
@pytest.mark.parametrize("merit", [0, 1])
def test_line_search_matches_port(port, merit, fused):
    """n_alphas = 8 (extension): parity with the plain-C port, which test_oracle_golden pins on reference functions — on the fused
    on-chip kernel (cilqr_fused_ls.cuh) and on the round kernels."""
    p = port.params_default(horizon=50)
    B, n_obs = 24, 10
    sc = scenarios.make_scenarios(p, 9000, B, n_obs, curved=True)
    s = make_solver(p, sc, B, n_obs)
    s.set_option(cilqr_b200.OPT_FUSED, fused)
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"], n_alphas=8, merit=merit)
    r = port.solve_batch(p, B, n_obs, sc["dim"], sc["pos"], sc["x0"], sc["U0"], sc["poly"], sc["xlp"], n_alphas=8, merit=merit)
    same = np.asarray(s.iterations) == r["iters"]
    assert same.all(), np.nonzero(~same)
    assert np.array_equal(np.asarray(s.status), r["status"])
    for b in range(B):
        assert rel(U[b], r["U"][b]) < TOL and abs(s.J_result[b] - r["J"][b]) <= TOL * abs(r["J"][b])
    s.close()


@pytest.mark.parametrize("merit", [0, 1])
def test_config3_at_size_matches_reference_functions(checker, port, merit):
    """BASELINE config 3 at its own size — 4 096 scenarios x 10 covariance-inflated obstacles, 8 step lengths, where the two-wave
    line search has non-trivial retry lists and empty wave-2 grids — on a 256-scenario sample (every 16th scenario).
    Tracking merit: against the alpha loop assembled from reference member functions only (iLQR::backward_pass / forward_pass
    with k scaled / Constraints::get_J; oracle ref_solve_alpha_batch).  Total merit (needs barrier VALUES, which the reference API
    does not return): against the plain-C port, whose derivative pieces are pinned on the reference."""
    import torch
    p = checker.params_default(horizon=50)
    B, n_obs, na, step = 4096, 10, 8, 16
    sc = scenarios.make_scenarios(p, 0, B, n_obs, dense=True)
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")}
    s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
    s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
    U = t["U0"].clone()
    s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"], n_alphas=na, merit=merit)
    torch.cuda.synchronize()
    Uh, Xh, Jh = U.cpu().numpy(), s.X_result.cpu().numpy(), s.J_result.cpu().numpy()
    ih, sh = s.iterations.cpu().numpy(), s.status.cpu().numpy()
    sel = np.arange(0, B, step)
    sub = {k: np.ascontiguousarray(sc[k][sel]) for k in ("dim", "pos", "x0", "U0", "poly", "xlp")}
    if merit == 0:
        if not hasattr(checker, "solve_alpha_batch"):
            pytest.skip("needs the compiled reference (oracle/_ref)")
        r = checker.solve_alpha_batch(p, sel.size, n_obs, sub["dim"], sub["pos"], sub["x0"], sub["U0"], sub["poly"], sub["xlp"], na, threads=os.cpu_count() or 1)
    else:
        r = port.solve_batch(p, sel.size, n_obs, sub["dim"], sub["pos"], sub["x0"], sub["U0"], sub["poly"], sub["xlp"], n_alphas=na, merit=merit,
                             threads=os.cpu_count() or 1)
    flips = [int(b) for i, b in enumerate(sel) if ih[b] != r["iters"][i] or sh[b] != r["status"][i]]
    assert flips == [], "iteration count / status differ in scenarios %s" % flips
    worst = 0.0
    for i, b in enumerate(sel):
        eu, ex = rel(Uh[b], r["U"][i]), rel(Xh[b], r["X"][i])
        ej = abs(Jh[b] - r["J"][i]) / abs(r["J"][i])
        assert eu < TOL and ex < TOL and ej < TOL, (int(b), eu, ex, ej)
        worst = max(worst, eu, ex, ej)
    hist = np.bincount(sh, minlength=4)
    print("config 3 (merit %d): %d of %d scenarios checked, worst relative difference %.3g, status histogram of the batch %s, mean iterations %.2f"
          % (merit, sel.size, B, worst, hist.tolist(), float(ih.mean())))
    s.close()


@pytest.mark.parametrize("B,n_obs,cv,na,merit", [(333, 6, True, 8, 1), (97, 3, False, 4, 0), (50, 0, True, 2, 1), (1000, 10, True, 8, 0)])
def test_line_search_fused_kernel_against_the_round_kernels(B, n_obs, cv, na, merit):
    """The two schedules of the line-search extension on the same inputs: equal iteration counts and status, U*, X*, J within 1e-9
    (they call the same device functions; whether the compiler contracts every expression identically in both kernels is not
    relied upon), with automatic and with 3 slots per CTA (slots refilled many times)."""
    import torch
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(__import__("ctypes").byref(p))
    p = p.replace(horizon=50)
    sc = scenarios.make_scenarios(p, 31337, B, n_obs, curved=True, dense=not cv)
    dev = torch.device("cuda:0")
    keys = ("x0", "U0", "poly", "xlp") + ((("obs_state0", "obs_LW", "obs_sigma") if cv else ("dim", "pos")) if n_obs else ())
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in keys}
    res = []
    for fused, slots in [(0, 0), (1, 0), (1, 3)]:
        s = cilqr_b200.iLQR(p, max_batch=B, max_obs=max(n_obs, 1), options={cilqr_b200.OPT_FUSED: fused, cilqr_b200.OPT_FUSED_SLOTS: slots})
        if n_obs:
            if cv:
                s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
            else:
                s.set_obstacles_dense(t["dim"], t["pos"])
        U = t["U0"].clone()
        s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"], n_alphas=na, merit=merit)
        torch.cuda.synchronize()
        res.append((U.cpu().numpy(), s.X_result.cpu().numpy(), s.J_result.cpu().numpy(), s.iterations.cpu().numpy(), s.status.cpu().numpy()))
        s.close()
    for got in res[1:]:
        assert np.array_equal(got[3], res[0][3]) and np.array_equal(got[4], res[0][4])
        for b in range(B):
            assert rel(got[0][b], res[0][0][b]) < TOL and rel(got[1][b], res[0][1][b]) < TOL and abs(got[2][b] - res[0][2][b]) <= TOL * abs(res[0][2][b]), b


def test_covariance_inflated_obstacles_on_device_match_host_expansion(checker):
    """cilqr_set_obstacles_cv (device-side constant-velocity prediction + covariance inflation) against the unmodified
    Obstacle class fed with host-inflated per-timestep dimensions."""
    p = checker.params_default(horizon=50)
    B, n_obs = 32, 20
    sc = scenarios.make_scenarios(p, 12000, B, n_obs)
    s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
    s.set_obstacles_cv(sc["obs_state0"], sc["obs_LW"], sc["obs_sigma"], sc["kappa"])
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    for b in range(B):
        r = checker.solve(p, sc["dim"][b], sc["pos"][b], sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        assert int(s.iterations[b]) == r["iters"], b
        assert rel(U[b], r["U"]) < TOL and abs(s.J_result[b] - r["J"]) <= TOL * abs(r["J"])
    s.close()


def test_obstacle_set_is_complete_whoever_uses_it_first():
    """cilqr_set_obstacles_cv leaves the per-timestep half of the expansion (the 1/b^2 table) to the first consumer of the
    obstacle set.  Whatever that consumer is — round kernels, on-chip kernel, line search (both schedules), a stage entry point —
    and even when the caller's device arrays are overwritten right after the call, the results are those of the dense form
    built on the host from the same obstacles."""
    import torch
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(__import__("ctypes").byref(p))
    p = p.replace(horizon=50)
    B, n_obs = 97, 7
    sc = scenarios.make_scenarios(p, 4242, B, n_obs, curved=True, dense=True)
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma", "dim", "pos")}

    def cv_set(s):
        st, lw, sg = t["obs_state0"].clone(), t["obs_LW"].clone(), t["obs_sigma"].clone()
        s.set_obstacles_cv(st, lw, sg, sc["kappa"])
        st.fill_(float("nan")); lw.fill_(float("nan")); sg.fill_(float("nan"))  # stream-ordered after the call: must not be seen

    def solve(s, **kw):
        U = t["U0"].clone()
        s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"], **kw)
        torch.cuda.synchronize()
        return U.cpu().numpy(), s.J_result.cpu().numpy(), s.iterations.cpu().numpy()

    ref = {}
    for fused in (0, 1):  # the dense form on both schedules is the reference of this test
        s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, options={cilqr_b200.OPT_FUSED: fused})
        s.set_obstacles_dense(t["dim"], t["pos"])
        ref[1] = solve(s)
        ref[8] = solve(s, n_alphas=8, merit=cilqr_b200.MERIT_TOTAL)
        X = s.get_nominal_trajectory(t["x0"], t["U0"])
        ref["lx"] = [a.cpu().numpy() for a in s.get_state_cost(X, t["poly"], t["xlp"])]
        s.close()
        for first in ("solve", "line_search", "stage", "solve_twice"):
            s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, options={cilqr_b200.OPT_FUSED: fused})
            cv_set(s)
            if first == "stage":
                got = [a.cpu().numpy() for a in s.get_state_cost(X, t["poly"], t["xlp"])]
                for a, b in zip(got, ref["lx"]):
                    assert rel(a, b) < TOL, (fused, first)
            elif first == "line_search":
                got = solve(s, n_alphas=8, merit=cilqr_b200.MERIT_TOTAL)
                assert np.array_equal(got[2], ref[8][2]) and rel(got[0], ref[8][0]) < TOL, (fused, first)
            got = solve(s)
            assert np.array_equal(got[2], ref[1][2]) and rel(got[0], ref[1][0]) < TOL and rel(got[1], ref[1][1]) < TOL, (fused, first)
            if first == "solve_twice":  # the same obstacle set again (no pending pass), then a fresh set on the same handle
                again = solve(s)
                assert all(np.array_equal(a, b) for a, b in zip(got, again))
                cv_set(s)
                again = solve(s)
                assert all(np.array_equal(a, b) for a, b in zip(got, again))
            s.close()


def test_warm_start_is_carried_unshifted_like_the_reference(checker):
    """run_step twice: the second solve starts from the first one's U (iLQR.cpp:201,222,253)."""
    p = checker.params_default(horizon=50)
    sc = scenarios.make_scenarios(p, 300, 4, 3)
    s = make_solver(p, sc, 4, 3)
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    x1 = np.asarray(s.X_result)[:, 4:8].copy()  # the state one step ahead
    U2 = U.copy()
    s.get_optimal_control_seq(x1, U2, sc["poly"], sc["xlp"])
    for b in range(4):
        r1 = checker.solve(p, sc["dim"][b], sc["pos"][b], sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        r2 = checker.solve(p, sc["dim"][b], sc["pos"][b], r1["X"][4:8], r1["U"], sc["poly"][b], sc["xlp"][b])
        assert int(s.iterations[b]) == r2["iters"] and rel(U2[b], r2["U"]) < TOL
    s.close()


def test_run_step_matches_reference_golden(checker, gold):
    p = checker.params_default(horizon=50)
    s = cilqr_b200.iLQR(p, max_batch=1, max_obs=0)
    s.set_global_plan(np.stack([np.arange(200, dtype=np.float64), np.zeros(200)], axis=1))
    X, U = s.run_step(np.array([0.3, 0.2, 4.0, 0.01]))
    assert rel(U[0], gold["run_step/U"]) < 1e-7  # looser: the polynomial fit feeds in at ~1e-12 and the solve amplifies it
    assert rel(X[0], gold["run_step/X"]) < 1e-7
    s.close()


# ------------------------------------------------------------------------------------------------ memory spaces, edge cases, errors
def test_device_resident_and_host_buffers_give_identical_bits():
    import torch
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(__import__("ctypes").byref(p))
    p = p.replace(horizon=50)
    B, n_obs = 257, 5  # not a multiple of the warps per CTA
    sc = scenarios.make_scenarios(p, 0, B, n_obs, curved=True)
    s = make_solver(p, sc, B, n_obs)
    Uh = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], Uh, sc["poly"], sc["xlp"])
    Xh, Jh, ih = s.X_result.copy(), s.J_result.copy(), s.iterations.copy()
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("x0", "U0", "poly", "xlp", "dim", "pos")}
    s.set_obstacles_dense(t["dim"], t["pos"])
    Ud = t["U0"].clone()
    s.get_optimal_control_seq(t["x0"], Ud, t["poly"], t["xlp"])
    torch.cuda.synchronize()
    assert np.array_equal(Ud.cpu().numpy(), Uh) and np.array_equal(s.X_result.cpu().numpy(), Xh)
    assert np.array_equal(s.J_result.cpu().numpy(), Jh) and np.array_equal(s.iterations.cpu().numpy(), ih)
    s.close()


@pytest.mark.parametrize("B,n_obs,cv,horizon", [(333, 6, True, 50), (97, 3, False, 50), (64, 0, True, 50), (41, 4, True, 33), (9, 2, False, 100)])
def test_fused_kernel_gives_the_bits_of_the_round_kernels(B, n_obs, cv, horizon):
    """cilqr_fused.cuh (scenario groups resident in shared memory from rollout to result) against the batch-synchronous round
    kernels: the same per-step device functions in the same order, so U*, X*, J, iteration counts, status and the pass tallies are
    identical — with 1 / 5 / 14 / automatic slots per CTA."""
    import torch
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(__import__("ctypes").byref(p))
    p = p.replace(horizon=horizon)
    sc = scenarios.make_scenarios(p, 2024, B, n_obs, curved=True, dense=not cv)
    dev = torch.device("cuda:0")
    keys = ("x0", "U0", "poly", "xlp") + (("obs_state0", "obs_LW", "obs_sigma") if cv else ("dim", "pos"))
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in keys if n_obs or not k.startswith(("obs", "dim", "pos"))}
    ref = None
    for r0, slots in [(0, 0), (1, 0), (1, 1), (1, 5), (1, 14)]:  # (CILQR_OPT_FUSED, CILQR_OPT_FUSED_SLOTS): round kernels first
        s = cilqr_b200.iLQR(p, max_batch=B, max_obs=max(n_obs, 1), options={cilqr_b200.OPT_FUSED: r0, cilqr_b200.OPT_FUSED_SLOTS: slots})
        if n_obs:
            if cv:
                s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
            else:
                s.set_obstacles_dense(t["dim"], t["pos"])
        for rep in range(3):  # direct launches, then (round kernels) the recorded graph
            U = t["U0"].clone()
            s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"])
            torch.cuda.synchronize()
            got = (U.cpu().numpy(), s.X_result.cpu().numpy(), s.J_result.cpu().numpy(), s.iterations.cpu().numpy(), s.status.cpu().numpy())
            stats = s.solve_stats()[:2]
            if ref is None:
                ref = got + (stats,)
            for a, b in zip(ref[:5], got):
                assert np.array_equal(a, b), (r0, slots, rep)
            assert stats == ref[5], (r0, slots, rep)
        # host buffers through the same plan
        Uh = sc["U0"].copy()
        if n_obs:
            if cv:
                s.set_obstacles_cv(sc["obs_state0"], sc["obs_LW"], sc["obs_sigma"], sc["kappa"])
            else:
                s.set_obstacles_dense(sc["dim"], sc["pos"])
        s.get_optimal_control_seq(sc["x0"], Uh, sc["poly"], sc["xlp"])
        for a, b in zip(ref[:5], (Uh, s.X_result, s.J_result, s.iterations, s.status)):
            assert np.array_equal(a, np.asarray(b)), (r0, slots, "host")
        s.close()


def test_two_live_handles_with_different_horizons(checker):
    """Dynamic shared-memory limits are per kernel and process-wide: a second handle with a shorter horizon must not lower them
    under a first one that needs more than 48 KB (N = 100, S = 200 -> 67 KB in the stage kernel)."""
    pa, pb = checker.params_default(horizon=100), checker.params_default(horizon=20)
    sa, sb_ = scenarios.make_scenarios(pa, 3, 5, 2), scenarios.make_scenarios(pb, 3, 5, 2)
    a = make_solver(pa, sa, 5, 2)
    b = make_solver(pb, sb_, 5, 2)  # created second, smaller
    for s, p, sc in ((a, pa, sa), (b, pb, sb_), (a, pa, sa)):
        X0 = s.get_nominal_trajectory(sc["x0"], sc["U0"])
        lx, lxx = s.get_state_cost(X0, sc["poly"], sc["xlp"])
        ok = s.backward_pass(X0, sc["U0"], sc["poly"], sc["xlp"], 1.0)
        assert np.all(ok == 1) and np.isfinite(lx).all()
        U = sc["U0"].copy()
        s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
        r = checker.solve(p, sc["dim"][0], sc["pos"][0], sc["x0"][0], sc["U0"][0], sc["poly"][0], sc["xlp"][0])
        assert int(s.iterations[0]) == r["iters"] and rel(U[0], r["U"]) < TOL
        Xr = checker.rollout(p, sc["x0"][0], sc["U0"][0])
        assert rel(X0[0], Xr) < 1e-12
    a.close()
    b.close()


@pytest.mark.parametrize("chunks", [2, 3, 8])
def test_chunked_streams_give_the_bits_of_one_stream(chunks):
    """cilqr_solve_batch cuts the batch into chunks on internal streams (host->device copy, rounds, device->host copy per
    chunk); ragged last chunk, host and device buffers, constant-velocity obstacles: same bits as the single-stream solve."""
    import torch
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(__import__("ctypes").byref(p))
    p = p.replace(horizon=50)
    B, n_obs = 203, 4  # chunks of 128/75, 96/96/11, 32 x 6 + 11
    sc = scenarios.make_scenarios(p, 11, B, n_obs, curved=True, dense=False)
    out = {}
    for k in (1, chunks):
        s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, options={cilqr_b200.OPT_CHUNKS: k, cilqr_b200.OPT_FUSED: 0 if k == chunks else 1})
        s.set_obstacles_cv(sc["obs_state0"], sc["obs_LW"], sc["obs_sigma"], sc["kappa"])
        U = sc["U0"].copy()
        s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
        host = (U, s.X_result.copy(), s.J_result.copy(), s.iterations.copy(), s.status.copy())
        stats = s.solve_stats()
        dev = torch.device("cuda:0")
        t = {n: torch.from_numpy(sc[n]).to(dev) for n in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")}
        s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
        Ud = t["U0"].clone()
        N = p.horizon
        od = {"X": torch.empty(B, 4 * (N + 1), dtype=torch.float64, device=dev), "J": torch.empty(B, dtype=torch.float64, device=dev),
              "iters": torch.empty(B, dtype=torch.int32, device=dev), "status": torch.empty(B, dtype=torch.int32, device=dev)}
        for rep in range(4):  # the second call with the same buffers records a CUDA graph, the later ones replay it
            Ud.copy_(t["U0"])
            for o in od.values():
                o.zero_()
            s.get_optimal_control_seq(t["x0"], Ud, t["poly"], t["xlp"], outputs=od)
            torch.cuda.synchronize()
            devr = (Ud.cpu().numpy(), od["X"].cpu().numpy(), od["J"].cpu().numpy(), od["iters"].cpu().numpy(), od["status"].cpu().numpy())
            for a, b in zip(host, devr):
                assert np.array_equal(a, b), rep
        assert s.solve_stats() == stats
        # pinned host buffers: recorded too, host<->device copies included
        pin = {n: torch.from_numpy(sc[n]).pin_memory() for n in ("x0", "U0", "poly", "xlp")}
        s.set_obstacles_cv(sc["obs_state0"], sc["obs_LW"], sc["obs_sigma"], sc["kappa"])
        Up = pin["U0"].clone().pin_memory().numpy()
        op = {n: torch.empty(o.shape, dtype=o.dtype).pin_memory().numpy() for n, o in od.items()}
        for rep in range(4):
            Up[...] = sc["U0"]
            for o in op.values():
                o[...] = 0
            s.get_optimal_control_seq(pin["x0"].numpy(), Up, pin["poly"].numpy(), pin["xlp"].numpy(), outputs=op)
            for a, b in zip(host, (Up, op["X"], op["J"], op["iters"], op["status"])):
                assert np.array_equal(a, b), rep
        for a, b in zip(host, devr):
            assert np.array_equal(a, b)
        out[k] = host + (stats,)
        s.close()
    for a, b in zip(out[1][:5], out[chunks][:5]):
        assert np.array_equal(a, b)
    assert out[1][5][:2] == out[chunks][5][:2]  # cost / derivative passes summed over the chunks


def test_solve_trace_lists_every_launch_in_order(checker):
    """cilqr_solve_trace (OPT_TRACE): one entry per launch of the last solve, rounds in order on each chunk's stream."""
    p = checker.params_default(horizon=50)
    B, n_obs = 96, 3
    sc = scenarios.make_scenarios(p, 77, B, n_obs, dense=False)
    s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, options={cilqr_b200.OPT_TRACE: 1, cilqr_b200.OPT_CHUNKS: 2})
    s.set_obstacles_cv(sc["obs_state0"], sc["obs_LW"], sc["obs_sigma"], sc["kappa"])
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    tl = s.solve_trace()
    M = p.max_iterations
    assert len(tl) == 2 * (1 + (M + 1) + M + M + 1)  # per chunk: init, cost x (M + 1), derivs x M, sweep x M, output
    for c in range(2):
        mine = [e for e in tl if e[0] == c]
        assert mine[0][2] == "init" and mine[-1][2] == "output"
        for a, b in zip(mine, mine[1:]):
            assert a[4] <= b[3] + 1000, (a, b)  # one stream per chunk: a launch starts after the previous one ended (1 us timer slack)
        order = [(e[1], e[2]) for e in mine[1:-1]]
        assert order[:4] == [(0, "cost"), (0, "derivs"), (0, "sweep"), (1, "cost")]
    s2 = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)  # a handle created without the switch has no trace
    with pytest.raises(cilqr_b200.CilqrError, match="CILQR_OPT_TRACE"):
        s2.solve_trace()
    s.close()
    s2.close()


@pytest.mark.parametrize("token", [False, True])
def test_two_host_threads_two_handles_give_the_single_thread_bits(monkeypatch, token):
    """Handles are independent: two host threads, one handle and one stream each, blocking host-buffer calls in a loop
    (the way bench.py's end-to-end leg hides PCIe time) — every result equals the one-thread result.  With the opt-in compute
    token (cilqr_set_compute_token) the two handles' kernels take turns; same bits."""
    import threading
    import torch
    monkeypatch.delenv("CILQR_EXCLUSIVE", raising=False)
    p = cilqr_b200.Parameters()
    cilqr_b200.load_library().cilqr_params_default(__import__("ctypes").byref(p))
    p = p.replace(horizon=50)
    B, n_obs, N = 4099, 6, 50
    sc = [scenarios.make_scenarios(p, 500 + 5000 * i, B, n_obs, curved=bool(i), dense=False) for i in range(2)]
    want = []
    for i in range(2):
        s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
        s.set_obstacles_cv(sc[i]["obs_state0"], sc[i]["obs_LW"], sc[i]["obs_sigma"], sc[i]["kappa"])
        U = sc[i]["U0"].copy()
        s.get_optimal_control_seq(sc[i]["x0"], U, sc[i]["poly"], sc[i]["xlp"])
        want.append((U, s.X_result.copy(), s.J_result.copy(), s.iterations.copy()))
        s.close()
    errors = []

    def worker(i):
        try:
            torch.cuda.set_device(0)
            s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, stream=torch.cuda.Stream(), compute_token=True if token else None)
            pin = {k: torch.from_numpy(sc[i][k]).pin_memory().numpy() for k in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")}
            U = torch.empty((B, 2 * N), dtype=torch.float64).pin_memory().numpy()
            out = {"X": torch.empty((B, 4 * (N + 1)), dtype=torch.float64).pin_memory().numpy(), "J": torch.empty(B, dtype=torch.float64).pin_memory().numpy(),
                   "iters": torch.empty(B, dtype=torch.int32).pin_memory().numpy()}
            for rep in range(6):  # the same buffers again and again: direct launches, then the recorded graph
                s.set_obstacles_cv(pin["obs_state0"], pin["obs_LW"], pin["obs_sigma"], sc[i]["kappa"])
                U[...] = pin["U0"]
                s.get_optimal_control_seq(pin["x0"], U, pin["poly"], pin["xlp"], outputs=out, want=())
                for a, b in zip(want[i], (U, out["X"], out["J"], out["iters"])):
                    assert np.array_equal(a, b), (i, rep)
            s.close()
        except Exception as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    th = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert errors == []


def test_edge_cases_and_error_codes(checker):
    p = checker.params_default(horizon=50)
    sc = scenarios.make_scenarios(p, 0, 4, 2)
    s = cilqr_b200.iLQR(p, max_batch=4, max_obs=2)
    # empty batch is a no-op
    U = np.zeros((0, 100))
    s.get_optimal_control_seq(np.zeros((0, 4)), U, np.zeros((0, 6)), np.zeros((0, 2)))
    # batch larger than the handle was created for
    big = scenarios.make_scenarios(p, 0, 5, 0)
    with pytest.raises(cilqr_b200.CilqrError, match="batch"):
        s.get_optimal_control_seq(big["x0"], big["U0"].copy(), big["poly"], big["xlp"])
    # more obstacles than max_obs
    sc3 = scenarios.make_scenarios(p, 0, 4, 3)
    with pytest.raises(cilqr_b200.CilqrError, match="n_obs"):
        s.set_obstacles_dense(sc3["dim"], sc3["pos"])
    # obstacles set for another batch size
    s.set_obstacles_dense(sc["dim"][:2], sc["pos"][:2])
    with pytest.raises(cilqr_b200.CilqrError, match="obstacles were set"):
        s.get_optimal_control_seq(sc["x0"], sc["U0"].copy(), sc["poly"], sc["xlp"])
    # clear_Obstacle() -> the obstacle-free solve
    s.clear_Obstacle()
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    r = checker.solve(p, None, None, sc["x0"][0], sc["U0"][0], sc["poly"][0], sc["xlp"][0])
    assert int(s.iterations[0]) == r["iters"] and rel(U[0], r["U"]) < TOL
    with pytest.raises(cilqr_b200.CilqrError, match="n_alphas"):
        s.get_optimal_control_seq(sc["x0"], sc["U0"].copy(), sc["poly"], sc["xlp"], n_alphas=9)
    s.close()


@pytest.mark.parametrize("horizon", [1, 7, 33, 64, 100])
def test_other_horizons(checker, horizon):
    p = checker.params_default(horizon=horizon)
    B, n_obs = 6, 2
    sc = scenarios.make_scenarios(p, 40, B, n_obs, curved=True)
    s = make_solver(p, sc, B, n_obs)
    U = sc["U0"].copy()
    s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
    for b in range(B):
        r = checker.solve(p, sc["dim"][b], sc["pos"][b], sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        assert int(s.iterations[b]) == r["iters"] and rel(U[b], r["U"]) < TOL and rel(s.X_result[b], r["X"]) < TOL
    s.close()


# ------------------------------------------------------------------------------------------------ full BASELINE sizes: size-independent properties
def test_full_size_properties_config4_shard(checker):
    """One rank's shard of config 4 (65 536 / 8 = 8 192 scenarios, 20 uncertain obstacles): determinism, independence of
    batch composition (the property rank-sharding relies on), status histogram, and a sample against the checker."""
    import torch
    p = checker.params_default(horizon=50)
    B, n_obs = 8192, 20
    sc = scenarios.make_scenarios(p, 0, B, n_obs, dense=False)
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")}
    s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
    s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
    U1 = t["U0"].clone()
    s.get_optimal_control_seq(t["x0"], U1, t["poly"], t["xlp"])
    J1, it1, st1 = s.J_result.clone(), s.iterations.clone(), s.status.clone()
    U2 = t["U0"].clone()
    s.get_optimal_control_seq(t["x0"], U2, t["poly"], t["xlp"])
    assert torch.equal(U1, U2) and torch.equal(J1, s.J_result) and torch.equal(it1, s.iterations)  # deterministic
    # a sub-batch taken from the middle gives the same bits as inside the full batch
    lo, hi = 3000, 3512
    s.set_obstacles_cv(t["obs_state0"][lo:hi].contiguous(), t["obs_LW"][lo:hi].contiguous(), t["obs_sigma"][lo:hi].contiguous(), sc["kappa"])
    U3 = t["U0"][lo:hi].clone()
    s.get_optimal_control_seq(t["x0"][lo:hi].contiguous(), U3, t["poly"][lo:hi].contiguous(), t["xlp"][lo:hi].contiguous())
    assert torch.equal(U3, U1[lo:hi]) and torch.equal(s.J_result, J1[lo:hi])
    hist = np.bincount(st1.cpu().numpy(), minlength=4)
    assert hist.sum() == B and hist[3] == 0
    assert torch.isfinite(U1).all() and torch.isfinite(J1).all()
    assert int(it1.min()) >= 1 and int(it1.max()) <= p.max_iterations
    Uh, Jh, ih = U1.cpu().numpy(), J1.cpu().numpy(), it1.cpu().numpy()
    flips = []
    for b in range(0, B, B // 256):  # 256 scenarios of the shard against the reference itself
        dim, pos = scenarios.expand_obstacles_cv(p, sc["obs_state0"][b], sc["obs_LW"][b], sc["obs_sigma"][b], sc["kappa"])
        r = checker.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        if int(ih[b]) != r["iters"]:
            flips.append(b)
            continue
        assert rel(Uh[b], r["U"]) < TOL and abs(Jh[b] - r["J"]) <= TOL * abs(r["J"]), b
    assert flips == [], "iteration counts differ from the reference in scenarios %s" % flips
    s.close()


def test_config4_parity_across_all_rank_shards(checker):
    """256 consecutive scenarios from each of the 8 rank shards of config 4 (ids g * 65 536 + 30 000 ...: 2 048 solves, 20
    uncertain obstacles, N = 50), every one against the checker on all host threads: U*, X* and J within 1e-9 relative, and
    no scenario may take a different accept / reject path (which would show up as a difference of order one)."""
    import torch
    if not hasattr(checker, "solve_batch"):
        pytest.skip("needs a batch checker")
    p = checker.params_default(horizon=50)
    per, n_obs = 256, 20
    dev = torch.device("cuda:0")
    s = cilqr_b200.iLQR(p, max_batch=per, max_obs=n_obs)
    worst = 0.0
    for g in range(8):
        sc = scenarios.make_scenarios(p, g * 65536 + 30000, per, n_obs, dense=True)
        t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")}
        s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
        U = t["U0"].clone()
        s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"])
        Uh, Xh, Jh = U.cpu().numpy(), s.X_result.cpu().numpy(), s.J_result.cpu().numpy()
        r = checker.solve_batch(p, per, n_obs, sc["dim"], sc["pos"], sc["x0"], sc["U0"], sc["poly"], sc["xlp"], threads=os.cpu_count() or 1)
        for b in range(per):
            eu, ex = rel(Uh[b], r["U"][b]), rel(Xh[b], r["X"][b])
            ej = abs(Jh[b] - r["J"][b]) / abs(r["J"][b])
            assert eu < TOL and ex < TOL and ej < TOL, (g, b, eu, ex, ej)
            worst = max(worst, eu, ex, ej)
    s.close()
    print("config 4, %d scenarios over 8 shards" % (8 * per) + ": worst relative difference %.3g" % worst)


# ------------------------------------------------------------------------------------------------ row N1: LocalPlanner on the device
@pytest.mark.parametrize("curved", [False, True])
def test_local_planner_on_device_matches_reference(checker, curved):
    """cilqr_local_plan_batch against the compiled reference LocalPlanner (LocalPlanner.cpp:24-116): the first / last abscissa
    of the local plan exactly, the fitted polynomial to 1e-9 on the plan's span (the coefficients themselves are conditioned by a
    degree-5 Vandermonde matrix; the reference's own two builds differ there by more than they differ in the fitted values)."""
    if not hasattr(checker, "local_plan"):
        pytest.skip("needs the compiled reference (oracle/_ref)")
    import torch
    p = checker.params_default(horizon=50)
    t = np.arange(200, dtype=np.float64) * 0.9
    plan = np.stack([t, 0.002 * t ** 2 if curved else np.zeros_like(t)], axis=1)
    rng = np.random.default_rng(7 + int(curved))
    B = 96
    s = rng.uniform(0.0, 175.0, B)
    ego = np.stack([s + rng.uniform(-0.4, 0.4, B), (0.002 * s ** 2 if curved else 0.0) + rng.uniform(-2.0, 2.0, B),
                    rng.uniform(2, 6, B), rng.uniform(-0.1, 0.1, B)], axis=1)
    ego[-1, 0] = 178.0  # fewer than num_of_local_wpts waypoints left (LocalPlanner.cpp:51-54)
    lp = cilqr_b200.LocalPlanner(p)
    lp.set_global_planner(plan)
    poly_h, xlp_h = lp.plan_batch_device(ego)  # host buffers
    poly_d, xlp_d = lp.plan_batch_device(torch.from_numpy(ego).cuda())  # device buffers
    assert np.array_equal(poly_d.cpu().numpy(), poly_h) and np.array_equal(xlp_d.cpu().numpy(), xlp_h)
    for b in range(B):
        pr, xr, _ = checker.local_plan(p, plan, ego[b])
        assert np.array_equal(xlp_h[b], xr), b
        xs = np.linspace(xr[0], xr[1], 40)
        y_mine = sum(poly_h[b, j] * xs ** j for j in range(6))
        y_ref = sum(pr[j] * xs ** j for j in range(6))
        assert np.max(np.abs(y_mine - y_ref)) <= 1e-9 * max(1.0, np.max(np.abs(y_ref))), b


# ------------------------------------------------------------------------------------------------ row N4: receding-horizon loop
def test_receding_horizon_loop_matches_chained_reference_calls(checker, tmp_path):
    """Four closed-loop ticks (no pose noise) of three vehicles against the same loop assembled from reference calls:
    LocalPlanner -> get_optimal_control_seq warm-started from the previous tick, un-shifted -> Model::forward_simulate with
    U_result(:, 0) (ilqr_uncertainty_node.cpp:111-128, iLQR.cpp:247-255)."""
    if not hasattr(checker, "local_plan"):
        pytest.skip("needs the compiled reference (oracle/_ref)")
    from cilqr_b200 import receding_horizon as rh
    p = checker.params_default(horizon=50)
    B, n_obs, ticks = 3, 3, 4
    sc = scenarios.make_scenarios(p, 500, B, n_obs)
    plan = np.stack([np.arange(200, dtype=np.float64), np.zeros(200)], axis=1)
    s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
    s.set_obstacles_dense(sc["dim"], sc["pos"])
    path = str(tmp_path / "loop.bin")
    loop = rh.RecedingHorizon(s, plan, sc["x0"], log_path=path)
    for _ in range(ticks):
        loop.tick()
    loop.close()
    N, Bl, rec = rh.read_log(path)
    assert (N, Bl, rec.size) == (50, B, ticks * B)
    for b in range(B):
        x = sc["x0"][b].copy()
        U = sc["U0"][b].copy()  # the constructor's control sequence (iLQR.cpp:9-15)
        for k in range(ticks):
            poly, xlp, _ = checker.local_plan(p, plan, x)
            r = checker.solve(p, sc["dim"][b], sc["pos"][b], x, U, poly, xlp)
            got = rec[k * B + b]
            assert rel(got["start_pos"], x) < 1e-9 or np.max(np.abs(got["start_pos"] - x)) < 1e-9
            assert rel(got["U"], r["U"]) < 1e-6 and rel(got["X"], r["X"]) < 1e-6, (b, k)  # the fit feeds in at ~1e-12, k solves amplify it
            U = r["U"]
            x = checker.forward_simulate(p, x, r["U"][:2])
    s.close()
