"""The plain-C restatement (oracle/cilqr_oracle.c) against the golden vectors that
tests/golden/make_golden.py generated from the UNMODIFIED reference (oracle/_ref).  CPU only."""
import os

import numpy as np
import pytest

import cilqr_b200
from cilqr_b200 import scenarios

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-9  # BASELINE.json north_star: optimal controls and cost within 1e-9 relative in FP64


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "solver_golden.npz"))


CASES = ["n50_obs3", "n50_obs0_curved", "n40_obs2_curved", "n50_obs10"]


def case_inputs(port, gold, name):
    N, n_obs, curved, B = (int(gold[name + "/" + k]) for k in ("horizon", "n_obs", "curved", "B"))
    p = port.params_default(horizon=N)
    sc = scenarios.make_scenarios(p, 0, B, n_obs, curved=bool(curved))
    return p, sc, n_obs, B


@pytest.mark.parametrize("name", CASES)
def test_port_solve_matches_reference_golden(port, gold, name):
    p, sc, n_obs, B = case_inputs(port, gold, name)
    for b in range(B):
        dim = sc["dim"][b] if n_obs else None
        pos = sc["pos"][b] if n_obs else None
        r = port.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        assert r["iters"] == int(gold[name + "/iters"][b])
        assert rel(r["U"], gold[name + "/U"][b]) < TOL
        assert rel(r["X"], gold[name + "/X"][b]) < TOL
        assert abs(r["J"] - gold[name + "/J"][b]) <= TOL * abs(gold[name + "/J"][b])


@pytest.mark.parametrize("name", CASES)
def test_port_stages_match_reference_golden(port, gold, name):
    p, sc, n_obs, B = case_inputs(port, gold, name)
    b = 0
    dim = sc["dim"][b] if n_obs else None
    pos = sc["pos"][b] if n_obs else None
    g = lambda k: gold[name + "/stage/" + k]  # noqa: E731
    X0 = port.rollout(p, sc["x0"][b], sc["U0"][b])
    assert rel(X0, g("X0")) < 1e-13
    lx, lxx = port.state_cost(p, dim, pos, g("X0"), sc["poly"][b], sc["xlp"][b])
    assert rel(lx, g("l_x")) < 1e-12 and rel(lxx, g("l_xx")) < 1e-12
    lu, luu = port.control_cost(p, g("X0"), sc["U0"][b])
    assert rel(lu, g("l_u")) < 1e-13 and rel(luu, g("l_uu")) < 1e-13
    assert abs(port.cost(p, g("X0"), sc["U0"][b], sc["poly"][b], sc["xlp"][b]) - float(g("J0"))) < 1e-12 * abs(float(g("J0")))
    ok, k, K = port.backward(p, dim, pos, g("X0"), sc["U0"][b], sc["poly"][b], sc["xlp"][b], 1.0)
    assert ok == 1 and rel(k, g("k")) < 1e-11 and rel(K, g("K")) < 1e-11
    Xn, Un = port.forward(p, g("X0"), sc["U0"][b], g("k"), g("K"), 1.0)
    assert rel(Xn, g("X_new")) < 1e-13 and rel(Un, g("U_new")) < 1e-13
    Xh, Uh = port.forward(p, g("X0"), sc["U0"][b], g("k"), g("K"), 0.25)
    assert rel(Xh, g("X_new_a025")) < 1e-13 and rel(Uh, g("U_new_a025")) < 1e-13


def test_port_alpha_extension_matches_reference_pinned_wrapper(port, gold):
    """n_alphas=8, merit=get_J: the oracle wrapper that only calls reference functions (RefOracle.solve_alpha)."""
    p = port.params_default(horizon=50)
    sc = scenarios.make_scenarios(p, 100, 3, 3)
    for b in range(3):
        r = port.solve(p, sc["dim"][b], sc["pos"][b], sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b], n_alphas=8, merit=0)
        assert r["iters"] == int(gold["alpha8/%d/iters" % b])
        assert rel(r["U"], gold["alpha8/%d/U" % b]) < TOL
        assert abs(r["J"] - float(gold["alpha8/%d/J" % b])) < TOL * abs(float(gold["alpha8/%d/J" % b]))


def test_default_parameters_are_the_reference_defaults(port):
    p = port.params_default()
    assert (p.horizon, p.max_iterations, p.num_of_local_wpts, p.poly_order) == (40, 20, 20, 5)
    assert (p.timestep, p.tolerance, p.w_pos, p.w_vel, p.w_acc, p.w_yawrate) == (0.1, 1e-4, 0.65, 3.0, 1.0, 4.0)
    assert (p.q1_front, p.q2_front, p.q1_rear, p.q2_rear) == (2.75, 2.75, 2.5, 2.5)
    assert (p.acc_max, p.acc_min, p.steer_angle_max, p.wheelbase) == (2.0, -5.5, 0.75, 2.94)
    assert p.ego_front == 1.47 + 0.925 and (p.lamb_factor, p.lamb_max) == (10.0, 10000.0)
    U = port.default_controls(p.replace(horizon=50))
    assert np.all(U[0::2] == 0.5) and np.all(U[1:50:2] == 0.0) and np.all(U[51::2] == 0.1)


def test_port_warp_matches_reference_golden(port):
    g = np.load(os.path.join(HERE, "golden", "warp_golden.npz"))
    G = cilqr_b200.CostmapGeom
    sg, tg = G(*g["small/src_geom"]), G(*g["small/tgt_geom"])
    rs, cs = port.geom_size(sg)
    src = scenarios.make_source_map(rs, cs)
    val, idx = port.warp(src, sg, tg, g["small/poses"])
    assert np.array_equal(idx, g["small/idx"])
    assert np.array_equal(val.view(np.uint32), g["small/val"].view(np.uint32))  # bit-for-bit, NaNs included
    assert (idx < 0).any(), "the golden set must cover out-of-map cells"
    # config 2 of BASELINE.json: reference geometry, 400x400 target
    sg2, tg2 = scenarios.reference_global_geom(), G(80.0, 80.0, 0.2, 35.0, 0.0)
    rs2, cs2 = port.geom_size(sg2)
    assert (rs2, cs2) == (1506, 1506) and port.geom_size(tg2) == (400, 400)
    src2 = scenarios.make_source_map(rs2, cs2)
    val2, idx2 = port.warp(src2, sg2, tg2, np.array([[93.14, -205.96, np.sin(0.7), np.cos(0.7)]]))
    g0, g1 = idx2[0] % rs2, idx2[0] // rs2
    assert int(np.sum(g0.astype(np.int64) * 1506 + g1.astype(np.int64))) == int(g["config2/index_checksum"]) == 149172896914
    assert int(np.sum(idx2[0] < 0)) == int(g["config2/oob"]) == 0
    assert int(np.sum(np.isnan(val2[0]))) == int(g["config2/nan_count"])
    assert np.array_equal(idx2[0][:400], g["config2/idx_row0"])
