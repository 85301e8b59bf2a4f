"""Generates the committed golden vectors from the UNMODIFIED reference compiled in oracle/_ref
(`make -C oracle ref`, needs /root/reference).  Run here, in the build container:

    python tests/golden/make_golden.py

Outputs (small, committed): tests/golden/solver_golden.npz, tests/golden/warp_golden.npz, tests/golden/propagate_golden.npz.
The GPU box has no /root/reference; its tests read these files (and the prebuilt oracle/_ref .so).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
import cilqr_b200  # noqa: E402
from cilqr_b200 import scenarios  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def solver_golden(R):
    out = {}
    cases = [("n50_obs3", dict(horizon=50), 3, False, 6), ("n50_obs0_curved", dict(horizon=50), 0, True, 3),
             ("n40_obs2_curved", dict(), 2, True, 3), ("n50_obs10", dict(horizon=50), 10, False, 3)]
    for name, kw, n_obs, curved, B in cases:
        p = R.params_default(**kw)
        sc = scenarios.make_scenarios(p, 0, B, n_obs, curved=curved)
        N = p.horizon
        Us, Xs, Js, its, lambs = [], [], [], [], []
        for b in range(B):
            dim = sc["dim"][b] if n_obs else None
            pos = sc["pos"][b] if n_obs else None
            r = R.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
            Us.append(r["U"]); Xs.append(r["X"]); Js.append(r["J"]); its.append(r["iters"]); lambs.append(r["lamb"])
        out[name + "/horizon"] = np.int32(N)
        out[name + "/n_obs"] = np.int32(n_obs)
        out[name + "/curved"] = np.int32(curved)
        out[name + "/B"] = np.int32(B)
        out[name + "/U"] = np.stack(Us); out[name + "/X"] = np.stack(Xs); out[name + "/J"] = np.array(Js)
        out[name + "/iters"] = np.array(its, dtype=np.int32); out[name + "/lamb"] = np.array(lambs)
        # stage vectors on scenario 0: nominal rollout, derivatives, gains, candidate
        b = 0
        dim = sc["dim"][b] if n_obs else None
        pos = sc["pos"][b] if n_obs else None
        X0 = R.rollout(p, sc["x0"][b], sc["U0"][b])
        lx, lxx = R.state_cost(p, dim, pos, X0, sc["poly"][b], sc["xlp"][b])
        lu, luu = R.control_cost(p, X0, sc["U0"][b])
        J0 = R.cost(p, X0, sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        ok, k, K = R.backward(p, dim, pos, X0, sc["U0"][b], sc["poly"][b], sc["xlp"][b], 1.0)
        Xn, Un = R.forward(p, X0, sc["U0"][b], k, K, 1.0)
        Xh, Uh = R.forward(p, X0, sc["U0"][b], k, K, 0.25)
        for key, val in dict(X0=X0, l_x=lx, l_xx=lxx, l_u=lu, l_uu=luu, J0=np.float64(J0), k=k, K=K, X_new=Xn, U_new=Un,
                             X_new_a025=Xh, U_new_a025=Uh).items():
            out[name + "/stage/" + key] = val
    # alpha line-search extension pinned on reference functions (oracle.RefOracle.solve_alpha)
    p = R.params_default(horizon=50)
    sc = scenarios.make_scenarios(p, 100, 3, 3)
    for b in range(3):
        r = R.solve_alpha(p, sc["dim"][b], sc["pos"][b], sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b], 8)
        out["alpha8/%d/U" % b] = r["U"]; out["alpha8/%d/J" % b] = np.float64(r["J"]); out["alpha8/%d/iters" % b] = np.int32(r["iters"])
    # LocalPlanner / run_step on the straight road of SURVEY.md §8(d)
    plan = np.stack([np.arange(200, dtype=np.float64), np.zeros(200)], axis=1)
    ego = np.array([0.3, 0.2, 4.0, 0.01])
    poly, xlp, n = R.local_plan(p, plan, ego)
    Xr, Ur = R.run_step(p, None, None, plan, ego)
    out["run_step/poly"] = poly; out["run_step/xlp"] = xlp; out["run_step/X"] = Xr; out["run_step/U"] = Ur
    plan2 = np.stack([np.arange(200, dtype=np.float64) * 0.9, 0.002 * (np.arange(200) * 0.9) ** 2], axis=1)
    poly2, xlp2, _ = R.local_plan(p, plan2, np.array([5.2, 0.4, 3.0, 0.05]))
    out["local_plan_curved/poly"] = poly2; out["local_plan_curved/xlp"] = xlp2
    np.savez_compressed(os.path.join(HERE, "solver_golden.npz"), **out)
    print("solver_golden.npz:", len(out), "arrays")


def warp_golden(R):
    out = {}
    G = cilqr_b200.CostmapGeom
    # small case: every index stored
    sg, tg = G(16.0, 12.0, 0.2, 3.0, -2.0), G(6.0, 4.4, 0.2, 1.0, 0.5)
    rs, cs = R.geom_size(sg)
    src = scenarios.make_source_map(rs, cs)
    th = np.array([0.0, 0.7, 2.0, 3.141592653589793, 4.4, 5.9])
    poses = np.stack([3.0 + 0.3 * np.cos(th), -2.0 + 0.25 * np.sin(th), np.sin(th), np.cos(th)], axis=1)
    poses[3, 0] += 9.0  # drives part of the frame out of the map (reference throws; -1 / NaN here)
    val, idx = R.warp(src, sg, tg, poses)
    out["small/src_geom"] = np.array([sg.length_x, sg.length_y, sg.resolution, sg.pos_x, sg.pos_y])
    out["small/tgt_geom"] = np.array([tg.length_x, tg.length_y, tg.resolution, tg.pos_x, tg.pos_y])
    out["small/poses"] = poses; out["small/idx"] = idx; out["small/val"] = val
    # config 2: reference geometry, 400x400 target, pose (93.14, -205.96, 0.7): checksum only
    sg2 = scenarios.reference_global_geom()
    tg2 = G(80.0, 80.0, 0.2, 35.0, 0.0)
    rs2, cs2 = R.geom_size(sg2)
    src2 = scenarios.make_source_map(rs2, cs2)
    pose2 = np.array([[93.14, -205.96, np.sin(0.7), np.cos(0.7)]])
    val2, idx2 = R.warp(src2, sg2, tg2, pose2)
    g0, g1 = idx2[0] % rs2, idx2[0] // rs2
    out["config2/index_checksum"] = np.int64(np.sum(g0.astype(np.int64) * 1506 + g1.astype(np.int64)))
    out["config2/linear_checksum"] = np.int64(np.sum(idx2[0].astype(np.int64)))
    out["config2/oob"] = np.int64(np.sum(idx2[0] < 0))
    out["config2/value_sum"] = np.float64(np.nansum(val2[0].astype(np.float64)))
    out["config2/nan_count"] = np.int64(np.sum(np.isnan(val2[0])))
    out["config2/idx_row0"] = idx2[0][:400].copy()
    np.savez_compressed(os.path.join(HERE, "warp_golden.npz"), **out)
    print("warp_golden.npz: config2 checksum", int(out["config2/index_checksum"]), "oob", int(out["config2/oob"]))


def propagate_golden(R):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_propagate as tp
    out = {}
    for name, geom, theta, sigma, index in tp.CASES:
        g = cilqr_b200.CostmapGeom(*geom)
        rows, cols = R.geom_size(g)
        c, v, e = R.propagate(tp.make_map(rows, cols), g, np.sin(theta), np.cos(theta), *sigma, index=index, want_ellipse=True)
        out[name + "/count"] = c; out[name + "/value"] = v; out[name + "/ellipse"] = e
    np.savez_compressed(os.path.join(HERE, "propagate_golden.npz"), **out)
    print("propagate_golden.npz:", len(out), "arrays")


if __name__ == "__main__":
    oracle.build(("ref",))
    R = oracle.RefOracle()
    if "--only-propagate" not in sys.argv:
        solver_golden(R)
        warp_golden(R)
    propagate_golden(R)
