"""GPU diagnostics: every stage and the fused solve against the CPU checker, printing error magnitudes
instead of asserting (one gpurun round trip should tell as much as possible).  Not a test, not the bench."""
import os
import sys
import time
import traceback

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import cilqr_b200  # noqa: E402
from cilqr_b200 import scenarios  # noqa: E402
import oracle  # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def section(name):
    print("\n=== %s" % name, flush=True)


def main():
    print("torch", torch.__version__, "cuda", torch.cuda.is_available(), torch.cuda.get_device_name(0) if torch.cuda.is_available() else "-")
    chk = oracle.best_oracle()
    port = oracle.PortOracle()
    print("checker:", chk.kind)
    p = chk.params_default(horizon=50)
    N = 50
    for n_obs, curved in [(3, False), (0, True), (10, True)]:
        section("stages n_obs=%d curved=%s" % (n_obs, curved))
        try:
            B = 5
            sc = scenarios.make_scenarios(p, 0, B, n_obs, curved=curved)
            s = cilqr_b200.iLQR(p, max_batch=64, max_obs=max(n_obs, 1))
            if n_obs:
                s.set_obstacles_dense(sc["dim"], sc["pos"])
            X = s.get_nominal_trajectory(sc["x0"], sc["U0"])
            Xr = np.stack([chk.rollout(p, sc["x0"][b], sc["U0"][b]) for b in range(B)])
            print("rollout      ", rel(X, Xr))
            lx, lxx = s.get_state_cost(Xr, sc["poly"], sc["xlp"])
            ref = [chk.state_cost(p, sc["dim"][b] if n_obs else None, sc["pos"][b] if n_obs else None, Xr[b], sc["poly"][b], sc["xlp"][b]) for b in range(B)]
            print("l_x          ", rel(lx, np.stack([r[0] for r in ref])), " l_xx", rel(lxx, np.stack([r[1] for r in ref])))
            lu, luu = s.get_control_cost(Xr, sc["U0"])
            ref = [chk.control_cost(p, Xr[b], sc["U0"][b]) for b in range(B)]
            print("l_u          ", rel(lu, np.stack([r[0] for r in ref])), " l_uu", rel(luu, np.stack([r[1] for r in ref])))
            J = s.get_J(Xr, sc["U0"], sc["poly"], sc["xlp"])
            Jr = np.array([chk.cost(p, Xr[b], sc["U0"][b], sc["poly"][b], sc["xlp"][b]) for b in range(B)])
            print("J            ", rel(J, Jr))
            for lamb in (1.0, 1e-5):
                ok = s.backward_pass(Xr, sc["U0"], sc["poly"], sc["xlp"], lamb)
                ref = [chk.backward(p, sc["dim"][b] if n_obs else None, sc["pos"][b] if n_obs else None, Xr[b], sc["U0"][b], sc["poly"][b], sc["xlp"][b], lamb) for b in range(B)]
                kr, Kr = np.stack([r[1] for r in ref]), np.stack([r[2] for r in ref])
                print("backward l=%g" % lamb, "ok", ok.tolist(), " k", rel(s.k, kr), " K", rel(s.K, Kr))
            Xn, Un = s.forward_pass(Xr, sc["U0"], kr, Kr)
            ref = [chk.forward(p, Xr[b], sc["U0"][b], kr[b], Kr[b], 1.0) for b in range(B)]
            print("forward a=1  ", rel(Xn, np.stack([r[0] for r in ref])), rel(Un, np.stack([r[1] for r in ref])))
            Xn, Un = s.forward_pass(Xr, sc["U0"], kr, Kr, alpha=0.25)
            ref = [chk.forward(p, Xr[b], sc["U0"][b], kr[b], Kr[b], 0.25) for b in range(B)]
            print("forward a=.25", rel(Xn, np.stack([r[0] for r in ref])), rel(Un, np.stack([r[1] for r in ref])))
            U = sc["U0"].copy()
            s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"])
            for b in range(B):
                r = chk.solve(p, sc["dim"][b] if n_obs else None, sc["pos"][b] if n_obs else None, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
                print("solve b=%d iters %d/%d status %d  U %.2e X %.2e J %.2e" % (b, s.iterations[b], r["iters"], s.status[b], rel(U[b], r["U"]),
                                                                              rel(s.X_result[b], r["X"]), abs(s.J_result[b] - r["J"]) / abs(r["J"])))
            for merit in (0, 1):
                U = sc["U0"].copy()
                s.get_optimal_control_seq(sc["x0"], U, sc["poly"], sc["xlp"], n_alphas=8, merit=merit)
                for b in range(min(B, 3)):
                    r = port.solve(p, sc["dim"][b] if n_obs else None, sc["pos"][b] if n_obs else None, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b], n_alphas=8, merit=merit)
                    print("alpha8 merit=%d b=%d iters %d/%d status %d/%d U %.2e J %.2e" % (merit, b, s.iterations[b], r["iters"], s.status[b], r["status"],
                                                                                       rel(U[b], r["U"]), abs(s.J_result[b] - r["J"]) / abs(r["J"])))
            s.close()
        except Exception:
            traceback.print_exc()

    section("batched solve parity + timing (device-resident)")
    try:
        for B, n_obs, na in [(1024, 3, 1), (4096, 10, 1), (4096, 10, 8), (8192, 20, 1)]:
            sc = scenarios.make_scenarios(p, 0, B, n_obs, dense=False)
            s = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs)
            dev = torch.device("cuda")
            t = {k: torch.from_numpy(sc[k]).to(dev) for k in ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")}
            s.set_obstacles_cv(t["obs_state0"], t["obs_LW"], t["obs_sigma"], sc["kappa"])
            U = t["U0"].clone()
            s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"], n_alphas=na, merit=1 if na > 1 else 0)
            torch.cuda.synchronize()
            ts = []
            for _ in range(3):
                U = t["U0"].clone()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                s.get_optimal_control_seq(t["x0"], U, t["poly"], t["xlp"], n_alphas=na, merit=1 if na > 1 else 0)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            it = s.iterations.cpu().numpy()
            st = s.status.cpu().numpy()
            print("B=%d n_obs=%d n_alphas=%d: %.2f ms -> %.0f solves/s; iters mean %.1f; status hist %s" % (
                B, n_obs, na, min(ts), B / min(ts) * 1e3, it.mean(), np.bincount(st, minlength=4).tolist()))
            # parity on a sample vs the checker (dense expansion on the host feeds the unmodified Obstacle)
            Uh = U.cpu().numpy()
            Jh = s.J_result.cpu().numpy()
            worst, bad_it = 0.0, 0
            sample = list(range(0, B, max(1, B // 24)))[:24]
            scd = scenarios.make_scenarios(p, 0, B, n_obs, dense=False)
            for b in sample:
                dim, pos = scenarios.expand_obstacles_cv(p, scd["obs_state0"][b], scd["obs_LW"][b], scd["obs_sigma"][b], scd["kappa"])
                if na == 1:
                    r = chk.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
                else:
                    r = port.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b], n_alphas=na, merit=1)
                e = rel(Uh[b], r["U"])
                worst = max(worst, e)
                bad_it += int(it[b] != r["iters"])
                if e > 1e-9:
                    print("   b=%d relU %.3e iters %d/%d J %.6f/%.6f" % (b, e, it[b], r["iters"], Jh[b], r["J"]))
            print("   parity sample %d: worst rel U %.2e, iteration mismatches %d" % (len(sample), worst, bad_it))
            s.close()
    except Exception:
        traceback.print_exc()

    section("costmap warp")
    try:
        G = cilqr_b200.CostmapGeom
        g = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "warp_golden.npz"))
        sg, tg = G(*g["small/src_geom"]), G(*g["small/tgt_geom"])
        rs, cs = chk.geom_size(sg)
        src = scenarios.make_source_map(rs, cs)
        VARS = (1, 2) if os.environ.get('FIRST_LIGHT_TMA', '1') == '1' else (1,)
        for variant in VARS:
            try:
                val, idx = cilqr_b200.warp_batch(src, sg, tg, g["small/poses"], variant=variant)
                print("small variant %d: idx mismatches %d, value bit mismatches %d" % (variant, int(np.sum(idx != g["small/idx"])),
                                                                                  int(np.sum(val.view(np.uint32) != g["small/val"].view(np.uint32)))))
            except Exception as e:
                print("small variant %d: %s" % (variant, e))
        sg2, tg2 = scenarios.reference_global_geom(), G(80.0, 80.0, 0.2, 35.0, 0.0)
        src2 = scenarios.make_source_map(1506, 1506)
        th = np.arange(64) * 2 * np.pi / 64
        poses = np.stack([93.14 + 15 * np.cos(3 * th), -205.96 + 15 * np.sin(5 * th), np.sin(th), np.cos(th)], axis=1)
        poses[0] = (93.14, -205.96, np.sin(0.7), np.cos(0.7))
        vr, ir = chk.warp(src2, sg2, tg2, poses, threads=0)
        val, idx = cilqr_b200.warp_batch(src2, sg2, tg2, poses, variant=1)
        print("config2 direct: idx mismatches %d / %d, value mismatches %d, checksum pose0 %d" % (
            int(np.sum(idx != ir)), idx.size, int(np.sum(val.view(np.uint32) != vr.view(np.uint32))),
            int(np.sum((idx[0] % 1506).astype(np.int64) * 1506 + idx[0] // 1506))))
        # sweep, device resident
        dev = torch.device("cuda")
        for T in (512, 1024, 2048, 4096):
            S = 2 * T
            sgT, tgT = G(S * 0.2, S * 0.2, 0.2, 0.0, 0.0), G(T * 0.2, T * 0.2, 0.2, 0.0, 0.0)
            srcT = torch.from_numpy(scenarios.make_source_map(S, S, nan_band=False)).to(dev)
            P = max(4, min(64, (1 << 28) // (T * T)))
            posesT = torch.from_numpy(scenarios.make_poses(P, (0.0, 0.0), 0.05 * T * 0.2)).to(dev)
            out = torch.empty((P, T * T), dtype=torch.float32, device=dev)
            idxT = torch.empty((P, T * T), dtype=torch.int32, device=dev)
            res = {}
            for variant in VARS:
                for with_idx in (False, True):
                    cilqr_b200.warp_batch(srcT, sgT, tgT, posesT, variant=variant, out=out, idx=idxT if with_idx else None, want_idx=False)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(3):
                        cilqr_b200.warp_batch(srcT, sgT, tgT, posesT, variant=variant, out=out, idx=idxT if with_idx else None, want_idx=False)
                    e1.record()
                    torch.cuda.synchronize()
                    ms = e0.elapsed_time(e1) / 3
                    by = P * T * T * (12 if with_idx else 8)
                    res[(variant, with_idx)] = by / ms / 1e6
                if variant == 1:
                    keep_idx, keep_val = idxT.clone(), out.clone()
                else:
                    print("   T=%d: tma vs direct idx mismatches %d, value mismatches %d" % (T, int((keep_idx != idxT).sum()),
                                                                                         int((keep_val.view(torch.int32) != out.view(torch.int32)).sum())))
            print("T=%d P=%d GB/s:" % (T, P), {k: round(v) for k, v in res.items()})
            del srcT, out, idxT
    except Exception:
        traceback.print_exc()
    print("\nlaunches:", cilqr_b200.launch_count())


if __name__ == "__main__":
    main()
