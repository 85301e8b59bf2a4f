"""CPU-only host logic: seeded scenario generator, rank sharding, the LocalPlanner restatement against the
reference's golden output, and the world-size-2 gather over gloo."""
import os
import subprocess
import sys

import numpy as np

import cilqr_b200
from cilqr_b200 import scenarios, dist

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_scenarios_are_independent_of_batching(port):
    p = port.params_default(horizon=50)
    a = scenarios.make_scenarios(p, 0, 16, 5, curved=True)
    b = scenarios.make_scenarios(p, 8, 4, 5, curved=True)
    for k in ("x0", "poly", "xlp", "obs_state0", "dim", "pos"):
        assert np.array_equal(a[k][8:12], b[k])
    assert a["x0"][:, 0].min() >= 0 and a["x0"][:, 0].max() < 2 and a["x0"][:, 2].min() >= 2 and a["x0"][:, 2].max() < 6
    assert a["obs_state0"][..., 0].min() >= 6 and a["obs_state0"][..., 0].max() < 30
    assert a["dim"].shape == (16, 5, 100) and a["pos"].shape == (16, 5, 200)
    # covariance inflation: width grows along the horizon, length does not
    assert np.all(np.diff(a["dim"][0, 0, 1::2]) >= 0) and np.all(a["dim"][0, 0, 0::2] == a["dim"][0, 0, 0])
    assert np.isclose(a["dim"][0, 0, 0], p.length + 2 * scenarios.KAPPA * 0.16)


def test_shard_range_partitions_contiguously():
    for total, world in [(65536, 8), (10, 3), (7, 8), (4096, 2), (1, 1)]:
        spans = [dist.shard_range(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


def test_local_planner_matches_reference_golden(port):
    g = np.load(os.path.join(HERE, "golden", "solver_golden.npz"))
    p = port.params_default(horizon=50)
    lp = cilqr_b200.LocalPlanner(p)
    lp.set_global_planner(np.stack([np.arange(200, dtype=np.float64), np.zeros(200)], axis=1))
    poly, xlp, _ = lp.plan_batch(np.array([[0.3, 0.2, 4.0, 0.01]]))
    assert np.array_equal(xlp[0], g["run_step/xlp"])
    xs = np.linspace(xlp[0, 0], xlp[0, 1], 50)
    y_mine = sum(poly[0, j] * xs ** j for j in range(6))
    y_ref = sum(g["run_step/poly"][j] * xs ** j for j in range(6))
    assert np.max(np.abs(y_mine - y_ref)) < 1e-9
    # a curved plan: coefficients agree to rounding because the same pivoted Householder QR is used
    t = np.arange(200, dtype=np.float64) * 0.9
    lp.set_global_planner(np.stack([t, 0.002 * t ** 2], axis=1))
    poly2, xlp2, _ = lp.plan_batch(np.array([[5.2, 0.4, 3.0, 0.05]]))
    assert np.array_equal(xlp2[0], g["local_plan_curved/xlp"])
    xs = np.linspace(xlp2[0, 0], xlp2[0, 1], 50)
    y_mine = sum(poly2[0, j] * xs ** j for j in range(6))
    y_ref = sum(g["local_plan_curved/poly"][j] * xs ** j for j in range(6))
    assert np.max(np.abs(y_mine - y_ref)) < 1e-9 * max(1.0, np.max(np.abs(y_ref)))


_GLOO = r"""
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as td
import cilqr_b200
from cilqr_b200 import dist, scenarios
import oracle
rank, local, world = dist.init_process_group("gloo")
port = oracle.PortOracle()
p = port.params_default(horizon=50)
TOTAL = 6
lo, hi = dist.shard_range(TOTAL, rank, world)
sc = scenarios.make_scenarios(p, lo, hi - lo, 2)
# each rank solves its shard (the CPU checker stands in for the GPU here: this test is about the
# sharding + gather plumbing, which is identical on NCCL)
r = port.solve_batch(p, hi - lo, 2, sc["dim"], sc["pos"], sc["x0"], sc["U0"], sc["poly"], sc["xlp"], threads=1)
U = dist.gather_trajectories(torch.from_numpy(r["U"]), counts=[h - l for l, h in (dist.shard_range(TOTAL, q, world) for q in range(world))])
if rank == 0:
    full = scenarios.make_scenarios(p, 0, TOTAL, 2)
    want = port.solve_batch(p, TOTAL, 2, full["dim"], full["pos"], full["x0"], full["U0"], full["poly"], full["xlp"], threads=1)["U"]
    assert U.shape == (TOTAL, 100), U.shape
    assert np.array_equal(U.numpy(), want), "sharded + gathered result differs from the unsharded solve"
    print("GLOO_OK")
# ragged shards (7 scenarios over 2 ranks: 4 + 3) with counts left to the ranks to exchange
lo7, hi7 = dist.shard_range(7, rank, world)
rows = torch.arange(lo7, hi7, dtype=torch.float64).reshape(-1, 1).repeat(1, 3)
G7 = dist.gather_trajectories(rows)
assert G7.shape == (7, 3) and torch.equal(G7[:, 0], torch.arange(7, dtype=torch.float64)), G7
if rank == 0:
    print("RAGGED_OK")
td.barrier()
td.destroy_process_group()
"""


def test_world_size_2_shard_and_gather_over_gloo(tmp_path, port):
    script = tmp_path / "gloo_job.py"
    script.write_text(_GLOO % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29611", str(script)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "GLOO_OK" in r.stdout and "RAGGED_OK" in r.stdout


# ------------------------------------------------------------------------------------------------ row N4: receding-horizon wrapper (host side)
class _FakeSolver:
    """Stands in for cilqr_b200.iLQR: the wrapper's bookkeeping (noise, warm-start shift, log) needs no GPU."""

    def __init__(self, p, B):
        self.params, self.B = p, B
        N = p.horizon
        self.control_seq = np.tile(np.arange(2 * N, dtype=np.float64), (B, 1))
        self.seen = []

    def set_global_plan(self, plan):
        self.plan = plan

    def run_step(self, ego):
        N = self.params.horizon
        self.seen.append((ego.copy(), self.control_seq.copy()))
        X = np.tile(ego, (1, N + 1)) + np.arange(4 * (N + 1))
        return X, self.control_seq[:ego.shape[0]] + 1.0

    def get_nominal_trajectory(self, x0, U):
        N = self.params.horizon
        X = np.tile(x0, (1, N + 1))
        X[:, 4:8] = x0 + 0.1 * U[:, [0, 1, 0, 1]]
        return X


def test_receding_horizon_noise_log_and_shift(port, tmp_path):
    from cilqr_b200 import receding_horizon as rh
    p = port.params_default(horizon=50)
    B = 3
    x0 = np.array([[0.0, 0.1, 4.0, 0.0], [5.0, -0.2, 3.0, 0.01], [9.0, 0.0, 5.0, -0.02]])
    path = str(tmp_path / "run.bin")
    s = _FakeSolver(p, B)
    loop = rh.RecedingHorizon(s, np.zeros((10, 2)), x0, sigma=(0.16, 0.16, 0.017), seed=11, shift=True, log_path=path)
    for _ in range(3):
        loop.tick()
    loop.close()
    # noise: x, y, theta perturbed, v untouched (ilqr_uncertainty_node.cpp:111); reproducible; roughly the right scale
    ego0 = s.seen[0][0]
    assert np.array_equal(ego0[:, 2], x0[:, 2]) and not np.array_equal(ego0[:, [0, 1, 3]], x0[:, [0, 1, 3]])
    assert np.allclose(rh.pose_noise(11, 0, B, (0.16, 0.16, 0.017)), ego0[:, [0, 1, 3]] - x0[:, [0, 1, 3]], rtol=0, atol=1e-12)
    big = rh.pose_noise(3, 7, 20000, (0.16, 0.16, 0.017))
    assert np.allclose(big.std(axis=0), (0.16, 0.16, 0.017), rtol=0.05) and np.all(np.abs(big.mean(axis=0)) < 0.01)
    # shift extension: tick k >= 1 starts from the previous controls advanced by one step, last control repeated
    N = p.horizon
    before = np.tile(np.arange(2 * N, dtype=np.float64), (B, 1)).reshape(B, N, 2)
    assert np.array_equal(s.seen[0][1].reshape(B, N, 2), before)
    assert np.array_equal(s.seen[1][1].reshape(B, N, 2)[:, :-1], before[:, 1:]) and np.array_equal(s.seen[1][1].reshape(B, N, 2)[:, -1], before[:, -1])
    # log: one Experiment record per (tick, scenario), start_pos = the noise-free state, X / U as returned
    Nr, Br, rec = rh.read_log(path)
    assert (Nr, Br, rec.size) == (N, B, 3 * B)
    assert np.array_equal(rec["tick"], np.repeat(np.arange(3), B)) and np.array_equal(rec["scenario"], np.tile(np.arange(B), 3))
    assert np.array_equal(rec["start_pos"][:B], x0) and np.all(rec["planning_time"] >= 0.0)
    assert np.array_equal(rec["X"][:B], np.tile(ego0, (1, N + 1)) + np.arange(4 * (N + 1)))
    assert np.array_equal(rec["start_pos"][B:2 * B], x0 + 0.1 * loop.commands[0][:, [0, 1, 0, 1]])


def test_bench_clock_sampler_window():
    """bench.py's nvidia-smi sampler: only samples stamped inside the timed region are reported; a region too short to hold
    one falls back to the samples under load and says so."""
    import datetime
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    text = "\n".join([
        "2026/11/05 00:00:00.100, 0, 345, 1965, 140.0, 0x0000000000000001, Not Active, Not Active, Not Active, Not Active",
        "2026/11/05 00:00:01.100, 0, 1965, 1965, 600.0, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
        "2026/11/05 00:00:01.120, 0, 1950, 1965, 640.0, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
        "2026/11/05 00:00:02.000, 0, 1200, 1965, 650.0, 0x0000000000000008, Active, Not Active, Not Active, Not Active",
        "garbage line",
    ])
    rows = bench.ClockSampler.parse(text)
    assert len(rows) == 4
    t = lambda s: datetime.datetime.strptime("2026/11/05 " + s, "%Y/%m/%d %H:%M:%S.%f")
    out = bench.ClockSampler.summarise(rows, t("00:00:01.000"), t("00:00:01.500"))
    assert out["samples"] == 2 and out["window"] == "timed region"
    assert out["sm_mhz"] == 1957.5 and out["sm_max_mhz"] == 1965.0 and out["reasons"] == ["sw_power_cap"]
    out = bench.ClockSampler.summarise(rows, t("00:00:03.000"), t("00:00:03.010"))
    assert out["samples"] == 3 and out["window"].startswith("no sample inside")
    assert "hw_slowdown" in out["reasons"]
    assert bench.ClockSampler.summarise([], None, None)["sm_mhz"] is None
