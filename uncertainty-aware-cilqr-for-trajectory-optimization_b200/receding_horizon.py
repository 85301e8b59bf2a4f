"""Receding-horizon wrapper around the batched solver: the caller side of the hot path without ROS.

What ilqr_uncertainty_node::odomCallback does every odometry tick (CILQR/src/ilqr/include/ilqr/ilqr_uncertainty_node.cpp:
82-130), for B ego vehicles at once:

  1. Gaussian noise on the measured pose: x + N(0, sigma_x), y + N(0, sigma_y), theta + N(0, sigma_theta), v unchanged
     (:82-90, :111); the reference seeds std::mt19937 from std::random_device, here the draws come from a counter-based
     generator keyed by (seed, tick, scenario) so that a run can be repeated;
  2. run_step on the noisy state (:119) — LocalPlanner + solve, warm-started from the previous tick's controls, NOT shifted
     (iLQR.cpp:201, :222, :253); `shift=True` is an extension that advances the warm start by one step;
  3. the first control U_result(:, 0) is the command (:127);
  4. an Experiment record {start_time, start_pos (the noise-free state), planning_time, X, U} is published
     (:128, publishExperimentData :243-283, vehiclepub/msg/Experiment.msg) — here appended to a flat binary log.

The plant between ticks is the reference's own kinematic model (Model::forward_simulate, Model.cpp:17-30 — CARLA is out of
scope): the next true state is the noise-free state advanced by the commanded control.
"""
import struct
import time

import numpy as np

MAGIC = b"CILQREXP"
VERSION = 1
_HEADER = struct.Struct("<8sIIII")  # magic, version, horizon N, batch B, reserved


def record_dtype(N):
    """One Experiment message (vehiclepub/msg/Experiment.msg) per (tick, scenario); X and U flattened like
    publishExperimentData does (column by column)."""
    return np.dtype([("tick", "<u4"), ("scenario", "<u4"), ("start_time", "<f8"), ("planning_time", "<f8"),
                     ("start_pos", "<f8", (4,)), ("X", "<f8", (4 * (N + 1),)), ("U", "<f8", (2 * N,))])


class ExperimentLog:
    """Append-only binary log of Experiment records."""

    def __init__(self, path, N, B):
        self.N, self.B = int(N), int(B)
        self.f = open(path, "wb")
        self.f.write(_HEADER.pack(MAGIC, VERSION, self.N, self.B, 0))
        self.dt = record_dtype(self.N)

    def append(self, tick, start_time, planning_time, start_pos, X, U):
        rec = np.zeros(self.B, dtype=self.dt)
        rec["tick"] = tick
        rec["scenario"] = np.arange(self.B)
        rec["start_time"] = start_time
        rec["planning_time"] = planning_time
        rec["start_pos"] = start_pos
        rec["X"] = X
        rec["U"] = U
        self.f.write(rec.tobytes())

    def close(self):
        self.f.close()


def read_log(path):
    """-> (N, B, records) with records a structured array of record_dtype(N)."""
    with open(path, "rb") as f:
        magic, version, N, B, _ = _HEADER.unpack(f.read(_HEADER.size))
        if magic != MAGIC or version != VERSION:
            raise ValueError("not a CILQR experiment log")
        return N, B, np.frombuffer(f.read(), dtype=record_dtype(N))


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    return z ^ (z >> np.uint64(31))


def pose_noise(seed, tick, B, sigma):
    """[B,3] draws N(0, sigma_k): Box-Muller on SplitMix64 streams keyed by (seed, tick, scenario, k)."""
    with np.errstate(over="ignore"):
        key = (np.uint64(seed) ^ (np.uint64(tick) << np.uint64(32))) + np.arange(B, dtype=np.uint64) * np.uint64(8)
        out = np.zeros((B, 3))
        for k in range(3):
            a = _splitmix64(key + np.uint64(2 * k))
            b = _splitmix64(key + np.uint64(2 * k + 1))
            u1 = ((a >> np.uint64(11)).astype(np.float64) + 1.0) / 9007199254740993.0
            u2 = (b >> np.uint64(11)).astype(np.float64) / 9007199254740992.0
            out[:, k] = sigma[k] * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    return out


class RecedingHorizon:
    """B closed loops on one shared global plan.

        loop = RecedingHorizon(solver, global_plan, x0 [B,4], sigma=(0.16, 0.16, 0.017), seed=1, log_path="run.bin")
        for _ in range(ticks): loop.tick()
    """

    def __init__(self, solver, global_plan, x0, sigma=(0.0, 0.0, 0.0), seed=0, shift=False, log_path=None):
        self.solver = solver
        self.solver.set_global_plan(global_plan)
        self.state = np.array(x0, dtype=np.float64).reshape(-1, 4)
        self.B = self.state.shape[0]
        self.sigma, self.seed, self.shift = tuple(float(s) for s in sigma), int(seed), bool(shift)
        self.k = 0
        self.commands = []
        self.log = ExperimentLog(log_path, solver.params.horizon, self.B) if log_path else None

    def tick(self):
        N = self.solver.params.horizon
        noisy = self.state.copy()
        if any(s != 0.0 for s in self.sigma):
            n = pose_noise(self.seed, self.k, self.B, self.sigma)
            noisy[:, 0] += n[:, 0]
            noisy[:, 1] += n[:, 1]
            noisy[:, 3] += n[:, 2]
        if self.shift and self.k > 0:  # extension: advance the warm start by one step, repeat the last control
            u = self.solver.control_seq[:self.B].reshape(self.B, N, 2)
            u[:, :-1] = u[:, 1:].copy()
        t_begin = time.time()
        t0 = time.perf_counter()
        X, U = self.solver.run_step(noisy)
        planning_time = time.perf_counter() - t0
        X, U = np.asarray(X), np.asarray(U)
        cmd = U[:, :2].copy()  # U_result(0, 0), U_result(1, 0)  (ilqr_uncertainty_node.cpp:127)
        self.commands.append(cmd)
        if self.log:
            self.log.append(self.k, t_begin, planning_time, self.state, X, U)
        # plant: Model::forward_simulate on the noise-free state with the commanded control
        u_step = np.zeros((self.B, 2 * N))
        u_step[:, :2] = cmd
        self.state = np.asarray(self.solver.get_nominal_trajectory(self.state, u_step))[:, 4:8].copy()
        self.k += 1
        return X, U

    def close(self):
        if self.log:
            self.log.close()
