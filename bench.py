"""bench.py — the driver's measurement contract for the CILQR hot path.

    [blocks of the JSON line added in round 2: `roofline` is the FP64 figure (bound "fp64"), `hbm` the HBM one; `strong_scaling`
    (N > 1: BASELINE config 4 as written, 65 536 scenarios in total sharded over the ranks, gather included); `config3`,
    `few_obstacles`; `costmap_warp.sweep` (all five config-5 sizes) and CPU baselines inside the two costmap blocks;
    `e2e.host_ceiling`, `e2e.u_only_value`.]

    python bench.py --gpus 1 --steps K --warmup W                  (this repo's CUDA path)
    torchrun ... bench.py --gpus N --steps K --warmup W            (one rank per GPU, NCCL gather of U*)
    python bench.py --impl reference --steps K --warmup W          (the reference's own CPU code, oracle/_ref)

A "step" is one pass of the hot path over one batch of synthetic scenarios of BASELINE.json config 4 (the
configuration the target is quoted on): 65 536 independent scenarios per GPU, horizon N = 50, 20 Gaussian-uncertain
constant-velocity obstacles each with covariance propagation over the horizon on the device
(cilqr_set_obstacles_cv), then the batched solve (cilqr_solve_batch, reference semantics n_alphas = 1), then
— for N > 1 — the single NCCL all-gather of the optimal control sequences.  Scenarios shard by batch index with no
data-path collective, so per-GPU work is fixed as N grows ("scaling": "weak").

`value`   whole-job solves/s with the inputs resident in HBM (CUDA events, max over ranks).
`e2e`     the same metric through the C ABI with HOST (pinned) buffers: host->device copies of every input and
          device->host copies of U*, X*, J, iterations, status are inside the timed region.
`roofline` the solve (its round kernels together): mandatory HBM bytes per solve of the batch / its CUDA-event duration against the
          measured copy bandwidth of MEASURED_PEAKS.json; the solver is FP64-issue bound, so `fp64` reports the
          algorithmic FP64 flop rate against an FMA micro-benchmark run on the same GPU.
`costmap_warp` the second half of BASELINE.json's metric: HBM GB/s of costmap_warp_batch (config 5, one grid size).
`cpu_baseline` the reference's own CPU solver (oracle/_ref, unmodified sources) on the host cores of this box, on a
          bounded sample of the same workload (rank 0, N = 1 only).
Nothing here reads /root/reference.  oracle/ is only executed by the cpu_baseline leg and by --impl reference.
"""
import argparse
import datetime
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "batched CILQR solves/sec (N=50)"
UNIT = "solves/s"
HORIZON = 50
N_OBS = 20
BATCH_PER_GPU = 65536
WARP_T = 2048          # config 5 grid edge timed by the default run (source 2T x 2T = 67 MB, poses x 16.8 MB targets)
WARP_POSES = 1024      # BASELINE config 5: 1024 poses, processed in chunks
WARP_CHUNK_BYTES = 2 << 30   # poses per costmap_warp_batch call: as many as fit 2 GiB of value + index maps (64 at T = 2048), at most all 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH_PER_GPU, help="scenarios per GPU")
    ap.add_argument("--n-obs", type=int, default=N_OBS)
    ap.add_argument("--n-alphas", type=int, default=1)
    ap.add_argument("--warp-T", type=int, default=0, help="time one costmap grid edge only (0 = the whole config-5 sweep)")
    ap.add_argument("--no-strong", action="store_true", help="skip the sharded (strong-scaling) block at N > 1")
    ap.add_argument("--no-extra", action="store_true", help="skip the config-3 and few-obstacle blocks")
    ap.add_argument("--no-warp", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-threads", type=int, default=4, help="host threads (one solver handle each) in the end-to-end measurement")
    ap.add_argument("--fused", type=int, default=-1, help="CILQR_OPT_FUSED (-1 automatic, 1 fused on-chip kernel, 0 round kernels)")
    ap.add_argument("--fused-slots", type=int, default=0, help="CILQR_OPT_FUSED_SLOTS (0 automatic)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="the U* exchange at N > 1: peer-to-peer copies on the copy engines "
                                                                                 "(cilqr_gather_copy) or NCCL's all-gather kernel")
    ap.add_argument("--chunks", type=int, default=0, help="CILQR_OPT_CHUNKS (0 automatic: one chunk for device buffers)")
    ap.add_argument("--no-graph", action="store_true", help="CILQR_OPT_GRAPH = 0: direct launches (for ncu launch lists)")
    return ap.parse_args()


def measured_peaks():
    """HBM GB/s from the driver-written MEASURED_PEAKS.json, else the fallback B200_PROFILING.md states."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture of this same command
    (profiles/traffic.json, written by tools/ncu_summary.py traffic); None when no capture is recorded."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            v = json.load(f).get(key)
        return float(v) if v is not None else None
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------ algorithmic counts
def solver_bytes_per_solve(N, n_obs):
    """Mandatory HBM traffic of one solve with the compact obstacle form:
    in: x0 32 + U 16N + poly 48 + xlp 16 + obstacles 72 n_obs (state0 4, LW 2, sigma 3 doubles);
    out: X 32(N+1) + U 16N + J 8 + iterations 4 + status 4."""
    return 32 + 16 * N + 48 + 16 + 72 * n_obs + 32 * (N + 1) + 16 * N + 16


def solver_flops(N, n_obs, cost_passes, deriv_passes):
    """Algorithmic FP64 flops of the work the solver actually ran (cilqr_solve_stats), each operation counted once;
    exp / sincos counted as 20 flops:
      cost pass (Constraints::get_J), per scenario:            N * (5 polyline samples * 22 + 14)
          (the bounded closest-point walk visits ~5 samples: 12 flops for the polynomial, 10 for offsets and bound)
      derivative + backward + forward pass, per scenario:
          obstacles       N * n_obs * 2 * (46 + 20)   (two ellipses per obstacle, one exp each)
          control cost    N * (4 * 22 + 16)
          Riccati sweep   N * 330          (sparse A, B; symmetric V_xx; closed-form 2x2 inverse)
          rollout         N * (40 + 20)    (gain application, bicycle step, one sincos)
    The reference's `iters` counter is NOT used: it also counts iterations in which a rejected nominal only waits for lamb
    to pass lamb_max, which cost nothing here."""
    per_cost = N * (5 * 22 + 14)
    per_deriv = N * n_obs * 2 * 66 + N * 104 + N * 330 + N * 60
    return cost_passes * per_cost + deriv_passes * per_deriv, per_cost, per_deriv


def propagation_flops(N, n_obs):
    """Covariance-inflated obstacle set of one scenario (cilqr_set_obstacles_cv; square root and division counted once each):
    per obstacle a sincos and the longitudinal axis (20 + 12), per (obstacle, timestep) the lever arm, the lateral sigma and
    1/b^2 (19)."""
    return n_obs * (32 + N * 19)


# ------------------------------------------------------------------------------------------------ clocks during the timed region
class ClockSampler:
    """nvidia-smi sampler (B200_PROFILING.md's clocks line).  Started before the warm-up — nvidia-smi needs a second or two to
    come up on an 8-GPU box, longer than a short timed region — and every sample carries nvidia-smi's timestamp, so that only
    the samples taken DURING the timed region (mark_begin / mark_end) are reported.  If the region was too short to hold a
    sample, the samples taken under load during the warm-up stand in and `window` says so."""
    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, gpu_index):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        self.t0 = self.t1 = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=self.tmp, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def mark_begin(self):
        self.t0 = datetime.datetime.now()

    def mark_end(self):
        self.t1 = datetime.datetime.now()

    @classmethod
    def parse(cls, text):
        """[(time or None, sm MHz, max MHz, watts, {reasons})] from the sampler's CSV."""
        rows = []
        for line in text.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 10:
                continue
            try:
                sm, mx, pw = float(f[2]), float(f[3]), float(f[4])
            except ValueError:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f")
            except ValueError:
                ts = None
            rows.append((ts, sm, mx, pw, {n for n, v in zip(cls.NAMES, f[6:10]) if v.lower().startswith("active")}))
        return rows

    @staticmethod
    def summarise(rows, t0, t1):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not rows:
            return out
        inside = [r for r in rows if r[0] is not None and t0 is not None and t1 is not None and t0 <= r[0] <= t1]
        if inside:
            use, out["window"] = inside, "timed region"
        else:  # a region shorter than the sampling period: the samples taken under load (warm-up included)
            top = max(r[3] for r in rows)
            use, out["window"] = [r for r in rows if r[3] >= 0.5 * top] or rows, "no sample inside the timed region; samples under load around it"
        out.update(sm_mhz=float(np.median([r[1] for r in use])), sm_max_mhz=float(max(r[2] for r in use)),
                   power_w_max=float(max(r[3] for r in use)), samples=len(use))
        out["reasons"] = sorted(set().union(*[r[4] for r in use]))
        return out

    def stop(self):
        if self.proc is None:
            return self.summarise([], None, None)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        rows = self.parse(self.tmp.read())
        try:
            os.unlink(self.tmp.name)
        except OSError:
            pass
        return self.summarise(rows, self.t0, self.t1)


# ------------------------------------------------------------------------------------------------ CPU legs (the only users of oracle/)
def cpu_solver():
    import oracle
    if oracle.ref_available():
        return oracle.RefOracle()
    oracle.build(("port",))
    return oracle.PortOracle()


def cpu_solve_sample(chk, p, sc, n, threads=0):
    """Times chk.solve_batch on the first n scenarios of `sc` with every host thread; returns (seconds, threads)."""
    if threads <= 0:  # every host thread, named explicitly: torchrun exports OMP_NUM_THREADS=1 to the ranks it launches
        threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    r = chk.solve_batch(p, n, sc["n_obs"], sc["dim"][:n], sc["pos"][:n], sc["x0"][:n], sc["U0"][:n], sc["poly"][:n], sc["xlp"][:n],
                        threads=threads, want_X=False)
    return time.perf_counter() - t0, int(r["threads"])


def cpu_sample_size(chk, p, sc, target_s, cap):
    """Pilot run -> number of scenarios that take about `target_s` seconds on all host threads."""
    cores = os.cpu_count() or 1
    pilot = min(cap, max(2 * cores, 8))
    dt, used = cpu_solve_sample(chk, p, sc, pilot)
    n = int(max(pilot, min(cap, pilot * target_s / max(dt, 1e-3))))
    return n, used


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation (oracle/_ref = unmodified sources, Eigen kernels, no MKL)
    on all host threads; each step is a bounded sample of the same workload.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import cilqr_b200
    from cilqr_b200 import scenarios
    chk = cpu_solver()
    p = chk.params_default(horizon=HORIZON)
    cap = 4096
    sc = scenarios.make_scenarios(p, 0, cap, args.n_obs, dense=True)
    n, used = cpu_sample_size(chk, p, sc, target_s=max(1.0, min(6.0, 90.0 / max(1, args.steps + args.warmup))), cap=cap)
    for _ in range(args.warmup):
        cpu_solve_sample(chk, p, sc, n)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_solve_sample(chk, p, sc, n)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = "%d of the %d scenarios per step (config 4 inputs: N=50, %d covariance-inflated obstacles, host-expanded)" % (n, args.batch, args.n_obs)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / max(1, args.steps), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(args.batch, args.n_obs, 1), "scenarios_per_gpu": args.batch, "horizon": HORIZON,
                   "n_obs": args.n_obs, "n_alphas": 1, "sample_per_step": n,
                   "sample": "each CPU step solves the first %d scenarios of the workload (obstacles host-expanded), all host threads" % n},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": chk.kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline_leg(p, args, chk):
    """cpu_baseline of the main arm: ~15 s of the reference CPU solver on all host threads (rank 0, N = 1)."""
    from cilqr_b200 import scenarios
    cap = 4096
    sc = scenarios.make_scenarios(p, 0, cap, args.n_obs, dense=True)
    n, used = cpu_sample_size(chk, p, sc, target_s=15.0, cap=cap)
    dt, used = cpu_solve_sample(chk, p, sc, n)
    out = {"value": n / dt, "unit": UNIT, "cores": used, "kind": chk.kind,
           "sample": "first %d scenarios of the step's %d (same inputs, obstacles host-expanded), %.1f s, all host threads, Eigen kernels without MKL"
                     % (n, args.batch, dt)}
    # single-thread figure for context (the reference itself is single-threaded)
    n1 = max(4, min(64, n // max(1, used)))
    dt1, _ = cpu_solve_sample(chk, p, sc, n1, threads=1)
    out["single_thread_value"] = n1 / dt1
    return out


# ------------------------------------------------------------------------------------------------ the CUDA arm
def event_pair():
    import torch
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def max_over_ranks(x, world, device):
    import torch
    import torch.distributed as dist
    if world == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def workload_name(B_total, n_obs, n_alphas):
    if (B_total, n_obs, n_alphas) == (BATCH_PER_GPU, N_OBS, 1):
        return "config4: 65536 scenarios x 20 Gaussian-uncertain obstacles (covariance propagated over the horizon), N=50, n_alphas=1"
    if (B_total, n_obs, n_alphas) == (4096, 10, 8):
        return "config3: 4096 scenarios x 10 Gaussian-uncertain obstacles, N=50, 8 parallel line-search alphas"
    return "custom: %d scenarios x %d Gaussian-uncertain obstacles, N=50, n_alphas=%d" % (B_total, n_obs, n_alphas)


FP64_PEAK = {}


def fp64_peak(lib):
    """DFMA micro-benchmark on this GPU (cilqr_measure_fp64_peak): builder-measured, there is no driver-written FP64 figure."""
    import ctypes as C
    if "v" not in FP64_PEAK:
        v = C.c_double(0.0)
        FP64_PEAK["v"] = v.value if lib.cilqr_measure_fp64_peak(C.byref(v)) != 0 else v.value
    return FP64_PEAK["v"]


def solve_case(args, p, device, rank, world, first_id, B, n_obs, n_alphas, steps, warmup, opts, gather, sampler=None):
    """Times `steps` solves of scenarios [first_id, first_id + B) on this rank (inputs resident in HBM): per step
    cilqr_set_obstacles_cv (covariance propagation on device), cilqr_solve_batch, and for world > 1 the NCCL all-gather of U* on
    its own stream.  Returns the timing, the work tallies and the status histogram; CUDA events, max over ranks."""
    import torch
    import cilqr_b200
    from cilqr_b200 import scenarios
    N = HORIZON
    sc = scenarios.make_scenarios(p, first_id, B, n_obs, dense=False)
    keys = ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")
    d = {k: torch.from_numpy(sc[k]).to(device) for k in keys}
    U = torch.empty_like(d["U0"])
    solver = cilqr_b200.iLQR(p, max_batch=B, max_obs=max(n_obs, 1), options=opts)
    merit = cilqr_b200.MERIT_TOTAL if n_alphas > 1 else cilqr_b200.MERIT_TRACKING
    res = {"X": torch.empty((B, 4 * (N + 1)), dtype=torch.float64, device=device), "J": torch.empty(B, dtype=torch.float64, device=device),
           "iters": torch.empty(B, dtype=torch.int32, device=device), "status": torch.empty(B, dtype=torch.int32, device=device)}
    kern_ev = []
    do_gather = gather and world > 1
    if do_gather:
        # The exchange step runs on its own stream from a second U buffer: step i's all-gather travels over
        # NVLink under step i+1's solve, which writes the other buffer (no snapshot copy).  The timed region ends with a
        # device-wide synchronize, so the last gather is inside it.
        U2 = torch.empty_like(U)
        peer = None
        if args.gather == "peer":  # copy-engine exchange (cilqr_gather_copy): no kernel competes with the next step's solve
            from cilqr_b200 import dist as cdist
            peer = cdist.PeerGather(solver, B, 2 * N, device)
            gathered = peer.gathered
        else:
            gathered = torch.empty((world * B, 2 * N), dtype=torch.float64, device=device)
        comm_stream = torch.cuda.Stream(device=device)
        solved, gathered_ev = torch.cuda.Event(), [torch.cuda.Event(), torch.cuda.Event()]
        for e in gathered_ev:
            e.record()
    bufs = [U, U2] if do_gather else [U]
    state = {"i": 0}

    def step(timed):
        i = state["i"]
        state["i"] += 1
        Ub = bufs[i % len(bufs)]
        cur = torch.cuda.current_stream()
        if do_gather:
            cur.wait_event(gathered_ev[i % 2])  # the gather that read this buffer two steps ago has finished
        Ub.copy_(d["U0"])
        if timed:  # kernel_ms: the library's device work of the step — covariance propagation (its per-timestep half runs as the
            ev = event_pair()  # first kernel of the solve) and the solve
            ev[0].record()
        solver.set_obstacles_cv(d["obs_state0"], d["obs_LW"], d["obs_sigma"], sc["kappa"])  # covariance propagation on device
        solver.get_optimal_control_seq(d["x0"], Ub, d["poly"], d["xlp"], n_alphas=n_alphas, merit=merit, outputs=res)
        if timed:
            ev[1].record()
            kern_ev.append(ev)
        if do_gather:
            solved.record(cur)
            with torch.cuda.stream(comm_stream):
                comm_stream.wait_event(solved)
                if peer is not None:
                    peer.gather(Ub)
                else:
                    torch.distributed.all_gather_into_tensor(gathered, Ub)
                gathered_ev[i % 2].record(comm_stream)
        return Ub

    last = None
    for _ in range(max(warmup, 0)):
        last = step(False)
    torch.cuda.synchronize()
    launches0 = cilqr_b200.launch_count()
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = event_pair()
    if sampler:
        sampler.mark_begin()
    e0.record()
    for _ in range(steps):
        last = step(True)
    e1.record()
    torch.cuda.synchronize()
    if sampler:
        sampler.mark_end()
    if world > 1:
        torch.distributed.barrier()
    total_ms = max_over_ranks(e0.elapsed_time(e1), world, device)
    if do_gather:
        assert torch.equal(gathered[rank * B:(rank + 1) * B], last), "all-gather of U*: this rank's slice differs from its own result"
        # every rank's slice, on every rank: the rows' sums agree with the owners' (exchanged with a plain NCCL all-gather)
        sums = torch.empty(world * B, dtype=torch.float64, device=device)
        torch.distributed.all_gather_into_tensor(sums, last.sum(dim=1))
        assert torch.equal(gathered.sum(dim=1), sums), "gathered U*: a peer's slice differs from what the peer solved"
        if peer is not None:
            peer.close()
    launches = cilqr_b200.launch_count() - launches0
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kern_ev]))
    iters = solver.iterations.cpu().numpy()
    status = solver.status.cpu().numpy()
    stats = solver.solve_stats() if n_alphas == 1 else (int(iters.sum()), int(iters.sum()), int(iters.max()))
    out = {"total_ms": total_ms, "ms_per_step": total_ms / steps, "kernel_ms": kernel_ms, "launches": int(launches), "stats": stats,
           "mean_iterations": float(iters.mean()), "U": last.cpu().numpy(), "sc": sc,
           "status_hist": {"tolerance": int((status == 0).sum()), "max_iter": int((status == 1).sum()), "lamb_max": int((status == 2).sum()),
                           "nonfinite": int((status == 3).sum())}}
    solver.close()
    return out


def fp64_block(lib, N, n_obs, B, stats, kernel_ms):
    flops_launch, per_cost, per_deriv = solver_flops(N, n_obs, stats[0], stats[1])
    flops_launch += B * propagation_flops(N, n_obs)
    peak = fp64_peak(lib)
    tf = flops_launch / (kernel_ms * 1e-3) / 1e12
    return {"bound": "fp64", "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak if peak else None,
            "peak_source": "builder-measured: DFMA micro-benchmark on this GPU (cilqr_measure_fp64_peak); MEASURED_PEAKS.json has no FP64 figure",
            "flops_per_cost_pass": per_cost, "flops_per_derivative_pass": per_deriv, "cost_passes_per_solve": stats[0] / B,
            "derivative_passes_per_solve": stats[1] / B, "rounds": stats[2], "kernel_ms": kernel_ms,
            "flops_covariance_propagation_per_solve": propagation_flops(N, n_obs),
            "covers": "cilqr_set_obstacles_cv + cilqr_solve_batch (every kernel of the step)",
            "counting": "each operation once, exp / sincos = 20 flops (DESIGN.md section 5)"}


def bench_warp(args, device, hbm_peak, peak_kind, chk):
    """costmap_warp_batch over BASELINE config 5: target T x T from a 2T x 2T source, T in {256 .. 4096}, 1024 poses in chunks,
    value + index map emitted (12 B per cell).  T = 4096 is the headline: its 268 MB source does not fit L2, so the gathers really
    come from HBM (at T <= 2048 the source is L2-resident and algorithmic bytes overstate what HBM moves)."""
    import torch
    import cilqr_b200
    from cilqr_b200 import scenarios
    G = cilqr_b200.CostmapGeom
    sizes = [args.warp_T] if args.warp_T else [256, 512, 1024, 2048, 4096]
    sweep = []
    for T in sizes:
        S = 2 * T
        sg, tg = G(S * 0.2, S * 0.2, 0.2, 0.0, 0.0), G(T * 0.2, T * 0.2, 0.2, 0.0, 0.0)
        src = torch.from_numpy(scenarios.make_source_map(S, S, nan_band=False)).to(device)
        chunk = max(1, min(WARP_POSES, WARP_CHUNK_BYTES // (T * T * 8)))
        poses = torch.from_numpy(scenarios.make_poses(WARP_POSES, (0.0, 0.0), 0.05 * T * 0.2)).to(device)
        out = torch.empty((chunk, T * T), dtype=torch.float32, device=device)
        idx = torch.empty((chunk, T * T), dtype=torch.int32, device=device)
        res = {}
        for name, with_idx in (("value_only", False), ("value_and_index", True)):
            reps = max(1, 2048 // T)  # small grids: several sweeps back to back, so that one launch's latency is not the measurement

            def sweep_once():
              for _ in range(reps):
                for c in range(0, WARP_POSES, chunk):
                    cilqr_b200.warp_batch(src, sg, tg, poses[c:c + chunk], out=out[:min(chunk, WARP_POSES - c)],
                                          idx=idx[:min(chunk, WARP_POSES - c)] if with_idx else None, want_idx=False)
            sweep_once()
            torch.cuda.synchronize()
            e0, e1 = event_pair()
            e0.record()
            sweep_once()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            by = WARP_POSES * T * T * (12 if with_idx else 8)
            res[name] = {"gbs": by / ms / 1e6, "ms_per_pose": ms / WARP_POSES}
        best = res["value_and_index"]
        sweep.append({"T": T, "source": "%dx%d (%.0f MB)" % (S, S, S * S * 4 / 1e6), "poses": WARP_POSES, "chunk": chunk, "gbs": best["gbs"],
                      "frac": best["gbs"] / hbm_peak, "value_only_gbs": res["value_only"]["gbs"], "ms_per_pose": best["ms_per_pose"],
                      "traffic": recorded_traffic("warp_dram_bytes_per_launch_T%d" % T), "launches_counted": 4 * reps * ((WARP_POSES + chunk - 1) // chunk),
                      "source_in_l2": S * S * 4 < 100e6})
        del src, out, idx
        torch.cuda.empty_cache()
    head = [e for e in sweep if e["T"] == 4096] or [sweep[-1]]
    head = head[0]
    blk = {"metric": "costmap warp HBM GB/s", "value": head["gbs"], "unit": "GB/s",
           "config": {"workload": "config5: target TxT from source 2Tx2T, %d poses in chunks, float value + int32 index map; headline T=%d "
                                  "(source larger than L2)" % (WARP_POSES, head["T"]), "l2": "working set per chunk larger than L2"},
           "roofline": {"bound": "hbm", "achieved": head["gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": head["gbs"] / hbm_peak,
                        "traffic": head["traffic"], "peak_source": peak_kind,
                        "algorithmic_bytes_per_cell": "4 B source read + 4 B value write + 4 B index write"},
           "sweep": sweep, "launches": sum(e["launches_counted"] for e in sweep)}
    # the same frame on the wire format (int8 occupancy data in and out, N2 fused into the warp): 1 B + 1 B (+ 4 B index) per cell
    occ_blk = []
    for T in ([args.warp_T] if args.warp_T else [1024, 4096]):
        S = 2 * T
        sg, tg = G(S * 0.2, S * 0.2, 0.2, 0.0, 0.0), G(T * 0.2, T * 0.2, 0.2, 0.0, 0.0)
        occ = torch.randint(-1, 101, (S * S,), device=device, dtype=torch.int32).to(torch.int8)
        chunk = max(1, min(WARP_POSES, WARP_CHUNK_BYTES // (T * T * 8)))
        poses = torch.from_numpy(scenarios.make_poses(WARP_POSES, (0.0, 0.0), 0.05 * T * 0.2)).to(device)
        out8 = torch.empty((chunk, T * T), dtype=torch.int8, device=device)

        def sweep8():
            for c in range(0, WARP_POSES, chunk):
                cilqr_b200.warp_occupancy_batch(occ, sg, tg, poses[c:c + chunk], out=out8[:min(chunk, WARP_POSES - c)])
        sweep8()
        torch.cuda.synchronize()
        e0, e1 = event_pair()
        e0.record()
        sweep8()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        occ_blk.append({"T": T, "ms_per_pose": ms / WARP_POSES, "gbs": WARP_POSES * T * T * 2 / ms / 1e6, "frac": WARP_POSES * T * T * 2 / ms / 1e6 / hbm_peak,
                        "cells_per_s": WARP_POSES * T * T / (ms * 1e-3), "bytes_per_cell": "1 B int8 read + 1 B int8 written"})
        del occ, out8
        torch.cuda.empty_cache()
    blk["occupancy_in_out"] = {"what": "costmap_warp_occupancy_batch: fromOccupancyGrid + warp + toOccupancyGrid in one gather", "sweep": occ_blk,
                               "float_path_ms_per_pose": {e["T"]: e["ms_per_pose"] for e in sweep},
                               "bound": "gather instruction rate (8 byte loads per thread), not HBM: 2 B per cell are algorithmic"}
    if chk is not None and hasattr(chk, "warp"):  # CPU baseline (iii) of SURVEY.md section 8(d): the reference loop, 1 thread and OpenMP over poses
        T = 1024
        S = 2 * T
        sg, tg = G(S * 0.2, S * 0.2, 0.2, 0.0, 0.0), G(T * 0.2, T * 0.2, 0.2, 0.0, 0.0)
        srcn = scenarios.make_source_map(S, S, nan_band=False)
        cores = os.cpu_count() or 1
        posn = scenarios.make_poses(max(2, cores), (0.0, 0.0), 0.05 * T * 0.2)
        t0 = time.perf_counter()
        chk.warp(srcn, sg, tg, posn[:2], want_idx=True, threads=1)
        t1 = time.perf_counter()
        chk.warp(srcn, sg, tg, posn, want_idx=True, threads=cores)
        t2 = time.perf_counter()
        blk["cpu_baseline"] = {"kind": chk.kind, "unit": "GB/s", "cores": cores, "value": len(posn) * T * T * 12 / (t2 - t1) / 1e9,
                               "single_thread_value": 2 * T * T * 12 / (t1 - t0) / 1e9,
                               "sample": "T=%d: 2 poses on one thread, %d poses over %d threads (local_costmap.cpp:242-264 on grid_map::GridMap objects)"
                                         % (T, len(posn), cores)}
    return blk


def propagate_flops(counts):
    """Algorithmic FP64 flops of costmap_propagate_uncertainty: per cell ~120 (covariance, float eigen-solve, ellipse set-up, the
    final division), per cell of its bounding box 12 (position, rotation, membership test), per cell inside the ellipse 14 + exp (20).
    The box is ~4/pi of the ellipse's area."""
    inside = float(np.sum(counts))
    return 120.0 * counts.size + 12.0 * inside * 4.0 / math.pi + 34.0 * inside


def bench_propagate(args, device, lib, chk):
    """costmap_propagate_uncertainty (rows W2-W4), the reference's experiment sigmas on BASELINE config 2's 400 x 400 local costmap
    at 0.2 m.  FP64 / divergence bound: the map (640 KB) is L2-resident, as in the reference's node."""
    import torch
    import cilqr_b200
    from cilqr_b200 import scenarios
    S = 400
    g = cilqr_b200.CostmapGeom(S * 0.2, S * 0.2, 0.2, 0.0, 0.0)
    srcn = scenarios.make_source_map(S, S, nan_band=False)
    src = torch.from_numpy(srcn).to(device)
    sig = scenarios.SIGMA_EXPERIMENT

    def once():
        return cilqr_b200.propagate_uncertainty(src, g, math.sin(0.7), math.cos(0.7), *sig)

    for _ in range(2):
        cnt, _ = once()
    torch.cuda.synchronize()
    reps = 20
    e0, e1 = event_pair()
    e0.record()
    for _ in range(reps):
        once()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    flops = propagate_flops(cnt.cpu().numpy())
    peak = fp64_peak(lib)
    blk = {"metric": "uncertainty propagation Mcells/s", "value": S * S / ms / 1e3, "unit": "Mcells/s", "ms_per_map": ms,
           "config": {"workload": "%dx%d map, 0.2 m cells, sigma (x, y, theta) = %s, all cells" % (S, S, tuple(sig)),
                      "l2": "L2-resident (2.6 MB per map), as in the reference's map node"},
           "roofline": {"bound": "fp64", "achieved": flops / (ms * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                        "frac": flops / (ms * 1e-3) / 1e12 / peak if peak else None, "traffic": None,
                        "peak_source": "builder-measured DFMA micro-benchmark (cilqr_measure_fp64_peak)", "flops_per_map": flops,
                        "note": "each cell averages the map over its own uncertainty ellipse (a data-dependent number of cells): divergence bound"},
           "launches": reps + 2}
    if chk is not None and hasattr(chk, "propagate"):
        cores = os.cpu_count() or 1
        t0 = time.perf_counter()
        chk.propagate(srcn, g, math.sin(0.7), math.cos(0.7), *sig, threads=cores)
        t1 = time.perf_counter()
        blk["cpu_baseline"] = {"kind": chk.kind, "unit": "Mcells/s", "cores": cores, "value": S * S / (t1 - t0) / 1e6,
                               "sample": "one 400x400 map, OpenMP over cells on %d threads (the reference uses 8, arbitrary_transformation.cu:102)" % cores}
    return blk


def host_ceiling(device, world, h2d, d2h, reps=6):
    """All ranks at once move only one step's pinned host->device and device->host volumes (two streams, no kernels): what the
    box's host memory / PCIe paths allow, in steps per second per rank."""
    import torch
    a = torch.empty(h2d, dtype=torch.uint8).pin_memory()
    b = torch.empty(d2h, dtype=torch.uint8).pin_memory()
    da = torch.empty(h2d, dtype=torch.uint8, device=device)
    db = torch.empty(d2h, dtype=torch.uint8, device=device)
    s1, s2 = torch.cuda.Stream(device=device), torch.cuda.Stream(device=device)

    def go(n):
        for _ in range(n):
            with torch.cuda.stream(s1):
                da.copy_(a, non_blocking=True)
            with torch.cuda.stream(s2):
                b.copy_(db, non_blocking=True)
        torch.cuda.synchronize()

    go(2)
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    go(reps)
    return reps / max_over_ranks(time.perf_counter() - t0, world, device)


def bench_e2e(args, p, device, world, sc, U_dev, opts, n_obs):
    """End to end through the C ABI with pinned HOST buffers (copies inside the timed region).  Every step is
    cilqr_set_obstacles_cv + cilqr_solve_batch on host arrays; the calls block until the results are in host memory.  Like a host
    application that keeps the GPU busy, --e2e-threads host threads, each with its own solver handle, CUDA stream and buffers,
    take the steps in turn (compute token on): one step's transfers run under another's kernels.  Every step has its own
    pre-filled in/out U buffer in pinned memory (the inputs of a step are in pinned host memory when the step starts)."""
    import threading
    import torch
    import cilqr_b200
    N, B = HORIZON, args.batch
    keys = ("x0", "U0", "poly", "xlp", "obs_state0", "obs_LW", "obs_sigma")
    pin = {k: torch.from_numpy(sc[k]).pin_memory() for k in keys}
    h2d = 8 * (B * (4 + 2 * N + (p.poly_order + 1) + 2) + B * n_obs * 9)
    d2h = 8 * B * (2 * N + 4 * (N + 1) + 1) + 4 * B * 2
    d2h_u = 8 * B * (2 * N + 1) + 4 * B * 2
    merit = cilqr_b200.MERIT_TRACKING
    n_thr = max(1, args.e2e_threads)
    ksteps = max(n_thr, min(args.steps, 48 if world == 1 else 24))  # (every step has its own 52 MB pinned in/out buffer on every rank)
    each = max(1, (ksteps + n_thr - 1) // n_thr)

    class Worker:
        def __init__(self, token, steps, want_x=True, reuse=False):
            self.reuse = reuse  # one in/out buffer restored by a host memcpy before every call (keeps the recorded graph's key)
            self.stream = torch.cuda.Stream(device=device)
            self.hs = cilqr_b200.iLQR(p, max_batch=B, max_obs=n_obs, stream=self.stream, compute_token=True if token else None, options=opts)
            self.Us = [pin["U0"].clone().pin_memory() for _ in range(steps + 3)]  # one in/out buffer per step, filled before timing
            self.k = 0
            self.outs = {"J": torch.empty(B, dtype=torch.float64).pin_memory().numpy(), "iters": torch.empty(B, dtype=torch.int32).pin_memory().numpy(),
                         "status": torch.empty(B, dtype=torch.int32).pin_memory().numpy()}
            if want_x:
                self.outs["X"] = torch.empty((B, 4 * (N + 1)), dtype=torch.float64).pin_memory().numpy()

        def step(self):
            self.hs.set_obstacles_cv(pin["obs_state0"].numpy(), pin["obs_LW"].numpy(), pin["obs_sigma"].numpy(), sc["kappa"])
            U = self.Us[0 if self.reuse else self.k % len(self.Us)]
            if self.reuse:
                U.copy_(pin["U0"])
            self.k += 1
            self.hs.get_optimal_control_seq(pin["x0"].numpy(), U.numpy(), pin["poly"].numpy(), pin["xlp"].numpy(), n_alphas=1, merit=merit,
                                            outputs=self.outs, want=())
            return U

    def run(workers, steps_each):
        """steps_each steps on every worker, all workers at once; wall time from the common start to the last result."""
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
        gate = threading.Barrier(len(workers) + 1)

        def loop(w):
            torch.cuda.set_device(device)
            gate.wait()
            for _ in range(steps_each):
                w.last = w.step()

        th = [threading.Thread(target=loop, args=(w,)) for w in workers]
        for t in th:
            t.start()
        gate.wait()
        t0 = time.perf_counter()
        for t in th:
            t.join()
        torch.cuda.synchronize()
        return max_over_ranks(time.perf_counter() - t0, world, device)

    out = {"unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "host_threads": n_thr}
    for name, want_x in (("full", True), ("u_only", False)):
        workers = [Worker(n_thr > 1, each, want_x) for _ in range(n_thr)]
        for w in workers:  # warm-up: allocations (the first three buffers)
            for _ in range(3):
                w.step()
        dt = run(workers, each)
        for w in workers:
            assert np.array_equal(w.last.numpy(), U_dev), "host-buffer path and device-resident path disagree"
            w.hs.close()
        v = world * B * each * n_thr / dt
        if name == "full":
            out.update(value=v, steps=each * n_thr)
        else:
            out.update(u_only_value=v, u_only_d2h_bytes_per_step=d2h_u)
        del workers
    solo = Worker(False, 0, reuse=True)  # one caller: the library's default host path (chunked streams, recorded graph)
    for _ in range(3):
        solo.step()
    dt1 = run([solo], ksteps)
    solo.hs.close()
    out["one_thread_value"] = world * B * ksteps / dt1
    ceil_steps = host_ceiling(device, world, h2d, d2h)
    out["host_ceiling"] = {"value": world * B * ceil_steps, "unit": UNIT,
                           "how": "all %d rank(s) at once copy one step's pinned H2D (%d B) and D2H (%d B) volumes on two streams, no kernels" % (world, h2d, d2h)}
    out["exchange"] = ("none (1 GPU)" if world == 1 else
                       "each rank's host receives its own shard's U*, X*, J, iterations, status (the contract of cilqr_solve_batch with "
                       "CILQR_MEM_HOST); the all-gather of U* is a device-side step and is measured in `value`, not here")
    out["path"] = ("cilqr_set_obstacles_cv + cilqr_solve_batch with CILQR_MEM_HOST pinned buffers, blocking calls; %d host thread(s), one solver "
                   "handle and stream each, compute token on; u_only_value: X = NULL (U*, J, iterations, status only); one_thread_value: "
                   "one handle on the default host path" % n_thr)
    return out


def run_b200_arm(args):
    import ctypes as C
    import torch
    import cilqr_b200
    from cilqr_b200 import dist as cdist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - this arm has no CPU path (use --impl reference for the CPU reference)")
    rank, local_rank, world = cdist.init_process_group("nccl")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    sampler = ClockSampler(local_rank) if rank == 0 else None  # started here: it is up and sampling well before the timed region
    lib = cilqr_b200.load_library()
    p = cilqr_b200.Parameters()
    lib.cilqr_params_default(C.byref(p))
    p = p.replace(horizon=HORIZON)
    N, B, n_obs = HORIZON, args.batch, args.n_obs
    hbm_peak, peak_kind = measured_peaks()
    opts = {cilqr_b200.OPT_FUSED: args.fused, cilqr_b200.OPT_FUSED_SLOTS: args.fused_slots,
            cilqr_b200.OPT_GRAPH: 0 if args.no_graph else 1, cilqr_b200.OPT_CHUNKS: args.chunks}

    # ---- the headline: every rank solves its own B scenarios (ids rank*B ...), weak scaling
    main = solve_case(args, p, device, rank, world, rank * B, B, n_obs, args.n_alphas, args.steps, args.warmup, opts, True, sampler)
    clocks = sampler.stop() if sampler else None
    value = world * B * args.steps / (main["total_ms"] * 1e-3)

    # ---- BASELINE config 4 as written: 65 536 scenarios in total, sharded over the ranks by contiguous id ranges, gather included
    strong = None
    if world > 1 and args.n_alphas == 1 and not args.no_strong:
        lo, hi = cdist.shard_range(B, rank, world)
        ksteps = max(10, min(args.steps, 50))
        st = solve_case(args, p, device, rank, world, lo, hi - lo, n_obs, 1, ksteps, max(3, args.warmup), opts, (hi - lo) * world == B)
        strong = {"value": B * ksteps / (st["total_ms"] * 1e-3), "unit": UNIT, "ms_per_step": st["ms_per_step"], "solver_ms": st["kernel_ms"],
                  "scenarios_total": B, "scenarios_per_gpu": hi - lo, "steps": ksteps, "scaling": "strong",
                  "vs_one_gpu_rate": None, "gpu_launches": st["launches"],
                  "workload": workload_name(B, n_obs, 1) + ", sharded over %d GPUs by contiguous scenario ranges, all-gather of U* included" % world}

    e2e = None
    if not args.no_e2e:
        e2e = bench_e2e(args, p, device, world, main["sc"], main["U"], opts, n_obs)

    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    # ---- roofline of the solve: FP64 (what bounds it), HBM beside it
    stats = main["stats"]
    kernel_ms = main["kernel_ms"]
    bytes_launch = B * solver_bytes_per_solve(N, n_obs)
    achieved = bytes_launch / (kernel_ms * 1e-3) / 1e9
    roof = fp64_block(lib, N, n_obs, B, stats, kernel_ms)
    roof["kernel"] = ("cilqr_set_obstacles_cv + cilqr_solve_batch: one 'launch' = one step of the batch (obstacle prep, then fused_solve_kernel for "
                      "batches up to 16 384 scenarios; batch_init + <= %d rounds of batch_cost / batch_derivs / batch_sweep + batch_output for "
                      "larger ones; batch_derivs_kernel dominates, see profiles/)" % (p.max_iterations + 1))
    roof["traffic"] = recorded_traffic("solver_dram_bytes_per_step")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(B, n_obs, args.n_alphas), "scenarios_per_gpu": B, "horizon": N, "n_obs": n_obs, "n_alphas": args.n_alphas,
                   "parallelism": "scenario shards x%d, every GPU its own %d scenarios" % (world, B),
                   "l2": "%s (%.2f GB of obstacle data re-generated every step + %.2f GB of solver state)"
                         % ("working set larger than L2" if B * (n_obs * 8 * (8 + N) + (18 + 18 * (N + 1)) * 8) > 126e6 else
                            "working set L2-resident: a parity-size configuration, not the headline", B * n_obs * 8 * (8 + N) / 1e9,
                            B * (18 + 8 * (N + 1) + 14 * N) * 8 / 1e9),
                   "collective": ("none" if world == 1 else "all-gather of U* by peer-to-peer copies on the copy engines (cilqr_gather_copy) + a one-element NCCL all-reduce "
                                  "as the barrier, on its own stream under the next step's solve" if args.gather == "peer" else
                                  "all_gather U* (NCCL), on its own stream under the next step's solve")},
        "clocks": clocks, "e2e": e2e, "gpu_launches": main["launches"],
        "roofline": roof,
        "hbm": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": recorded_traffic("solver_dram_bytes_per_step"), "peak_source": peak_kind,
                "algorithmic_bytes_per_solve": solver_bytes_per_solve(N, n_obs), "note": "the solve is FP64-issue / latency bound, not HBM bound"},
        "solver": {"mean_iterations": main["mean_iterations"], "status_hist": main["status_hist"],
                   "note": "status as the reference would stop (iLQR.cpp:217-239, lagged accept test): lamb_max = a nominal was rejected and lamb "
                           "grew past lamb_max; the iterations after the rejection change nothing and cost nothing here"},
        "strong_scaling": strong,
    }
    if world > 1:
        torch.distributed.destroy_process_group()
    chk = None
    if world == 1 and not args.no_cpu:
        chk = cpu_solver()
    if world == 1 and not args.no_extra and args.n_alphas == 1:
        # config 3 (line search) and a batch whose scenarios mostly run to the iteration limit instead of stopping on a rejection
        c3 = solve_case(args, p, device, 0, 1, 0, 4096, 10, 8, 20, 3, opts, False)
        line["config3"] = {"workload": workload_name(4096, 10, 8), "value": 4096 * 20 / (c3["total_ms"] * 1e-3), "unit": UNIT,
                           "ms_per_step": c3["ms_per_step"], "mean_iterations": c3["mean_iterations"], "status_hist": c3["status_hist"],
                           "gpu_launches": c3["launches"], "merit": "total (tracking + obstacle + control barriers)"}
        c1 = solve_case(args, p, device, 0, 1, 0, B, 3, 1, 10, 3, opts, False)
        line["few_obstacles"] = {"workload": workload_name(B, 3, 1), "value": B * 10 / (c1["total_ms"] * 1e-3), "unit": UNIT,
                                 "ms_per_step": c1["ms_per_step"], "mean_iterations": c1["mean_iterations"], "status_hist": c1["status_hist"],
                                 "roofline": fp64_block(lib, N, 3, B, c1["stats"], c1["kernel_ms"])}
    if not args.no_warp:
        line["costmap_warp"] = bench_warp(args, device, hbm_peak, peak_kind, chk)
        line["costmap_propagate"] = bench_propagate(args, device, lib, chk)
    if chk is not None:
        line["cpu_baseline"] = cpu_baseline_leg(p, args, chk)
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    # stdout is reserved for the one JSON line: whatever native libraries write to file descriptor 1 (NCCL's version banner,
    # for one) goes to stderr instead; Python's own sys.stdout keeps the real stdout
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
