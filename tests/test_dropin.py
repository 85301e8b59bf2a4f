"""The header-only C++ drop-in (include/cilqr_dropin.hpp): a host program written against the reference's class API
(tests/cpp/dropin_demo.cpp) compiles against it with the Eigen the reference vendors, links libcilqr_b200.so, and —
on the GPU box — reproduces the reference's run_step output.  The binary is built here (needs the vendored Eigen under
/root/reference) into build/ and travels to the GPU box."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "build")
DEMO = os.path.join(BUILD, "dropin_demo")
STAGES = os.path.join(BUILD, "dropin_stages")
MAPDEMO = os.path.join(BUILD, "map_dropin_demo")
GATHER = os.path.join(BUILD, "gather_demo")
EIGEN = "/root/reference/CILQR/src/map_engine/include/map_engine"
GMC = "/root/reference/CILQR/src/grid_map/grid_map_core"
PKG = "uncertainty-aware-cilqr-for-trajectory-optimization_b200"


def build_demo():
    """The three host programs of tests/cpp/, compiled against include/ with the Eigen and grid_map_core the reference vendors
    (Eigen 3.2.10 has no unsupported/ tree: the Tensor header comes from the oracle's shim; grid_map_core's objects are the ones
    oracle/Makefile compiled from the unmodified sources)."""
    import glob
    import oracle
    os.makedirs(BUILD, exist_ok=True)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    common = [cxx, "-std=gnu++14", "-O2", "-w", "-include", "cstdint", "-I" + os.path.join(ROOT, "include"), "-I" + EIGEN, "-I" + os.path.join(ROOT, "oracle", "shims")]
    link = ["-L" + os.path.join(ROOT, PKG), "-l:libcilqr_b200.so", "-Wl,-rpath,$ORIGIN/../" + PKG]
    for src, out in (("dropin_demo.cpp", DEMO), ("dropin_stages.cpp", STAGES)):
        subprocess.run(common + [os.path.join(ROOT, "tests", "cpp", src), "-o", out] + link, check=True)
    nvcc = "/usr/local/cuda/bin/nvcc"
    if os.path.exists(nvcc) and os.path.exists("/usr/include/nccl.h"):  # one host thread, one handle per GPU, cilqr_gather on the application's NCCL communicators
        subprocess.run([nvcc, "-ccbin", cxx, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I" + os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "tests", "cpp", "gather_demo.cpp"), "-o", GATHER, "-L" + os.path.join(ROOT, PKG), "-l:libcilqr_b200.so", "-lnccl",
                        "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../" + PKG], check=True)
    oracle.build(("ref",))
    objs = sorted(glob.glob(os.path.join(ROOT, "oracle", "_ref", "obj", "gmc*_*.o")))
    subprocess.run(common + ["-I" + os.path.join(GMC, "include"), "-I" + os.path.join(ROOT, "oracle", "_ref", "inc"),
                             os.path.join(ROOT, "tests", "cpp", "map_dropin_demo.cpp")] + objs + ["-o", MAPDEMO] + link, check=True)


def parse_named(out):
    d = {}
    for line in out.splitlines():
        f = line.split()
        if len(f) >= 2 and f[1].lstrip("-").isdigit() and len(f) == 2 + int(f[1]):
            d[f[0]] = np.array([float(x) for x in f[2:]])
    return d


def parse(out):
    d = {}
    for line in out.splitlines():
        f = line.split()
        if f and f[0] in ("U", "X"):
            d[f[0]] = np.array([float(x) for x in f[1:]])
        elif f and f[0] == "iterations":
            d["iterations"], d["status"], d["J"] = int(f[1]), int(f[3]), float(f[5])
    return d


def test_dropin_compiles_links_and_fails_loudly_without_a_gpu(lib):
    import torch
    if not os.path.isdir(EIGEN):
        pytest.skip("the vendored Eigen is only available where /root/reference is mounted")
    build_demo()
    for cmd in ([DEMO], [MAPDEMO]):
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
        if torch.cuda.is_available():
            assert r.returncode == 0, r.stdout
        else:
            assert r.returncode == 3 and "no CUDA device" in r.stdout, r.stdout


@pytest.mark.gpu
def test_dropin_run_step_matches_reference(checker):
    if not os.path.exists(DEMO):
        pytest.skip("build/dropin_demo was not built (needs the vendored Eigen at build time)")
    g = np.load(os.path.join(ROOT, "tests", "golden", "solver_golden.npz"))
    r = subprocess.run([DEMO], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    d = parse(r.stdout)
    rel = lambda a, b: float(np.max(np.abs(a - b)) / np.max(np.abs(b)))  # noqa: E731
    assert rel(d["U"], g["run_step/U"]) < 1e-7 and rel(d["X"], g["run_step/X"]) < 1e-7
    if checker.kind != "reference":
        return
    # with obstacles, against the compiled reference's own run_step
    p = checker.params_default(horizon=50)
    N, n_obs = 50, 2
    dim, pos = np.zeros((n_obs, N, 2)), np.zeros((n_obs, N, 4))
    for j in range(n_obs):
        t = np.arange(N)
        dim[j, :, 0], dim[j, :, 1] = p.length, p.width
        pos[j, :, 0], pos[j, :, 1], pos[j, :, 2], pos[j, :, 3] = 12.0 + 6.0 * j + 1.5 * t * p.timestep, -1.0 + 1.2 * j, 1.5, 0.0
    plan = np.stack([np.arange(200, dtype=np.float64), np.zeros(200)], axis=1)
    Xr, Ur = checker.run_step(p, dim.reshape(n_obs, -1), pos.reshape(n_obs, -1), plan, np.array([0.3, 0.2, 4.0, 0.01]))
    r = subprocess.run([DEMO, str(n_obs)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    d = parse(r.stdout)
    assert rel(d["U"], Ur) < 1e-7 and rel(d["X"], Xr) < 1e-7


@pytest.mark.gpu
@pytest.mark.parametrize("n_obs,first", [(3, 40), (0, 41), (12, 42)])
def test_dropin_classes_match_reference_stage_by_stage(checker, tmp_path, n_obs, first):
    """Constraints, Model, iLQR::backward_pass / forward_pass / get_optimal_control_seq through the header (public `constraints`
    and `model` members, set_uncertainty_map with a foreign type) against the compiled reference, fed the SAME poly / xlp —
    the solve at 1e-9: the looser tolerance of the run_step tests is the polynomial fit's conditioning (row N1), not the solver."""
    if not os.path.exists(STAGES):
        pytest.skip("build/dropin_stages was not built (needs the vendored Eigen at build time)")
    from cilqr_b200 import scenarios
    p = checker.params_default(horizon=50)
    N = p.horizon
    sc = scenarios.make_scenarios(p, first, 1, n_obs, curved=True)
    lamb = 1.0
    f = tmp_path / "scenario.txt"
    rows = {"horizon": [N], "n_obs": [n_obs], "x0": sc["x0"][0], "U": sc["U0"][0], "poly": sc["poly"][0], "xlp": sc["xlp"][0], "lamb": [lamb],
            "dim": sc["dim"][0].ravel() if n_obs else [], "pos": sc["pos"][0].ravel() if n_obs else []}
    f.write_text("".join("%s %d %s\n" % (k, len(v), " ".join(repr(float(x)) for x in v)) for k, v in rows.items()))
    r = subprocess.run([STAGES, str(f)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    d = parse_named(r.stdout)
    rel = lambda a, b: float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))  # noqa: E731
    dim = sc["dim"][0] if n_obs else None
    pos = sc["pos"][0] if n_obs else None
    x0, U, poly, xlp = sc["x0"][0], sc["U0"][0], sc["poly"][0], sc["xlp"][0]
    X0 = checker.rollout(p, x0, U)
    assert rel(d["X0"], X0) < 1e-12
    lx, lxx = checker.state_cost(p, dim, pos, X0, poly, xlp)
    assert rel(d["l_x"], lx) < 1e-11 and rel(d["l_xx"], lxx) < 1e-11
    lu, luu = checker.control_cost(p, X0, U)
    assert rel(d["l_u"], lu) < 1e-12 and rel(d["l_uu"], luu) < 1e-12
    assert rel(d["J0"], [checker.cost(p, X0, U, poly, xlp)]) < 1e-12
    assert rel(d["closest"], checker.closest_point(p, X0[12:16], poly, xlp)) < 1e-12
    e = 2.75 * np.exp(2.75 * -0.3)
    cd = np.array([0.5, -1.0])
    assert rel(d["barrier"], np.concatenate([[e], 2.75 * e * cd, (2.75 ** 2 * e * np.outer(cd, cd)).ravel()])) < 1e-14
    assert d["l_ux_abs"][0] == 0.0
    assert rel(d["step"], checker.forward_simulate(p, x0, U[:2])) < 1e-14
    if hasattr(checker, "linearize"):
        Xm = X0.reshape(N + 1, 4)
        A, B = checker.linearize(p, Xm[1:, 2], Xm[1:, 3], U.reshape(N, 2)[:, 0])
        assert rel(d["A_t"], A) < 1e-13 and rel(d["B_t"], B) < 1e-13
    ok, k, K = checker.backward(p, dim, pos, X0, U, poly, xlp, lamb)
    assert d["ok"][0] == float(ok) and rel(d["k"], k) < 1e-10 and rel(d["K"], K) < 1e-10
    Xn, Un = checker.forward(p, X0, U, k, K, 1.0)
    assert rel(d["X_new"], Xn) < 1e-9 and rel(d["U_new"], Un) < 1e-9  # (the gains fed in are the drop-in's own, equal to 1e-10)
    s = checker.solve(p, dim, pos, x0, U, poly, xlp)
    assert int(d["J_iters_status"][1]) == s["iters"]
    assert rel(d["U_result"], s["U"]) < 1e-9 and rel(d["X_result"], s["X"]) < 1e-9
    assert abs(d["J_iters_status"][0] - s["J"]) <= 1e-9 * abs(s["J"])


@pytest.mark.gpu
@pytest.mark.parametrize("index", [0, 777])
def test_map_dropin_matches_reference(checker, index):
    """thrust_propagateUncertainty(grid_map::GridMap, int, double x 5) (ARBIT.cuh:109) and the warp loop through
    include/cilqr_map_dropin.hpp, on real grid_map::GridMap objects, against the compiled reference."""
    if not os.path.exists(MAPDEMO):
        pytest.skip("build/map_dropin_demo was not built (needs grid_map_core at build time)")
    if checker.kind != "reference":
        pytest.skip("needs the compiled reference (oracle/_ref)")
    import cilqr_b200
    r = subprocess.run([MAPDEMO, str(index)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    d = parse_named(r.stdout)
    G = cilqr_b200.CostmapGeom
    sg, tg = G(48.0, 40.0, 0.2, 10.0, -5.0), G(12.0, 8.0, 0.2, 2.0, 0.0)
    rs, cs = checker.geom_size(sg)
    i0, i1 = np.meshgrid(np.arange(rs), np.arange(cs), indexing="ij")
    src = np.where((i0 // 7 + i1 // 5) % 3 == 0, 100.0, 0.0).astype(np.float32).ravel(order="F")
    rt, ct = checker.geom_size(tg)
    bbox = np.zeros((rt, ct), np.float32)
    bbox[3, 4] = 100.0
    pose = np.array([[12.0, -4.0, np.sin(0.7), np.cos(0.7)]])
    val, _ = checker.warp(src, sg, tg, pose, bbox=bbox.ravel(order="F")[None, :])
    assert np.array_equal(d["warp"].astype(np.float32), val[0]), "warped layer differs from the reference loop"
    cr, vr = checker.propagate(val[0], tg, np.sin(0.7), np.cos(0.7), 0.16, 0.16, 0.017, index=index)
    flips = int(np.sum(d["count"].astype(np.int64) != cr))
    assert flips <= 1e-4 * cr.size, flips
    same = d["count"].astype(np.int64) == cr
    ok = same & np.isfinite(vr)
    assert np.max(np.abs(d["value"][ok] - vr[ok]) / np.maximum(np.abs(vr[ok]), 1e-300)) < 1e-12


@pytest.mark.gpu
def test_one_thread_drives_a_handle_per_gpu_and_gathers_over_nccl():
    """cilqr_create_on + cilqr_gather from a C++ host (tests/cpp/gather_demo.cpp): shards solved on every device of the box, the
    all-gathered U* identical on every device and bit-identical to the one-device solve.  Needs at least two GPUs."""
    import torch
    if not os.path.exists(GATHER):
        pytest.skip("build/gather_demo was not built")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two GPUs (run under gpurun --gpus 2)")
    r = subprocess.run([GATHER], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 mismatching devices" in r.stdout
