"""The header-only C++ drop-in (include/cilqr_dropin.hpp): a host program written against the reference's class API
(tests/cpp/dropin_demo.cpp) compiles against it with the Eigen the reference vendors, links libcilqr_b200.so, and —
on the GPU box — reproduces the reference's run_step output.  The binary is built here (needs the vendored Eigen under
/root/reference) into build/ and travels to the GPU box."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "build", "dropin_demo")
EIGEN = "/root/reference/CILQR/src/map_engine/include/map_engine"
PKG = "uncertainty-aware-cilqr-for-trajectory-optimization_b200"


def build_demo():
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=gnu++14", "-O2", "-w", "-I" + os.path.join(ROOT, "include"), "-I" + EIGEN, os.path.join(ROOT, "tests", "cpp", "dropin_demo.cpp"),
           "-o", DEMO, "-L" + os.path.join(ROOT, PKG), "-l:libcilqr_b200.so", "-Wl,-rpath,$ORIGIN/../" + PKG]
    subprocess.run(cmd, check=True)


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
    r = subprocess.run([DEMO], capture_output=True, text=True, timeout=120)
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
