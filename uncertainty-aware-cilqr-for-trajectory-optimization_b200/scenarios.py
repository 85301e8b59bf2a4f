"""Seeded synthetic workloads — generated once on the host and handed identically to
the CPU oracle and to the GPU path.  Counter-based SplitMix64 keyed by ``seed ^ scenario_id`` so that a
scenario's inputs do not depend on the batch it is generated in (rank-sharding safe).

No road network / CARLA: a straight road y = 0 with waypoints every metre, the shape the reference's
LocalPlanner would fit (CILQR/src/ilqr/include/ilqr/LocalPlanner.cpp:66-116) reduced to its output
(poly, x_first, x_last).  ``curved=True`` adds a gentle cubic so that the closest-point search
(Constraints.cpp:24-59) is exercised off-axis.
"""
import numpy as np

SEED = 0xC11A5EED
KAPPA = 2.4477  # chisquare_val, CILQR/src/map_engine/include/ARBIT.cuh:87
SIGMA_EXPERIMENT = (0.16, 0.16, 0.017)  # CILQR/src/ilqr/launch/Experiment.launch:9-11

_G = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def _mix(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def uniforms(scenario_ids, n, seed=SEED):
    """[len(ids), n] doubles in [0,1): draw k of scenario s = mix(mix(seed ^ s) + (k+1)*golden)."""
    ids = np.asarray(scenario_ids, dtype=np.uint64)
    with np.errstate(over="ignore"):
        key = _mix((np.uint64(seed) ^ ids) * _G + _G)
        ctr = (np.arange(1, n + 1, dtype=np.uint64) * _G)[None, :]
        z = _mix(key[:, None] + ctr)
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def make_scenarios(params, first, count, n_obs, curved=False, seed=SEED, sigma=SIGMA_EXPERIMENT, dense=True):
    """Scenarios ``first .. first+count-1``.

    Returns a dict of C-contiguous float64 arrays:
      x0 [B,4], U0 [B,2N] (iLQR ctor default, iLQR.cpp:9-15), poly [B,poly_order+1], xlp [B,2],
      obs_state0 [B,n_obs,4], obs_LW [B,n_obs,2], obs_sigma [B,n_obs,3], kappa,
      and when ``dense``: dim [B,n_obs,2N], pos [B,n_obs,4N] — the per-timestep matrices the reference's
      Obstacle(Parameters, dimension, relative_pos_array) takes (Obstacle.cpp:8-13), with the
      covariance-inflated dimensions of SURVEY.md §8(d).
    """
    N = params.horizon
    B = int(count)
    ids = np.arange(first, first + B, dtype=np.uint64)
    u = uniforms(ids, 8 + 4 * max(n_obs, 0), seed)
    x0 = np.stack([2.0 * u[:, 0], -0.5 + u[:, 1], 2.0 + 4.0 * u[:, 2], -0.05 + 0.1 * u[:, 3]], axis=1)
    npoly = params.poly_order + 1
    poly = np.zeros((B, npoly))
    if curved:
        poly[:, 0] = -0.3 + 0.6 * u[:, 4]
        poly[:, 1] = -0.03 + 0.06 * u[:, 5]
        if npoly > 2:
            poly[:, 2] = -0.004 + 0.008 * u[:, 6]
        if npoly > 3:
            poly[:, 3] = -1e-4 + 2e-4 * u[:, 7]
    # nearest waypoint of the 1 m grid, then the next num_of_local_wpts waypoints (LocalPlanner.cpp:46-60)
    i0 = np.floor(x0[:, 0] + 0.5)
    xlp = np.stack([i0, i0 + (params.num_of_local_wpts - 1)], axis=1)
    U0 = np.zeros((B, 2 * N))
    U0[:, 0::2] = 0.5
    U0[:, 1::2] = np.where(np.arange(N) < N // 2, 0.0, 0.1)[None, :]
    out = dict(x0=np.ascontiguousarray(x0), U0=U0, poly=poly, xlp=np.ascontiguousarray(xlp), kappa=KAPPA, n_obs=n_obs, ids=ids)
    if n_obs > 0:
        o = u[:, 8:].reshape(B, n_obs, 4)
        st = np.stack([6.0 + 24.0 * o[..., 0], -3.5 + 7.0 * o[..., 1], 4.0 * o[..., 2], -0.2 + 0.4 * o[..., 3]], axis=-1)
        LW = np.empty((B, n_obs, 2))
        LW[..., 0], LW[..., 1] = params.length, params.width
        sg = np.empty((B, n_obs, 3))
        sg[...] = np.asarray(sigma, dtype=np.float64)
        out.update(obs_state0=np.ascontiguousarray(st), obs_LW=LW, obs_sigma=sg)
        if dense:
            dim, pos = expand_obstacles_cv(params, st, LW, sg, KAPPA)
            out.update(dim=dim, pos=pos)
    else:
        out.update(obs_state0=None, obs_LW=None, obs_sigma=None, dim=None, pos=None)
    return out


def expand_obstacles_cv(params, state0, LW, sigma, kappa):
    """Host-side expansion of the compact obstacle form into Obstacle()'s (dimension 2xN, pos 4xN):
    constant-velocity prediction, sigma_long = sx, sigma_lat(i) = sqrt(sy^2 + sth^2 (v i dt)^2)
    (lever-arm form of ARBIT.cuh:58-68), dimension = size + 2*kappa*sigma."""
    N = params.horizon
    t = np.arange(N, dtype=np.float64) * params.timestep
    x0, y0, v, th = (state0[..., i][..., None] for i in range(4))
    c, s = np.cos(th), np.sin(th)
    pos = np.empty(state0.shape[:-1] + (N, 4))
    pos[..., 0] = x0 + v * c * t
    pos[..., 1] = y0 + v * s * t
    pos[..., 2] = v
    pos[..., 3] = th
    arm = v * t
    dim = np.empty(state0.shape[:-1] + (N, 2))
    dim[..., 0] = LW[..., 0:1] + 2.0 * kappa * sigma[..., 0:1]
    dim[..., 1] = LW[..., 1:2] + 2.0 * kappa * np.sqrt(sigma[..., 1:2] ** 2 + sigma[..., 2:3] ** 2 * arm * arm)
    return (np.ascontiguousarray(dim.reshape(state0.shape[:-1] + (2 * N,))),
            np.ascontiguousarray(pos.reshape(state0.shape[:-1] + (4 * N,))))


# ------------------------------------------------------------------------------------------ costmap
def reference_global_geom():
    """global_map_.setGeometry(Length(301.2, 301.2), 0.2, Position(93.14, -205.96)) — local_costmap.cpp:119."""
    from .params import CostmapGeom
    return CostmapGeom(301.2, 301.2, 0.2, 93.14, -205.96)


def make_source_map(rows, cols, nan_band=True):
    """Deterministic pattern of SURVEY.md §8(d) config 2: src(i,j) = ((i/7 + j/5) % 3 == 0) ? 100 : 0 plus a
    NaN band; returned flat in grid_map's column-major order (index0 fastest)."""
    i = np.arange(rows, dtype=np.int64)[:, None]
    j = np.arange(cols, dtype=np.int64)[None, :]
    m = np.where(((i // 7 + j // 5) % 3) == 0, 100.0, 0.0).astype(np.float32)
    if nan_band:
        m[rows // 3: rows // 3 + 4, :] = np.nan
    return np.ascontiguousarray(m.T).reshape(-1)  # column-major flat


def make_poses(count, centre, spread, seed=SEED, first=0):
    """[count,4] = (Vx, Vy, sin(theta), cos(theta)); theta ~ U(0, 2pi); sin/cos from host libm, as
    local_costmap.cpp:201-202 computes them once per frame."""
    u = uniforms(np.arange(first, first + count, dtype=np.uint64) + np.uint64(1 << 40), 3, seed)
    th = 2.0 * np.pi * u[:, 0]
    return np.ascontiguousarray(np.stack([centre[0] + spread * (2 * u[:, 1] - 1), centre[1] + spread * (2 * u[:, 2] - 1),
                                          np.sin(th), np.cos(th)], axis=1))
