"""Pins the plain-C restatement against the compiled reference on fresh seeded inputs (skipped where
oracle/_ref could not be built, i.e. never in this container).  CPU only."""
import numpy as np
import pytest

import cilqr_b200
from cilqr_b200 import scenarios


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@pytest.mark.parametrize("n_obs,curved", [(0, False), (3, True), (20, False)])
def test_full_solves(ref, port, n_obs, curved):
    p = ref.params_default(horizon=50)
    B = 6
    sc = scenarios.make_scenarios(p, 1000, B, n_obs, curved=curved)
    for b in range(B):
        dim = sc["dim"][b] if n_obs else None
        pos = sc["pos"][b] if n_obs else None
        r = ref.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        o = port.solve(p, dim, pos, sc["x0"][b], sc["U0"][b], sc["poly"][b], sc["xlp"][b])
        assert r["iters"] == o["iters"]
        assert rel(o["U"], r["U"]) < 1e-9 and rel(o["X"], r["X"]) < 1e-9 and abs(o["J"] - r["J"]) < 1e-9 * abs(r["J"])


def test_stage_functions_on_random_states(ref, port):
    p = ref.params_default(horizon=50)
    rng = np.random.default_rng(7)
    sc = scenarios.make_scenarios(p, 2000, 4, 5, curved=True)
    for b in range(4):
        U = sc["U0"][b] + rng.normal(0, 0.3, 100)
        X = ref.rollout(p, sc["x0"][b], U)
        assert rel(port.rollout(p, sc["x0"][b], U), X) < 1e-13
        a, c = ref.state_cost(p, sc["dim"][b], sc["pos"][b], X, sc["poly"][b], sc["xlp"][b]), port.state_cost(p, sc["dim"][b], sc["pos"][b], X, sc["poly"][b], sc["xlp"][b])
        assert rel(c[0], a[0]) < 1e-12 and rel(c[1], a[1]) < 1e-12
        a, c = ref.control_cost(p, X, U), port.control_cost(p, X, U)
        assert rel(c[0], a[0]) < 1e-13 and rel(c[1], a[1]) < 1e-13
        for lamb in (1.0, 1e-6, 1e3):
            ok_r, k_r, K_r = ref.backward(p, sc["dim"][b], sc["pos"][b], X, U, sc["poly"][b], sc["xlp"][b], lamb)
            ok_o, k_o, K_o = port.backward(p, sc["dim"][b], sc["pos"][b], X, U, sc["poly"][b], sc["xlp"][b], lamb)
            assert ok_r == ok_o == 1 and rel(k_o, k_r) < 1e-10 and rel(K_o, K_r) < 1e-10
        pt_r = ref.closest_point(p, X[8:12], sc["poly"][b], sc["xlp"][b])
        assert np.allclose(port.closest_point(p, X[8:12], sc["poly"][b], sc["xlp"][b]), pt_r, rtol=0, atol=1e-13)
        vx_r, mx_r = ref.obstacle_cost(p, sc["dim"][b][0], sc["pos"][b][0], 7, X[28:32])
        vx_o, mx_o = port.obstacle_cost(p, sc["dim"][b][0], sc["pos"][b][0], 7, X[28:32])
        assert rel(vx_o, vx_r) < 1e-12 and rel(mx_o, mx_r) < 1e-12


def test_warp_index_maps_bit_exact(ref, port):
    G = cilqr_b200.CostmapGeom
    sg, tg = G(40.0, 30.2, 0.2, 10.0, -5.0), G(12.0, 9.0, 0.2, 2.0, 0.0)
    rs, cs = ref.geom_size(sg)
    assert (rs, cs) == port.geom_size(sg)
    src = scenarios.make_source_map(rs, cs)
    poses = scenarios.make_poses(24, (10.0, -5.0), 16.0)
    rt, ct = ref.geom_size(tg)
    bbox = np.where(np.random.default_rng(3).random((24, rt * ct)) > 0.97, 100.0, 0.0).astype(np.float32)
    v_r, i_r = ref.warp(src, sg, tg, poses, bbox=bbox)
    v_o, i_o = port.warp(src, sg, tg, poses, bbox=bbox)
    assert np.array_equal(i_r, i_o) and np.array_equal(v_r.view(np.uint32), v_o.view(np.uint32))
    assert (i_r < 0).any() and (i_r >= 0).any()


def test_expand_obstacles_matches_numpy_generator(port):
    p = port.params_default(horizon=50)
    sc = scenarios.make_scenarios(p, 5, 3, 4)
    dim, pos = port.expand_obstacles_cv(p, sc["obs_state0"][1], sc["obs_LW"][1], sc["obs_sigma"][1], sc["kappa"])
    assert rel(dim, sc["dim"][1]) < 1e-15 and rel(pos, sc["pos"][1]) < 1e-15


# ------------------------------------------------------------------------------------------------ rows N2 / N3: known answers
def test_occupancy_wire_format_known_answers(ref):
    """GridMapRosConverter.cpp:259-266 / :291-306 restated in oracle/ref_harness.cpp: reverse order, -1 <-> NaN, truncation."""
    occ = np.array([0, 50, -1, 100, 59], dtype=np.int8)
    data = ref.from_occupancy(occ)
    assert np.array_equal(np.isnan(data), [False, False, True, False, False])
    assert np.array_equal(data[[0, 1, 3, 4]], [59.0, 100.0, 50.0, 0.0])
    back = ref.to_occupancy(data, 0.0, 100.0)
    assert back.tolist() == [0, 50, -1, 100, 58]  # (59 / 100) * 100 = 58.99999 in float: the reference truncates
    assert ref.to_occupancy(np.array([-7.0, 250.0, 12.5], np.float32), 0.0, 50.0).tolist() == [25, 100, 0]


def test_bounding_box_rasterisation_known_answer(ref):
    """bondingBoxHandle (local_costmap.cpp:860-922) on the reference's Polygon / PolygonIterator: an axis-aligned 4.0 x 2.0 m
    vehicle (+0.2 m) right ahead of the ego covers 21 x 11 cell centres of a 0.2 m grid; a vehicle 150 m away is skipped."""
    import cilqr_b200
    g = cilqr_b200.CostmapGeom(40.0, 40.0, 0.2, 0.0, 0.0)
    rows, cols = ref.geom_size(g)
    boxes = np.array([[10.1, -0.1, 0.0, 4.0, 2.0], [150.0, 0.0, 0.3, 4.0, 2.0]])
    layer = ref.rasterize_boxes(g, (0.0, 0.0, 0.0), boxes, np.zeros(rows * cols, np.float32))
    filled = np.flatnonzero(layer == 100.0)
    assert filled.size == 21 * 11 and set(np.unique(layer)) == {0.0, 100.0}
    i, j = filled % rows, filled // rows  # column-major linear index
    x = 0.0 + (0.5 * 40.0 - 0.5 * 0.2) - 0.2 * i
    y = 0.0 + (0.5 * 40.0 - 0.5 * 0.2) - 0.2 * j
    assert x.min() > 10.1 - 2.1 - 1e-9 and x.max() < 10.1 + 2.1 + 1e-9 and y.min() > 0.1 - 1.1 - 1e-9 and y.max() < 0.1 + 1.1 + 1e-9


def _move_unwrapped_layer():
    """OccupancyGridConversion.withMove (grid_map_ros/test/GridMapRosTest.cpp:116-139): an 8.0 x 5.0 m map at 0.5 m (16 x 10
    cells) filled with 1.0, then moved by (-1, -1) m: the two rows / columns that enter on the low-coordinate side (the highest
    indices) are unobserved (NaN).  Returned in unwrapped linear (column-major) order, as toOccupancyGrid walks it."""
    rows, cols = 16, 10
    m = np.ones((rows, cols), np.float32)
    before = m.ravel(order="F").copy()
    m[rows - 2:, :] = np.nan
    m[:, cols - 2:] = np.nan
    return before, m.ravel(order="F")


def test_occupancy_known_answers_transcribed_from_grid_map_ros_tests(ref):
    """The reference's own tests of the wire format, as known answers for the checker's restated loops (GridMapRosConverter.cpp
    itself needs ROS headers): GridMapRosTest.cpp:116-139 withMove (cell 0 is 100 before the move and -1 after it) and :141-177
    roundTrip (fromOccupancyGrid then toOccupancyGrid(-1, 100) returns every value of [-1, 100] unchanged)."""
    for chk in (ref,):
        before, after = _move_unwrapped_layer()
        occ = chk.to_occupancy(before, 0.0, 1.0)
        assert occ[0] == 100 and np.all(occ == 100)
        occ_new = chk.to_occupancy(after, 0.0, 1.0)
        assert occ_new[0] == -1
        assert np.array_equal(occ_new == -1, np.isnan(after)[::-1]) and np.all(occ_new[occ_new != -1] == 100)
        rng = np.random.default_rng(1234)
        grid = (rng.integers(0, 102, 50 * 100) - 1).astype(np.int8)  # rand() % 102 - 1: [-1, 100]
        grid[:102] = np.arange(-1, 101, dtype=np.int8)               # every value at least once
        layer = chk.from_occupancy(grid)
        assert np.array_equal(np.isnan(layer), (grid == -1)[::-1])
        assert np.array_equal(chk.to_occupancy(layer, -1.0, 100.0), grid)
