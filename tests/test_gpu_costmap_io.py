"""GPU parity tests of the formats either side of the costmap warp, through the C ABI.

N2  costmap_from_occupancy / costmap_to_occupancy against the restated loops of GridMapRosConverter.cpp:259-266 / :291-306
    (oracle/ref_harness.cpp; grid_map_ros itself needs ROS headers) — bit-exact, plus the round-trip property at full size.
N3  costmap_rasterize_boxes against LocalCostmap::bondingBoxHandle restated on the reference's own grid_map::Polygon /
    PolygonIterator (compiled unmodified into oracle/_ref) — every cell identical, including boxes whose edges run exactly
    through cell centres."""
import numpy as np
import pytest

import cilqr_b200

pytestmark = pytest.mark.gpu
G = cilqr_b200.CostmapGeom


def need_ref(checker):
    if not hasattr(checker, "lib") or getattr(checker, "kind", "") != "reference":
        pytest.skip("needs the compiled reference (oracle/_ref)")


@pytest.mark.parametrize("n", [1, 15, 16, 17, 400 * 400, 1506 * 1506, 1000003])
def test_occupancy_conversions_match_reference_loops(checker, n):
    need_ref(checker)
    rng = np.random.default_rng(n)
    occ = rng.integers(-1, 101, n).astype(np.int8)
    occ[rng.integers(0, n, max(1, n // 7))] = -1
    data = cilqr_b200.from_occupancy(occ)
    ref = checker.from_occupancy(occ)
    assert np.array_equal(data.view(np.uint32), ref.view(np.uint32))
    vals = rng.uniform(-20.0, 130.0, n).astype(np.float32)
    vals[rng.integers(0, n, max(1, n // 9))] = np.nan
    vals[rng.integers(0, n, max(1, n // 11))] = 100.0
    for lo, hi in ((0.0, 100.0), (-5.0, 37.5)):
        assert np.array_equal(cilqr_b200.to_occupancy(vals, lo, hi), checker.to_occupancy(vals, lo, hi))
    # the wire format is NOT idempotent in the reference either ((59 / 100) * 100 truncates to 58): the composition has to
    # reproduce that, and unknown cells have to survive it
    back = cilqr_b200.from_occupancy(cilqr_b200.to_occupancy(data, 0.0, 100.0))
    ref_back = checker.from_occupancy(checker.to_occupancy(ref, 0.0, 100.0))
    assert np.array_equal(back.view(np.uint32), ref_back.view(np.uint32))
    assert np.array_equal(np.isnan(back), np.isnan(data))


def test_occupancy_known_answers_transcribed_from_grid_map_ros_tests():
    """grid_map_ros/test/GridMapRosTest.cpp:116-139 (withMove) and :141-177 (roundTrip) against the CUDA kernels: known answers
    held by the reference's own tests, no checker involved."""
    rows, cols = 16, 10
    m = np.ones((rows, cols), np.float32)
    assert np.all(cilqr_b200.to_occupancy(m.ravel(order="F"), 0.0, 1.0) == 100)
    m[rows - 2:, :] = np.nan
    m[:, cols - 2:] = np.nan
    flat = m.ravel(order="F")
    occ = cilqr_b200.to_occupancy(flat, 0.0, 1.0)
    assert occ[0] == -1 and np.array_equal(occ == -1, np.isnan(flat)[::-1]) and np.all(occ[occ != -1] == 100)
    rng = np.random.default_rng(1234)
    grid = (rng.integers(0, 102, 50 * 100) - 1).astype(np.int8)
    grid[:102] = np.arange(-1, 101, dtype=np.int8)
    layer = cilqr_b200.from_occupancy(grid)
    assert np.array_equal(np.isnan(layer), (grid == -1)[::-1])
    assert np.array_equal(cilqr_b200.to_occupancy(layer, -1.0, 100.0), grid)


@pytest.mark.parametrize("rows_t,cols_t,with_bbox", [(400, 400, True), (64, 40, False), (150, 100, True), (37, 21, False)])
def test_warp_on_the_wire_format_equals_the_three_step_frame(checker, rows_t, cols_t, with_bbox):
    """costmap_warp_occupancy_batch (N2 fused into W1: int8 in, int8 out) against the reference's frame assembled from its
    pieces — fromOccupancyGrid, the warp loop on grid_map::GridMap objects, toOccupancyGrid — byte for byte, index map int for
    int; sizes with and without 8-cell alignment, poses that leave the global map (unknown cells: -1)."""
    need_ref(checker)
    import torch
    sg = G(60.0, 48.0, 0.2, 10.0, -5.0)
    tg = G(rows_t * 0.2, cols_t * 0.2, 0.2, 2.0, 0.0)
    rs, cs = checker.geom_size(sg)
    rng = np.random.default_rng(rows_t * 1000 + cols_t)
    occ = rng.integers(-1, 101, rs * cs).astype(np.int8)
    poses = np.array([[12.0, -4.0, np.sin(0.7), np.cos(0.7)], [30.0, 10.0, np.sin(2.9), np.cos(2.9)], [10.0, -5.0, 0.0, 1.0],
                      [-15.0, -25.0, np.sin(5.5), np.cos(5.5)]])
    bbox = None
    if with_bbox:
        bbox = np.zeros((len(poses), rows_t * cols_t), np.float32)
        bbox[:, rng.integers(0, rows_t * cols_t, 50)] = 100.0
    layer = checker.from_occupancy(occ)
    val, idx_ref = checker.warp(layer, sg, tg, poses, bbox=bbox)
    want = np.stack([checker.to_occupancy(v, 0.0, 100.0) for v in val])
    got, idx = cilqr_b200.warp_occupancy_batch(occ, sg, tg, poses, bbox=bbox, want_idx=True)
    assert np.array_equal(idx, idx_ref)
    assert np.array_equal(got, want)
    assert (got == -1).any() and (got >= 0).any()
    # device buffers, other scaling
    want2 = np.stack([checker.to_occupancy(v, -5.0, 37.5) for v in val])
    got2, _ = cilqr_b200.warp_occupancy_batch(torch.from_numpy(occ).cuda(), sg, tg, torch.from_numpy(poses).cuda(),
                                              bbox=None if bbox is None else torch.from_numpy(bbox).cuda(), data_min=-5.0, data_max=37.5)
    assert np.array_equal(got2.cpu().numpy(), want2)
    # and against this library's own three kernels
    three = cilqr_b200.to_occupancy(cilqr_b200.warp_batch(cilqr_b200.from_occupancy(occ), sg, tg, poses[:1], bbox=None if bbox is None else bbox[:1],
                                                          want_idx=False)[0][0], 0.0, 100.0)
    assert np.array_equal(three, got[0])


def test_occupancy_device_buffers_and_full_size_round_trip():
    import torch
    n = 4096 * 4096
    g = torch.Generator(device="cuda").manual_seed(5)
    occ = torch.randint(-1, 101, (n,), generator=g, device="cuda", dtype=torch.int32).to(torch.int8)
    data = cilqr_b200.from_occupancy(occ)
    assert data.dtype == torch.float32 and data.shape == (n,)
    unknown = occ.flip(0) == -1
    assert torch.equal(torch.isnan(data), unknown)
    assert torch.equal(data[~unknown], occ.flip(0)[~unknown].to(torch.float32))
    # toOccupancyGrid's float arithmetic restated with numpy float32 ops (IEEE division, multiplication, truncation)
    d = data.cpu().numpy()
    with np.errstate(invalid="ignore"):
        v = (d - np.float32(0.0)) / (np.float32(100.0) - np.float32(0.0))
        want = np.where(np.isnan(v), np.float32(-1.0), np.float32(0.0) + np.clip(v, np.float32(0.0), np.float32(1.0)) * np.float32(100.0))
    want = want.astype(np.int32).astype(np.int8)[::-1]
    assert np.array_equal(cilqr_b200.to_occupancy(data, 0.0, 100.0).cpu().numpy(), want)


def boxes_case(seed, n_boxes, aligned):
    rng = np.random.default_rng(seed)
    own = np.array([rng.uniform(-50, 50), rng.uniform(-50, 50), 0.0 if aligned else rng.uniform(-3.1, 3.1)])
    b = np.zeros((n_boxes, 5))
    b[:, 0] = own[0] + rng.uniform(-45, 45, n_boxes)
    b[:, 1] = -(own[1] + rng.uniform(-45, 45, n_boxes))  # the message's y is negated by the reference
    b[:, 2] = 0.0 if aligned else rng.uniform(-3.1, 3.1, n_boxes)
    b[:, 3] = rng.uniform(1.5, 6.0, n_boxes)
    b[:, 4] = rng.uniform(1.0, 2.6, n_boxes)
    if aligned:  # edges exactly through cell centres: centres sit at odd multiples of 0.1 m
        b[:, 0] = own[0] + np.round(rng.uniform(-30, 30, n_boxes) * 5) / 5 + 0.1
        b[:, 1] = -(own[1] + np.round(rng.uniform(-30, 30, n_boxes) * 5) / 5 + 0.1)
        b[:, 3], b[:, 4] = 4.6, 2.2  # + 0.2 -> half sizes 2.4 and 1.2
    b[-1, 0] = own[0] + 500.0  # farther than 100 m: skipped (local_costmap.cpp:879)
    return own, b


@pytest.mark.parametrize("seed,n_boxes,aligned", [(1, 6, False), (2, 12, False), (3, 8, True), (4, 3, False)])
def test_rasterize_boxes_matches_reference_polygon_iterator(checker, seed, n_boxes, aligned):
    need_ref(checker)
    g = G(80.0, 80.0, 0.2, 35.0, 0.0)  # the vehicle map of local_costmap.cpp:213
    rows, cols = cilqr_b200.geom_size(g)
    own, b = boxes_case(seed, n_boxes, aligned)
    rng = np.random.default_rng(seed)
    layer0 = np.where(rng.uniform(size=rows * cols) < 0.02, 55.0, 0.0).astype(np.float32)
    ref = checker.rasterize_boxes(g, own, b, layer0)
    mine = cilqr_b200.rasterize_boxes(g, own[None], b[None], layer0.copy()[None])[0]
    assert np.array_equal(mine, ref)
    assert (ref == 100.0).sum() > 0 and (mine[layer0 == 55.0] >= 55.0).all()


def test_rasterize_boxes_batched_frames_on_device(checker):
    need_ref(checker)
    import torch
    g = G(51.2, 38.4, 0.2, 5.0, -3.0)
    rows, cols = cilqr_b200.geom_size(g)
    P, n_boxes = 5, 7
    cases = [boxes_case(10 + f, n_boxes, False) for f in range(P)]
    own = np.stack([c[0] for c in cases])
    bx = np.stack([c[1] for c in cases])
    layer = torch.zeros((P, rows * cols), dtype=torch.float32, device="cuda")
    cilqr_b200.rasterize_boxes(g, own, bx, layer)
    got = layer.cpu().numpy()
    for f in range(P):
        assert np.array_equal(got[f], checker.rasterize_boxes(g, own[f], bx[f], np.zeros(rows * cols, np.float32))), f
    # the rasterised layer is the warp's overlay input (local_costmap.cpp:260-263): values > 90 win
    sg = G(120.0, 120.0, 0.2, 5.0, -3.0)
    rs, cs = cilqr_b200.geom_size(sg)
    src = torch.zeros(rs * cs, dtype=torch.float32, device="cuda")
    poses = torch.tensor([[5.0, -3.0, 0.0, 1.0]] * P, dtype=torch.float64, device="cuda")
    out, _ = cilqr_b200.warp_batch(src, sg, g, poses, bbox=layer)
    assert torch.equal(out == 100.0, layer == 100.0)
