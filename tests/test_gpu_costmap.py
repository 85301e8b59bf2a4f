"""GPU parity tests of the costmap warp (run on the B200 box: pytest -m gpu), through the C ABI.

Bar: source-index maps identical int-for-int and values identical bit-for-bit to the reference's grid_map_core
arithmetic (oracle/_ref = the unmodified reference sources; else the plain-C port), on both kernels (direct gather
and TMA-staged)."""
import os

import numpy as np
import pytest

import cilqr_b200
from cilqr_b200 import scenarios
from cilqr_b200.costmap import VARIANT_AUTO, VARIANT_DIRECT, VARIANT_TMA

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
G = cilqr_b200.CostmapGeom


def bits(a):
    return np.asarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "warp_golden.npz"))


@pytest.mark.parametrize("variant", [VARIANT_DIRECT, VARIANT_TMA, VARIANT_AUTO])
def test_small_golden_every_index(gold, variant):
    sg, tg = G(*gold["small/src_geom"]), G(*gold["small/tgt_geom"])
    rs, cs = cilqr_b200.geom_size(sg)
    src = scenarios.make_source_map(rs, cs)
    val, idx = cilqr_b200.warp_batch(src, sg, tg, gold["small/poses"], variant=variant)
    assert np.array_equal(idx, gold["small/idx"])
    assert np.array_equal(bits(val), bits(gold["small/val"]))
    assert (idx < 0).any()  # out-of-map cells are covered: -1 / NaN where the reference throws


def test_config2_reference_geometry(checker, gold):
    """BASELINE config 2: 1506x1506 global map (local_costmap.cpp:119), 400x400 ego frame, 64 poses."""
    sg, tg = scenarios.reference_global_geom(), G(80.0, 80.0, 0.2, 35.0, 0.0)
    assert cilqr_b200.geom_size(sg) == (1506, 1506) and cilqr_b200.geom_size(tg) == (400, 400)
    src = scenarios.make_source_map(1506, 1506)
    th = np.arange(64) * 2 * np.pi / 64
    poses = np.stack([93.14 + 15 * np.cos(3 * th), -205.96 + 15 * np.sin(5 * th), np.sin(th), np.cos(th)], axis=1)
    poses[0] = (93.14, -205.96, np.sin(0.7), np.cos(0.7))
    val, idx = cilqr_b200.warp_batch(src, sg, tg, poses)
    g0, g1 = idx[0] % 1506, idx[0] // 1506
    assert int(np.sum(g0.astype(np.int64) * 1506 + g1.astype(np.int64))) == int(gold["config2/index_checksum"]) == 149172896914
    assert int(np.sum(idx[0] < 0)) == int(gold["config2/oob"]) == 0
    assert int(np.sum(np.isnan(val[0]))) == int(gold["config2/nan_count"])
    assert np.array_equal(idx[0][:400], gold["config2/idx_row0"])
    vr, ir = checker.warp(src, sg, tg, poses, threads=0)
    assert np.array_equal(idx, ir) and np.array_equal(bits(val), bits(vr))


@pytest.mark.parametrize("variant", [VARIANT_DIRECT, VARIANT_TMA])
@pytest.mark.parametrize("shape", [(200, 152, 60, 44), (256, 256, 128, 128), (400, 300, 129, 67), (64, 64, 90, 90), (1024, 512, 301, 77)])
def test_ragged_shapes_with_overlay_match_checker(checker, variant, shape):
    """Tile edges that are not multiples of 64, target rows that are not multiples of 4 (scalar-store path), targets
    larger than the source (mostly out of map), and the bounding-box overlay (local_costmap.cpp:260-263)."""
    rs, cs, rt, ct = shape
    sg, tg = G(rs * 0.2, cs * 0.2, 0.2, 10.0, -5.0), G(rt * 0.2, ct * 0.2, 0.2, 2.0, 0.5)
    assert cilqr_b200.geom_size(sg) == (rs, cs) and cilqr_b200.geom_size(tg) == (rt, ct)
    src = scenarios.make_source_map(rs, cs)
    poses = scenarios.make_poses(12, (10.0, -5.0), 0.3 * min(rs, cs) * 0.2)
    poses[0] = (10.0, -5.0, 0.0, 1.0)
    poses[1] = (10.0, -5.0, 1.0, 0.0)
    poses[2] = (10.1, -5.1, np.sin(np.pi), np.cos(np.pi))
    rng = np.random.default_rng(5)
    bbox = np.where(rng.random((12, rt * ct)) > 0.97, 100.0, np.where(rng.random((12, rt * ct)) > 0.5, 50.0, 0.0)).astype(np.float32)
    vr, ir = checker.warp(src, sg, tg, poses, bbox=bbox)
    val, idx = cilqr_b200.warp_batch(src, sg, tg, poses, bbox=bbox, variant=variant)
    assert np.array_equal(idx, ir)
    assert np.array_equal(bits(val), bits(vr))


def test_cell_border_positions_follow_the_truncation_rule(checker):
    """grid_map_core/test/GridMapMathTest.cpp:116-137 (IndexFromPosition.EdgeCases) at kernel level: ego offsets that
    put every cell centre exactly on (or one ulp either side of) a source cell border."""
    sg, tg = G(51.2, 51.2, 0.2, 0.0, 0.0), G(12.8, 12.8, 0.2, 0.0, 0.0)
    src = scenarios.make_source_map(256, 256, nan_band=False)
    offs = [0.0, 0.1, -0.1, 0.1 + 2e-16, 0.1 - 2e-16, 0.3, np.nextafter(0.1, 1), np.nextafter(0.1, 0), 0.2, 1e-17]
    poses = np.array([[o, -o, 0.0, 1.0] for o in offs] + [[o, o, 1.0, 0.0] for o in offs] + [[o, 0.0, -1.0, 0.0] for o in offs])
    vr, ir = checker.warp(src, sg, tg, poses)
    for variant in (VARIANT_DIRECT, VARIANT_TMA):
        val, idx = cilqr_b200.warp_batch(src, sg, tg, poses, variant=variant)
        assert np.array_equal(idx, ir) and np.array_equal(bits(val), bits(vr))


def test_known_answers_identity_and_whole_cell_shift():
    """Known-answer properties at a BASELINE sweep size (2048^2 target, 4096^2 source): the identity pose maps the
    centred target window onto itself, and a shift by k whole cells shifts every index by k (GridMapTest.cpp:441-463
    ValueAtPosition.NearestNeighbor / GridMapMathTest.cpp:85-114 at scale)."""
    import torch
    T, S = 2048, 4096
    sg, tg = G(S * 0.25, S * 0.25, 0.25, 0.0, 0.0), G(T * 0.25, T * 0.25, 0.25, 0.0, 0.0)  # 0.25: exactly representable
    dev = torch.device("cuda:0")
    src = torch.arange(S * S, dtype=torch.float32, device=dev) % 65536.0
    k = 7
    poses = torch.tensor([[0.0, 0.0, 0.0, 1.0], [k * 0.25, -3 * k * 0.25, 0.0, 1.0]], dtype=torch.float64, device=dev)
    for variant in (VARIANT_DIRECT, VARIANT_TMA):
        val, idx = cilqr_b200.warp_batch(src, sg, tg, poses, variant=variant)
        i = torch.arange(T, device=dev, dtype=torch.int64)
        base = (S - T) // 2
        want0 = ((base + i)[:, None] * S + (base + i)[None, :]).reshape(-1)  # [j][i] -> g0 + S*g1 = (base+i) + S*(base+j)
        assert torch.equal(idx[0].to(torch.int64), want0)
        want1 = ((base + 3 * k + i)[:, None] * S + (base - k + i)[None, :]).reshape(-1)
        assert torch.equal(idx[1].to(torch.int64), want1)
        assert torch.equal(val[0], src[want0]) and torch.equal(val[1], src[want1])


def test_full_sweep_size_tma_equals_direct_and_sample_matches_checker(checker):
    """BASELINE config 5 at its largest grid (4096^2 target from 8192^2 source): both kernels agree bit-for-bit on
    every cell, rotating by theta then sampling agrees with the checker on one pose, and out-of-map cells are -1."""
    import torch
    T, S = 4096, 8192
    sg, tg = G(S * 0.2, S * 0.2, 0.2, 0.0, 0.0), G(T * 0.2, T * 0.2, 0.2, 0.0, 0.0)
    dev = torch.device("cuda:0")
    src_h = scenarios.make_source_map(S, S, nan_band=True)
    src = torch.from_numpy(src_h).to(dev)
    poses_h = scenarios.make_poses(4, (0.0, 0.0), 0.05 * T * 0.2)
    poses_h[3, 0] += 0.45 * S * 0.2  # drives part of the frame out of the map
    poses = torch.from_numpy(poses_h).to(dev)
    v1, i1 = cilqr_b200.warp_batch(src, sg, tg, poses, variant=VARIANT_DIRECT)
    v2, i2 = cilqr_b200.warp_batch(src, sg, tg, poses, variant=VARIANT_TMA)
    assert torch.equal(i1, i2) and torch.equal(v1.view(torch.int32), v2.view(torch.int32))
    assert int((i1[3] < 0).sum()) > 0 and int((i1[:3] < 0).sum()) == 0
    assert torch.isnan(v1[3][i1[3] < 0]).all()
    vr, ir = checker.warp(src_h, sg, tg, poses_h[1:2], threads=0)
    assert np.array_equal(i1[1].cpu().numpy(), ir[0]) and np.array_equal(bits(v1[1].cpu().numpy()), bits(vr[0]))


def test_empty_and_bad_arguments():
    sg, tg = G(12.8, 12.8, 0.2, 0.0, 0.0), G(6.4, 6.4, 0.2, 0.0, 0.0)
    src = np.zeros(64 * 64, np.float32)
    val, idx = cilqr_b200.warp_batch(src, sg, tg, np.zeros((0, 4)))
    assert val.shape == (0, 32 * 32) and idx.shape == (0, 32 * 32)
    with pytest.raises(cilqr_b200.CilqrError, match="variant"):
        cilqr_b200.warp_batch(src, sg, tg, np.array([[0.0, 0.0, 0.0, 1.0]]), variant=7)
    sg_odd = G(12.6, 12.8, 0.2, 0.0, 0.0)  # 63 rows: pitch not a multiple of 16 bytes -> TMA refused, auto falls back to direct
    src_odd = np.zeros(63 * 64, np.float32)
    with pytest.raises(cilqr_b200.CilqrError, match="TMA"):
        cilqr_b200.warp_batch(src_odd, sg_odd, tg, np.array([[0.0, 0.0, 0.0, 1.0]]), variant=VARIANT_TMA)
    val, idx = cilqr_b200.warp_batch(src_odd, sg_odd, tg, np.array([[0.0, 0.0, 0.0, 1.0]]))
    assert (idx >= 0).all()
