"""Loop-closure support operations through the C ABI against the CPU oracle: voxel-overlap indices
(computeIndicesOfOverlappingPoints, helpers.cpp:320-345) bit-exact as sets; the 6x6 information matrix
(GetInformationMatrixFromPointClouds) within 1e-9 relative — the correspondences come from the path's fp32 nearest-neighbour
search and the oracle's from an fp64 brute force, so a near-tie may pick another target point a few centimetres away."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def submap_pair(seed, n=20000):
    sc = scenes.Scene(seed, size=(30.0, 24.0, 6.0), n_boxes=8, n_cyl=5)
    a, _ = sc.sample_surface(n, seed + 1)
    b, _ = sc.sample_surface(n, seed + 2)
    rng = np.random.default_rng(seed)
    a = a[(a[:, 0] < 20.0)]                      # partial overlap: source covers x < 20, target x > 10
    b = b[(b[:, 0] > 10.0)]
    T = scenes.rigid(scenes.rot_axis_angle([0.1, 0.2, 1.0], np.deg2rad(3.0)), [0.4, -0.3, 0.1])
    src = (np.linalg.inv(T) @ np.concatenate([a, np.ones((a.shape[0], 1))], axis=1).T).T[:, :3] + rng.normal(0, 0.005, a.shape)
    return src, b, T


@pytest.mark.parametrize("voxel,min_pts", [(0.5, 1), (0.9, 3), (0.2, 1), (2.0, 25)])
def test_overlap_indices_match_oracle(ctx, oracle, voxel, min_pts):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    src, tgt, T = submap_pair(61)
    so, to = oracle.overlap(src, tgt, T, voxel, min_pts)
    sg, tg = ctx.overlap_indices(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), T, voxel, min_pts)
    assert np.array_equal(sg, so) and np.array_equal(tg, to)
    assert 0 < sg.size < src.shape[0] and 0 < tg.size < tgt.shape[0]          # a partial overlap, as constructed
    # disjoint clouds, empty clouds, parameter errors
    far = np.eye(4); far[:3, 3] = 500.0
    sg, tg = ctx.overlap_indices(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), far, voxel, min_pts)
    assert sg.size == 0 and tg.size == 0
    sg, tg = ctx.overlap_indices(capi.Cloud64.upload(ctx, np.zeros((0, 3))), capi.Cloud64.upload(ctx, tgt), T, voxel, min_pts)
    assert sg.size == 0 and tg.size == 0
    with pytest.raises(capi.InvalidParameter):
        ctx.overlap_indices(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), T, 0.0, 1)
    with pytest.raises(capi.InvalidParameter):
        ctx.overlap_indices(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), T, voxel, 0)
    with pytest.raises(capi.InvalidParameter):
        ctx.overlap_indices(capi.Cloud64.upload(ctx, src * 1e6), capi.Cloud64.upload(ctx, tgt), T, 1e-3, 1)   # voxel index beyond 2^20


@pytest.mark.parametrize("max_dist", [0.2, 0.6, 2.0])
def test_information_matrix_matches_oracle(ctx, oracle, max_dist):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    src, tgt, T = submap_pair(62, n=6000)
    info_o, cnt_o = oracle.information_matrix(src, tgt, max_dist, T)
    info_g, cnt_g = ctx.information_matrix(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), max_dist, T)
    assert abs(cnt_g - cnt_o) <= max(2, cnt_o // 2000), (cnt_g, cnt_o)          # radius boundary in fp32 vs fp64
    assert cnt_o > 100
    assert np.allclose(info_g, info_g.T, rtol=0, atol=0)                         # built from 10 moments: exactly symmetric
    assert np.abs(info_g - info_o).max() <= 2e-3 * np.abs(info_o).max() * max(1, abs(cnt_g - cnt_o)) + 1e-9 * np.abs(info_o).max()
    if cnt_g == cnt_o:
        assert np.abs(info_g - info_o).max() <= 1e-6 * np.abs(info_o).max()
    assert info_g[3, 3] == info_g[4, 4] == info_g[5, 5] == cnt_g                 # the translation block is count * I
    # nothing within reach -> zero matrix; empty source -> zero matrix; bad radius -> InvalidParameter
    far = np.eye(4); far[:3, 3] = 500.0
    z, c = ctx.information_matrix(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), max_dist, far)
    assert c == 0 and not z.any()
    z, c = ctx.information_matrix(capi.Cloud64.upload(ctx, np.zeros((0, 3))), capi.Cloud64.upload(ctx, tgt), max_dist, T)
    assert c == 0 and not z.any()
    with pytest.raises(capi.InvalidParameter):
        ctx.information_matrix(capi.Cloud64.upload(ctx, src), capi.Cloud64.upload(ctx, tgt), 0.0, T)
