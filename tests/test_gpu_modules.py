"""The remaining chain modules of SURVEY 8f rank 3 through the C ABI against the CPU oracle: VarTrimmedDistOutlierFilter (ratio, limit
and weights bit-exact: the fp32 running sum is taken in the reference's order on the device), PointToPointErrorMinimizer (transform
within 1e-5 of the oracle's fp64 lane), the PointToPlaneWithCov covariance (1e-9 relative) and the compacting DataPointsFilters
MaxDist / MinDist / BoundingBox / RandomSampling (membership and order bit-exact, RandomSampling for a given seed)."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def f4(a):
    return np.concatenate([np.asarray(a, np.float32), np.ones((len(a), 1), np.float32)], axis=1)


@pytest.mark.parametrize("lam", [0.0, 0.5, 0.95, 2.0])
def test_var_trimmed_dist_matches_oracle(ctx, oracle, lam):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    rng = np.random.default_rng(31)
    for n, lo, hi in ((5, 0.0000001, 1.0), (64, 0.2, 0.9), (1000, 0.05, 0.99), (5000, 0.6, 0.8), (131072, 0.05, 0.99)):
        inl = int(0.7 * n)
        d2 = np.concatenate([np.abs(rng.normal(0, 0.1, inl)) ** 2, rng.random(n - inl) * 4 + 1]).astype(np.float32)
        rng.shuffle(d2)
        if n > 10:
            d2[::37] = np.inf          # unmatched points
            d2[3] = 0.0                # exact hits are left out of the sorted list (:190) but count in points_nbr
            d2[10:14] = d2[14]         # ties
        st, ratio_o, tau_o, w_o = oracle.var_trimmed_dist(d2, lo, hi, lam)
        ratio_g, tau_g, w_g = ctx.var_trimmed_dist(d2, lo, hi, lam)
        assert st == 0
        assert np.float32(ratio_g).view(np.uint32) == np.float32(ratio_o).view(np.uint32), (n, lam, ratio_g, ratio_o)
        assert np.float32(tau_g).view(np.uint32) == np.float32(tau_o).view(np.uint32) and np.array_equal(w_g, w_o)
    # the reference's known answers (utest/ui/Outliers.cpp:126-151)
    d2 = np.array([4, 5, 5, 5, 5], dtype=np.float32)
    assert ctx.var_trimmed_dist(d2, 0.0000001, 1.0, 0.0)[2].tolist() == [1, 0, 0, 0, 0]
    assert ctx.var_trimmed_dist(d2, 0.0000001, 1.0, 1.0)[2].tolist() == [1, 1, 1, 1, 1]
    with pytest.raises(capi.InvalidParameter):
        ctx.var_trimmed_dist(d2, 0.8, 0.6, 1.0)
    with pytest.raises(capi.ConvergenceError):
        ctx.var_trimmed_dist(np.array([np.inf, 0.0], np.float32), 0.1, 0.9, 1.0)
    with pytest.raises(capi.ConvergenceError):
        ctx.var_trimmed_dist(np.zeros(0, np.float32), 0.1, 0.9, 1.0)


def matched_pair(seed, n=20000, k=1):
    sc = scenes.Scene(seed, size=(30.0, 24.0, 6.0), n_boxes=8, n_cyl=5)
    ref, nrm = sc.sample_surface(n, seed + 1)
    rng = np.random.default_rng(seed)
    T = scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(4.0)), rng.uniform(-0.3, 0.3, 3))
    sel = rng.permutation(n)[: n // 2]
    reading = (np.linalg.inv(T) @ np.concatenate([ref[sel] + rng.normal(0, 0.01, (sel.size, 3)), np.ones((sel.size, 1))], axis=1).T).T[:, :3]
    ids = np.stack([np.roll(sel, j) if j else sel for j in range(k)], axis=1).astype(np.int32)   # slot 0 = the true partner
    d2 = rng.random(ids.shape).astype(np.float32)
    d2[::19, 0] = np.inf
    w = (rng.random(ids.shape) < 0.8).astype(np.float32)
    if k > 1:
        w[:, 1:] = 0.0
        w[::5, 1] = 0.25       # a few fractional-weight second neighbours
    return f4(reading), f4(ref), nrm.astype(np.float32), ids, d2, w, T


@pytest.mark.parametrize("k", [1, 3])
def test_point_to_point_matches_oracle(ctx, oracle, k):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    reading, ref, nrm, ids, d2, w, T = matched_pair(71 + k, k=k)
    st, T_o, kept_o = oracle.point_to_point(reading, ref, ids, d2, w)
    cr, cf = ctx.upload(reading), ctx.upload(ref)
    T_g, stats = ctx.point_to_point(cr, cf, ids, d2, w)
    assert st == 0 and stats.kept == kept_o
    assert np.abs(T_g - T_o).max() < 1e-5, np.abs(T_g - T_o).max()          # 1e-5 m / 1e-5 rad (north_star tolerance)
    if k == 1:
        assert np.abs(T_g.astype(np.float64) - T).max() < 5e-3              # and it is the transform the pairs were made with
    assert abs(np.linalg.det(T_g[:3, :3].astype(np.float64)) - 1.0) < 1e-5
    # mirrored reference: the reflection branch still returns a rotation, equal to the oracle's
    mir = ref.copy(); mir[:, 2] *= -1
    st, T_o, _ = oracle.point_to_point(reading, mir, ids, d2, w)
    T_g, _ = ctx.point_to_point(cr, ctx.upload(mir), ids, d2, w)
    assert st == 0 and np.linalg.det(T_g[:3, :3].astype(np.float64)) > 0.99 and np.abs(T_g - T_o).max() < 1e-4
    with pytest.raises(capi.ConvergenceError):
        ctx.point_to_point(cr, cf, ids, d2, np.zeros_like(w))


def test_point_to_plane_covariance_matches_oracle(ctx, oracle):
    reading, ref, nrm, ids, d2, w, T = matched_pair(81)
    w = (w != 0).astype(np.float32)
    cr, cf = ctx.upload(reading), ctx.upload(ref, nrm)
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    A, b, stats = ctx.point_to_plane(cr, cf, ids, d2, w)
    Tstep, x = capi.solve_point_to_plane(A, b, stats.mean_reading, stats.mean_reference)
    cov_g = ctx.point_to_plane_covariance(cr, cf, ids, d2, w, stats.mean_reading, stats.mean_reference, Tstep, 0.03)
    st, cov_o = oracle.point_to_plane_covariance(reading, ref, nrm, ids, d2, w, Tstep, 0.03)
    assert st == 0 and np.isfinite(cov_g).all()
    scale = np.abs(cov_o).max()
    assert np.abs(cov_g - cov_o).max() < 1e-9 * scale, np.abs(cov_g - cov_o).max() / scale
    assert np.allclose(cov_g, cov_g.T, rtol=0, atol=1e-12 * scale) and (np.linalg.eigvalsh(cov_g) > -1e-12 * scale).all()
    # a reference without normals: max * Identity (PointToPlaneWithCov.cpp:91-92)
    none = ctx.point_to_plane_covariance(cr, ctx.upload(ref), ids, d2, w, stats.mean_reading, stats.mean_reference, Tstep, 0.03)
    assert np.array_equal(none, np.eye(6) * np.finfo(np.float32).max)


@pytest.mark.parametrize("n", [0, 1, 1000, 131072])
def test_compacting_filters_match_oracle(ctx, oracle, n):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    rng = np.random.default_rng(91 + n)
    pts = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    nrm = rng.normal(size=(n, 3)).astype(np.float32)
    if n >= 1000:
        pts[:8] = [[3, 4, 0], [0, 0, 5], [5, 0, 0], [0, -5, 0], [2, 2, 2], [-2, 2, 2], [2, 2, 6], [2, 6, 2]]   # exactly on radius 5 / box faces
    cases = [dict(kind=capi.FILTER_MAX_DIST, dim=-1, dist=-5.0), dict(kind=capi.FILTER_MAX_DIST, dim=0, dist=2.0),
             dict(kind=capi.FILTER_MIN_DIST, dim=-1, dist=5.0), dict(kind=capi.FILTER_MIN_DIST, dim=2, dist=-1.5),
             dict(kind=capi.FILTER_BOUNDING_BOX, box=(-2, 2, -2, 6, -3, 6), remove_inside=True),
             dict(kind=capi.FILTER_BOUNDING_BOX, box=(-2, 2, -2, 6, -3, 6), remove_inside=False),
             dict(kind=capi.FILTER_RANDOM_SAMPLING, prob=0.35, seed=12345, random_method=0),
             dict(kind=capi.FILTER_RANDOM_SAMPLING, prob=0.35, seed=987654321987, random_method=1),
             dict(kind=capi.FILTER_RANDOM_SAMPLING, prob=1.0, seed=5), dict(kind=capi.FILTER_RANDOM_SAMPLING, prob=0.0, seed=5)]
    for c in cases:
        keep = oracle.cloud_filter(f4(pts), **c) if n else np.zeros(0, bool)
        cl = ctx.upload(f4(pts), nrm).filter(**c)
        assert cl.size == int(keep.sum()), (c, cl.size, int(keep.sum()))
        got, got_nrm = cl.download()
        assert np.array_equal(got[:, :3], pts[keep]) and np.array_equal(got_nrm, nrm[keep]), c      # same points, same order
        cl.free()
    with pytest.raises(capi.InvalidParameter):
        ctx.upload(f4(pts)).filter(capi.FILTER_MAX_DIST, dim=3)
    with pytest.raises(capi.InvalidParameter):
        ctx.upload(f4(pts)).filter(capi.FILTER_RANDOM_SAMPLING, prob=1.5)
