"""GPU parity tests (run on the B200 with -m gpu): CUDA path through the C ABI vs the CPU oracle on the same inputs.

Bars (north_star): kNN indices and squared distances bit-exact; trimmed-distance limit bit-exact; kept counts exact;
A, b vs the fp64 oracle at 1e-9 relative; final transforms within 1e-5 m / 1e-5 rad.
"""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu

TOL_T = 1e-5     # metres   (north_star)
TOL_R = 1e-5     # radians  (north_star)


def rot_angle(Ra, Rb):
    R = Ra.T.astype(np.float64) @ Rb.astype(np.float64)
    return float(np.arccos(np.clip((np.trace(R) - 1) / 2, -1, 1)))


def assert_same_T(Ta, Tb, tol_t=TOL_T, tol_r=TOL_R):
    assert np.abs(Ta[:3, 3].astype(np.float64) - Tb[:3, 3]).max() <= tol_t, (Ta, Tb)
    # arccos loses precision near 0: use the Frobenius bound |R_a - R_b|_F ~= sqrt(2) * angle for small angles
    ang = np.linalg.norm(Ta[:3, :3].astype(np.float64) - Tb[:3, :3]) / np.sqrt(2)
    assert ang <= tol_r, (ang, Ta, Tb)


def small_T(seed, t=0.2, deg=3.0):
    rng = np.random.default_rng(seed)
    R = scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(deg))
    return scenes.rigid(R, rng.uniform(-t, t, 3)).astype(np.float32)


def random_cloud(n, seed, scale=10.0):
    rng = np.random.default_rng(seed)
    return (rng.random((n, 3)) * scale - scale / 2).astype(np.float32)


# ---------------------------------------------------------------------------------------------------------------
# kNN: ids and d2 bit patterns vs the oracle k-d tree (and brute force on a sub-sample)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("k,max_dist", [(1, np.inf), (1, 0.5), (2, 1.0), (3, 0.5), (10, 1.5), (20, np.inf)])
def test_knn_car_clouds_bit_exact(ctx, oracle, golden, k, max_dist):
    ref = oracle.feat4(golden["car"]["ref_xyz"])
    q = oracle.feat4(golden["car"]["data_xyz"])
    tree = oracle.KdTree(ref)
    ids_o, d2_o, _ = tree.knn(q, k=k, max_dist=max_dist, nthreads=8)
    cref, cq = ctx.upload(ref), ctx.upload(q)
    index = ctx.build_index(cref)
    ids_g, d2_g, visits = ctx.knn(index, cq, k=k, max_dist=max_dist)
    assert np.array_equal(ids_g, ids_o)
    assert np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))
    assert visits > 0


@pytest.mark.parametrize("cell", [0.0, 0.05, 0.37, 2.5])
def test_knn_cell_size_does_not_change_results(ctx, oracle, golden, cell):
    ref = oracle.feat4(golden["car"]["ref_xyz"])
    q = oracle.feat4(golden["car"]["data_xyz"][::7])
    ids_o, d2_o, _ = oracle.KdTree(ref).knn(q, k=4, max_dist=0.8, nthreads=8)
    index = ctx.build_index(ctx.upload(ref), cell=cell)
    ids_g, d2_g, _ = ctx.knn(index, ctx.upload(q), k=4, max_dist=0.8)
    assert np.array_equal(ids_g, ids_o)
    assert np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))


def test_knn_fused_transform_matches_oracle_transform(ctx, oracle, golden):
    ref = oracle.feat4(golden["car"]["ref_xyz"])
    q = oracle.feat4(golden["car"]["data_xyz"])
    T = small_T(3)
    st, qT, _ = oracle.transform(T, q)
    assert st == 0
    ids_o, d2_o, _ = oracle.KdTree(ref).knn(qT, k=1, max_dist=1.0, nthreads=8)
    index = ctx.build_index(ctx.upload(ref))
    ids_g, d2_g, _ = ctx.knn(index, ctx.upload(q), T=T, k=1, max_dist=1.0)
    assert np.array_equal(ids_g, ids_o)
    assert np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))
    # the standalone transform entry point is bit-exact too
    c = ctx.upload(q)
    c.transform(T)
    f, _ = c.download()
    assert np.array_equal(f.view(np.uint32), qT.view(np.uint32))


def test_knn_ties_smallest_index_and_padding(ctx, oracle):
    # lattice => massive exact distance ties; queries at cell centres and far away (padding)
    g = np.stack(np.meshgrid(np.arange(12), np.arange(12), np.arange(12), indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    ref = oracle.feat4(np.concatenate([g, g[:200]]))   # duplicates: ties on identical points too
    q = oracle.feat4(np.concatenate([g[::5] + 0.5, g[::11], np.array([[100, 100, 100], [-50, 3, 3]], dtype=np.float32)]))
    for k, md in [(1, np.inf), (8, 1.0), (16, 2.0), (5, 0.4)]:
        ids_o, d2_o = oracle.knn_brute(ref, q, k=k, max_dist=md)
        index = ctx.build_index(ctx.upload(ref))
        ids_g, d2_g, _ = ctx.knn(index, ctx.upload(q), k=k, max_dist=md)
        assert np.array_equal(ids_g, ids_o), (k, md)
        assert np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))
    ids_g, d2_g, _ = ctx.knn(index, ctx.upload(q[-2:]), k=3, max_dist=1.0)
    assert (ids_g == -1).all() and np.isinf(d2_g).all()


def test_knn_self_match_and_no_self(ctx, oracle):
    pts = oracle.feat4(random_cloud(5000, 5))
    c = ctx.upload(pts)
    index = ctx.build_index(c)
    ids, d2, _ = ctx.knn(index, c, k=3, allow_self=True)
    assert np.array_equal(ids[:, 0], np.arange(5000)) and (d2[:, 0] == 0).all()
    ids_o, d2_o, _ = oracle.KdTree(pts).knn(pts, k=3, allow_self=False, nthreads=8)
    ids_n, d2_n, _ = ctx.knn(index, c, k=3, allow_self=False)
    assert np.array_equal(ids_n, ids_o) and np.array_equal(d2_n.view(np.uint32), d2_o.view(np.uint32))


def test_knn_synthetic_scan_vs_map(ctx, oracle):
    c = scenes.config2(m=400_000, n_beams=32, n_az=512)
    ref = oracle.feat4(c["map_xyz"])
    q = oracle.feat4(c["scan"])
    T = c["T_init"].astype(np.float32)
    _, qT, _ = oracle.transform(T, q)
    ids_o, d2_o, _ = oracle.KdTree(ref).knn(qT, k=1, max_dist=1.0, nthreads=8)
    index = ctx.build_index(ctx.upload(ref))
    ids_g, d2_g, _ = ctx.knn(index, ctx.upload(q), T=T, k=1, max_dist=1.0)
    assert np.array_equal(ids_g, ids_o)
    assert np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))


@pytest.mark.parametrize("k", [2, 8, 9, 16, 17, 32, 33, 48, 64])
def test_knn_list_kernel_equals_one_thread_kernel_and_oracle(ctx, oracle, k):
    """The W-lanes-per-query list kernel (2 <= k <= 32; two entries per lane for 32 < k <= 64) against the one-thread kernel
    (B200REG_KNN_LANES=1, k <= 32) and the oracle k-d tree on a raw LiDAR scan (density varies by orders of magnitude: crowded cells and
    many-ring walks), queries outside the grid, k > reachable points."""
    import os
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    c = scenes.config2(m=150_000, n_beams=32, n_az=512)
    ref = oracle.feat4(np.concatenate([c["scan"], c["map_xyz"][:30_000]]))
    rng = np.random.default_rng(k)
    far = (rng.random((64, 3)) * 400 - 200).astype(np.float32)                       # mostly outside the grid
    q = oracle.feat4(np.concatenate([c["scan"][::7] + rng.normal(0, 0.05, (c["scan"][::7].shape[0], 3)).astype(np.float32), far]))
    os.environ["B200REG_KNN_LANES"] = "1"
    try:
        ctx1 = capi.Context(0)
    finally:
        del os.environ["B200REG_KNN_LANES"]
    tree = oracle.KdTree(ref)
    for md, allow in ((np.inf, True), (0.8, True), (0.15, False)):
        ids_l, d2_l, vis = ctx.knn(ctx.build_index(ctx.upload(ref)), ctx.upload(q), k=k, max_dist=md, allow_self=allow)
        if k <= 32:
            ids_1, d2_1, _ = ctx1.knn(ctx1.build_index(ctx1.upload(ref)), ctx1.upload(q), k=k, max_dist=md, allow_self=allow)
            assert np.array_equal(ids_l, ids_1), (k, md)
            assert np.array_equal(d2_l.view(np.uint32), d2_1.view(np.uint32))
        else:
            with pytest.raises(capi.InvalidParameter):
                ctx1.knn(ctx1.build_index(ctx1.upload(ref)), ctx1.upload(q), k=k, max_dist=md, allow_self=allow)
        ids_o, d2_o, _ = tree.knn(q, k=k, max_dist=md, allow_self=allow, nthreads=8)
        assert np.array_equal(ids_l, ids_o), (k, md)
        assert np.array_equal(d2_l.view(np.uint32), d2_o.view(np.uint32))
        assert vis > 0
    # fewer points than k: padding (-1, inf) after the real neighbours
    tiny = oracle.feat4(random_cloud(max(k - 1, 1), 77))
    ids_t, d2_t, _ = ctx.knn(ctx.build_index(ctx.upload(tiny)), ctx.upload(q[:100]), k=k)
    ids_o, d2_o = oracle.knn_brute(tiny, q[:100], k=k)
    assert np.array_equal(ids_t, ids_o) and np.array_equal(d2_t.view(np.uint32), d2_o.view(np.uint32))
    assert (ids_t[:, -1] == -1).all()
    # a one-point reference far from the queries: the search starts at a ring ~1e6 cells out (k = 1 group kernel and list kernel)
    one = oracle.feat4(random_cloud(1, 78))
    for kk in (1, k):
        ids_t, d2_t, _ = ctx.knn(ctx.build_index(ctx.upload(one)), ctx.upload(q[:100]), k=kk)
        ids_o, d2_o = oracle.knn_brute(one, q[:100], k=kk)
        assert np.array_equal(ids_t, ids_o) and np.array_equal(d2_t.view(np.uint32), d2_o.view(np.uint32))


def test_knn_empty_and_errors(ctx, oracle, golden):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    ref = ctx.upload(oracle.feat4(golden["car"]["ref_xyz"][:100]))
    index = ctx.build_index(ref)
    empty = ctx.upload(np.zeros((0, 4), dtype=np.float32))
    ids, d2, v = ctx.knn(index, empty, k=1)
    assert ids.shape == (0, 1) and v == 0
    with pytest.raises(capi.NoPoints):
        ctx.build_index(empty)
    with pytest.raises(capi.InvalidParameter):
        ctx.knn(index, ref, k=0)
    with pytest.raises(capi.InvalidParameter):
        ctx.knn(index, ref, k=1, epsilon=-1.0)


# ---------------------------------------------------------------------------------------------------------------
# one ICP iteration: tau bit-exact, kept exact, A / b vs the fp64 oracle
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("k,trim,angle", [(1, 0.85, -1.0), (1, -1.0, -1.0), (3, 0.8, -1.0), (1, 0.9, 1.2), (10, 0.8, 0.42), (1, 1.0, -1.0)])
def test_icp_step_matches_oracle(ctx, oracle, golden, k, trim, angle):
    car = golden["car"]
    ref, rn = oracle.feat4(car["ref_xyz"]), car["ref_normals"]
    q = oracle.feat4(car["data_xyz"])
    # reading normals for the SurfaceNormalOutlierFilter gate: any deterministic unit vectors do
    qn = oracle.surface_normals(q, 8, nthreads=8)["normals"] if angle >= 0 else None
    T = small_T(11, t=0.05, deg=2.0)
    _, qT, qnT = oracle.transform(T, q, qn)
    md = 1.0
    ids, d2, _ = oracle.KdTree(ref).knn(qT, k=k, max_dist=md, nthreads=8)
    w = None
    tau_o = np.inf
    if trim > 0:
        st, tau_o, w = oracle.trimmed_dist(d2, trim)
        assert st == 0
    if angle >= 0:
        w2 = oracle.surface_normal_outlier(qnT, rn, ids, angle)
        w = w2 if w is None else w * w2
    if w is None:
        w = oracle.no_filter_weights(d2)
    o = oracle.point_to_plane(qT, ref, rn, ids, d2, w, precision=oracle.ORC_MIXED)
    assert o["status"] == 0

    cref, cq = ctx.upload(ref, rn), ctx.upload(q, qn)
    index = ctx.build_index(cref)
    A, b, st = ctx.icp_step(index, cq, T=T, k=k, max_dist=md, trim_ratio=trim, max_angle=angle)
    if trim > 0:
        assert np.float32(st.tau).view(np.uint32) == np.float32(tau_o).view(np.uint32)
    assert st.kept == o["kept"]
    assert st.n_finite == int(np.isfinite(d2).sum())
    assert st.point_used_ratio == pytest.approx(o["point_used_ratio"], rel=1e-6)
    scale = np.abs(o["A"]).max()
    assert np.abs(A - o["A"]).max() <= 1e-9 * scale
    assert np.abs(b - o["b"]).max() <= 1e-9 * max(np.abs(o["b"]).max(), 1e-3 * scale)
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    Tg, x = capi.solve_point_to_plane(A, b, list(st.mean_reading), list(st.mean_reference))
    assert_same_T(Tg, o["T"], 1e-6, 1e-6)


@pytest.mark.parametrize("kind,param", [(0, 0.35), (0, 2.0), (1, 0.35), (2, 3.0), (2, 0.5), (2, 1.0)])
def test_distance_outlier_filters_match_oracle(ctx, oracle, kind, param):
    """MaxDist / MinDist / MedianDist (b200reg_dist_outlier) against the oracle: limit and weights bit-exact, incl. +inf padding,
    exact ties at the limit and the median's float-multiply rank."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    rng = np.random.default_rng(17 + kind)
    for n in (1, 7, 1000, 131072):
        d2 = (rng.random(n).astype(np.float32) ** 2) * np.float32(1.5)
        d2[rng.random(n) < 0.1] = np.inf
        if n > 10:
            d2[:5] = np.float32(param) * np.float32(param)          # values exactly at a MaxDist / MinDist limit
            d2[5:9] = d2[9]                                          # ties
        if not np.isfinite(d2).any():
            d2[0] = 0.25
        st, lim_o, w_o = oracle.dist_outlier(kind, d2, param)
        lim_g, w_g = ctx.dist_outlier(kind, d2, param)
        assert st == 0 and np.float32(lim_g).view(np.uint32) == np.float32(lim_o).view(np.uint32), (n, lim_g, lim_o)
        assert np.array_equal(w_g, w_o), n
    # k > 1 shaped input (n x k matrix) is just a longer array
    d2 = rng.random((500, 3)).astype(np.float32)
    assert np.array_equal(ctx.dist_outlier(kind, d2, param)[1], oracle.dist_outlier(kind, d2, param)[2])
    # errors of the CPU twins: parameter bounds -> InvalidParameter; a median without any finite match -> ConvergenceError
    with pytest.raises(capi.InvalidParameter):
        ctx.dist_outlier(kind, d2, 0.0)
    if kind == 2:
        with pytest.raises(capi.ConvergenceError):
            ctx.dist_outlier(2, np.full(16, np.inf, dtype=np.float32), 3.0)
        with pytest.raises(capi.ConvergenceError):
            ctx.dist_outlier(2, np.zeros(0, dtype=np.float32), 3.0)
    else:
        assert ctx.dist_outlier(kind, np.zeros(0, dtype=np.float32), 1.0)[1].size == 0


def test_trimmed_quantile_rank_semantics(ctx, oracle):
    # n = 131072, ratio 0.8 -> rank 104857 (float multiply + truncation); heavy ties at the limit
    rng = np.random.default_rng(1)
    ref = oracle.feat4(random_cloud(50_000, 2, scale=4.0))
    rn = np.tile(np.array([[0, 0, 1]], dtype=np.float32), (ref.shape[0], 1))
    q = oracle.feat4(np.round(random_cloud(131072, 3, scale=4.0) * 8) / 8)   # quantised => tied distances
    ids, d2, _ = oracle.KdTree(ref).knn(q, k=1, max_dist=np.inf, nthreads=8)
    index = ctx.build_index(ctx.upload(ref, rn))
    cq = ctx.upload(q)
    for ratio in (0.8, 0.5, 0.999, 1.0, 1e-6):
        _, tau_o, w = oracle.trimmed_dist(d2, ratio)
        A, b, st = ctx.icp_step(index, cq, k=1, trim_ratio=ratio)
        assert np.float32(st.tau).view(np.uint32) == np.float32(tau_o).view(np.uint32), ratio
        assert st.kept == int(w.sum())


def test_icp_step_errors(ctx, oracle, golden):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    car = golden["car"]
    ref = oracle.feat4(car["ref_xyz"])
    q = ctx.upload(oracle.feat4(car["data_xyz"] + 1000.0))
    with_n = ctx.build_index(ctx.upload(ref, car["ref_normals"]))
    without_n = ctx.build_index(ctx.upload(ref))
    with pytest.raises(capi.InvalidField):          # PointToPlane.cpp:115
        ctx.icp_step(without_n, q, k=1)
    with pytest.raises(capi.ConvergenceError):      # Matches.cpp:76-77
        ctx.icp_step(with_n, q, k=1, max_dist=0.5, trim_ratio=0.8)
    with pytest.raises(capi.ConvergenceError):      # ErrorMinimizer.cpp:76-77
        ctx.icp_step(with_n, q, k=1, max_dist=0.5)


# ---------------------------------------------------------------------------------------------------------------
# whole registrations
# ---------------------------------------------------------------------------------------------------------------
def _gpu_icp(ctx, ref, rn, q, qn=None, T_init=None, **kw):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    icp = capi.ICP(ctx, **kw)
    icp.set_reference(ctx.upload(ref, rn))
    T, st = icp.register(ctx.upload(q, qn), T_init)
    icp.close()
    return T, st


def test_icp_car_clouds_config1(ctx, oracle, golden):
    """BASELINE config #1 data (car_cloud400 <- car_cloud401), default chain minus the two random samplers:
    KDTree(k1) + TrimmedDist(0.85) + PointToPlane + Counter(40) + Differential(1e-3, 1e-2, 3)."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8), ref, rn, q)
    assert code == 0
    T_g, st_g = _gpu_icp(ctx, ref, rn, q)
    assert st_g.iterations == st_o.iterations
    assert_same_T(T_g, T_o)
    assert st_g.kept_last == st_o.kept_last
    assert np.float32(st_g.last_tau).view(np.uint32) == np.float32(st_o.last_tau).view(np.uint32)
    # reference's own acceptance (utest.h:65-86): same basin as validT3d
    V = golden["T"]["validT3d"]
    assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(T_g[:3, 3])) < 0.1
    assert rot_angle(V[:3, :3], T_g[:3, :3]) < 0.1
    # and the pure-fp32 restatement stays inside the tolerance too
    code, T_32, _ = oracle.icp(oracle.icp_config(precision=oracle.ORC_REF32, nthreads=8), ref, rn, q)
    assert_same_T(T_g, T_32, 1e-5, 1e-5)


@pytest.mark.parametrize("knn,eps,maxd", [(1, 0.0, 1.0), (2, 0.0, 0.5), (3, 0.2, 1.0)])
def test_icp_matcher_param_grid(ctx, oracle, golden, knn, eps, maxd):
    """utest/ui/Matcher.cpp:68-92 shape: KDTreeMatcher x knn x epsilon x maxDist inside the default chain."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    code, T_o, st_o = oracle.icp(oracle.icp_config(knn=knn, epsilon=eps, max_dist=maxd, precision=oracle.ORC_MIXED, nthreads=8), ref, rn, q)
    assert code == 0
    T_g, st_g = _gpu_icp(ctx, ref, rn, q, knn=knn, epsilon=eps, max_dist=maxd)
    assert st_g.iterations == st_o.iterations
    assert_same_T(T_g, T_o)
    V = golden["T"]["validT3d"]
    assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(T_g[:3, 3])) < 0.1 and rot_angle(V[:3, :3], T_g[:3, :3]) < 0.1


def test_icp_open3d_slam_chain(ctx, oracle, golden):
    """param/icp.yaml chain: maxDist 0.5, TrimmedDist 0.9, SurfaceNormalOutlierFilter 1.57, Differential + Counter 15."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    qn = oracle.surface_normals(q, 10, nthreads=8, orient_origin=[0, 0, 0])["normals"]
    kw = dict(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15)
    code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8, **kw), ref, rn, q, qn)
    assert code == 0
    T_g, st_g = _gpu_icp(ctx, ref, rn, q, qn, **kw)
    assert st_g.iterations == st_o.iterations
    assert st_g.max_iterations_reached == st_o.max_iterations_reached
    assert_same_T(T_g, T_o)


def test_icp_synthetic_scan_to_map(ctx, oracle):
    """BASELINE config #2 at a size the oracle finishes in seconds."""
    c = scenes.config2(m=500_000, n_beams=64, n_az=512)
    ref, rn, q = oracle.feat4(c["map_xyz"]), c["map_nrm"].astype(np.float32), oracle.feat4(c["scan"])
    Ti = c["T_init"].astype(np.float32)
    kw = dict(knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40)
    code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8, **kw), ref, rn, q, None, Ti)
    assert code == 0
    T_g, st_g = _gpu_icp(ctx, ref, rn, q, None, Ti, **kw)
    assert st_g.iterations == st_o.iterations
    assert_same_T(T_g, T_o)
    E = np.linalg.inv(c["T_true"]) @ T_g.astype(np.float64)
    assert np.abs(E[:3, 3]).max() < 0.01     # converged to the true pose (1 cm noise)


def test_icp_singular_planar_grid(ctx, oracle):
    """utest/ui/icp/GeneralTests.cpp:152-188 (icpSingular): planar grid shifted 1 m in z -> pure z translation."""
    nX = 10
    d = 0.1
    xs = np.array([[d * x - nX * d / 2, d * y - nX * d / 2, 0.0] for x in range(nX) for y in range(nX)], dtype=np.float32)
    ref = xs.copy()
    ref[:, 2] = 1.0
    rn = np.tile(np.array([[0, 0, 1]], dtype=np.float32), (ref.shape[0], 1))
    kw = dict(knn=1, trim_ratio=1.0, max_iter=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    T_g, st = _gpu_icp(ctx, oracle.feat4(ref), rn, oracle.feat4(xs), **kw)
    expected = np.eye(4)
    expected[2, 3] = 1
    assert np.allclose(T_g, expected, atol=1e-5)
    code, T_o, _ = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, **kw), oracle.feat4(ref), rn, oracle.feat4(xs))
    assert code == 0 and np.allclose(T_o, expected, atol=1e-5)


def test_icp_identity_same_cloud(ctx, oracle, golden):
    """GeneralTests.cpp:190-210 (icpIdentity): same cloud twice -> identity within 1e-4 (NaN guard, PointToPlane.cpp:326-332)."""
    pts = oracle.feat4(golden["vtk"]["cloud0"])
    nrm = oracle.surface_normals(pts, 10, nthreads=8)["normals"]
    kw = dict(knn=1, trim_ratio=1.0, max_iter=40, smooth_length=4)
    T_g, st = _gpu_icp(ctx, pts, nrm, pts, **kw)
    assert np.allclose(T_g, np.eye(4), atol=1e-4)


def test_icp_errors(ctx, oracle, golden):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    icp = capi.ICP(ctx)
    with pytest.raises(capi.NoPoints):
        icp.set_reference(ctx.upload(np.zeros((0, 4), dtype=np.float32)))
    icp.set_reference(ctx.upload(ref, rn))
    with pytest.raises(capi.NoPoints):                       # ICP.cpp:357-359
        icp.register(ctx.upload(np.zeros((0, 4), dtype=np.float32)))
    bad = np.eye(4, dtype=np.float32) * 2
    bad[3, 3] = 1
    with pytest.raises(capi.TransformationError):            # TransformationsImpl.cpp:74
        icp.register(ctx.upload(q), bad)
    far = np.eye(4, dtype=np.float32)
    far[:3, 3] = 1000
    icp2 = capi.ICP(ctx, max_dist=0.5)
    icp2.set_reference(ctx.upload(ref, rn))
    with pytest.raises(capi.ConvergenceError):
        icp2.register(ctx.upload(q), far)
    with pytest.raises(capi.InvalidParameter):
        capi.ICP(ctx, knn=0)
    with pytest.raises(capi.InvalidParameter):
        capi.ICP(ctx, trim_ratio=1.5)


# ---------------------------------------------------------------------------------------------------------------
# device-resident loop, batch, streaming upload
# ---------------------------------------------------------------------------------------------------------------
def test_graph_loop_equals_host_loop(ctx, oracle, golden):
    """The CUDA-graph while loop (match, select, reduce, tail on the device) and the host-driven loop share their
    arithmetic (__host__ __device__, no FMA contraction): same iteration count, transforms within one fp32 ulp-scale
    step (device sin/cos vs glibc), both within tolerance of the oracle."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    kw = dict(knn=1, max_dist=1.0, trim_ratio=0.85, max_iter=40)
    Ti = small_T(3, t=0.1, deg=2.0)
    icp = capi.ICP(ctx, **kw)
    icp.set_reference(ctx.upload(ref, rn))
    cq = ctx.upload(q)
    T_graph, st_graph = icp.register(cq, Ti)
    ctx.set_profiling(True)          # profiling brackets force the host-driven loop
    T_host, st_host = icp.register(cq, Ti)
    ms = ctx.profile_read()
    ctx.set_profiling(False)
    assert len(ms) == st_host.iterations and (ms > 0).all()
    assert st_graph.iterations == st_host.iterations
    assert st_graph.kept_last == st_host.kept_last and st_graph.last_tau == st_host.last_tau
    assert st_graph.points_touched > 0 and st_host.points_touched > 0   # (the count depends on the lanes-per-query decomposition)
    assert np.abs(T_graph.astype(np.float64) - T_host).max() <= 2e-6
    code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8, **kw), ref, rn, q, None, Ti)
    assert code == 0 and st_o.iterations == st_graph.iterations
    assert_same_T(T_graph, T_o)
    assert_same_T(T_host, T_o)
    icp.close()


def test_graph_loop_trace_and_reuse_across_sizes(ctx, oracle, golden):
    """One instantiated graph serves every reading size / reference; the device trace shows the body's kernels per iteration."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    kw = dict(knn=1, max_dist=1.0, trim_ratio=0.85, max_iter=40)
    icp = capi.ICP(ctx, **kw)
    icp.set_reference(ctx.upload(ref, rn))
    ctx.set_trace(True)
    for n in (q.shape[0], 5000, 777, q.shape[0]):
        T_g, st = icp.register(ctx.upload(q[:n]))
        tr = ctx.trace_read()
        tags = [t for t, _ in tr]
        assert tags.count(1) == st.iterations and tags.count(5) == st.iterations and tags.count(7) == st.iterations
        code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8, **kw), ref, rn, q[:n])
        assert code == 0 and st_o.iterations == st.iterations
        assert_same_T(T_g, T_o)
    ctx.set_trace(False)
    icp.set_reference(ctx.upload(ref[::2], rn[::2]))     # new reference, same graph
    T_g, st = icp.register(ctx.upload(q))
    code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8, **kw), ref[::2], rn[::2], q)
    assert code == 0 and st_o.iterations == st.iterations
    assert_same_T(T_g, T_o)
    icp.close()


def test_register_batch_matches_individual(ctx, oracle):
    """b200reg_icp_register_batch (BASELINE config #4 shape, small): every pair equals its stand-alone registration and the
    oracle; a failing pair reports its status and keeps T_init without stopping the batch."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    kw = dict(knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40)
    cfg = capi.icp_config(**kw)
    refs, reads, Tis, expect = [], [], [], []
    for i in range(3):
        c = scenes.config2(m=60_000, n_beams=16, n_az=256, scene_seed=3000 + i)
        ref, rn, q = oracle.feat4(c["map_xyz"]), c["map_nrm"].astype(np.float32), oracle.feat4(c["scan"])
        refs.append(ctx.upload(ref, rn))
        reads.append(ctx.upload(q))
        Tis.append(c["T_init"].astype(np.float32))
        code, T_o, st_o = oracle.icp(oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=8, **kw), ref, rn, q, None, Tis[-1])
        assert code == 0
        expect.append((T_o, st_o.iterations))
    far = np.eye(4, dtype=np.float32)
    far[:3, 3] = 1000.0
    refs.append(refs[0]); reads.append(reads[0]); Tis.append(far)
    T, stats, status = capi.register_batch(ctx, cfg, refs, reads, np.stack(Tis))
    assert list(status[:3]) == [0, 0, 0] and status[3] == capi.E_NOT_CONVERGED
    assert np.array_equal(T[3], far)
    for i in range(3):
        assert stats[i].iterations == expect[i][1]
        assert_same_T(T[i], expect[i][0])
        icp = capi.ICP(ctx, **kw)
        icp.set_reference(refs[i])
        T_one, _ = icp.register(reads[i], Tis[i])
        assert np.array_equal(T_one, T[i])
        icp.close()


def test_cloud_update_streaming(ctx, oracle, golden):
    """b200reg_cloud_update_f32: re-fill (grow, shrink, async) == a fresh upload."""
    car = golden["car"]
    q = oracle.feat4(car["data_xyz"])
    c = ctx.upload(q[:100])
    for n, wait in ((5000, True), (50, True), (q.shape[0], False), (1234, False)):
        c.update(q[:n], wait=wait)
        f, _ = c.download()
        assert c.size == n and np.array_equal(f, q[:n])
    nrm = car["ref_normals"][:1234]
    c.update(q[:1234, :3], normals=nrm)        # stride 3 + normals through the pack kernel
    f, nn = c.download(normals=True)
    assert np.array_equal(f[:, :3], q[:1234, :3]) and np.array_equal(nn, nrm)
