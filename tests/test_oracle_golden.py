"""CPU tests (-m "not gpu"): pin the oracle against the reference's own fixtures and known answers, and
cross-check its pieces against independent numpy / scipy computations.  The oracle is the checker for the GPU tests, so it
has to be right first."""
import numpy as np
import pytest
from scipy.spatial import cKDTree

import scenes


def rot_angle(Ra, Rb):
    R = Ra.T.astype(np.float64) @ Rb.astype(np.float64)
    sk = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]]) / 2     # robust near identity (arccos is not)
    return float(np.arctan2(np.linalg.norm(sk), (np.trace(R) - 1) / 2))


# ---- reference fixture: car_cloud400 <- car_cloud401, validT3d (utest/utest.cpp:74-89, tolerance utest.h:65-86) ----------
@pytest.mark.parametrize("precision", [0, 1])
def test_car_clouds_reach_validT3d(oracle, golden, precision):
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"], oracle.feat4(car["data_xyz"])
    assert ref.shape[0] == 24989 and q.shape[0] == 25193
    code, T, st = oracle.icp(oracle.icp_config(precision=precision, nthreads=4), ref, rn, q)
    assert code == 0 and 3 <= st.iterations < 40 and not st.max_iterations_reached
    V = golden["T"]["validT3d"]
    assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(T[:3, 3])) < 0.1      # utest.h:83
    assert rot_angle(V[:3, :3], T[:3, :3]) < 0.1                               # utest.h:84


@pytest.mark.parametrize("knn,eps,maxd", [(1, 0.0, 1.0), (2, 0.2, 1.0), (3, 0.0, 0.5)])
def test_matcher_grid_reaches_validT3d(oracle, golden, knn, eps, maxd):
    """utest/ui/Matcher.cpp:68-92"""
    car = golden["car"]
    code, T, st = oracle.icp(oracle.icp_config(knn=knn, epsilon=eps, max_dist=maxd, nthreads=4), oracle.feat4(car["ref_xyz"]),
                             car["ref_normals"], oracle.feat4(car["data_xyz"]))
    V = golden["T"]["validT3d"]
    assert code == 0
    assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(T[:3, 3])) < 0.1 and rot_angle(V[:3, :3], T[:3, :3]) < 0.1


def test_trimmed_outlier_filter_ratio_085(oracle, golden):
    """utest/ui/Outliers.cpp:106-111"""
    car = golden["car"]
    code, T, st = oracle.icp(oracle.icp_config(trim_ratio=0.85, nthreads=4), oracle.feat4(car["ref_xyz"]), car["ref_normals"],
                             oracle.feat4(car["data_xyz"]))
    assert code == 0 and abs(st.weighted_point_used_ratio - 0.85) < 1e-3


def test_golden_ref_trans_point_to_plane_chain(oracle, golden):
    """GeneralTests.cpp:70-150 acceptance (median relative point error < 5 %) for the on-path golden config
    defaultPointToPlaneMinDistDataPointsFilter (KDTree k1 + TrimmedDist 0.75 + PointToPlane + Counter 40 + Differential
    1e-3/1e-2/4).  Its two DataPointsFilters are out of scope: MinDist 1.0 is applied here in numpy, and the reference
    normals come from the oracle's SurfaceNormal (knn 10) instead of SamplingSurfaceNormal — allowed by the 5 % bar."""
    c0, c1 = golden["vtk"]["cloud0"], golden["vtk"]["cloud1"]
    ref = oracle.feat4(c0)
    rn = oracle.surface_normals(ref, 10, nthreads=4)["normals"]
    reading = c1[np.linalg.norm(c1, axis=1) > 1.0]                             # MinDistDataPointsFilter minDist 1
    cfg = oracle.icp_config(trim_ratio=0.75, smooth_length=4, max_iter=40, nthreads=4)
    code, T, st = oracle.icp(cfg, ref, rn, oracle.feat4(reading))
    assert code == 0
    G = golden["T"]["defaultPointToPlaneMinDistDataPointsFilter"]
    P = np.concatenate([c1.astype(np.float64), np.ones((c1.shape[0], 1))], axis=1).T
    err = np.linalg.norm((T.astype(np.float64) @ P - G @ P)[:3], axis=0)
    mag = np.linalg.norm((T.astype(np.float64) @ P)[:3], axis=0)
    assert np.median(err) / np.median(mag) < 0.05                              # GeneralTests.cpp:136-148


def test_icp_singular_and_identity(oracle, golden):
    """GeneralTests.cpp:152-210"""
    nX, d = 10, 0.1
    xs = np.array([[d * x - nX * d / 2, d * y - nX * d / 2, 0.0] for x in range(nX) for y in range(nX)], dtype=np.float32)
    ref = xs.copy()
    ref[:, 2] = 1.0
    rn = np.tile(np.array([[0, 0, 1]], dtype=np.float32), (100, 1))
    expected = np.eye(4)
    expected[2, 3] = 1
    for prec in (0, 1):
        code, T, _ = oracle.icp(oracle.icp_config(precision=prec, trim_ratio=1.0, smooth_length=4), oracle.feat4(ref), rn, oracle.feat4(xs))
        assert code == 0 and np.allclose(T, expected, atol=1e-5)
    pts = oracle.feat4(golden["vtk"]["cloud0"])
    nrm = oracle.surface_normals(pts, 10, nthreads=4)["normals"]
    code, T, _ = oracle.icp(oracle.icp_config(precision=0, trim_ratio=1.0, smooth_length=4, nthreads=4), pts, nrm, pts)
    assert code == 0 and np.allclose(T, np.eye(4), atol=1e-4)


@pytest.mark.parametrize("shift", [(0, 0, 0), (100.0, -30.0, 35.0)])
def test_conditioning_mirror_matcher(oracle, shift):
    """utest/ui/icp/Conditioning.cpp:354-372 shape: icp_point_to_plane.yaml (MirrorMatcher, no filters, PointToPlane,
    Differential 1e-5/1e-4/3, Counter 30) on a 10 000-point 1x3x5 box; same cloud displaced -> error < 1e-5 relative scale."""
    rng = np.random.default_rng(4)
    # box surface samples with analytic normals
    n = 10_000
    face = rng.integers(0, 6, n)
    u, v = rng.random(n), rng.random(n)
    dims = np.array([1.0, 3.0, 5.0])
    pts = np.zeros((n, 3)); nrm = np.zeros((n, 3))
    for f in range(6):
        ax = f // 2
        o = [a for a in range(3) if a != ax]
        m = face == f
        pts[m, ax] = dims[ax] * (f % 2) - dims[ax] / 2
        pts[m, o[0]] = (u[m] - 0.5) * dims[o[0]]
        pts[m, o[1]] = (v[m] - 0.5) * dims[o[1]]
        nrm[m, ax] = 1.0 if f % 2 else -1.0
    pts += np.array(shift)
    T_true = scenes.rigid(scenes.rot_axis_angle([0.3, -1, 0.5], 0.05), [0.05, -0.02, 0.03])
    reading = (np.linalg.inv(T_true) @ np.concatenate([pts, np.ones((n, 1))], axis=1).T).T[:, :3]
    cfg = oracle.icp_config(matcher_kind=1, trim_ratio=-1.0, max_iter=30, min_diff_rot=1e-5, min_diff_trans=1e-4, smooth_length=3, precision=1)
    code, T, st = oracle.icp(cfg, oracle.feat4(pts), nrm.astype(np.float32), oracle.feat4(reading))
    assert code == 0
    E = np.linalg.inv(T_true) @ T.astype(np.float64)
    tol = 2e-5 if shift == (0, 0, 0) else 2e-4      # fp32 coordinates: ulp grows with the 100 m offset
    assert np.abs(E[:3, 3]).max() < tol and rot_angle(np.eye(3), E[:3, :3]) < tol


# ---- known answers / independent cross-checks ------------------------------------------------------------------------
def test_quantile_rank_is_float_multiply_truncation(oracle):
    """Matches.cpp:61-87: rank = size_t(float(n) * ratio).  KAT values from utest/ui/Outliers.cpp:126-151 (dists 4,5,5,5,5)."""
    d = np.array([4, 5, 5, 5, 5], dtype=np.float32)
    st, tau, w = oracle.trimmed_dist(d, 0.1)
    assert st == 0 and tau == 4.0 and w.tolist() == [1, 0, 0, 0, 0]
    st, tau, w = oracle.trimmed_dist(d, 0.5)
    assert tau == 5.0 and w.sum() == 5
    n = 131072
    rng = np.random.default_rng(0)
    x = rng.random(n).astype(np.float32)
    st, tau, w = oracle.trimmed_dist(x, 0.8)
    rank = int(np.float32(n) * np.float32(0.8))
    assert rank == 104857
    assert tau == np.sort(x)[rank]
    x2 = x.copy()
    x2[:10] = np.inf
    st, tau2, w2 = oracle.trimmed_dist(x2, 1.0)
    assert tau2 == x2[np.isfinite(x2)].max() and w2[:10].sum() == 0
    st, _, _ = oracle.trimmed_dist(np.full(4, np.inf, dtype=np.float32), 0.5)
    assert st == -1      # ConvergenceError "No matches available"


def test_kdtree_vs_brute_and_scipy(oracle):
    rng = np.random.default_rng(3)
    ref = oracle.feat4(rng.normal(size=(20000, 3)).astype(np.float32))
    q = oracle.feat4(rng.normal(size=(3000, 3)).astype(np.float32))
    tree = oracle.KdTree(ref)
    for k, md in [(1, np.inf), (5, 0.3), (12, np.inf)]:
        ids, d2, visits = tree.knn(q, k=k, max_dist=md, nthreads=4)
        ids_b, d2_b = oracle.knn_brute(ref, q, k=k, max_dist=md, nthreads=4)
        assert np.array_equal(ids, ids_b) and np.array_equal(d2.view(np.uint32), d2_b.view(np.uint32))
        assert visits < 0.2 * ref.shape[0] * q.shape[0]
    # fp64 scipy agrees except on fp32 rounding ties
    ids, d2, _ = tree.knn(q, k=1)
    dd, ii = cKDTree(ref[:, :3].astype(np.float64)).query(q[:, :3].astype(np.float64), k=1)
    differ = ids[:, 0] != ii
    assert differ.mean() < 1e-3
    assert np.allclose(np.sqrt(d2[:, 0]), dd, rtol=1e-5, atol=1e-6)
    # padding
    ids, d2, _ = tree.knn(q[:5] + 100, k=3, max_dist=1.0)
    assert (ids == -1).all() and np.isinf(d2).all()


def test_point_to_plane_system_vs_numpy(oracle):
    rng = np.random.default_rng(8)
    n = 4000
    ref = rng.normal(size=(n, 3)) * 3
    nrm = rng.normal(size=(n, 3)); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    T = scenes.rigid(scenes.rot_axis_angle([1, 2, 3], 0.01), [0.01, 0.02, -0.01])
    reading = (np.linalg.inv(T) @ np.concatenate([ref, np.ones((n, 1))], 1).T).T[:, :3].astype(np.float32)
    ref32, nrm32 = ref.astype(np.float32), nrm.astype(np.float32)
    ids = np.arange(n, dtype=np.int32).reshape(n, 1)
    d2 = np.zeros((n, 1), dtype=np.float32)
    w = np.ones((n, 1), dtype=np.float32)
    w[::7] = 0
    d2[5::11] = np.inf
    o = oracle.point_to_plane(oracle.feat4(reading), oracle.feat4(ref32), nrm32, ids, d2, w, precision=1)
    keep = (w[:, 0] != 0) & np.isfinite(d2[:, 0])
    assert o["kept"] == keep.sum()
    p = reading[keep].astype(np.float64); q = ref32[keep].astype(np.float64); nn = nrm32[keep].astype(np.float64)
    p = p - p.mean(0); qc = q - q.mean(0)
    G = np.concatenate([np.cross(p, nn), nn], axis=1).T
    h = np.sum((p - qc) * nn, axis=1)
    A = G @ G.T
    b = -G @ h
    assert np.allclose(o["A"], A, rtol=1e-10, atol=1e-9) and np.allclose(o["b"], b, rtol=1e-9, atol=1e-9)
    assert np.allclose(o["x"], np.linalg.solve(A, b), rtol=1e-8)
    E = np.linalg.inv(T) @ o["T"].astype(np.float64)
    assert np.abs(E - np.eye(4)).max() < 5e-4            # one linearised step from 0.01 rad: second-order residual
    o32 = oracle.point_to_plane(oracle.feat4(reading), oracle.feat4(ref32), nrm32, ids, d2, w, precision=0)
    assert np.abs(o32["T"] - o["T"]).max() < 1e-5
    # force4DOF: rows/cols {gamma, tx, ty, tz}
    o4 = oracle.point_to_plane(oracle.feat4(reading), oracle.feat4(ref32), nrm32, ids, d2, w, force4DOF=True, precision=1)
    assert np.allclose(o4["A"], A[2:, 2:], rtol=1e-10, atol=1e-9)
    assert np.allclose(o4["T"][2, :3], [0, 0, 1], atol=1e-7)


def test_voxelgrid_membership_vs_numpy(oracle):
    rng = np.random.default_rng(2)
    pts = (rng.random((20000, 3)) * 8 - 4).astype(np.float32)
    v = np.float32(0.37)
    m, f, _, cells = oracle.voxelgrid(oracle.feat4(pts), None, v, v, v)
    minB = pts.min(0) / v
    ijk = np.floor(pts / v - minB).astype(np.int64)
    nd = (np.float32(1) + pts.max(0) / v - minB).astype(np.int64)
    lin = ijk[:, 0] + ijk[:, 1] * nd[0] + ijk[:, 2] * nd[0] * nd[1]
    assert np.array_equal(lin, cells.astype(np.int64))
    uniq, first = np.unique(lin, return_index=True)
    assert m == uniq.size
    order = np.sort(first)
    # centroids: fp32 accumulation; compare with fp64 means loosely and with order of first appearance exactly
    for o_idx in (0, m // 2, m - 1):
        members = np.nonzero(lin == lin[order[o_idx]])[0]
        assert np.allclose(f[o_idx, :3], pts[members].astype(np.float64).mean(0), atol=1e-5)


def test_surface_normals_plane_and_descriptors(oracle):
    rng = np.random.default_rng(6)
    n = 3000
    xy = rng.random((n, 2)) * 4
    pts = np.stack([xy[:, 0], xy[:, 1], 0.3 * xy[:, 0] + 0.001 * rng.normal(size=n)], axis=1).astype(np.float32)
    o = oracle.surface_normals(oracle.feat4(pts), 12, nthreads=4, keep=("normals", "densities", "eig_values", "eig_vectors", "matched_ids", "mean_dists", "lin_plan_curv"))
    expect = np.array([-0.3, 0, 1.0]) / np.linalg.norm([-0.3, 0, 1.0])
    cosang = np.abs(o["normals"] @ expect)
    assert np.median(cosang) > 0.9999
    assert (np.diff(o["eig_values"], axis=1) >= 0).all()                       # ascending (our documented order)
    assert np.array_equal(o["matched_ids"][:, 0], np.arange(n, dtype=np.float32))   # self is the first neighbour
    assert (o["densities"] > 0).all() and o["degenerate"] == 0
    assert (o["lin_plan_curv"][:, 1] > 0.5).mean() > 0.9                       # planar


def test_open3d_slam_ops_vs_numpy(oracle):
    rng = np.random.default_rng(12)
    xyz = rng.normal(size=(50000, 3)) * 4
    nrm = rng.normal(size=(50000, 3))
    c = np.array([0.5, -0.3, 0.2])
    d = np.linalg.norm(xyz - c, axis=1)
    for kind, kw, expect in [(0, dict(rmax=3.0), d <= 3.0), (1, dict(rmin=3.0), d >= 3.0), (2, dict(rmin=1.0, rmax=3.0), (d <= 3.0) & (d >= 1.0)),
                             (3, dict(rmax=2.0, zmin=-1.0, zmax=1.0), (xyz[:, 2] >= -1) & (xyz[:, 2] <= 1) & (np.linalg.norm((xyz - c)[:, :2], axis=1) <= 2.0))]:
        xo, no, mask = oracle.crop(oracle.crop_spec(kind, c, **kw), xyz, nrm)
        assert np.array_equal(mask, expect) and np.array_equal(xo, xyz[expect]) and np.array_equal(no, nrm[expect])
        xo2, _, mask2 = oracle.crop(oracle.crop_spec(kind, c, invert=True, **kw), xyz, nrm)
        assert np.array_equal(mask2, ~expect)
    # voxelize within volume: membership + means vs numpy
    spec = oracle.crop_spec(0, c, rmax=3.0)
    xo, no, npass = oracle.voxelize_within_volume(0.5, spec, xyz, nrm)
    inside = d <= 3.0
    assert npass == (~inside).sum() and np.array_equal(xo[:npass], xyz[~inside])
    key = np.floor(xyz[inside] * (1.0 / 0.5)).astype(np.int64)
    uniq, inv = np.unique(key, axis=0, return_inverse=True)
    assert xo.shape[0] == npass + uniq.shape[0]
    means = np.zeros((uniq.shape[0], 3)); np.add.at(means, inv.ravel(), xyz[inside]); means /= np.bincount(inv.ravel())[:, None]
    assert np.allclose(xo[npass:], means, atol=1e-12)                          # np.unique(axis=0) sorts lexicographically like our key order
    assert np.allclose(np.linalg.norm(no[npass:], axis=1), 1.0)
    # Open3D VoxelDownSample convention: origin = min_bound - voxel/2
    xd, nd = oracle.voxel_down_sample(0.5, xyz, nrm)
    key = np.floor((xyz - (xyz.min(0) - 0.25)) / 0.5).astype(np.int64)
    assert xd.shape[0] == np.unique(key, axis=0).shape[0]
    f, n32 = oracle.o3d_to_pm(xyz, nrm)
    assert np.array_equal(f[:, :3], xyz.astype(np.float32)) and (f[:, 3] == 1).all() and np.array_equal(n32, nrm.astype(np.float32))


def test_transform_and_determinant_check(oracle):
    rng = np.random.default_rng(1)
    pts = oracle.feat4(rng.normal(size=(1000, 3)).astype(np.float32))
    T = scenes.rigid(scenes.rot_axis_angle([0, 1, 1], 0.4), [1, 2, 3]).astype(np.float32)
    st, out, _ = oracle.transform(T, pts)
    assert st == 0 and np.allclose(out[:, :3], pts[:, :3] @ T[:3, :3].T + T[:3, 3], atol=1e-5) and (out[:, 3] == 1).all()
    bad = T.copy()
    bad[:3, :3] *= 1.01
    st, _, _ = oracle.transform(bad, pts)
    assert st == -2      # TransformationError (TransformationsImpl.cpp:74)


def test_space_carving_known_answers(oracle):
    """orc_carve against hand-derived answers for getIdxsOfCarvedPoints (helpers.cpp:245-281): a wall the rays end on (kept: rays
    stop truncation_distance short of it), a ghost plane in free space facing the sensor (removed), a ghost plane seen edge-on
    (|ray . normal| <= minDot: kept), points outside the carving volume (never touched), and a map without normals (everything
    a ray passes through goes)."""
    O = oracle
    yy, zz = np.meshgrid(np.arange(-1.0, 1.0001, 0.05), np.arange(-1.0, 1.0001, 0.05), indexing="ij")
    plane = lambda x: np.stack([np.full(yy.size, x), yy.ravel(), zz.ravel()], axis=1)  # noqa: E731
    wall, ghost, edge, outside = plane(5.0), plane(2.03), plane(3.03), plane(2.53) + np.array([0.0, 30.0, 0.0])
    nrm = np.concatenate([np.tile([-1.0, 0, 0], (len(wall), 1)), np.tile([-1.0, 0, 0], (len(ghost), 1)), np.tile([0, 0, 1.0], (len(edge), 1)),
                          np.tile([-1.0, 0, 0], (len(outside), 1))])
    map_xyz = np.concatenate([wall, ghost, edge, outside])
    scan = np.concatenate([wall, plane(2.53) + np.array([0.0, 30.0, 0.0]) * 0 + np.array([2.47, 30.0, 0.0])])   # rays to the wall and toward `outside`
    sensor = np.zeros(3)
    vol = O.crop_spec(O.CROP_MAX_RADIUS, sensor, rmax=10.0)
    m = O.carve(scan, sensor, map_xyz, nrm, vol, voxel=0.1, max_ray_length=20.0, truncation=0.1, min_dot=0.5)
    a, b, c = len(wall), len(wall) + len(ghost), len(wall) + len(ghost) + len(edge)
    fan = lambda pts, x: (np.abs(pts[:, 1]) <= 0.85 * x / 5.0) & (np.abs(pts[:, 2]) <= 0.85 * x / 5.0)      # well inside the ray fan  # noqa: E731
    off = lambda pts, x: (np.abs(pts[:, 1]) >= 1.25 * x / 5.0) | (np.abs(pts[:, 2]) >= 1.25 * x / 5.0)      # clearly outside it  # noqa: E731
    assert not m[:a].any()                                   # the surface the rays end on survives
    assert m[a:b][fan(ghost, 2.0)].all() and not m[a:b][off(ghost, 2.1)].any()   # free-space ghost facing the sensor: carved where rays pass
    assert not m[b:c].any()                                  # edge-on surface: |dir . n| = |dir_z| <= 0.28 < 0.5
    assert not m[c:].any()                                   # outside the cropping volume (30 m away > rmax 10)
    m2 = O.carve(scan, sensor, map_xyz, None, vol, voxel=0.1)   # no normals: everything a ray passes through goes
    assert m2[a:b][fan(ghost, 2.0)].all() and m2[b:c][fan(edge, 3.0)].all() and not m2[:a].any() and not m2[c:].any()
    # ray length: maximalPathTraveled = max(voxel, min(length - truncation, maxRaytracingLength)) — a 1 m cap stops before the ghost
    assert not O.carve(scan, sensor, map_xyz, nrm, vol, voxel=0.1, max_ray_length=1.0).any()
    # brute-force restatement in numpy (fp64, same sampling) on a random case
    rng = np.random.default_rng(3)
    mp = rng.uniform(-4, 4, (3000, 3)); mn = rng.normal(size=(3000, 3))
    sc = rng.uniform(-6, 6, (200, 3)); s0 = np.array([0.3, -0.2, 0.1])
    vol2 = O.crop_spec(O.CROP_MAX_RADIUS, s0, rmax=5.0)
    got = O.carve(sc, s0, mp, mn, vol2, voxel=0.25, max_ray_length=4.0, truncation=0.2, min_dot=0.3)
    inside = np.linalg.norm(mp - s0, axis=1) <= 5.0
    keys = {}
    for i in np.nonzero(inside)[0]:
        keys.setdefault(tuple(np.floor(mp[i] * (1.0 / 0.25)).astype(int)), []).append(i)
    want = np.zeros(3000, dtype=bool)
    for p in sc:
        d = p - s0
        L = np.sqrt((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2])
        u = d / L
        dist, mx = 0.0, max(0.25, min(L - 0.2, 4.0))
        while dist < mx:
            for i in keys.get(tuple(np.floor((dist * u + s0) * (1.0 / 0.25)).astype(int)), []):
                n = mn[i] / np.sqrt((mn[i][0] * mn[i][0] + mn[i][1] * mn[i][1]) + mn[i][2] * mn[i][2])
                if abs((u[0] * n[0] + u[1] * n[1]) + u[2] * n[2]) > 0.3:
                    want[i] = True
            dist += 0.25
    assert np.array_equal(got, want) and want.sum() > 20


def test_distance_outlier_filters_known_answers(oracle):
    """MaxDist / MinDist / MedianDist (OutlierFiltersImpl.cpp:64-125) on squared distances, incl. the +inf padding of missing matches:
    inf <= limit is false, inf >= limit is true; the median is getDistsQuantile(0.5) = the element of rank uint(float(n_finite) * 0.5)."""
    d2 = np.array([0.04, 0.25, np.inf, 1.0, 0.0, 4.0, 9.0, 0.09], dtype=np.float32)
    st, lim, w = oracle.dist_outlier(0, d2, 0.5)
    assert st == 0 and lim == np.float32(0.25) and w.tolist() == [1, 1, 0, 0, 1, 0, 0, 1]          # the boundary value is kept (<=)
    st, lim, w = oracle.dist_outlier(1, d2, 0.5)
    assert st == 0 and w.tolist() == [0, 1, 1, 1, 0, 1, 1, 0]
    st, lim, w = oracle.dist_outlier(2, d2, 3.0)
    finite = np.sort(d2[np.isfinite(d2)])                                                           # 7 finite values, rank uint(7 * 0.5) = 3
    assert st == 0 and lim == np.float32(3.0) * finite[3] and w.tolist() == [1, 1, 0, 0, 1, 0, 0, 1]
    assert oracle.dist_outlier(2, np.full(4, np.inf, dtype=np.float32), 3.0)[0] != 0                # no finite match: ConvergenceError
    x = np.float32(0.3)                                                                             # pow(float, 2) in double, rounded once
    assert oracle.dist_outlier(0, d2, float(x))[1] == np.float32(np.float64(x) ** 2) == x * x


def _pc2_message(n, fields, point_step, seed, dtype_override=None):
    """A sensor_msgs/PointCloud2 data buffer: `fields` = {name: byte offset}; every listed field holds a FLOAT32; the rest of each
    point record is random bytes (intensity, ring, time ... of a real driver)."""
    rng = np.random.default_rng(seed)
    buf = rng.integers(0, 256, size=(n, point_step), dtype=np.uint8)
    vals = {}
    for name, off in fields.items():
        v = rng.normal(size=n).astype(np.float32) * 20
        if name == "x" and n > 3:
            v[3] = np.nan          # invalid returns stay in the cloud (is_dense = false)
        vals[name] = v
        buf[:, off:off + 4] = v.view(np.uint8).reshape(n, 4)
    return buf.reshape(-1), vals


@pytest.mark.parametrize("fields,step", [({"x": 0, "y": 4, "z": 8}, 12), ({"x": 0, "y": 4, "z": 8, "normal_x": 12, "normal_y": 16, "normal_z": 20}, 24),
                                         ({"x": 0, "y": 4, "z": 8, "normal_x": 16, "normal_y": 20, "normal_z": 24}, 48),
                                         ({"z": 1, "x": 9, "y": 5, "normal_x": 13, "normal_y": 17}, 23)])
def test_pointcloud2_deserialisation_layouts(oracle, fields, step):
    """orc_pointcloud2 against numpy on the layouts of RosPointCloud2DeserializerTest.cpp:195-399 (points; points + normals) plus an
    Ouster-like 48-byte record and a packed, unaligned one: value = bytes at i * point_step + offset as float32; features get
    pad = 1; normals only when all three fields exist; the fp64 side is the exact widening of the floats."""
    n = 257
    data, vals = _pc2_message(n, fields, step, 11)
    st, f4, n3, x64, n64 = oracle.pointcloud2(data, n, step, fields)
    assert st == 0
    want = np.stack([vals["x"], vals["y"], vals["z"]], axis=1)
    assert np.array_equal(f4[:, :3].view(np.uint32), want.view(np.uint32)) and (f4[:, 3] == 1).all()
    assert np.array_equal(x64, want.astype(np.float64), equal_nan=True)
    if "normal_z" in fields:
        wn = np.stack([vals["normal_x"], vals["normal_y"], vals["normal_z"]], axis=1)
        assert np.array_equal(n3, wn) and np.array_equal(n64, wn.astype(np.float64))
    else:
        assert n3 is None and n64 is None
    assert oracle.pointcloud2(data, n, step, {"x": 0, "y": 4})[0] != 0   # "Field z does not exist"


def test_loop_closure_support_ops_known_answers(oracle):
    """computeIndicesOfOverlappingPoints (helpers.cpp:320-345) and GetInformationMatrixFromPointClouds (Open3D v0.15.1) on cases with
    hand-derived answers."""
    O = oracle
    # two unit-spaced rows of points; voxel 1.0: target occupies voxels x = 0..4, source (after a +2.5 shift) voxels 2..6
    tgt = np.array([[i + 0.5, 0.5, 0.5] for i in range(5)] + [[0.6, 0.5, 0.5]])        # voxel 0 holds two target points
    src = np.array([[i + 0.5 - 2.5 + 2.0, 0.5, 0.5] for i in range(5)])                 # x = 0, 1, 2, 3, 4 before the shift
    T = np.eye(4); T[0, 3] = 2.5                                                          # -> 2.5 .. 6.5: voxels 2..6
    si, ti = O.overlap(src, tgt, T, 1.0, 1)
    assert si.tolist() == [0, 1, 2] and ti.tolist() == [2, 3, 4]                          # shared voxels 2, 3, 4
    si, ti = O.overlap(src, tgt, np.eye(4), 1.0, 2)                                       # min 2 points per voxel on BOTH sides: none
    assert si.size == 0 and ti.size == 0
    si, ti = O.overlap(np.concatenate([src, src[:1] + 0.1]), tgt, np.eye(4), 1.0, 2)      # source x = 0 and 0.1 share voxel 0 with 0.5, 0.6
    assert si.tolist() == [0, 5] and ti.tolist() == [0, 5]
    # information matrix: one correspondence at q = (1, 2, 3): G^T G with G = [-[q]x | I]
    info, cnt = O.information_matrix(np.array([[1.0, 2.0, 3.05]]), np.array([[1.0, 2.0, 3.0], [9.0, 9.0, 9.0]]), 0.1, np.eye(4))
    q = np.array([1.0, 2.0, 3.0])
    qx = np.array([[0, -q[2], q[1]], [q[2], 0, -q[0]], [-q[1], q[0], 0]])
    G = np.hstack([-qx, np.eye(3)])
    assert cnt == 1 and np.allclose(info, G.T @ G, rtol=0, atol=1e-12)
    assert O.information_matrix(np.array([[1.0, 2.0, 3.2]]), np.array([[1.0, 2.0, 3.0]]), 0.1, np.eye(4))[1] == 0   # beyond the radius
    # many points: equals the numpy restatement (correspondence = nearest target within the radius, d2 < r2)
    rng = np.random.default_rng(4)
    tg = rng.uniform(-5, 5, (400, 3)); sr = tg[rng.permutation(400)[:300]] + rng.normal(0, 0.02, (300, 3))
    Tm = np.eye(4); Tm[:3, 3] = [0.01, -0.02, 0.005]
    info, cnt = O.information_matrix(sr, tg, 0.05, Tm)
    moved = sr + Tm[:3, 3]
    d = np.linalg.norm(moved[:, None] - tg[None], axis=2)
    j = d.argmin(axis=1); ok = d.min(axis=1) < 0.05
    want = np.zeros((6, 6))
    for t in tg[j[ok]]:
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        Gm = np.hstack([-tx, np.eye(3)])
        want += Gm.T @ Gm
    assert cnt == int(ok.sum()) and cnt > 100 and np.allclose(info, want, rtol=1e-12)


def test_var_trimmed_dist_reference_known_answers(oracle):
    """utest/ui/Outliers.cpp:126-151 (VarTrimmedDistOutlierFilterParameters): dists 4 5 5 5 5, minRatio 1e-7, maxRatio 1."""
    O = oracle
    d2 = np.array([4, 5, 5, 5, 5], dtype=np.float32)
    st, ratio, tau, w = O.var_trimmed_dist(d2, 0.0000001, 1.0, 0.0)      # lambda 0: everything but the minimum is rejected
    assert st == 0 and w.tolist() == [1, 0, 0, 0, 0] and ratio == 0.0 and tau == 4.0
    st, ratio, tau, w = O.var_trimmed_dist(d2, 0.0000001, 1.0, 1.0)      # lambda 1: all kept
    assert st == 0 and w.tolist() == [1, 1, 1, 1, 1] and ratio == np.float32(0.8)
    assert O.var_trimmed_dist(d2, 0.8, 0.6, 1.0)[0] == O.ORC_E_PARAM                      # minRatio >= maxRatio (:160-163)
    assert O.var_trimmed_dist(np.array([np.inf, 0.0], np.float32), 0.1, 0.9, 1.0)[0] == O.ORC_E_CONVERGENCE   # no valid match (:193)
    # numpy restatement (float32 cumulative sum in order, FRMS, first argmin) on a contaminated distribution
    rng = np.random.default_rng(5)
    d2 = np.concatenate([rng.random(700) ** 2 * 0.01, rng.random(300) * 4 + 1]).astype(np.float32)
    rng.shuffle(d2)
    d2[::50] = np.inf
    d2[7] = 0.0
    prev = 0.0
    for lam in (0.5, 0.95, 2.0):
        st, ratio, tau, w = O.var_trimmed_dist(d2, 0.05, 0.99, lam)
        srt = np.sort(d2[np.isfinite(d2) & (d2 > 0)])
        cum = np.cumsum(srt, dtype=np.float32)
        n = d2.size
        lo, hi = int(np.floor(np.float32(0.05) * n)), min(int(np.floor(np.float32(0.99) * n)), srt.size)
        ids = np.arange(lo + 1, hi + 1, dtype=np.float32)
        deno = ((ids / np.float32(n)).astype(np.float64) ** lam).astype(np.float32)
        frms = (cum[lo:hi] * (np.float32(1) / ids)) * ((np.float32(1) / deno) * (np.float32(1) / deno))
        want = np.float32(int(frms.argmin()) + lo) / np.float32(n)
        assert st == 0 and ratio == want, (lam, ratio, want)
        assert ratio >= prev                                                              # a larger lambda rewards a larger inlier set
        prev = ratio
        st2, tau2, w2 = O.trimmed_dist(d2, ratio)
        assert tau == tau2 and np.array_equal(w, w2)


def test_point_to_point_minimizer_known_answers(oracle):
    """PointToPoint.cpp:55-94: exact pairs under a known rigid transform are recovered; a mirrored set yields a rotation (the
    reflection branch :80-85); weights and invalid matches act as in ErrorElements."""
    O = oracle
    rng = np.random.default_rng(6)
    p = rng.uniform(-5, 5, (200, 3))
    T = scenes.rigid(scenes.rot_axis_angle([0.3, -0.2, 1.0], np.deg2rad(25.0)), [0.7, -1.2, 0.4])
    q = p @ T[:3, :3].T + T[:3, 3]
    ids = np.arange(200, dtype=np.int32).reshape(-1, 1)
    d2 = np.ones((200, 1), np.float32)
    w = np.ones((200, 1), np.float32)
    for prec in (O.ORC_REF32, O.ORC_MIXED):
        st, To, kept = O.point_to_point(O.feat4(p), O.feat4(q), ids, d2, w, prec)
        assert st == 0 and kept == 200 and np.abs(To - T).max() < (2e-5 if prec == O.ORC_REF32 else 2e-6)
    # invalid matches and zero weights drop out: corrupt those pairs, the answer must not move
    q2 = q.copy(); q2[:20] += 3.0
    d2b = d2.copy(); d2b[:10] = np.inf
    wb = w.copy(); wb[10:20] = 0.0
    st, To, kept = O.point_to_point(O.feat4(p), O.feat4(q2), ids, d2b, wb)
    assert st == 0 and kept == 180 and np.abs(To - T).max() < 2e-6
    # fractional weights: equals the numpy weighted Kabsch
    wf = rng.random((200, 1)).astype(np.float32) + 0.1
    qn = q + rng.normal(0, 0.05, q.shape)
    st, To, _ = O.point_to_point(O.feat4(p), O.feat4(qn), ids, d2, wf)
    pf, qf = O.feat4(p)[:, :3].astype(np.float64), O.feat4(qn)[:, :3].astype(np.float64)
    ww = wf[:, 0].astype(np.float64)
    mp, mq = (pf * ww[:, None]).sum(0) / ww.sum(), (qf * ww[:, None]).sum(0) / ww.sum()
    U, _, Vt = np.linalg.svd(((qf - mq) * ww[:, None]).T @ (pf - mp))
    R = U @ Vt
    assert np.linalg.det(R) > 0 and np.abs(To[:3, :3] - R).max() < 2e-6 and np.abs(To[:3, 3] - (mq - R @ mp)).max() < 2e-6
    # a mirrored copy: U V^T is a reflection, the minimiser must still answer with a proper rotation
    qm = p * np.array([1.0, 1.0, -1.0])
    st, To, _ = O.point_to_point(O.feat4(p), O.feat4(qm), ids, d2, w)
    assert st == 0 and abs(np.linalg.det(To[:3, :3].astype(np.float64)) - 1.0) < 1e-5
    assert O.point_to_point(O.feat4(p), O.feat4(q), ids, d2, np.zeros_like(w))[0] == O.ORC_E_CONVERGENCE


def test_point_to_plane_covariance_equals_numpy_restatement(oracle):
    """PointToPlaneWithCov.cpp:70-160 written with numpy arrays, against the oracle's loop (fp64 lane), plus the no-normals answer."""
    O = oracle
    rng = np.random.default_rng(7)
    n = 500
    p = rng.uniform(-8, 8, (n, 3)); q = p + rng.normal(0, 0.02, (n, 3))
    nr = rng.normal(size=(n, 3)); nr /= np.linalg.norm(nr, axis=1, keepdims=True)
    ids = rng.permutation(n).astype(np.int32).reshape(-1, 1)
    d2 = np.ones((n, 1), np.float32); d2[::17] = np.inf
    w = np.ones((n, 1), np.float32); w[::13] = 0
    T = scenes.rigid(scenes.rot_axis_angle([0.2, 0.5, 1.0], np.deg2rad(0.8)), [0.01, -0.02, 0.015]).astype(np.float32)
    pf, qf, nf = O.feat4(p), O.feat4(q), nr.astype(np.float32)
    st, cov = O.point_to_plane_covariance(pf, qf, nf, ids, d2, w, T, 0.05)
    keep = np.isfinite(d2[:, 0]) & (w[:, 0] != 0)
    r = pf[keep, :3].astype(np.float64); s = qf[ids[keep, 0], :3].astype(np.float64); nn = nf[ids[keep, 0]].astype(np.float64)
    r -= r.mean(0); s -= s.mean(0)
    Td = T.astype(np.float64)
    beta = -np.arcsin(Td[2, 0]); alpha = np.arctan2(Td[2, 1], Td[2, 2]); gamma = np.arctan2(Td[1, 0] / np.cos(beta), Td[0, 0] / np.cos(beta))
    t = Td[:3, 3]
    rr = np.linalg.norm(r, axis=1); fr = np.linalg.norm(s, axis=1)
    rd = r / rr[:, None]; fd = s / fr[:, None]
    nabg = np.stack([nn[:, 2] * rd[:, 1] - nn[:, 1] * rd[:, 2], nn[:, 0] * rd[:, 2] - nn[:, 2] * rd[:, 0], nn[:, 1] * rd[:, 0] - nn[:, 0] * rd[:, 1]], 1)
    Rlin = np.array([[1, -gamma, beta], [gamma, 1, -alpha], [-beta, alpha, 1]])
    E = (nn * (r @ Rlin.T + t - s)).sum(1)
    Nr = (nn * (rd @ Rlin.T)).sum(1)
    Nf = -(nn * fd).sum(1)
    V = np.hstack([nn, rr[:, None] * nabg])
    A = np.hstack([nn * Nr[:, None], nabg * (E + rr * Nr)[:, None]])
    B = np.hstack([nn * Nf[:, None], nabg * (fr * Nf)[:, None]])
    Hi = np.linalg.inv(V.T @ V)
    want = float(np.float32(0.05)) ** 2 * Hi @ (A.T @ A + B.T @ B) @ Hi                           # sensorStdDev is a float parameter
    assert st == 0 and np.allclose(cov, want, rtol=1e-8, atol=1e-18)
    assert np.allclose(cov, cov.T, rtol=1e-9, atol=1e-20) and (np.linalg.eigvalsh(cov) > -1e-18).all()
    st, cov32 = O.point_to_plane_covariance(pf, qf, nf, ids, d2, w, T, 0.05, O.ORC_REF32)        # the reference's fp32 lane
    assert st == 0 and np.allclose(cov32, want, rtol=5e-2, atol=1e-12)
    st, none = O.point_to_plane_covariance(pf, qf, None, ids, d2, w, T, 0.05)                    # :91-92
    assert st == 0 and np.array_equal(none, np.eye(6) * np.finfo(np.float32).max)


def test_compacting_filters_known_answers(oracle):
    """MaxDist / MinDist / BoundingBox / RandomSampling DataPointsFilters (DataPointsFilters/*.cpp) on small clouds."""
    O = oracle
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3], [1, 1, 1], [-2, -2, -2], [3, 4, 0]], dtype=np.float32)
    f = O.feat4(pts)
    assert O.cloud_filter(f, O.FILTER_MAX_DIST, dim=-1, dist=-2.0).tolist() == [True, True, False, False, True, False, False]   # |maxDist|, strict
    assert O.cloud_filter(f, O.FILTER_MAX_DIST, dim=1, dist=2.0).tolist() == [True, True, False, True, True, True, False]
    assert O.cloud_filter(f, O.FILTER_MIN_DIST, dim=-1, dist=2.0).tolist() == [False, False, False, True, False, True, True]
    assert O.cloud_filter(f, O.FILTER_MIN_DIST, dim=2, dist=-1.0).tolist() == [True, True, True, True, True, False, True]
    box = (-0.5, 1.5, -0.5, 1.5, -0.5, 1.5)
    assert O.cloud_filter(f, O.FILTER_BOUNDING_BOX, box=box, remove_inside=True).tolist() == [False, False, True, True, False, True, True]
    assert O.cloud_filter(f, O.FILTER_BOUNDING_BOX, box=box, remove_inside=False).tolist() == [True, True, False, False, True, False, False]
    assert O.cloud_filter(f, O.FILTER_BOUNDING_BOX, box=(0, 1, -1, 1, -1, 1), remove_inside=False).tolist() == [False] * 7     # strict bounds
    # RandomSampling: minstd_rand(seed) draws against prob, at most floor(n * prob) + 1 points, deterministic per seed
    big = O.feat4(np.random.default_rng(8).uniform(-1, 1, (20000, 3)))
    for method in (0, 1):
        k1 = O.cloud_filter(big, O.FILTER_RANDOM_SAMPLING, prob=0.3, seed=42, random_method=method)
        k2 = O.cloud_filter(big, O.FILTER_RANDOM_SAMPLING, prob=0.3, seed=42, random_method=method)
        k3 = O.cloud_filter(big, O.FILTER_RANDOM_SAMPLING, prob=0.3, seed=43, random_method=method)
        assert np.array_equal(k1, k2) and not np.array_equal(k1, k3)
        assert abs(k1.mean() - 0.3) < 0.015 and k1.sum() <= int(20000 * 0.3) + 1
    # the first draws of minstd_rand(1): 48271, 182605794 -> / 2^31
    k = O.cloud_filter(big[:2], O.FILTER_RANDOM_SAMPLING, prob=0.05, seed=1)
    assert k.tolist() == [True, False]     # 48271 / 2^31 = 2.2e-5 < 0.05 ; 182605794 / 2^31 = 0.085 > 0.05
    assert O.cloud_filter(big, O.FILTER_RANDOM_SAMPLING, prob=1.0, seed=3).all() and not O.cloud_filter(big, O.FILTER_RANDOM_SAMPLING, prob=0.0, seed=3).any()


def test_pointcloud2_reference_test_layouts_oracle(oracle):
    """The oracle's deserialiser against the message layouts of the reference's RosPointCloud2DeserializerTest.cpp:195-430 (written by
    pointMatcherCloudToRosMsg<float>, restated in pc2_reference_layouts.py): deserialize(serialize(cloud)) == cloud."""
    import pc2_reference_layouts as R
    for k, case in enumerate(R.CASES):
        if case[1] > 200_000:
            case = (case[0], 50_000) + case[2:]        # the million-point cases run at full size in the GPU suite
        data, fields, step, wh, feat, nrm = R.build(case, 700 + k)
        st, f4, n3, x64, n64 = oracle.pointcloud2(data, feat.shape[0], step, fields)
        assert st == 0
        assert np.array_equal(f4.view(np.uint32), feat.view(np.uint32)), case[0]
        assert (n3 is None) == (nrm is None) and (nrm is None or np.array_equal(n3, nrm)), case[0]
        assert np.array_equal(x64, feat[:, :3].astype(np.float64))


def test_residual_error_equals_numpy_restatement(oracle):
    """PointToPlaneErrorMinimizer::computeResidualError (PointToPlane.cpp:372-407) over the ErrorElements pairs (finite distance, non-zero
    weight, ErrorMinimizer.cpp:103-108): sum of w ((p - q) . n)^2, z dropped under force2D — against a direct numpy evaluation."""
    rng = np.random.default_rng(5)
    n, m, k = 400, 300, 3
    q = np.concatenate([rng.normal(size=(n, 3)), np.ones((n, 1))], axis=1).astype(np.float32)
    ref = np.concatenate([rng.normal(size=(m, 3)), np.ones((m, 1))], axis=1).astype(np.float32)
    nrm = rng.normal(size=(m, 3))
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    ids = rng.integers(0, m, size=(n, k)).astype(np.int32)
    d2 = rng.uniform(0, 1, size=(n, k)).astype(np.float32)
    miss = rng.random((n, k)) < 0.2
    d2[miss] = np.inf
    ids[miss] = -1
    w = rng.choice(np.array([0.0, 0.5, 1.0], np.float32), size=(n, k))
    for force2D in (False, True):
        dims = 2 if force2D else 3
        keep = np.isfinite(d2) & (w != 0)
        p = np.repeat(q[:, None, :3].astype(np.float64), k, axis=1)[keep]
        r = ref[ids[keep], :3].astype(np.float64)
        nn = nrm[ids[keep]].astype(np.float64)
        want = float((w[keep].astype(np.float64) * ((p - r)[:, :dims] * nn[:, :dims]).sum(1) ** 2).sum())
        got = oracle.point_to_plane_residual(q, ref, nrm, ids, d2, w, force2D=force2D)
        assert abs(got - want) <= 1e-12 * max(1.0, want), (force2D, got, want)
        got32 = oracle.point_to_plane_residual(q, ref, nrm, ids, d2, w, force2D=force2D, precision=oracle.ORC_REF32)
        assert abs(got32 - want) <= 1e-4 * max(1.0, want)          # the reference's own float accumulation


def test_smooth_normals_equals_the_sequential_sweep(oracle):
    """smoothNormals (SurfaceNormal.cpp:299-323): an IN-PLACE sweep in index order — point i averages the (already smoothed, for j < i)
    normals of its matched neighbours, each flipped onto its own current normal — restated literally in float32 Python."""
    rng = np.random.default_rng(9)
    n, knn, md = 250, 6, 0.9
    pts = np.concatenate([rng.uniform(-1, 1, size=(n, 2)), 0.05 * rng.normal(size=(n, 1)), np.ones((n, 1))], axis=1).astype(np.float32)
    nrm = rng.normal(size=(n, 3)) * np.array([0.3, 0.3, 1.0])
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    ids, d2, _ = oracle.KdTree(pts).knn(pts, k=knn, max_dist=md)
    want = nrm.copy()
    for i in range(n):
        cur = want[i].copy()
        mean = np.zeros(3, np.float32)
        cnt = 0
        for j in range(knn):
            if d2[i, j] == np.inf:
                continue
            v = want[ids[i, j]]
            dot = np.float32(np.float32(np.float32(cur[0] * v[0]) + np.float32(cur[1] * v[1])) + np.float32(cur[2] * v[2]))
            mean = (mean + v if dot > 0 else mean - v).astype(np.float32)
            cnt += 1
        want[i] = mean / np.float32(cnt)
    got = oracle.smooth_normals(pts, nrm, knn, max_dist=md)
    assert np.abs(got - want).max() <= 2e-7, np.abs(got - want).max()
    assert np.abs(got - nrm).max() > 1e-2                              # and it did something


def test_estimate_normals_f64_equals_numpy_pca(oracle):
    """Open3D EstimateNormals(KDTreeSearchParamHybrid(radius, knn)) + OrientNormalsTowardsCameraLocation in fp64: covariance of the (at most
    knn, within radius) nearest neighbours incl. the point itself, eigenvector of the smallest eigenvalue, flipped towards the origin —
    against numpy (brute-force neighbours, eigh); the candidate-list mode used above 20 000 points gives the same normals."""
    rng = np.random.default_rng(13)
    n, knn, radius = 1500, 12, 0.35
    uv = rng.uniform(-1, 1, size=(n, 2))
    xyz = np.stack([uv[:, 0], uv[:, 1], 0.3 * np.sin(2 * uv[:, 0]) + 0.002 * rng.normal(size=n)], axis=1) + np.array([0.0, 0.0, 3.0])
    origin = np.zeros(3)
    got, deg = oracle.estimate_normals_f64(xyz, knn, radius, origin=origin, exact=True)
    got_c, _ = oracle.estimate_normals_f64(xyz, knn, radius, origin=origin, exact=False)
    D = ((xyz[:, None, :] - xyz[None, :, :]) ** 2).sum(-1)
    worst = 0.0
    for i in range(0, n, 7):
        order = np.argsort(D[i], kind="stable")[:knn]
        nb = order[D[i, order] <= radius * radius]
        if nb.size < 3:
            continue
        P = xyz[nb]
        C = np.cov(P.T, bias=True)
        wv, V = np.linalg.eigh(C)
        v = V[:, 0]
        if np.dot(v, origin - xyz[i]) < 0:
            v = -v
        worst = max(worst, float(np.linalg.norm(got[i] - v)), float(np.linalg.norm(got_c[i] - v)))
    assert deg == 0 and worst < 1e-7, (deg, worst)


@pytest.mark.parametrize("k", [33, 48, 64])
def test_kdtree_large_k_vs_brute(oracle, k):
    """k above 32 (served on the GPU by the two-entries-per-lane list kernel): the oracle's k-d tree against its brute-force scan."""
    rng = np.random.default_rng(k)
    ref = np.concatenate([rng.uniform(-2, 2, size=(3000, 3)), np.ones((3000, 1))], axis=1).astype(np.float32)
    q = np.concatenate([rng.uniform(-2.5, 2.5, size=(200, 3)), np.ones((200, 1))], axis=1).astype(np.float32)
    for md in (np.inf, 0.6):
        ids_t, d2_t, _ = oracle.KdTree(ref).knn(q, k=k, max_dist=md)
        ids_b, d2_b = oracle.knn_brute(ref, q, k=k, max_dist=md)
        assert np.array_equal(ids_t, ids_b) and np.array_equal(d2_t.view(np.uint32), d2_b.view(np.uint32))


def test_surface_normal_extra_descriptors_vs_numpy(oracle):
    """The descriptors added in round 2 (SurfaceNormal.cpp:256-259,294-296; utils.h:91-103): eigVectors = the eigenvector matrix serialised
    ROW-major (row k = k-th components of the three eigenvectors, columns in ascending-eigenvalue order), meanDists = |p - mean(neighbours)|,
    matchedIds = the neighbour ids as floats — against numpy on the neighbours the oracle itself reports."""
    rng = np.random.default_rng(21)
    n, knn = 800, 9
    pts = np.concatenate([rng.uniform(-1, 1, size=(n, 2)), 0.1 * rng.normal(size=(n, 1))], axis=1).astype(np.float32)
    feat = oracle.feat4(pts)
    o = oracle.surface_normals(feat, knn, nthreads=4, keep=("normals", "eig_values", "eig_vectors", "matched_ids", "mean_dists"))
    ids, d2, _ = oracle.KdTree(feat).knn(feat, k=knn)
    assert np.array_equal(o["matched_ids"], ids.astype(np.float32))
    worst_md = worst_ev = worst_vec = 0.0
    for i in range(0, n, 5):
        P = pts[ids[i]].astype(np.float64)
        mean = P.mean(0)
        worst_md = max(worst_md, abs(float(o["mean_dists"][i]) - float(np.linalg.norm(pts[i].astype(np.float64) - mean))))
        C = (P - mean).T @ (P - mean) / P.shape[0]          # SurfaceNormal.cpp:201-204: NN NN^T / realKnn
        wv, V = np.linalg.eigh(C)
        E = o["eig_vectors"][i].astype(np.float64).reshape(3, 3)      # row-major: E[k, c] = component k of eigenvector c
        worst_ev = max(worst_ev, float(np.abs(o["eig_values"][i] - wv).max() / max(wv.max(), 1e-12)))
        for c in range(3):
            worst_vec = max(worst_vec, 1.0 - abs(float(E[:, c] @ V[:, c])))
        assert 1.0 - abs(float(o["normals"][i].astype(np.float64) @ V[:, 0])) < 1e-5     # the normal is the first column
    assert worst_md < 1e-6 and worst_ev < 1e-5 and worst_vec < 1e-4, (worst_md, worst_ev, worst_vec)


def test_counter_checker_semantics(oracle, golden):
    """CounterTransformationChecker (TransformationCheckersImpl.cpp:58-77): the counter is compared with maxIterationCount AFTER the
    increment, so 0 and 1 both stop after one iteration and 2 after two, with maxNumIterationsReached set; a negative count means the
    checker is not in the chain and the Differential checker alone ends the loop."""
    car = golden["car"]
    ref, rn, q = oracle.feat4(car["ref_xyz"]), car["ref_normals"].astype(np.float32), oracle.feat4(car["data_xyz"])
    Ti = np.eye(4, dtype=np.float32)
    for mi, want in ((0, 1), (1, 1), (2, 2), (3, 3)):
        cfg = oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=2, knn=1, max_dist=np.inf, trim_ratio=0.8, max_iter=mi, smooth_length=-1)
        code, T, st = oracle.icp(cfg, ref, rn, q, None, Ti)
        assert code == 0 and st.iterations == want and st.max_iterations_reached == 1, (mi, st.iterations)
    cfg = oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=2, knn=1, max_dist=np.inf, trim_ratio=0.8, max_iter=-1, smooth_length=3,
                            min_diff_rot=1e-3, min_diff_trans=1e-2)
    code, T, st = oracle.icp(cfg, ref, rn, q, None, Ti)
    assert code == 0 and st.max_iterations_reached == 0 and 3 < st.iterations < 200
    cfg40 = oracle.icp_config(precision=oracle.ORC_MIXED, nthreads=2, knn=1, max_dist=np.inf, trim_ratio=0.8, max_iter=1000, smooth_length=3,
                              min_diff_rot=1e-3, min_diff_trans=1e-2)
    code2, T2, st2 = oracle.icp(cfg40, ref, rn, q, None, Ti)
    assert code2 == 0 and st2.iterations == st.iterations and np.array_equal(T, T2)      # an unreached Counter changes nothing
