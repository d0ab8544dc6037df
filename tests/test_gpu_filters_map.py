"""GPU parity tests for the DataPointsFilters and the open3d_slam voxel-map crop/insert (through the C ABI) vs the oracle.

Bars: voxel membership + output order bit-exact, centroids bit-exact (the device reproduces the reference's fp32 accumulation
order); normals within 1e-5 rad up to sign; crop membership bit-exact; map means bit-exact (fp64, same accumulation order).
"""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def random_cloud(n, seed, scale=10.0):
    rng = np.random.default_rng(seed)
    return (rng.random((n, 3)) * scale - scale / 2).astype(np.float32)


# ---------------------------------------------------------------------------------------------------------------
# VoxelGridDataPointsFilter
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,v,centroid,avg", [(100, (0.5, 0.5, 0.5), 1, 1), (50_000, (0.3, 0.45, 0.2), 1, 1), (50_000, (1.0, 1.0, 1.0), 1, 0),
                                              (20_000, (0.25, 0.25, 0.25), 0, 1), (1, (1.0, 1.0, 1.0), 1, 1)])
def test_voxelgrid_random_bit_exact(ctx, oracle, n, v, centroid, avg):
    """utest/ui/DataFilters.cpp:616-673 shape (random cloud, fewer points out) + exact membership/order/centroids."""
    pts = oracle.feat4(random_cloud(n, 7 + n, scale=2.0 if n <= 100 else 10.0))
    nrm = random_cloud(n, 8 + n, scale=2.0)
    m, f_o, d_o, cells = oracle.voxelgrid(pts, nrm, *v, use_centroid=bool(centroid), average_descriptors=bool(avg))
    c = ctx.upload(pts, nrm)
    c.voxelgrid(*v, use_centroid=bool(centroid), average_descriptors=bool(avg))
    f_g, d_g = c.download()
    assert c.size == m
    assert n == 1 or m < n
    assert np.array_equal(f_g.view(np.uint32), f_o.view(np.uint32))
    assert np.array_equal(d_g.view(np.uint32), d_o.view(np.uint32))


def test_voxelgrid_car_cloud_and_scan(ctx, oracle, golden):
    for xyz, v in [(golden["car"]["ref_xyz"], 0.2), (scenes.config2(m=10_000, n_beams=64, n_az=1024)["scan"], 0.1)]:
        pts = oracle.feat4(xyz)
        m, f_o, _, _ = oracle.voxelgrid(pts, None, v, v, v)
        c = ctx.upload(pts)
        c.voxelgrid(v, v, v)
        f_g, _ = c.download()
        assert c.size == m and m < pts.shape[0]
        assert np.array_equal(f_g.view(np.uint32), f_o.view(np.uint32))


def test_voxelgrid_idempotent_at_scale(ctx, oracle):
    """Size-independent property at full scan size: filtering twice with the same grid changes nothing the second time
    only if every voxel already holds one point — here we check monotonicity and determinism instead."""
    pts = oracle.feat4(scenes.config2(m=10_000, n_beams=128, n_az=1024)["scan"])
    a = ctx.upload(pts); a.voxelgrid(0.1, 0.1, 0.1)
    b = ctx.upload(pts); b.voxelgrid(0.1, 0.1, 0.1)
    fa, _ = a.download(); fb, _ = b.download()
    assert np.array_equal(fa.view(np.uint32), fb.view(np.uint32))
    assert 0 < a.size < pts.shape[0]
    n1 = a.size
    a.voxelgrid(0.1, 0.1, 0.1)
    assert a.size <= n1


def test_voxelgrid_errors(ctx, oracle):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    c = ctx.upload(oracle.feat4(random_cloud(10, 1)))
    with pytest.raises(capi.InvalidParameter):     # VoxelGrid.h:69 lower bound 0.001
        c.voxelgrid(0.0001, 1.0, 1.0)
    bad = oracle.feat4(random_cloud(10, 1))
    bad[3, 0] = np.nan
    e = ctx.upload(np.zeros((0, 4), dtype=np.float32))
    e.voxelgrid(1.0, 1.0, 1.0)
    assert e.size == 0


# ---------------------------------------------------------------------------------------------------------------
# SurfaceNormalDataPointsFilter
# ---------------------------------------------------------------------------------------------------------------
def _angle_up_to_sign(a, b):
    dot = np.abs(np.sum(a.astype(np.float64) * b.astype(np.float64), axis=1))
    cross = np.linalg.norm(np.cross(a.astype(np.float64), b.astype(np.float64)), axis=1)
    return np.arctan2(cross, dot)


@pytest.mark.parametrize("knn,maxd", [(5, np.inf), (12, np.inf), (20, 1.0), (10, 0.3)])
def test_surface_normals_vs_oracle(ctx, oracle, golden, knn, maxd):
    """utest/ui/DataFilters.cpp:219-240 configuration (knn 5, all keep*) plus the configs' k = 10/12/20."""
    pts = oracle.feat4(golden["car"]["data_xyz"])
    o = oracle.surface_normals(pts, knn, max_dist=maxd, precision=oracle.ORC_MIXED, nthreads=8,
                               keep=("normals", "densities", "eig_values", "lin_plan_curv"))
    c = ctx.upload(pts)
    g = c.surface_normals(knn, max_dist=maxd, keep=("densities", "eig_values", "lin_plan_curv"))
    _, n_g = c.download()
    assert g["degenerate"] == o["degenerate"]
    deg = np.all(o["normals"] == 0, axis=1)
    assert np.array_equal(deg, np.all(n_g == 0, axis=1))           # degenerate classification identical
    ang = _angle_up_to_sign(n_g[~deg], o["normals"][~deg])
    assert ang.max() <= 1e-5, ang.max()
    assert np.allclose(g["eig_values"], o["eig_values"], rtol=1e-5, atol=1e-9)
    assert np.allclose(g["densities"], o["densities"], rtol=1e-5)
    ok = np.isfinite(o["lin_plan_curv"]).all(axis=1)
    assert np.allclose(g["lin_plan_curv"][ok], o["lin_plan_curv"][ok], rtol=1e-3, atol=1e-5)


def test_surface_normals_orientation_and_analytic(ctx, oracle):
    c2 = scenes.config2(m=150_000, n_beams=16, n_az=256)
    pts = oracle.feat4(c2["map_xyz"])
    cl = ctx.upload(pts)
    origin = np.array([60.0, 60.0, 2.0], dtype=np.float32)
    cl.surface_normals(20, max_dist=1.0, orient_origin=origin)
    _, n_g = cl.download()
    o = oracle.surface_normals(pts, 20, max_dist=1.0, nthreads=8, orient_origin=origin)
    deg = np.all(o["normals"] == 0, axis=1)
    assert np.array_equal(deg, np.all(n_g == 0, axis=1))
    # oriented: the sign is pinned, compare directly
    d = np.linalg.norm(n_g[~deg].astype(np.float64) - o["normals"][~deg], axis=1)
    assert d.max() <= 2e-5
    # sanity vs the analytic normals of the planar faces (up to sign), away from edges
    an = c2["map_nrm"]
    ang = _angle_up_to_sign(n_g[~deg], an[~deg].astype(np.float32))
    assert np.median(ang) < 1e-3


def test_surface_normals_degenerate_and_errors(ctx, oracle):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    # collinear points: rank(C) + 1 < 3 -> zero normal (SurfaceNormal.cpp:229-250)
    line = np.zeros((50, 3), dtype=np.float32)
    line[:, 0] = np.arange(50) * 0.1
    pts = oracle.feat4(np.concatenate([line, random_cloud(200, 3, scale=3.0) + 10]))
    o = oracle.surface_normals(pts, 5, nthreads=4)
    c = ctx.upload(pts)
    g = c.surface_normals(5)
    _, n_g = c.download()
    assert g["degenerate"] == o["degenerate"] and g["degenerate"] >= 40
    assert np.array_equal(np.all(n_g == 0, axis=1), np.all(o["normals"] == 0, axis=1))
    with pytest.raises(capi.InvalidParameter):     # SurfaceNormal.h:69: knn >= 3
        c.surface_normals(2)


# ---------------------------------------------------------------------------------------------------------------
# croppers + map insert / crop
# ---------------------------------------------------------------------------------------------------------------
def _spec_pair(oracle, kind, center, **kw):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    return oracle.crop_spec(kind, center, **kw), capi.crop_spec(kind, center, **kw)


@pytest.mark.parametrize("kind,kw", [(0, dict(rmax=7.0)), (1, dict(rmin=3.0)), (2, dict(rmin=2.0, rmax=80.0)), (2, dict(rmin=2.0, rmax=6.0, invert=True)),
                                     (3, dict(rmax=5.0, zmin=-1.0, zmax=2.0))])
def test_crop_membership_bit_exact(ctx, oracle, kind, kw):
    rng = np.random.default_rng(5)
    xyz = rng.normal(size=(200_000, 3)) * 5.0
    nrm = rng.normal(size=(200_000, 3))
    center = [0.7, -1.2, 0.4]
    # points exactly on the boundary (inclusive tests, croppers.cpp:121-167)
    xyz[:3] = np.array(center) + np.array([[kw.get("rmax", 1.0), 0, 0], [0, kw.get("rmin", 1.0), 0], [0, 0, kw.get("rmax", 1.0)]])
    so, sg = _spec_pair(oracle, kind, center, **kw)
    xo, no, mask = oracle.crop(so, xyz, nrm)
    xg, ng = ctx.crop_f64(sg, xyz, nrm)
    assert xg.shape == xo.shape and 0 < xo.shape[0] < xyz.shape[0]
    assert np.array_equal(xg, xo) and np.array_equal(ng, no)


def test_map_insert_sequence_matches_oracle(ctx, oracle):
    """Submap::insertScan x 6 along a short trajectory: transform, append, re-voxelise inside the volume; the device map
    must equal the oracle's map after every insert (pass-through first, then voxels by ascending key): fp64 bit-exact."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    sc = scenes.Scene(77, size=(40.0, 30.0, 6.0), n_boxes=10, n_cyl=6, keepout=[[10 + i, 12.0] for i in range(8)])
    voxel = 0.15
    m = capi.VoxelMap(ctx, voxel)
    map_o = np.zeros((0, 3)); nrm_o = np.zeros((0, 3))
    for i in range(6):
        pose = scenes.rigid(scenes.rot_axis_angle([0, 0, 1], 0.2 * i), [10.0 + 0.8 * i, 12.0 + 0.1 * i, 1.5])
        scan = scenes.lidar_scan(sc, pose, 32, 256, 0.01, seed=100 + i)
        rng = np.random.default_rng(i)
        sn = rng.normal(size=scan.shape)
        sn /= np.linalg.norm(sn, axis=1, keepdims=True)
        if i == 2:
            sn[::50] = np.nan          # NaN normals are skipped in the sum (helpers.cpp:35-37)
        so, sg = _spec_pair(oracle, 2 if i % 2 == 0 else 0, pose[:3, 3], rmin=0.0, rmax=12.0)
        m.insert(scan, sn, pose, sg)
        tx, tn = oracle.transform_f64(pose, scan, sn)
        map_o, nrm_o, npass = oracle.voxelize_within_volume(voxel, so, np.concatenate([map_o, tx]), np.concatenate([nrm_o, tn]))
        xg, ng = m.download()
        assert m.size == map_o.shape[0], i
        assert np.array_equal(xg, map_o), i
        assert np.array_equal(ng, nrm_o, equal_nan=True), i
    # ScanToMapRegistration::cropSubmap + open3dToPointmatcher: order-preserving crop into a float cloud
    so, sg = _spec_pair(oracle, 2, [12.0, 12.0, 1.5], rmin=2.0, rmax=8.0)
    cx, cn, _ = oracle.crop(so, map_o, nrm_o)
    f_o, n_o = oracle.o3d_to_pm(cx, cn)
    cl = m.crop(sg)
    f_g, n_g = cl.download()
    assert np.array_equal(f_g.view(np.uint32), f_o.view(np.uint32))
    assert np.array_equal(n_g.view(np.uint32), n_o.view(np.uint32))
    # idempotence: re-inserting nothing new leaves per-voxel means of single-point voxels untouched
    m.close()


def test_map_insert_without_normals_and_upload_f64(ctx, oracle):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    rng = np.random.default_rng(9)
    xyz = rng.normal(size=(5000, 3)) * 3
    m = capi.VoxelMap(ctx, 0.5)
    so, sg = _spec_pair(oracle, 0, [0, 0, 0], rmax=4.0)
    T = scenes.rigid(scenes.rot_axis_angle([1, 1, 0], 0.3), [0.5, 0.2, -0.1])
    m.insert(xyz, None, T, sg)
    tx, _ = oracle.transform_f64(T, xyz)
    mo, _, _ = oracle.voxelize_within_volume(0.5, so, tx)
    xg, _ = m.download()
    assert np.array_equal(xg, mo)
    # fp64 upload path == open3dToPointmatcher rounding
    f_o, _ = oracle.o3d_to_pm(xyz)
    f_g, _ = ctx.upload(xyz).download()
    assert np.array_equal(f_g.view(np.uint32), f_o.view(np.uint32))


@pytest.mark.parametrize("with_normals", [True, False])
def test_map_carve_matches_oracle(ctx, oracle, with_normals):
    """Space carving on the resident map (Submap::carve -> getIdxsOfCarvedPoints): a map built from two scans, with a block of
    'ghost' points dropped into free space, carved by a third scan through b200reg_map_carve (host scan) and
    b200reg_map_carve_cloud64 (device scan): the surviving map must equal the oracle's mask applied to the same map, bit for bit
    and in order; carving twice with the same scan removes nothing new."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    sc = scenes.Scene(91, size=(40.0, 30.0, 6.0), n_boxes=10, n_cyl=6, keepout=[[10 + i, 12.0] for i in range(6)])
    voxel = 0.15
    rng = np.random.default_rng(9)
    maps = []
    for mode in ("host", "cloud64"):
        m = capi.VoxelMap(ctx, voxel)
        for i in range(2):
            pose = scenes.rigid(scenes.rot_axis_angle([0, 0, 1], 0.3 * i), [10.0 + 1.5 * i, 12.0, 1.5])
            scan = scenes.lidar_scan(sc, pose, 32, 256, 0.01, seed=300 + i)
            c32 = ctx.upload(scan.astype(np.float64))
            c32.surface_normals(10, max_dist=1.0, orient_origin=np.zeros(3, np.float32))
            sn = c32.download()[1].astype(np.float64) if with_normals else None
            _, sg = _spec_pair(oracle, 0, pose[:3, 3], rmax=25.0)
            m.insert(scan, sn, pose, sg)
        ghost = np.array([13.0, 12.5, 1.0]) + np.random.default_rng(10).uniform(-0.4, 0.4, (400, 3))   # a departed object in free space
        gn = np.tile([-1.0, 0.0, 0.0], (400, 1)) if with_normals else None
        _, far_vol = _spec_pair(oracle, 0, [1e3, 0, 0], rmax=1.0)                                          # nothing inside: plain append + no re-voxelisation
        m.insert(ghost, gn, np.eye(4), far_vol)
        before_x, before_n = m.download()
        pose = scenes.rigid(scenes.rot_axis_angle([0, 0, 1], 0.1), [11.0, 12.2, 1.5])
        raw = scenes.lidar_scan(sc, pose, 32, 256, 0.01, seed=310).astype(np.float64)
        so, sg = _spec_pair(oracle, 0, [10.5, 12.0, 1.5], rmax=12.0)                                        # cropper still at the previous insertion's pose
        kw = dict(voxel_size=0.1, max_raytracing_length=15.0, truncation_distance=0.2, min_dot=0.4)
        removed = m.carve(raw if mode == "host" else capi.Cloud64.upload(ctx, raw), pose, sg, **kw)
        tx, _ = oracle.transform_f64(pose, raw, None)
        mask = oracle.carve(tx, pose[:3, 3], before_x, before_n, so, voxel=0.1, max_ray_length=15.0, truncation=0.2, min_dot=0.4)
        after_x, after_n = m.download()
        assert removed == int(mask.sum()) and removed > 100, (mode, removed, int(mask.sum()))
        assert np.array_equal(after_x, before_x[~mask])
        if with_normals:
            assert np.array_equal(after_n, before_n[~mask])
        ghost_rows = (np.abs(before_x - np.array([13.0, 12.5, 1.0])) <= 0.4).all(axis=1)
        assert mask[ghost_rows].mean() > 0.5                      # most of the departed object is gone
        assert m.carve(raw, pose, sg, **kw) == 0                   # idempotent
        maps.append(after_x)
        m.close()
    assert np.array_equal(maps[0], maps[1])
    # parameter check and empty inputs
    m = capi.VoxelMap(ctx, voxel)
    assert m.carve(raw, pose, sg) == 0                             # empty map: nothing to do (Submap.cpp:132)
    m.insert(raw[:100], None, pose, sg)
    assert m.carve(np.zeros((0, 3)), pose, sg) == 0
    with pytest.raises(capi.InvalidParameter):
        m.carve(raw, pose, sg, voxel_size=0.0)


@pytest.mark.parametrize("fields,step", [({"x": 0, "y": 4, "z": 8}, 12), ({"x": 0, "y": 4, "z": 8, "normal_x": 12, "normal_y": 16, "normal_z": 20}, 24),
                                         ({"x": 0, "y": 4, "z": 8, "normal_x": 16, "normal_y": 20, "normal_z": 24}, 48),
                                         ({"z": 1, "x": 9, "y": 5, "normal_x": 13, "normal_y": 17}, 23)])
def test_pointcloud2_to_device_clouds(ctx, oracle, fields, step):
    """sensor_msgs/PointCloud2 bytes -> fp32 cloud (RosPointCloud2Deserializer) and fp64 cloud (rosToOpen3d) on the device, bit for
    bit against the oracle, incl. NaN returns, unaligned records, storage reuse, empty messages and the error cases."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from test_oracle_golden import _pc2_message
    c32 = c64 = None
    for n, (w, h) in ((64, (64, 1)), (131072, (1024, 128))):          # second pass reuses and grows the clouds
        data, _ = _pc2_message(n, fields, step, 5 + n)
        st, f4, n3, x64, n64 = oracle.pointcloud2(data, n, step, fields)
        L = capi.pc2_layout(w, h, step, fields)
        c32 = ctx.cloud_from_pointcloud2(data, L, out=c32)
        g4, gn = c32.download()
        assert np.array_equal(g4.view(np.uint32), f4.view(np.uint32))
        assert (gn is None) == (n3 is None) and (gn is None or np.array_equal(gn.view(np.uint32), n3.view(np.uint32)))
        c64 = ctx.cloud64_from_pointcloud2(data, L, out=c64)
        gx, gn64 = c64.download()
        assert np.array_equal(gx.view(np.uint64), x64.view(np.uint64))
        assert (gn64 is None) == (n64 is None) and (gn64 is None or np.array_equal(gn64, n64))
    assert ctx.cloud_from_pointcloud2(np.zeros(0, np.uint8), capi.pc2_layout(0, 0, step, fields)).size == 0
    assert ctx.cloud64_from_pointcloud2(np.zeros(0, np.uint8), capi.pc2_layout(0, 1, step, fields)).size == 0
    with pytest.raises(capi.InvalidField):
        ctx.cloud_from_pointcloud2(data, capi.pc2_layout(w, h, step, {"x": 0, "y": 4}))
    with pytest.raises(capi.InvalidParameter):
        ctx.cloud_from_pointcloud2(data[:-1], L)                        # buffer shorter than width * height * point_step
    with pytest.raises(capi.InvalidParameter):
        ctx.cloud64_from_pointcloud2(data, capi.pc2_layout(w, h, step, dict(fields, x=step - 2)))
