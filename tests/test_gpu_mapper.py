"""Scan-to-map mapping loop (BASELINE config #3 shape at a size the oracle finishes in seconds): the C++ Mapper driver over
the C ABI (host/mapper.h, mirror of o3d_slam::Mapper::addRangeMeasurement, Mapper.cpp:168-504) against the same control flow
composed from the CPU oracle's pieces.  Bars: processed-scan / map sizes equal at every step (voxel and crop membership is
bit-exact), poses within 1e-4 m / 1e-4 rad of the oracle trajectory (normals come from two fp32 implementations whose
eigen-solvers differ in the last bits), trajectory error vs ground truth bounded."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def trajectory(n, step=0.25):
    poses = []
    for i in range(n):
        s = i * step
        yaw = 0.35 * np.sin(0.25 * s)
        T = scenes.rigid(scenes.rot_axis_angle(np.array([0.0, 0.0, 1.0]), yaw), np.array([-6.0 + s, 1.5 * np.sin(0.3 * s), 1.6]))
        poses.append(T)
    return poses


def drifting_odometry(poses, seed):
    """external odometry = truth + random-walk drift"""
    rng = np.random.default_rng(seed)
    D = np.eye(4)
    out = []
    for T in poses:
        d = scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(0.05) * rng.normal()), rng.normal(size=3) * 0.004)
        D = D @ d
        out.append(T @ D)
    return out


class OracleMapper:
    """The same control flow as host/mapper.cpp, composed from oracle calls (CPU)."""

    def __init__(self, O, wide, narrow, icp_kw, scan_voxel, knn, radius, map_voxel, min_move, period, carving=None, carve_every=10):
        self.O, self.wide, self.narrow, self.icp_kw = O, wide, narrow, icp_kw
        self.carving, self.carve_every, self.n_inserted, self.cropper_pose, self.carved = carving, carve_every, 0, None, 0
        self.scan_voxel, self.knn, self.radius, self.map_voxel, self.min_move, self.period = scan_voxel, knn, radius, map_voxel, min_move, period
        self.pose = np.eye(4)
        self.prev = np.eye(4)
        self.last_insert = np.eye(4)
        self.odom_prev = None
        self.map_xyz = np.zeros((0, 3))
        self.map_nrm = np.zeros((0, 3))
        self.ref = None
        self.last_ref = 0.0

    def spec(self, s, T=None):
        O = self.O
        c = np.zeros(3) if T is None else T[:3, 3]
        return O.crop_spec(s["kind"], c, s.get("rmin", 0.0), s.get("rmax", 0.0), s.get("zmin", 0.0), s.get("zmax", 0.0))

    def process(self, xyz):
        O = self.O
        cx, _, _ = O.crop(self.spec(self.wide), xyz)
        mx, _ = O.voxel_down_sample(self.scan_voxel, cx)
        nrm, _ = O.estimate_normals_f64(mx, self.knn, self.radius, origin=np.zeros(3), nthreads=8)   # Open3D EstimateNormals: fp64
        nx, nn, _ = O.crop(self.spec(self.narrow), mx, nrm)
        return mx, nrm, nx, nn

    def insert(self, mx, mn):
        O = self.O
        self.carved = 0
        if self.carving is not None and self.map_xyz.shape[0] > 0 and self.n_inserted % self.carve_every == 0:   # Submap.cpp:56-60,130-144
            sx, _ = O.transform_f64(self.pose, self.raw, None)
            mask = O.carve(sx, self.pose[:3, 3], self.map_xyz, self.map_nrm, self.spec(self.wide, self.cropper_pose), **self.carving)
            self.map_xyz, self.map_nrm, self.carved = self.map_xyz[~mask], self.map_nrm[~mask], int(mask.sum())
        self.cropper_pose = self.pose.copy()
        self.n_inserted += 1
        tx, tn = O.transform_f64(self.pose, mx, mn)
        allx = np.concatenate([self.map_xyz, tx])
        alln = np.concatenate([self.map_nrm, tn])
        self.map_xyz, self.map_nrm, _ = O.voxelize_within_volume(self.map_voxel, self.spec(self.wide, self.pose), allx, alln)

    def add(self, xyz, time, odom):
        O = self.O
        self.raw = xyz
        st = dict(inserted=0, refreshed=0, iterations=0, code=0)
        if self.map_xyz.shape[0] == 0:
            self.prev = self.pose.copy()
            mx, mn, _, _ = self.process(xyz)
            self.insert(mx, mn)
            self.odom_prev = odom.copy()
            st.update(inserted=1, merge=mx.shape[0], map=self.map_xyz.shape[0])
            return True, st
        est = self.prev @ (np.linalg.inv(self.odom_prev) @ odom)
        mx, mn, nx, nn = self.process(xyz)
        rf, rn = O.o3d_to_pm(nx, nn)
        px, pn, _ = O.crop(self.spec(self.narrow, self.pose), self.map_xyz, self.map_nrm)
        if px.shape[0] == 0:
            return False, st
        cfg = O.icp_config(precision=O.ORC_MIXED, nthreads=8, **self.icp_kw)
        if self.ref is None or (time - self.last_ref) >= self.period:
            self.last_ref = time
            pf, pnn = O.o3d_to_pm(px, pn)
            self.ref = O.IcpReference(cfg, pf, pnn)
            st["refreshed"] = 1
        guess = est.astype(np.float32)
        code, T, ist = self.ref.compute(rf, rn, guess)
        st.update(code=code, iterations=ist.iterations, reading=rf.shape[0], merge=mx.shape[0], patch=px.shape[0])
        self.pose = (T if code == 0 else guess).astype(np.float64)
        moved = np.linalg.norm((np.linalg.inv(self.last_insert) @ self.pose)[:3, 3])
        if not (moved < self.min_move):
            self.insert(mx, mn)
            self.last_insert = self.pose.copy()
            st["inserted"] = 1
        self.prev = self.pose.copy()
        self.odom_prev = odom.copy()
        st["map"] = self.map_xyz.shape[0]
        return True, st


def test_mapping_loop_matches_oracle_composition(ctx, oracle):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper
    O = oracle
    scene = scenes.Scene(2001, size=(40.0, 24.0, 8.0), n_boxes=14, n_cyl=8)
    poses = trajectory(12)
    odom = drifting_odometry(poses, 2003)
    wide = dict(kind=capi.CROP_MINMAX_RADIUS, rmin=1.0, rmax=30.0)
    narrow = dict(kind=capi.CROP_MAX_RADIUS, rmax=25.0)
    icp_kw = dict(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
    kw = dict(scan_voxel=0.25, knn=10, radius=1.0, map_voxel=0.3, min_move=0.2, period=0.35)
    om = OracleMapper(O, wide, narrow, icp_kw, **kw)
    gm = Mapper(ctx, capi.crop_spec(wide["kind"], np.zeros(3), rmin=wide["rmin"], rmax=wide["rmax"]),
                capi.crop_spec(narrow["kind"], np.zeros(3), rmax=narrow["rmax"]), capi.icp_config(**icp_kw),
                scan_voxel_size=kw["scan_voxel"], normals_knn=kw["knn"], normals_radius=kw["radius"], map_voxel_size=kw["map_voxel"],
                min_movement=kw["min_move"], reference_period=kw["period"])
    om.pose = poses[0].copy()
    gm.set_pose(poses[0])
    refreshed = inserted = 0
    for i, (T, od) in enumerate(zip(poses, odom)):
        scan = scenes.lidar_scan(scene, T, n_beams=24, n_az=360, noise_sigma=0.01, seed=2002 + i).astype(np.float64)
        t = 0.1 * i
        ok_o, so = om.add(scan, t, od)
        ok_g, pose_g, sg, err = gm.add_range_measurement(scan, t, od)
        assert ok_o and ok_g, (i, err)
        assert sg.merge_points == so["merge"] and sg.map_points == so["map"], (i, sg.merge_points, so)
        assert sg.inserted == so["inserted"]
        if i > 0:
            assert sg.reading_points == so["reading"] and sg.patch_points == so["patch"]
            assert sg.icp_status == so["code"] == 0 and sg.reference_refreshed == so["refreshed"]
            assert abs(sg.iterations - so["iterations"]) <= 1, (i, sg.iterations, so["iterations"])
            refreshed += sg.reference_refreshed
            inserted += sg.inserted
        assert np.abs(pose_g[:3, 3] - om.pose[:3, 3]).max() < 1e-4, (i, pose_g, om.pose)
        assert np.linalg.norm(pose_g[:3, :3] - om.pose[:3, :3]) / np.sqrt(2) < 1e-4
        E = np.linalg.inv(T) @ pose_g
        assert np.abs(E[:3, 3]).max() < 0.05, (i, E)          # stays on the true trajectory despite the drifting odometry
    assert 2 <= refreshed < len(poses) - 1 and inserted >= 5   # both the stale-reference and the skip-insertion branches ran
    mx, mn = gm.map_points()
    assert mx.shape[0] == om.map_xyz.shape[0]
    # pass-through points keep their order, voxel means follow in ascending key order in both implementations
    assert np.abs(mx - om.map_xyz).max() < 2e-4
    gm.close()


def test_voxel_down_sample_o3d_convention(ctx, oracle):
    """o3d_slam::voxelize -> Open3D VoxelDownSample (a17): membership and emission order bit-exact, means bit-exact."""
    rng = np.random.default_rng(5)
    xyz = rng.uniform(-20, 20, (60_000, 3))
    nrm = rng.normal(size=(60_000, 3))
    nrm[::97] = np.nan
    for v in (0.1, 0.35, 2.0):
        gx, gn = ctx.voxel_down_sample(v, xyz, nrm)
        ox, on = oracle.voxel_down_sample(v, xyz, nrm)
        assert gx.shape == ox.shape
        assert np.array_equal(gx, ox) and np.array_equal(np.nan_to_num(gn), np.nan_to_num(on))
    gx, gn = ctx.voxel_down_sample(0.5, xyz)          # no normals
    ox, _ = oracle.voxel_down_sample(0.5, xyz)
    assert gn is None and np.array_equal(gx, ox)
    gx2, _ = ctx.voxel_down_sample(0.5, gx)           # idempotent on membership count (means move inside their voxels)
    assert gx2.shape[0] <= gx.shape[0]
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    with pytest.raises(capi.InvalidParameter):
        ctx.voxel_down_sample(0.0, xyz)
    e, _ = ctx.voxel_down_sample(0.5, np.zeros((0, 3)))
    assert e.shape[0] == 0


def test_cloud64_chain_equals_host_array_chain(ctx, oracle):
    """The device-resident scan-processing chain (b200reg_cloud64_*) gives bit-identical clouds to the host-array entry points it
    shares its kernels with (crop_f64 -> voxel_down_sample_f64 -> crop_f64 -> upload_f64), its fp64 normals (Open3D
    EstimateNormals is fp64, so the points are NOT rounded to fp32 first) are within 1e-5 rad of the fp64 oracle with brute-force
    neighbours, and the map built from it equals the map built from host arrays."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    scene = scenes.Scene(31, size=(60.0, 40.0, 8.0), n_boxes=25, n_cyl=15)
    T = scenes.rigid(scenes.rot_axis_angle(np.array([0.0, 0.0, 1.0]), 0.2), np.array([1.0, -2.0, 1.7]))
    scan = scenes.lidar_scan(scene, T, n_beams=48, n_az=600, noise_sigma=0.01, seed=5).astype(np.float64)
    wide = capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=1.5, rmax=40.0)
    narrow = capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=25.0)
    # host-array chain
    cx, _ = ctx.crop_f64(wide, scan)
    vx, _ = ctx.voxel_down_sample(0.12, cx)
    vn, deg_o = oracle.estimate_normals_f64(vx, 12, 0.8, origin=np.zeros(3), exact=True)
    c32 = ctx.upload(vx)
    c32.surface_normals(12, max_dist=0.8, orient_origin=np.zeros(3, np.float32))
    _, n32 = c32.download()
    # device chain, with storage reuse on a second pass (first pass on a smaller cloud so every buffer has to grow)
    raw = crop = merge = match = None
    for pts in (scan[:5000], scan):
        raw = capi.Cloud64.upload(ctx, pts, out=raw)
        crop = raw.crop(wide, out=crop)
        merge = crop.voxel_down_sample(0.12, out=merge)
        assert not merge.has_normals
        deg_g = merge.estimate_normals(12, 0.8, origin=np.zeros(3))
        match = merge.crop(narrow, out=match)
    dx, dn = merge.download()
    assert np.array_equal(dx, vx) and deg_g == deg_o
    live = np.linalg.norm(vn, axis=1) > 0
    assert np.array_equal(live, np.linalg.norm(dn, axis=1) > 0)
    ang = np.arccos(np.clip(np.sum(dn[live] * vn[live], axis=1), -1.0, 1.0))
    chord = np.linalg.norm(dn[live] - vn[live], axis=1)              # arccos loses precision near 0: bound the chord instead
    assert chord.max() <= 1e-5, (chord.max(), ang.max())
    # what rounding the points to fp32 first costs at these ranges (the round-1 path): above the 1e-5 rad bar
    chord32 = np.linalg.norm(n32[live].astype(np.float64) - vn[live], axis=1)
    assert (chord32 > 1e-5).any()
    vn = dn
    mx, mn = ctx.crop_f64(narrow, vx, vn)
    qx, qn = match.download()
    assert np.array_equal(qx, mx) and np.array_equal(qn, mn)
    f, fn = match.to_f32().download()
    assert np.array_equal(f[:, :3], mx.astype(np.float32)) and (f[:, 3] == 1).all() and np.array_equal(fn, mn.astype(np.float32))
    # a cloud with normals keeps them through crop and voxelisation; without normals to_f32 has none
    withn = capi.Cloud64.upload(ctx, vx, vn)
    hx, hn = ctx.voxel_down_sample(0.3, vx, vn)
    gx, gn = withn.voxel_down_sample(0.3).download()
    assert np.array_equal(gx, hx) and np.array_equal(gn, hn)
    assert not capi.Cloud64.upload(ctx, vx).to_f32().has_normals
    # map insertion from the device cloud == from host arrays
    vol = capi.crop_spec(capi.CROP_MAX_RADIUS, T[:3, 3], rmax=40.0)
    ma, mb = capi.VoxelMap(ctx, 0.2), capi.VoxelMap(ctx, 0.2)
    for Tm in (T, T @ scenes.rigid(np.eye(3), np.array([0.3, 0.1, 0.0]))):
        ma.insert(vx, vn, Tm, vol)
        mb.insert_cloud64(merge, Tm, vol)
    xa, na = ma.download()
    xb, nb = mb.download()
    assert np.array_equal(xa, xb) and np.array_equal(na, nb)
    # empty results and reuse of an emptied cloud
    none = merge.crop(capi.crop_spec(capi.CROP_MAX_RADIUS, np.array([1e4, 0, 0]), rmax=1.0), out=match)
    assert none.size == 0 and none.download()[0].shape == (0, 3)


def test_mapping_loop_with_space_carving(ctx, oracle):
    """The mapping loop with carving every 2nd inserted scan (param_replay.lua enables it): a block of ghost points is planted into
    both maps' first scan; carved counts, map sizes and poses must follow the oracle composition step by step."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper
    O = oracle
    scene = scenes.Scene(2001, size=(40.0, 24.0, 8.0), n_boxes=14, n_cyl=8)
    poses = trajectory(9)
    odom = drifting_odometry(poses, 2004)
    wide = dict(kind=capi.CROP_MINMAX_RADIUS, rmin=1.0, rmax=30.0)
    narrow = dict(kind=capi.CROP_MAX_RADIUS, rmax=25.0)
    icp_kw = dict(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
    kw = dict(scan_voxel=0.25, knn=10, radius=1.0, map_voxel=0.3, min_move=0.0, period=0.35)
    carve_o = dict(voxel=0.3, max_ray_length=12.0, truncation=0.3, min_dot=0.3)
    om = OracleMapper(O, wide, narrow, icp_kw, carving=carve_o, carve_every=2, **kw)
    gm = Mapper(ctx, capi.crop_spec(wide["kind"], np.zeros(3), rmin=wide["rmin"], rmax=wide["rmax"]),
                capi.crop_spec(narrow["kind"], np.zeros(3), rmax=narrow["rmax"]), capi.icp_config(**icp_kw),
                scan_voxel_size=kw["scan_voxel"], normals_knn=kw["knn"], normals_radius=kw["radius"], map_voxel_size=kw["map_voxel"],
                min_movement=kw["min_move"], reference_period=kw["period"], carving=capi.CarveParams(0.3, 12.0, 0.3, 0.3), carve_every=2)
    om.pose = poses[0].copy()
    gm.set_pose(poses[0])
    total_carved = 0
    for i, (T, od) in enumerate(zip(poses, odom)):
        scan = scenes.lidar_scan(scene, T, n_beams=32, n_az=360, noise_sigma=0.01, seed=700 + i).astype(np.float64)
        if i == 0:   # a parked object 3 m ahead of the first pose that is gone afterwards (sensor-frame points)
            obj = np.array([3.0, 0.0, -0.6]) + np.random.default_rng(1).uniform(-0.5, 0.5, (600, 3)) * np.array([0.1, 1.0, 1.0])
            scan = np.concatenate([scan, obj])
        ok_o, st_o = om.add(scan, 0.1 * i, od)
        ok_g, pose_g, st_g, err = gm.add_range_measurement(scan, 0.1 * i, od)
        assert ok_o and ok_g, err
        assert st_g.carved_points == om.carved, (i, st_g.carved_points, om.carved)
        assert st_g.map_points == st_o["map"], (i, st_g.map_points, st_o["map"])
        total_carved += om.carved
        d = np.linalg.inv(om.pose) @ pose_g
        assert np.abs(d[:3, 3]).max() < 1e-4 and np.linalg.norm(d[:3, :3] - np.eye(3)) < 1e-4, i
    assert total_carved > 20
    gx, _ = gm.map_points()
    assert gx.shape[0] == om.map_xyz.shape[0] and np.abs(gx - om.map_xyz).max() < 2e-4
    gm.close()


def test_mapper_pointcloud2_entry_equals_array_entry(ctx):
    """Feeding the mapping step with sensor_msgs/PointCloud2 bytes (an Ouster-like 48-byte record, x y z as FLOAT32) gives exactly the
    poses, counts and map of feeding it the same points as an fp64 array (rosToOpen3d widens the floats exactly)."""
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper
    scene = scenes.Scene(2001, size=(40.0, 24.0, 8.0), n_boxes=14, n_cyl=8)
    poses = trajectory(5)
    odom = drifting_odometry(poses, 2005)
    icp = capi.icp_config(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
    mk = lambda: Mapper(ctx, capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=1.0, rmax=30.0),  # noqa: E731
                        capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=25.0), icp, scan_voxel_size=0.25, normals_knn=10,
                        normals_radius=1.0, map_voxel_size=0.3, min_movement=0.0, reference_period=0.25)
    ma, mb = mk(), mk()
    ma.set_pose(poses[0]); mb.set_pose(poses[0])
    fields, step = {"x": 0, "y": 4, "z": 8}, 48
    for i, (T, od) in enumerate(zip(poses, odom)):
        scan = scenes.lidar_scan(scene, T, n_beams=32, n_az=360, noise_sigma=0.01, seed=900 + i).astype(np.float32)
        n = scan.shape[0]
        buf = np.random.default_rng(i).integers(0, 256, size=(n, step), dtype=np.uint8)
        buf[:, :12] = scan.view(np.uint8).reshape(n, 12)
        ok_a, pose_a, st_a, err_a = ma.add_range_measurement(scan.astype(np.float64), 0.1 * i, od)
        ok_b, pose_b, st_b, err_b = mb.add_pointcloud2(buf.reshape(-1), capi.pc2_layout(n, 1, step, fields), 0.1 * i, od)
        assert ok_a and ok_b, (err_a, err_b)
        assert np.array_equal(pose_a, pose_b), i
        assert (st_a.iterations, st_a.reading_points, st_a.merge_points, st_a.map_points) == (st_b.iterations, st_b.reading_points, st_b.merge_points, st_b.map_points)
    xa, na = ma.map_points()
    xb, nb = mb.map_points()
    assert np.array_equal(xa, xb) and np.array_equal(na, nb)
    ok, _, _, err = mb.add_pointcloud2(buf.reshape(-1)[:100], capi.pc2_layout(n, 1, step, fields), 1.0, odom[-1])
    assert not ok and "point_step" in err
    ma.close(); mb.close()


def test_mapper_prefetch_gives_the_same_trajectory(ctx):
    """Mapper.prefetch (worker threads with their own contexts run the scan-processing chain ahead; the sets of device clouds rotate)
    changes WHEN a scan is uploaded and processed, never what the step computes: poses, counts and the final map equal the plain
    add_range_measurement run bit for bit — also when prefetched scans are never consumed or are pushed out by later ones, when a
    prefetched scan fails in the chain, and when the scan sizes differ (storage of the sets is recycled and grown)."""
    import torch
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper
    scene = scenes.Scene(2001, size=(40.0, 24.0, 8.0), n_boxes=14, n_cyl=8)
    poses = trajectory(7)
    odom = drifting_odometry(poses, 2005)
    icp = capi.icp_config(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
    mk = lambda: Mapper(ctx, capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=1.0, rmax=30.0),  # noqa: E731
                        capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=25.0), icp, scan_voxel_size=0.25, normals_knn=10,
                        normals_radius=1.0, map_voxel_size=0.3, min_movement=0.0, reference_period=0.25)
    scans = [scenes.lidar_scan(scene, T, n_beams=24 + 4 * (i % 3), n_az=300 + 40 * i, noise_sigma=0.01, seed=900 + i).astype(np.float64) for i, T in enumerate(poses)]
    pinned = [torch.from_numpy(s).pin_memory().numpy() for s in scans]
    ma, mb = mk(), mk()
    ma.set_pose(poses[0]); mb.set_pose(poses[0])
    mb.prefetch(pinned[0])
    decoys = [np.ascontiguousarray(scans[3][k:1000 + k]) for k in range(3)]
    for i, od in enumerate(odom):
        if i + 1 < len(pinned):
            mb.prefetch(pinned[i + 1])          # uploads while scan i is processed
        if i == 4:                              # more prefetches than sets: the oldest pending scans are replaced and then processed inline
            for d in decoys:
                mb.prefetch(d)
        ok_a, pose_a, st_a, err_a = ma.add_range_measurement(scans[i], 0.1 * i, od)
        ok_b, pose_b, st_b, err_b = mb.add_range_measurement(pinned[i], 0.1 * i, od)
        assert ok_a and ok_b, (err_a, err_b)
        assert np.array_equal(pose_a, pose_b), i
        assert (st_a.iterations, st_a.reading_points, st_a.merge_points, st_a.map_points) == (st_b.iterations, st_b.reading_points, st_b.merge_points, st_b.map_points)
    xa, na = ma.map_points()
    xb, nb = mb.map_points()
    assert np.array_equal(xa, xb) and np.array_equal(na, nb)
    # a scan the wide crop empties (everything inside the minimum radius): the worker's failure is reported by the step that consumes it,
    # with the message of the inline path, and the next scan goes through
    near = np.ascontiguousarray(scans[0] * (0.5 / np.linalg.norm(scans[0], axis=1, keepdims=True)))
    mb.prefetch(near)
    ok_a, _, _, err_a = ma.add_range_measurement(near, 0.8, odom[-1])
    ok_b, _, _, err_b = mb.add_range_measurement(near, 0.8, odom[-1])
    assert not ok_a and not ok_b and err_a == err_b and "cropped size is zero" in err_b
    ok_a, pose_a, _, _ = ma.add_range_measurement(scans[-1], 0.9, odom[-1])
    ok_b, pose_b, _, _ = mb.add_range_measurement(pinned[-1], 0.9, odom[-1])
    assert ok_a and ok_b and np.array_equal(pose_a, pose_b)
    ma.close(); mb.close()
    # the cloud-level entry points
    c = capi.Cloud64.prefetch(ctx, pinned[2])
    c.acquire()
    assert np.array_equal(c.download()[0], scans[2])
    capi.Cloud64.prefetch(ctx, pinned[1], out=c)
    capi.Cloud64.prefetch(ctx, pinned[3], out=c)     # a second prefetch into a pending cloud waits for the first
    assert np.array_equal(c.acquire().download()[0], scans[3])
    c.free()
