import sys, os, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import scenes
from open3d_slam_advanced_rss_2024_public_b200 import capi
from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper
from test_gpu_mapper import trajectory, drifting_odometry
ctx = capi.Context(0)
scene = scenes.Scene(2001, size=(40.0, 24.0, 8.0), n_boxes=14, n_cyl=8)
poses = trajectory(5); odom = drifting_odometry(poses, 2005)
icp = capi.icp_config(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
m = Mapper(ctx, capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=1.0, rmax=30.0), capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=25.0), icp,
           scan_voxel_size=0.25, normals_knn=10, normals_radius=1.0, map_voxel_size=0.3, min_movement=0.0, reference_period=0.25)
m.set_pose(poses[0])
mk = lambda: Mapper(ctx, capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=1.0, rmax=30.0), capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=25.0), icp,
           scan_voxel_size=0.25, normals_knn=10, normals_radius=1.0, map_voxel_size=0.3, min_movement=0.0, reference_period=0.25)
mb = mk(); mb.set_pose(poses[0])
fields, step = {"x": 0, "y": 4, "z": 8}, 48
for i,(T,od) in enumerate(zip(poses,odom)):
    scan = scenes.lidar_scan(scene, T, n_beams=32, n_az=360, noise_sigma=0.01, seed=900+i).astype(np.float32)
    n = scan.shape[0]
    buf = np.random.default_rng(i).integers(0, 256, size=(n, step), dtype=np.uint8)
    buf[:, :12] = scan.view(np.uint8).reshape(n, 12)
    ok, pose, st, err = m.add_range_measurement(scan.astype(np.float64), 0.1*i, od)
    print("A", i, ok, "icp_status", st.icp_status, "iters", st.iterations, "read", st.reading_points, "patch", st.patch_points, "map", st.map_points, "refreshed", st.reference_refreshed, err, flush=True)
    try:
        ctx.sync(); print("  sync ok after A", flush=True)
    except Exception as e:
        print("  sync FAILED after A", e, flush=True)
    if os.environ.get("ARRAY_B"):
        okb, poseb, stb, errb = mb.add_range_measurement(scan.astype(np.float64), 0.1*i, od)
    else:
        okb, poseb, stb, errb = mb.add_pointcloud2(buf.reshape(-1), capi.pc2_layout(n, 1, step, fields), 0.1 * i, od)
    print("B", i, okb, "icp_status", stb.icp_status, "iters", stb.iterations, "read", stb.reading_points, "patch", stb.patch_points, "map", stb.map_points, "refreshed", stb.reference_refreshed, errb, flush=True)
    try:
        ctx.sync(); print("  sync ok after B", flush=True)
    except Exception as e:
        print("  sync FAILED after B", e, flush=True)
    if not (ok and okb): break
