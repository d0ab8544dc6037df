#!/usr/bin/env python3
"""profiles/mapping_bench.py — BASELINE config #3 shape: open3d_slam scan-to-map mapping loop (C++ Mapper driver over the C ABI) on a
synthetic trajectory; full-size 128x1024 scans.  Prints scans/s (host wall), iterations, map size, trajectory error.
  python profiles/mapping_bench.py [n_scans]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper  # noqa: E402

n_scans = int(sys.argv[1]) if len(sys.argv) > 1 else 40
scene = scenes.Scene(2001, size=(200.0, 60.0, 10.0), n_boxes=60, n_cyl=40)
poses = []
for i in range(n_scans):                       # figure-eight, 0.2 m spacing, yaw tangent
    s = 0.2 * i
    a = 2 * np.pi * s / 120.0
    x, y = 40.0 * np.sin(a), 12.0 * np.sin(2 * a)
    dx, dy = 40.0 * np.cos(a), 24.0 * np.cos(2 * a)
    poses.append(scenes.rigid(scenes.rot_axis_angle(np.array([0, 0, 1.0]), np.arctan2(dy, dx)), np.array([x, y, 1.8])))
rng = np.random.default_rng(2003)
D = np.eye(4)
odom = []
for T in poses:                                # external odometry = truth + random-walk drift (2 mm, 0.02 deg per step)
    D = D @ scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(0.02) * rng.normal()), rng.normal(size=3) * 0.002)
    odom.append(T @ D)
scans = [scenes.lidar_scan(scene, T, n_beams=128, n_az=1024, noise_sigma=0.01, seed=2002 + i).astype(np.float64) for i, T in enumerate(poses)]
ctx = capi.Context(0)
icp = capi.icp_config(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
m = Mapper(ctx, capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=2.0, rmax=80.0), capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=80.0),
           icp, scan_voxel_size=0.1, normals_knn=20, normals_radius=1.0, map_voxel_size=0.15, min_movement=0.0, reference_period=2.0,
           carving=capi.CarveParams(0.1, 20.0, 0.1, 0.5) if os.environ.get("B200MAP_CARVE") else None, carve_every=10)
m.set_pose(poses[0])
ts, its, errs, stages, carved = [], [], [], [], []
for i, (sc, od, T) in enumerate(zip(scans, odom, poses)):
    ctx.sync()
    t0 = time.perf_counter()
    ok, pose, st, err = m.add_range_measurement(sc, 0.1 * i, od)
    ts.append(time.perf_counter() - t0)
    assert ok, err
    its.append(st.iterations)
    stages.append((st.reference_refreshed, list(st.stage_ms)))
    if st.carved_points or (os.environ.get("B200MAP_CARVE") and i % 10 == 0 and i > 0):
        carved.append((i, st.carved_points, ts[-1] * 1e3 if ts else 0.0))
    errs.append(np.abs((np.linalg.inv(T) @ pose)[:3, 3]).max())
    if i in (1, n_scans - 1):
        print(f"scan {i}: {ts[-1]*1e3:.2f} ms, reading {st.reading_points}, merge {st.merge_points}, patch {st.patch_points}, map {st.map_points}, "
              f"iterations {st.iterations}, refreshed {st.reference_refreshed}")
ts = np.array(ts[1:])
print(f"{n_scans} scans: {1.0 / ts.mean():.1f} scans/s (mean {ts.mean()*1e3:.2f} ms, median {np.median(ts)*1e3:.2f} ms), mean iterations {np.mean(its[1:]):.1f}, "
      f"max translation error vs truth {max(errs):.4f} m")
from open3d_slam_advanced_rss_2024_public_b200.mapper import MapperStats  # noqa: E402
for tag, sel in (("reference refreshed", 1), ("reference kept", 0)):
    rows = np.array([v for r, v in stages[1:] if r == sel])
    if len(rows):
        print(f"stages, {tag} ({len(rows)} scans), median ms: " + ", ".join(f"{n} {x:.2f}" for n, x in zip(MapperStats.STAGES, np.median(rows, axis=0)))
              + f" | sum {np.median(rows, axis=0).sum():.2f}")

# ---- CPU arm: the same control flow composed from the oracle's pieces (tests/test_gpu_mapper.py::OracleMapper), first N scans
n_cpu = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if n_cpu > 0:
    from oracle import oracle_py as O  # noqa: E402  (profiles/ is measurement code: the oracle is the timed CPU baseline here)
    from test_gpu_mapper import OracleMapper  # noqa: E402
    O.lib()
    om = OracleMapper(O, dict(kind=capi.CROP_MINMAX_RADIUS, rmin=2.0, rmax=80.0), dict(kind=capi.CROP_MAX_RADIUS, rmax=80.0),
                      dict(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3),
                      scan_voxel=0.1, knn=20, radius=1.0, map_voxel=0.15, min_move=0.0, period=2.0)
    om.pose = poses[0].copy()
    tc = []
    for i in range(min(n_cpu, n_scans)):
        t0 = time.perf_counter()
        ok, st = om.add(scans[i], 0.1 * i, odom[i])
        tc.append(time.perf_counter() - t0)
        assert ok
    tc = np.array(tc[1:])
    print(f"CPU oracle composition (8 threads in kNN / normals / ICP), {len(tc)} scans: {1.0 / tc.mean():.2f} scans/s (mean {tc.mean()*1e3:.0f} ms)")
if carved:
    print("space carving (every 10th inserted scan; raw 131 072-ray scan, 0.1 m steps up to 20 m): " +
          ", ".join(f"scan {i}: {c} points carved" for i, c, _ in carved))
