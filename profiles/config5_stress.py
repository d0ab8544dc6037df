#!/usr/bin/env python3
"""profiles/config5_stress.py — BASELINE config #5: a 50 M-point voxel map, k = 1 matching (the ICP matcher) and k = 20 normal
estimation over 10 synthetic 128x1024 scans, 1 GPU.

Reports: map upload + index build time, per-scan normal estimation (k = 20, radius 1 m) and registration time, k-NN queries/s
of the standalone matcher against the 50 M map, and checks that do not need a CPU pass over the whole map:
  * a random sample of queries is compared bit for bit with the oracle's brute force over all 50 M points;
  * map points queried against their own index return themselves at distance 0 (k = 1, allow_self);
  * every registration ends within 2 cm / 0.1 deg of the ground-truth pose.
Usage: python profiles/config5_stress.py [map_points=50000000] [n_scans=10]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402
from oracle import oracle_py as O  # noqa: E402  (checker only)

M = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
N_SCANS = int(sys.argv[2]) if len(sys.argv) > 2 else 10
side = float(np.sqrt(M / 50e6) * 420.0)   # surface area grows with the point count (0.05 m de-duplication voxels)
t0 = time.perf_counter()
scene = scenes.Scene(5001, size=(side, side, 12.0), n_boxes=int(400 * M / 50e6) + 20, n_cyl=int(300 * M / 50e6) + 10)
map_xyz, map_nrm = scene.make_map(M, 5002)
print(f"map: {M} points over {side:.0f} x {side:.0f} x 12 m generated in {time.perf_counter() - t0:.1f} s", flush=True)
f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)  # noqa: E731
ref_host = f4(map_xyz)
nrm_host = map_nrm.astype(np.float32)
del map_xyz, map_nrm

ctx = capi.Context(0)
t0 = time.perf_counter()
ref = ctx.upload(ref_host, nrm_host)
ctx.sync()
t_up = time.perf_counter() - t0
icp = capi.ICP(ctx, knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=30, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
t0 = time.perf_counter()
icp.set_reference(ref)
ctx.sync()
t_build = time.perf_counter() - t0
print(f"upload {t_up*1e3:.0f} ms ({ref_host.nbytes + nrm_host.nbytes >> 20} MiB), initReference (mean, centring, index build over {M} points) "
      f"{t_build*1e3:.0f} ms = {M / t_build / 1e6:.0f} M points/s, cell {icp.index.cell:.3f} m", flush=True)

# ---- scans along a straight path through the middle of the scene
rng = np.random.default_rng(5003)
res = []
for i in range(N_SCANS):
    c = np.array([side * (0.3 + 0.4 * i / max(N_SCANS - 1, 1)), side * 0.5 + 3.0 * np.sin(i), 1.8])
    T_true = scenes.rigid(scenes.rot_axis_angle([0, 0, 1], np.deg2rad(10.0 * i)), c)
    scan = scenes.lidar_scan(scene, T_true, 128, 1024, 0.01, 5100 + i)
    T_init = (T_true @ scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(1.0)), rng.uniform(-0.1, 0.1, 3))).astype(np.float32)
    reading = ctx.upload(f4(scan))
    if i == 0:   # untimed warm-up: scratch allocation, graph instantiation
        warm = ctx.upload(f4(scan))
        warm.surface_normals(20, max_dist=1.0, orient_origin=np.zeros(3, np.float32))
        icp.register(warm, T_init)
        warm.free()
    ctx.timer_start()
    reading.surface_normals(20, max_dist=1.0, orient_origin=np.zeros(3, np.float32))
    t_nrm = ctx.timer_stop()
    ctx.flush_l2()
    ctx.timer_start()
    T, st = icp.register(reading, T_init)
    t_reg = ctx.timer_stop()
    # standalone matcher at the converged pose: k-NN queries/s against the 50 M map
    ctx.flush_l2()
    Tc = T.copy()
    Tc[:3, 3] -= icp.reference_mean   # the index is built over the centred reference
    ctx.timer_start()
    ctx.knn(icp.index, reading, T=Tc, k=1, max_dist=1.0, want=False)
    t_knn = ctx.timer_stop()
    d = np.linalg.inv(T_true) @ T.astype(np.float64)
    et = np.abs(d[:3, 3]).max()
    er = np.degrees(np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1)))
    res.append((t_nrm, t_reg, st.iterations, t_knn, et, er))
    print(f"scan {i}: normals k=20 {t_nrm:.3f} ms, registration {t_reg:.3f} ms ({st.iterations} iterations), k=1 match {t_knn*1e3:.1f} us "
          f"({scan.shape[0] / t_knn / 1e3:.0f} M queries/s), pose error {et*100:.2f} cm / {er:.3f} deg", flush=True)
    assert et < 0.02 and er < 0.1, (et, er)
    if i == 0:   # bit-exact spot check against brute force over the whole map (oracle, all host threads)
        sel = rng.choice(scan.shape[0], 256, replace=False)
        _, qT, _ = O.transform(Tc, f4(scan[sel]))
        ids_g, d2_g, _ = ctx.knn(icp.index, ctx.upload(f4(scan[sel])), T=Tc, k=1, max_dist=1.0)
        mean = icp.reference_mean
        centred = ref_host.copy()
        centred[:, :3] -= mean   # fl(p - mean), the same single rounding as the index build (ICP.cpp:320)
        tb = time.perf_counter()
        ids_o, d2_o = O.knn_brute(centred, qT, k=1, max_dist=1.0, nthreads=os.cpu_count() or 8)
        print(f"  brute force over {M} points for 256 queries: {time.perf_counter() - tb:.1f} s on the host; "
              f"ids equal {np.array_equal(ids_g, ids_o)}, d2 bits equal {np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))}", flush=True)
        assert np.array_equal(ids_g, ids_o) and np.array_equal(d2_g.view(np.uint32), d2_o.view(np.uint32))
        del centred
    reading.free()

# ---- self-match property over a slice of the map: every point finds itself at d2 = 0 (ties between duplicates -> smaller index)
sl = slice(0, 2_000_000)
q = ctx.upload(ref_host[sl])
ix = ctx.build_index(ref)
ctx.timer_start()
ids, d2, _ = ctx.knn(ix, q, k=1)
t_self = ctx.timer_stop()
assert (d2[:, 0] == 0).all() and (ids[:, 0] <= np.arange(2_000_000)).all()
same = ids[:, 0] == np.arange(2_000_000)
assert np.array_equal(ref_host[ids[~same, 0], :3], ref_host[sl][~same, :3])
print(f"self-match of 2 M map points against the {M}-point index: all at d2 = 0 ({(~same).sum()} exact duplicates -> smaller index), {t_self:.2f} ms incl. D2H")
r = np.array(res)
print(f"config #5 over {N_SCANS} scans: normals k=20 {r[:, 0].mean():.3f} ms/scan, registration {r[:, 1].mean():.3f} ms/scan "
      f"({r[:, 2].mean():.1f} iterations), k=1 match {r[:, 3].mean()*1e3:.1f} us = {131072 / r[:, 3].mean() / 1e3:.0f} M queries/s, "
      f"{536 * 131072 / r[:, 3].mean() / 1e6:.0f} GB/s algorithmic")
