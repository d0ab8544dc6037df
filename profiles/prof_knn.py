#!/usr/bin/env python3
"""profiles/prof_knn.py — small driver for ncu captures: config #2 map + scan, one registration (to get the converged
pose), then a few k = 1 match launches at the initial and the converged pose with the L2 flushed in between.
  ncu --set full --clock-control none --import-source on -k regex:knn1 -c 4 -o gpurun_out/knn1 python profiles/prof_knn.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402

m = int(os.environ.get("PROF_MAP", 5_000_000))
c = scenes.config2(m=m, n_beams=128, n_az=1024, scene_seed=1001)
f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)  # noqa: E731
ctx = capi.Context(0)
icp = capi.ICP(ctx, knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
icp.set_reference(ctx.upload(f4(c["map_xyz"]), c["map_nrm"].astype(np.float32)))
scan = ctx.upload(f4(c["scan"]))
Ti = c["T_init"].astype(np.float32)
T, st = icp.register(scan, Ti)
Tpre = np.eye(4, dtype=np.float32)
Tpre[:3, 3] = -icp.reference_mean
for pose in (Ti, T, T):
    Tq = (Tpre.astype(np.float64) @ pose.astype(np.float64)).astype(np.float32)
    ctx.flush_l2()
    ctx.timer_start()
    _, _, visits = ctx.knn(icp.index, scan, T=Tq, k=1, max_dist=1.0, want=False)
    print("knn ms", ctx.timer_stop(), "visits/query", visits / scan.n if hasattr(scan, "n") else visits)
print("iterations", st.iterations)
