#!/usr/bin/env python3
"""profiles/knn_sweep.py — A/B driver: lanes-per-query x index cell size on BASELINE config #2; prints the kernel-only
(event-bracketed) duration of every match launch inside a registration, L2 flushed before each registration."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402

c = scenes.config2(m=int(os.environ.get("PROF_MAP", 5_000_000)), n_beams=128, n_az=1024, scene_seed=1001)
f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)  # noqa: E731
ref, rn, scan_h, Ti = f4(c["map_xyz"]), c["map_nrm"].astype(np.float32), f4(c["scan"]), c["T_init"].astype(np.float32)
lanes_list = [int(x) for x in os.environ.get("SWEEP_LANES", "8,4,1").split(",")]
cells = [float(x) for x in os.environ.get("SWEEP_CELLS", "0,0.12,0.15,0.18,0.25,0.3").split(",")]
T0 = None
for lanes in lanes_list:
    os.environ["B200REG_KNN_LANES"] = str(lanes)
    ctx = capi.Context(0)
    cref = ctx.upload(ref, rn)
    scan = ctx.upload(scan_h)
    for cell in cells:
        icp = capi.ICP(ctx, cell=cell, knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
        icp.set_reference(cref)
        for _ in range(3):
            T, st = icp.register(scan, Ti)
        if T0 is None:
            T0 = T
        assert np.array_equal(T, T0), "result depends on lanes/cell"
        ctx.set_profiling(True)
        per = []
        for _ in range(10):
            ctx.flush_l2()
            ctx.timer_start()
            icp.register(scan, Ti)
            tot = ctx.timer_stop()
            per.append(ctx.profile_read())
        ctx.set_profiling(False)
        per = np.array(per) * 1e3
        print(f"lanes {lanes} cell {icp.index.cell:.3f} iters {st.iterations} knn us/launch {np.round(per.mean(0), 1)} mean {per.mean():.1f} "
              f"touched/query/iter {st.points_touched / scan.size / st.iterations:.1f} reg_ms {tot:.3f}", flush=True)
        icp.close()
    ctx.close()
