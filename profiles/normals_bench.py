#!/usr/bin/env python3
"""profiles/normals_bench.py — SurfaceNormal k = 20 / k = 10 on a voxel-filtered 128x1024 scan and on a map patch (config #3 / #5 shapes):
device time of index build + self-kNN + covariance/eigen kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402

c = scenes.config2(m=int(os.environ.get("PROF_MAP", 2_000_000)), n_beams=128, n_az=1024, scene_seed=1001)
f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)  # noqa: E731
ctx = capi.Context(0)
for name, pts in (("scan 131k", c["scan"]), ("map 2M", c["map_xyz"])):
    for k in (10, 20):
        cl = ctx.upload(f4(pts))
        for rep in range(3):
            cl2 = ctx.upload(f4(pts))
            ctx.timer_start()
            cl2.surface_normals(k, max_dist=1.0)
            ms = ctx.timer_stop()
            cl2.free()
        ix = ctx.build_index(cl)
        ctx.timer_start()
        ctx.knn(ix, cl, k=k, max_dist=1.0, want=False)
        kms = ctx.timer_stop()
        n = pts.shape[0]
        print(f"{name} k={k}: surface_normals {ms:.3f} ms ({n / ms / 1e3:.1f} M pts/s, {3072 * n / ms / 1e6:.1f} GB/s algorithmic at 3072 B/pt), "
              f"self-kNN alone {kms:.3f} ms ({(16 + 8 * k + 16 * (96 if k == 10 else 160)) * n / kms / 1e6:.1f} GB/s)")
