#!/usr/bin/env python3
"""profiles/knn_cell_sweep.py — how the k = 20 neighbour search (SurfaceNormal's self-kNN) depends on the index cell, per cloud shape:
a voxel-filtered scan (config #3's normals stage), a raw 128x1024 scan (config #5's per-scan normals) and map patches of 2 M / 8 M points.
Prints the automatic cell, points per occupied cell, and kNN time / visits per query over a sweep of cells (device time, L2 flushed)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402

K = int(os.environ.get("PROF_K", 22))
f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)  # noqa: E731
ctx = capi.Context(0)
c = scenes.config2(m=2_000_000, n_beams=128, n_az=1024, scene_seed=1001)
raw = c["scan"].astype(np.float64)
vox, _ = ctx.voxel_down_sample(0.1, raw[np.linalg.norm(raw, axis=1) > 2.0])
big = scenes.config2(m=8_000_000, n_beams=8, n_az=64, scene_seed=1001)["map_xyz"]
for name, pts in (("voxelized scan 0.1 m", vox), ("raw scan", raw), ("map 2M", c["map_xyz"]), ("map 8M", big)):
    cl = ctx.upload(f4(pts))
    n = pts.shape[0]
    auto = ctx.build_index(cl)
    a = auto.cell
    auto.free()
    print(f"== {name}: {n} points, automatic cell {a:.4f} m")
    for f in (0.25, 0.35, 0.5, 0.7, 1.0, 1.4):
        cell = a * f
        ix = ctx.build_index(cl, cell)
        occ = len(np.unique(np.floor(pts / ix.cell).astype(np.int64) @ np.array([1, 1 << 20, 1 << 40], dtype=np.int64))) if n <= 2_100_000 else 0
        ts = []
        for rep in range(3):
            ctx.flush_l2()
            ctx.timer_start()
            _, _, vis = ctx.knn(ix, cl, k=K, max_dist=1.0, want=False)
            ts.append(ctx.timer_stop())
        print(f"   cell {ix.cell:.4f} ({f:.2f} x auto)  pts/occupied cell {n / occ if occ else float('nan'):7.2f}  kNN k={K}: {min(ts):8.3f} ms  {n / min(ts) / 1e3:8.1f} M q/s  visits/query {vis / n:8.1f}")
        ix.free()
    cl.free()
