#!/usr/bin/env python3
"""profiles/index_build.py — the index build (initReference: stats -> grid -> zero -> count -> scan -> scatter, csrc/index.cu) timed
with CUDA events on the ctx stream: cold (first build: the stream-ordered pool grows) and warm (allocations come back from the pool),
against 156 B/pt.  Usage: python profiles/index_build.py [points ...]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402

sizes = [int(a) for a in sys.argv[1:]] or [200_000, 5_000_000]
peak = 6531.6
p = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(p):
    peak = float(json.load(open(p))["hbm_gbs"])
ctx = capi.Context(0)
for m in sizes:
    side = max(40.0, float(np.sqrt(m / 5e6) * 120.0))
    sc = scenes.Scene(1001, size=(side, side, 12.0))
    xyz, nrm = sc.make_map(m, 1011)
    f4 = np.concatenate([xyz.astype(np.float32), np.ones((m, 1), np.float32)], axis=1)
    ref = ctx.upload(f4, nrm.astype(np.float32))
    icp = capi.ICP(ctx, knn=1, max_dist=1.0, trim_ratio=0.8)
    ms = []
    for rep in range(6):
        ctx.flush_l2()
        ctx.timer_start()
        icp.set_reference(ref)
        ms.append(ctx.timer_stop())
    warm = float(np.median(ms[2:]))
    print(json.dumps({"points": m, "cell_m": icp.index.cell, "cold_ms": ms[0], "second_ms": ms[1], "warm_ms_median": warm, "all_ms": ms,
                      "warm_gbs_156B": 156 * m / (warm * 1e-3) / 1e9, "warm_frac": 156 * m / (warm * 1e-3) / 1e9 / peak}), flush=True)
    icp.close()
    ref.free()
