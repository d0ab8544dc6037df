#!/usr/bin/env python3
"""profiles/loop_trace.py — per-kernel timeline of the graph-resident registration loop on BASELINE config #2, from the
device-side %globaltimer stamps (b200reg_ctx_set_trace); L2 flushed before each registration."""
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
ctx = capi.Context(0)
icp = capi.ICP(ctx, knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3,
              cell=float(os.environ.get("PROF_CELL", 0.0)))
icp.set_reference(ctx.upload(f4(c["map_xyz"]), c["map_nrm"].astype(np.float32)))
Ti = c["T_init"].astype(np.float32)
pts = c["scan"]
if os.environ.get("PROF_SORT"):   # experiment: queries in the index's cell order (x fastest) under T_init instead of sensor order
    cell = float(os.environ.get("PROF_SORT_CELL", 0.208))
    w = (pts.astype(np.float64) @ Ti[:3, :3].T.astype(np.float64) + Ti[:3, 3]) / cell
    ijk = np.floor(w).astype(np.int64)
    ijk -= ijk.min(0)
    mode = os.environ["PROF_SORT"]
    if mode == "tile":   # the [128 beams][1024 azimuths] grid re-ordered into tiles of tb x ta rays (a warp's queries form a patch, not a line)
        tb, ta = (int(v) for v in os.environ.get("PROF_TILE", "4x8").split("x"))
        idx = np.arange(128 * 1024).reshape(128, 1024)
        key = np.empty(128 * 1024, dtype=np.int64)
        key[idx.reshape(128 // tb, tb, 1024 // ta, ta).transpose(0, 2, 1, 3).reshape(-1)] = np.arange(128 * 1024)
    elif mode.startswith("stride"):   # chunks of 8 consecutive queries dealt out in a strided order: runs of far queries are spread over the launch
        ch = int(os.environ.get("PROF_CHUNK", 8))
        nchunk = pts.shape[0] // ch
        order = (np.arange(nchunk, dtype=np.int64) * 40503) % nchunk
        assert len(np.unique(order)) == nchunk
        key = np.empty(pts.shape[0], dtype=np.int64)
        key[:nchunk * ch] = np.repeat(np.argsort(order), ch) * ch + np.tile(np.arange(ch), nchunk)
        key[nchunk * ch:] = np.arange(nchunk * ch, pts.shape[0])
    elif mode == "morton":
        def spread(v):
            r = np.zeros_like(v)
            for b in range(12):
                r |= ((v >> b) & 1) << (3 * b)
            return r
        key = spread(ijk[:, 0]) | (spread(ijk[:, 1]) << 1) | (spread(ijk[:, 2]) << 2)
    else:
        key = ijk[:, 0] + (ijk[:, 1] << 14) + (ijk[:, 2] << 28)
    pts = pts[np.argsort(key, kind="stable")]
scan = ctx.upload(f4(pts))
for _ in range(3):
    icp.register(scan, Ti)
ctx.set_trace(True)
NAMES = {1: "match", 2: "select0", 3: "select1", 4: "select2", 5: "reduce", 6: "tail", 7: "tail_end"}
acc = {}
for rep in range(10):
    ctx.flush_l2()
    ctx.timer_start()
    T, st = icp.register(scan, Ti)
    tot = ctx.timer_stop()
    tr = ctx.trace_read()
    t0 = tr[0][1]
    line = []
    for i, (tag, t) in enumerate(tr):
        nxt = tr[i + 1][1] if i + 1 < len(tr) else t
        acc.setdefault((i, NAMES.get(tag, tag)), []).append((nxt - t) / 1e3)
    if rep == 9:
        print(f"registration {tot*1e3:.1f} us event-timed, {st.iterations} iterations, trace span {(tr[-1][1]-t0)/1e3:.1f} us")
for (i, name), v in acc.items():
    print(f"  {i:3d} {name:9s} start->next start {np.mean(v):7.1f} us")
print(f"host: reading_preprocessing {st.reading_preprocessing_s*1e6:.1f} us, convergence (graph launch -> result on host) {st.convergence_s*1e6:.1f} us, "
      f"device loop {st.device_ms*1e3:.1f} us")
import time
ts = []
for _ in range(20):
    ctx.sync()
    t0 = time.perf_counter()
    icp.register(scan, Ti)
    ts.append(time.perf_counter() - t0)
print(f"host wall per register (L2 warm): min {min(ts)*1e6:.1f} us median {sorted(ts)[10]*1e6:.1f} us")
