#!/usr/bin/env python3
"""profiles/config4_multi.py — BASELINE config #4 through the C++ multi-GPU host (b200reg_multi: ONE process, one host thread and one
b200reg_batch per visible GPU, pair i -> device i mod G, one ncclAllGather of the 96-byte records).  Prints one JSON line.
    python profiles/config4_multi.py [pairs=256] [points=200000] [steps=3]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "profiles"))
import bench_workloads as W  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import capi  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
N = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else 3
import torch  # noqa: E402  (pinned host memory only)
data = [W.config4_pair(i, N) for i in range(P)]
pin = [(torch.from_numpy(r).pin_memory(), torch.from_numpy(rn).pin_memory(), torch.from_numpy(q).pin_memory()) for r, rn, q, _ in data]
m = capi.Multi(None, **W.C4_KW)
G = m.devices
refs = capi.Batch.host_clouds([(p[0].numpy(), p[1].numpy()) for p in pin])
reads = capi.Batch.host_clouds([(p[2].numpy(), None) for p in pin])
Ti = np.stack([np.eye(4, dtype=np.float32)] * P)
ms, gms = [], []
for s in range(2 + STEPS):
    t0 = time.perf_counter()
    T, stats, status, gather_ms = m.register_host(refs, reads, Ti)
    dt = 1000.0 * (time.perf_counter() - t0)
    if s >= 2:
        ms.append(dt)
        gms.append(gather_ms)
assert not status.any()
worst_t = worst_r = 0.0
for j, (_, _, _, T_true) in enumerate(data):
    E = np.linalg.inv(T_true) @ T[j].astype(np.float64)
    worst_t, worst_r = max(worst_t, float(np.abs(E[:3, 3]).max())), max(worst_r, W._rot_err_deg(E))
assert worst_t < 0.02 and worst_r < 0.1, (worst_t, worst_r)
print(json.dumps({"workload": f"config4: {P} pairs x {N} pts, ONE process driving {G} GPU(s) through b200reg_multi_register_host (pinned host clouds in, results out of the NCCL-gathered block)",
                  "n_gpus": G, "nccl_version": m.nccl_version, "pairs_per_s": P * 1000.0 / float(np.mean(ms)), "batch_ms": {"min": min(ms), "mean": float(np.mean(ms)), "max": max(ms)},
                  "gather_ms": float(np.mean(gms)), "mean_iterations": float(np.mean([s.iterations for s in stats])), "worst_pose_error_m": worst_t,
                  "worst_pose_error_deg": worst_r, "gpu_launches": m.launch_count}))
