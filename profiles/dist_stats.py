import os, sys
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import scenes
from open3d_slam_advanced_rss_2024_public_b200 import capi
c = scenes.config2(m=5_000_000, n_beams=128, n_az=1024, scene_seed=1001)
f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)
ctx = capi.Context(0)
ix = ctx.build_index(ctx.upload(f4(c["map_xyz"])), 0.0)
scan = ctx.upload(f4(c["scan"]))
print("cell", ix.cell)
for name, T in (("init", c["T_init"]), ("true", c["T_true"])):
    ids, d2, v = ctx.knn(ix, scan, T=T.astype(np.float32), k=1, max_dist=1.0)
    d = np.sqrt(d2[:, 0])
    fin = np.isfinite(d)
    print(name, "no match", 1 - fin.mean(), "visits/q", v / len(d))
    for thr in (0.05, 0.1, 0.2, 0.3, 0.4, 0.6, 0.8, 1.0):
        print(f"   d<{thr}: {(d[fin] < thr).sum() / len(d):.3f}")
