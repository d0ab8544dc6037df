#!/usr/bin/env python3
"""profiles/config4_batch.py — BASELINE config #4: batched loop-closure registration, 256 submap pairs of 200 k points each,
sharded over the GPUs of one box (pair i -> rank i mod G, one all_gather of the results; batch.py).

    python profiles/config4_batch.py [pairs=256] [points=200000]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29517 profiles/config4_batch.py

Each pair = two independent surface samplings (+ 1 cm noise) of the same procedural scene, the second moved by a random rigid
transform (<= 0.3 m per axis, <= 2 deg: what is left after the coarse global registration of PlaceRecognition.cpp:71-174).  Per pair, end to
end: upload both clouds from host memory, initReference (index build), ICP, result to the host.  Every rank checks its poses against
ground truth; rank 0 prints pairs/s (time = max over ranks between two barriers, gather included).
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from open3d_slam_advanced_rss_2024_public_b200 import batch, capi  # noqa: E402

PAIRS = int(sys.argv[1]) if len(sys.argv) > 1 else 256
POINTS = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
dist = None
device = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=device)

f4 = lambda a: np.concatenate([a.astype(np.float32), np.ones((a.shape[0], 1), np.float32)], axis=1)  # noqa: E731
N_SCENES = 4
scn = [scenes.Scene(4000 + s, size=(40.0, 40.0, 8.0), n_boxes=12, n_cyl=8) for s in range(N_SCENES)]
mine = batch.shard(PAIRS, rank, world)
data = {}
for i in mine:   # host-side inputs of this rank's pairs (pinned memory would be the production choice; these are pageable numpy arrays)
    rng = np.random.default_rng(4100 + i)
    sc = scn[i % N_SCENES]
    a, an = sc.sample_surface(POINTS, 4200 + 2 * i)
    b, _ = sc.sample_surface(POINTS, 4201 + 2 * i)
    a = a + rng.normal(0, 0.01, a.shape)
    b = b + rng.normal(0, 0.01, b.shape)
    T_true = scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3) * [0.2, 0.2, 1.0], np.deg2rad(rng.uniform(-2, 2))), rng.uniform(-0.3, 0.3, 3))
    b_local = (np.linalg.inv(T_true) @ np.concatenate([b, np.ones((b.shape[0], 1))], axis=1).T).T[:, :3]   # reading in its own frame
    data[i] = (f4(a), an.astype(np.float32), f4(b_local), T_true)

ctx = capi.Context(local)
cfg = capi.icp_config(knn=1, max_dist=1.5, trim_ratio=0.8, max_iter=40, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
icp = capi.ICP(ctx, cfg)
iters = {}


def register(i):
    ref, refn, rd, _ = data[i]
    cr = ctx.upload(ref, refn)
    cq = ctx.upload(rd)
    icp.set_reference(cr)
    T, st = icp.register(cq, np.eye(4, dtype=np.float32))
    iters[i] = st.iterations
    cr.free()
    cq.free()
    return T, 0, st.iterations


def barrier():
    ctx.sync()
    if dist is not None:
        dist.barrier()


register(mine[0])   # warm-up: scratch, graph instantiation, NCCL communicator
if dist is not None:
    batch.register_sharded(lambda i: (np.eye(4, dtype=np.float32), 0, 0), world, dist, device)
barrier()
t0 = time.perf_counter()
T, status, its, owner = batch.register_sharded(register, PAIRS, dist, device)
barrier()
dt = time.perf_counter() - t0
if dist is not None:
    import torch
    t = torch.tensor([dt], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
worst_t = worst_r = 0.0
bad = []
for i in mine:
    d = np.linalg.inv(data[i][3]) @ T[i].astype(np.float64)
    et, er = np.abs(d[:3, 3]).max(), np.degrees(np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1)))
    if et > 0.02 or er > 0.1:
        bad.append((i, round(float(et), 3), round(float(er), 3), int(its[i])))
    worst_t, worst_r = max(worst_t, et), max(worst_r, er)
assert not bad, (rank, bad)
assert (status == 0).all() and (owner == np.arange(PAIRS) % world).all()
if rank == 0:
    print(f"config #4: {PAIRS} pairs x {POINTS} points on {world} GPU(s): {dt*1e3:.1f} ms = {PAIRS / dt:.1f} pairs/s "
          f"({dt / len(mine) * 1e3:.2f} ms per pair and rank, upload + index build + ICP + gather), mean iterations {its.mean():.1f}, "
          f"rank-0 worst pose error {worst_t*100:.2f} cm / {worst_r:.3f} deg")
if dist is not None:
    dist.destroy_process_group()
