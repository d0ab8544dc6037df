"""N > 1 host logic on CPU: world_size-2 `gloo` run of the sharding + gather of independent registrations.
The per-pair registration inside the test is the CPU oracle (the GPU path is exercised by test_gpu_icp.py); what is checked
here is the partition (pair i -> rank i mod G), the fixed-size records, the all_gather and the re-ordering."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from open3d_slam_advanced_rss_2024_public_b200 import batch  # noqa: E402

N_PAIRS = 5


def _pairs():
    """Small independent registration problems: a random rigid motion of a noisy box sampling."""
    rng = np.random.default_rng(7)
    out = []
    for i in range(N_PAIRS):
        m = 3000
        face = rng.integers(0, 6, m)
        p = rng.uniform(-1, 1, (m, 3)) * np.array([2.0, 1.5, 1.0])
        nrm = np.zeros((m, 3))
        ext = np.array([2.0, 1.5, 1.0])
        for a in range(3):
            for sgn, f in ((-1, 2 * a), (1, 2 * a + 1)):
                sel = face == f
                p[sel, a] = sgn * ext[a]
                nrm[sel, a] = sgn
        ang = 0.03 * (i + 1)
        c, s = np.cos(ang), np.sin(ang)
        T = np.eye(4)
        T[:3, :3] = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
        T[:3, 3] = [0.02 * i, -0.01, 0.015]
        q = (p[::2] - T[:3, 3]) @ T[:3, :3]      # reading = T^-1 * reference subset
        out.append((p.astype(np.float32), nrm.astype(np.float32), q.astype(np.float32), T))
    return out


def _register(i, pairs):
    from oracle import oracle_py as O
    ref, nrm, q, _ = pairs[i]
    cfg = O.icp_config(precision=O.ORC_MIXED, nthreads=1, knn=1, max_dist=1.0, trim_ratio=0.9, max_iter=30)
    code, T, st = O.icp(cfg, O.feat4(ref), nrm, O.feat4(q), None, np.eye(4, dtype=np.float32))
    return T.astype(np.float32), code, st.iterations


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pairs = _pairs()
    T, status, iters, owner = batch.register_sharded(lambda i: _register(i, pairs), N_PAIRS, dist=dist)
    q.put((rank, T, status, iters, owner))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_is_round_robin_and_covers_everything():
    for world in (1, 2, 3, 8):
        seen = sorted(i for r in range(world) for i in batch.shard(11, r, world))
        assert seen == list(range(11))
        assert batch.shard(11, 1 % world, world)[0] == 1 % world
    assert batch.shard(0, 0, 4) == []


def test_records_round_trip_with_ragged_shards():
    n, world = 7, 4
    blocks = []
    for r in range(world):
        items = batch.shard(n, r, world)
        T = [np.full((4, 4), 10 * i, dtype=np.float32) for i in items]
        blocks.append(batch.pack_records(items, T, [-(i % 2) for i in items], [i + 1 for i in items], r, (n + world - 1) // world))
    T, status, iters, owner = batch.unpack_records(blocks, n, world)
    for i in range(n):
        assert T[i, 0, 0] == 10 * i and status[i] == -(i % 2) and iters[i] == i + 1 and owner[i] == i % world


@pytest.mark.timeout(300)
def test_world_size_2_gloo_matches_serial():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(2)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    pairs = _pairs()
    T_ser, st_ser, it_ser, _ = batch.register_sharded(lambda i: _register(i, pairs), N_PAIRS)
    for rank, T, status, iters, owner in got:
        assert np.array_equal(T, T_ser) and np.array_equal(status, st_ser) and np.array_equal(iters, it_ser)
        assert list(owner) == [i % 2 for i in range(N_PAIRS)]
    for i, (_, _, _, T_true) in enumerate(pairs):      # and the registrations themselves make sense
        assert st_ser[i] == 0
        assert np.abs(T_ser[i] - T_true).max() < 5e-3
