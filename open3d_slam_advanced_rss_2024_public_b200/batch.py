"""Sharding of independent registrations over the GPUs of one box: the only multi-GPU structure this path has.

A single registration does not shard (a short chain of tiny global reductions over a working set that fits one GPU), so
N GPUs run N independent registration streams: pair i goes to rank i mod G, every rank registers its pairs on its own
device through the C ABI, and ONE all_gather of the 4x4 results (+ per-pair status and iteration count) is the path's
only exchange step — NCCL over NVLink on the GPUs, gloo in the CPU tests of this host logic.
"""
import numpy as np

RECORD = 20   # floats per pair in the gathered record: 16 (T, row-major) + status + iterations + rank + valid


def shard(n_items, rank, world):
    """Static round-robin of SURVEY 8e: the items of `rank`."""
    return list(range(rank, n_items, world))


def pack_records(items, T, status, iterations, rank, slots):
    """Fixed-size block (slots x RECORD, float32) for the gather; unused slots have valid = 0."""
    rec = np.zeros((slots, RECORD), dtype=np.float32)
    for j, _ in enumerate(items):
        rec[j, :16] = np.asarray(T[j], dtype=np.float32).reshape(16)
        rec[j, 16] = float(status[j])
        rec[j, 17] = float(iterations[j])
        rec[j, 18] = float(rank)
        rec[j, 19] = 1.0
    return rec


def unpack_records(blocks, n_items, world):
    """Inverse of the round-robin: blocks[r][j] is item r + j * world."""
    T = np.zeros((n_items, 4, 4), dtype=np.float32)
    status = np.zeros(n_items, dtype=np.int32)
    iterations = np.zeros(n_items, dtype=np.int32)
    owner = np.full(n_items, -1, dtype=np.int32)
    for r in range(world):
        for j, i in enumerate(shard(n_items, r, world)):
            rec = blocks[r][j]
            assert rec[19] == 1.0, f"item {i} missing from rank {r}"
            T[i] = rec[:16].reshape(4, 4)
            status[i], iterations[i], owner[i] = int(rec[16]), int(rec[17]), int(rec[18])
    return T, status, iterations, owner


def register_sharded(register_fn, n_items, dist=None, device=None):
    """Run register_fn(i) -> (T 4x4, status, iterations) for this rank's items and gather every rank's results.

    dist: an initialised torch.distributed module (None = single process).  device: torch device of the gather buffers
    (cuda:<local_rank> under NCCL, cpu under gloo).  Returns (T n x 4 x 4, status, iterations, owner) on every rank."""
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist is not None else (0, 1)
    mine = shard(n_items, rank, world)
    out = [register_fn(i) for i in mine]
    slots = (n_items + world - 1) // world
    rec = pack_records(mine, [o[0] for o in out], [o[1] for o in out], [o[2] for o in out], rank, slots)
    if dist is None or world == 1:
        return unpack_records([rec], n_items, 1)
    import torch
    t = torch.from_numpy(rec)
    if device is not None:
        t = t.to(device)
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    return unpack_records([g.cpu().numpy() for g in gathered], n_items, world)
