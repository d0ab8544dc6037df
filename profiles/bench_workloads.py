"""profiles/bench_workloads.py — the BASELINE configurations other than #2 as `bench.py --workload {config3,config4,config5}`:
each prints ONE JSON line in bench.py's format (roofline, cpu_baseline, e2e) on the inputs SURVEY.md §8d defines.  These are not
the driver's bench line (that is config #2); their raw outputs are committed under profiles/r2/.

  config3  open3d_slam mapping loop over a 500-scan figure-eight (C++ Mapper driver over the C ABI)            scans/s
  config4  batched loop-closure registration, 256 pairs x 200 k points, pair i -> rank i mod G, one all_gather   pairs/s
  config5  50 M-point map: index build, k = 1 matching + k = 20 normals over 10 scans, normals over the map      scans/s, queries/s
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench as B  # noqa: E402  (helpers: peaks, clock sampler, byte constants)


def run(args, rank, local_rank, world):
    fn = {"config3": run_config3, "config4": run_config4, "config5": run_config5}[args.workload]
    return fn(args, rank, local_rank, world)


def _rot_err_deg(E):
    return float(np.degrees(np.arccos(np.clip((np.trace(E[:3, :3]) - 1) / 2, -1, 1))))


# =====================================================================================================================
# config #4 — 256 submap pairs x 200 k points sharded over the GPUs
# =====================================================================================================================
C4_KW = dict(knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)


def config4_pair(i, points):
    """Pair i of SURVEY 8d row 4: scene seed 3000 + i (40 x 40 x 8 m room), two area-uniform samplings with analytic normals, the
    second expressed in a frame moved by a random rigid T (|t| <= 0.5 m, angle <= 5 deg); T_init = I."""
    import scenes
    rng = np.random.default_rng(3500 + i)
    sc = scenes.Scene(3000 + i, size=(40.0, 40.0, 8.0), n_boxes=12, n_cyl=8)
    a, an = sc.sample_surface(points, 3100 + 2 * i)
    b, _ = sc.sample_surface(points, 3101 + 2 * i)
    t = rng.normal(size=3)
    t *= rng.uniform(0, 0.5) / np.linalg.norm(t)
    T_true = scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(rng.uniform(0, 5.0))), t)
    b_local = (b - T_true[:3, 3]) @ T_true[:3, :3]      # reading in its own frame: p_read = R^T (p - t)
    return B.feat4(a), np.ascontiguousarray(an, dtype=np.float32), B.feat4(b_local), T_true


def reference_config4(args, world):
    from oracle import oracle_py as O
    pairs = [config4_pair(i, args.pair_points) for i in range(min(args.warmup + args.steps, 8))]
    cores = B.pick_threads(O, pairs[0][0], pairs[0][2])
    cfg = O.icp_config(precision=O.ORC_REF32, nthreads=cores, **C4_KW)
    ts, its = [], []
    for s in range(args.warmup + args.steps):
        ref, rn, rd, _ = pairs[s % len(pairs)]
        t0 = time.time()
        code, T, st = O.icp(cfg, ref, rn, rd, None, np.eye(4, dtype=np.float32))    # kd-tree build + ICP, like initReference + compute
        if s >= args.warmup:
            ts.append(time.time() - t0)
            its.append(st.iterations)
    val = 1.0 / float(np.mean(ts))
    return {"metric": "batched loop-closure registration (256 submap pairs x 200k pts)", "value": val, "unit": "pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 / val, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"config4: {args.pairs} submap pairs x {args.pair_points} pts", "step": "one pair (kd-tree build + ICP)", "iterations": float(np.mean(its))},
            "cpu_baseline": {"value": val, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": f"{len(ts)} pairs, kd-tree build + ICP each"},
            "e2e": {"value": val, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_config4(args, rank, local_rank, world):
    if args.impl == "reference":
        return reference_config4(args, world) if rank == 0 else None
    torch, dist = B.dist_setup(world, local_rank)
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    P, N = args.pairs, args.pair_points
    mine = list(range(rank, P, world))
    t0 = time.time()
    data = [config4_pair(i, N) for i in mine]
    gen_s = time.time() - t0
    # pinned host staging (what a loop-closure module that owns its submaps would keep them in)
    pin = [(torch.from_numpy(r).pin_memory(), torch.from_numpy(rn).pin_memory(), torch.from_numpy(q).pin_memory()) for r, rn, q, _ in data]
    ctx = capi.Context(local_rank)
    batch = capi.Batch(ctx, lanes=args.lanes, **C4_KW)
    refs_h = batch.host_clouds([(p[0].numpy(), p[1].numpy()) for p in pin])
    reads_h = batch.host_clouds([(p[2].numpy(), None) for p in pin])
    Ti = np.stack([np.eye(4, dtype=np.float32)] * len(mine))
    peak, peak_src = B.hbm_peak()

    def barrier():
        ctx.sync()
        if dist is not None:
            dist.barrier()

    def gather(T, status, iters):
        """the path's only exchange step: ONE all_gather of the fixed-size records (16 + status + iterations + owner + valid)"""
        slots = (P + world - 1) // world
        rec = np.zeros((slots, 20), dtype=np.float32)
        rec[:len(mine), :16] = T.reshape(-1, 16)
        rec[:len(mine), 16] = status
        rec[:len(mine), 17] = iters
        rec[:len(mine), 18] = rank
        rec[:len(mine), 19] = 1.0
        if dist is None:
            return rec[None]
        t = torch.from_numpy(rec).to(torch.device("cuda", local_rank))
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return np.stack([o.cpu().numpy() for o in out])

    # ---- warm-up: lanes, graphs, pool, communicator
    nw = min(len(mine), 2 * batch.lanes)
    batch.register_host(batch.host_clouds([(p[0].numpy(), p[1].numpy()) for p in pin[:nw]]), batch.host_clouds([(p[2].numpy(), None) for p in pin[:nw]]), Ti[:nw])
    gather(np.zeros((len(mine), 4, 4), np.float32), np.zeros(len(mine)), np.zeros(len(mine)))
    # ---- resident arm (`value`): this rank's clouds already on the device
    refs_d = [ctx.upload(r, rn) for r, rn, _, _ in data]
    reads_d = [ctx.upload(q) for _, _, q, _ in data]
    sampler = B.ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    res_ms, e2e_ms = [], []
    launches0 = ctx.launch_count
    for s in range(args.warmup + args.steps):
        ctx.flush_l2()
        barrier()
        t0 = time.perf_counter()
        T, stats, status = batch.register(refs_d, reads_d, Ti)
        rec = gather(T, status, [st.iterations for st in stats])
        barrier()
        if s >= args.warmup:
            res_ms.append(1000.0 * (time.perf_counter() - t0))
    launches = ctx.launch_count - launches0
    # ---- end-to-end arm: host (pinned) clouds in, transforms out; upload + index build + ICP + gather inside the timed region
    for s in range(args.warmup + args.steps):
        ctx.flush_l2()
        barrier()
        t0 = time.perf_counter()
        T_e, stats_e, status_e = batch.register_host(refs_h, reads_h, Ti)
        rec = gather(T_e, status_e, [st.iterations for st in stats_e])
        barrier()
        if s >= args.warmup:
            e2e_ms.append(1000.0 * (time.perf_counter() - t0))
    clocks = sampler.stop() if sampler else None
    assert np.array_equal(T_e, T) and not status.any() and not status_e.any()
    iters = np.array([st.iterations for st in stats])
    worst_t = worst_r = 0.0
    for j, (_, _, _, T_true) in enumerate(data):
        E = np.linalg.inv(T_true) @ T[j].astype(np.float64)
        worst_t, worst_r = max(worst_t, float(np.abs(E[:3, 3]).max())), max(worst_r, _rot_err_deg(E))
    assert worst_t < 0.02 and worst_r < 0.1, (rank, worst_t, worst_r)
    # every rank holds every pair's record
    assert (rec[:, :, 19].sum() == P) and all(int(rec[i % world, i // world, 18]) == i % world for i in range(P))
    per_rank = B.gather_floats(torch, dist, local_rank, [float(np.mean(res_ms)), float(np.mean(e2e_ms)), float(iters.sum()), worst_t, worst_r, len(mine)])
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return None
    ms_res, ms_e2e = float(per_rank[:, 0].max()), float(per_rank[:, 1].max())
    total_iters = float(per_rank[:, 2].sum())
    alg_bytes = total_iters * N * B.B_ITER + P * N * B.B_BUILD
    line = {"metric": "batched loop-closure registration (256 submap pairs x 200k pts)", "value": P * 1000.0 / ms_res, "unit": "pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"config4: {P} submap pairs x {N} pts (SURVEY 8d row 4: scenes 3000+i, |t| <= 0.5 m, angle <= 5 deg, T_init = I, maxDist 1.0, TrimmedDist 0.8)",
                       "step": f"the whole batch: pair i -> rank i mod {world}, {batch.lanes} lanes in flight per GPU, one all_gather of the records",
                       "mean_iterations": total_iters / P, "worst_pose_error_m": float(per_rank[:, 3].max()), "worst_pose_error_deg": float(per_rank[:, 4].max()),
                       "l2": "flushed between steps", "per_rank_ms": B.spread(per_rank[:, 0]), "per_rank_e2e_ms": B.spread(per_rank[:, 1]),
                       "step_ms_rank0": B.spread(res_ms), "e2e_step_ms_rank0": B.spread(e2e_ms), "gen_s": gen_s},
            "roofline": {"bound": "hbm", "kernel": "whole pair pipeline (index build 156 B/pt + 600 B/pt per ICP iteration), all lanes",
                         "achieved": alg_bytes / (ms_res * 1e-3) / 1e9 / world, "peak": peak, "unit": "GB/s",
                         "frac": alg_bytes / (ms_res * 1e-3) / 1e9 / world / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_batch": alg_bytes, "note": "per GPU: total algorithmic bytes / batch time / n_gpus"},
            "e2e": {"value": P * 1000.0 / ms_e2e, "unit": "pairs/s", "h2d_bytes_per_step": int(P * (N * 16 * 2 + N * 12 + 2 * 3072)), "d2h_bytes_per_step": P * 144,
                    "ms_per_step": ms_e2e},
            "gpu_launches": int(launches), "clocks": clocks}
    if not args.no_cpu_baseline:
        from oracle import oracle_py as O
        cores = B.pick_threads(O, data[0][0], data[0][2])
        cfg = O.icp_config(precision=O.ORC_REF32, nthreads=cores, **C4_KW)
        ts, dT = [], 0.0
        for j in range(min(4, len(data))):
            t0 = time.time()
            code, T_o, st_o = O.icp(cfg, data[j][0], data[j][1], data[j][2], None, np.eye(4, dtype=np.float32))
            ts.append(time.time() - t0)
            dT = max(dT, float(np.abs(T_o[:3, 3] - T[j][:3, 3]).max()))
        line["cpu_baseline"] = {"value": 1.0 / float(np.mean(ts)), "unit": "pairs/s", "cores": cores, "kind": "port",
                                "sample": f"{len(ts)} pairs of the batch, kd-tree build + ICP each; mean", "gpu_vs_cpu_translation_diff_m": dT}
    if dist is not None:
        dist.destroy_process_group()
    return line


# =====================================================================================================================
# config #3 — 500-scan mapping loop
# =====================================================================================================================
C3_ICP = dict(knn=1, max_dist=0.5, trim_ratio=0.9, max_angle=1.57, max_iter=15, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
_C3_SCENE = None


def _c3_scan(job):
    import scenes
    global _C3_SCENE
    if _C3_SCENE is None:
        _C3_SCENE = scenes.Scene(2001, size=(200.0, 60.0, 10.0), n_boxes=60, n_cyl=40)
    i, T = job
    return scenes.lidar_scan(_C3_SCENE, T, n_beams=128, n_az=1024, noise_sigma=0.01, seed=2002 + i).astype(np.float64)


def config3_inputs(n_scans):
    import multiprocessing as mp

    import scenes
    poses = []
    for i in range(n_scans):                       # figure-eight, 0.2 m spacing, yaw tangent
        s = 0.2 * i
        a = 2 * np.pi * s / 120.0
        x, y = 40.0 * np.sin(a), 12.0 * np.sin(2 * a)
        dx, dy = 40.0 * np.cos(a), 24.0 * np.cos(2 * a)
        poses.append(scenes.rigid(scenes.rot_axis_angle(np.array([0, 0, 1.0]), np.arctan2(dy, dx)), np.array([x, y, 1.8])))
    rng = np.random.default_rng(2003)
    D = np.eye(4)
    odom = []
    for T in poses:                                # external odometry = truth + random-walk drift (2 mm, 0.02 deg per step)
        D = D @ scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(0.02) * rng.normal()), rng.normal(size=3) * 0.002)
        odom.append(T @ D)
    with mp.get_context("fork").Pool(min(os.cpu_count() or 1, 32)) as pool:   # before any CUDA call in this process
        scans = pool.map(_c3_scan, list(enumerate(poses)), chunksize=4)
    return poses, odom, scans


def run_config3(args, rank, local_rank, world):
    if rank != 0:
        return None
    n_scans = args.scans
    t0 = time.time()
    poses, odom, scans = config3_inputs(n_scans)
    gen_s = time.time() - t0
    if args.impl == "reference":
        return _config3_cpu(args, poses, odom, scans, world, full_line=True)
    import torch
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from open3d_slam_advanced_rss_2024_public_b200.mapper import Mapper, MapperStats
    ctx = capi.Context(local_rank)
    peak, peak_src = B.hbm_peak()

    def one_pass(inputs, prefetch=False):
        m = Mapper(ctx, capi.crop_spec(capi.CROP_MINMAX_RADIUS, np.zeros(3), rmin=2.0, rmax=80.0), capi.crop_spec(capi.CROP_MAX_RADIUS, np.zeros(3), rmax=80.0),
                   capi.icp_config(**C3_ICP), scan_voxel_size=0.1, normals_knn=20, normals_radius=1.0, map_voxel_size=0.15, min_movement=0.0,
                   reference_period=2.0)
        m.set_pose(poses[0])
        ts, traj, sts = [], [], []
        launches0 = ctx.launch_count
        for i, (sc, od) in enumerate(zip(inputs, odom)):
            ctx.sync()
            t0 = time.perf_counter()
            if prefetch and i == 0 and len(inputs) > 1:
                m.prefetch(inputs[1])
            if prefetch and i + 2 < len(inputs):
                m.prefetch(inputs[i + 2])                      # the Mapper's two workers upload and process scans i + 1 and i + 2 while scan i is registered
            ok, pose, st, err = m.add_range_measurement(sc, 0.1 * i, od)
            ts.append(time.perf_counter() - t0)
            assert ok, err
            traj.append(pose)
            sts.append((st.reference_refreshed, list(st.stage_ms), st.iterations, st.reading_points, st.merge_points, st.patch_points, st.map_points))
        n_launch = ctx.launch_count - launches0
        m.close()
        return np.array(ts), traj, sts, n_launch

    one_pass(scans[:8])                                            # warm-up: graphs, lazy module loading
    ts_cold, _, _, _ = one_pass(scans)                             # first full pass: the stream-ordered pool grows with the map (reported, not `value`)
    sampler = B.ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    ts, traj, sts, launches = one_pass(scans)                      # `value`: pageable numpy scans, pool warm (every block size of the loop cached)
    pinned = [torch.from_numpy(s).pin_memory().numpy() for s in scans]
    ts_e, traj_e, _, _ = one_pass(pinned, prefetch=True)           # `e2e`: pinned scans, the next one uploaded and processed ahead by the Mapper's worker
    clocks = sampler.stop()
    assert all(np.array_equal(a, b) for a, b in zip(traj, traj_e))
    err = [float(np.linalg.norm((np.linalg.inv(T) @ p)[:3, 3])) for T, p in zip(poses, traj)]
    ate = float(np.sqrt(np.mean(np.square(err))))
    # per-stage medians (reference kept) and their algorithmic-bytes fractions
    rows = np.array([v for r, v, *_ in sts[1:] if r == 0])
    med = np.median(rows, axis=0)
    info = np.array([[it, rd, mg, pt, mp] for _, _, it, rd, mg, pt, mp in sts[1:]], dtype=np.float64).mean(0)
    raw_n = scans[0].shape[0]
    stage_bytes = {"wide_crop": raw_n * 24 + info[2] * 24 * 1.3, "voxelize": raw_n * 0.8 * B.B_VOXEL, "normals": info[2] * B.B_NORMALS,
                   "register": info[0] * info[1] * (B.B_ITER + 16), "insert": info[2] * (48 + 12 + 16 * 8) + info[3] * 96}
    stages = {}
    for name, ms in zip(MapperStats.STAGES, med):
        b = stage_bytes.get(name)
        stages[name] = {"median_ms": float(ms), "algorithmic_bytes": None if b is None else float(b),
                        "frac": None if b is None or ms <= 0 else float(b / (ms * 1e-3) / 1e9 / peak)}
    refresh = np.array([v for r, v, *_ in sts[1:] if r == 1])
    val = 1.0 / float(ts[1:].mean())
    val_e = 1.0 / float(ts_e[1:].mean())
    line = {"metric": "open3d_slam scan-to-map mapping loop (500-scan synthetic trajectory)", "value": val, "unit": "scans/s", "n_gpus": 1,
            "steps": n_scans, "warmup": 8, "ms_per_step": 1000.0 / val, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64 map / f32 ICP", "data": "synthetic",
            "config": {"workload": f"config3: {n_scans}-scan figure-eight (scene 2001, 200 x 60 x 10 m, 0.2 m spacing), 128x1024 scans: crop 2-80 m, voxel 0.1, normals k=20 r=1, ICP (maxDist 0.5, TrimmedDist 0.9, SurfaceNormal 1.57, Counter 15), insert at 0.15 m, reference every 20 scans",
                       "step": "one Mapper::addRangeMeasurement (host scan in, pose out)", "scan_ms": B.spread(ts[1:] * 1e3), "ate_rmse_m": ate, "max_translation_error_m": max(err),
                       "mean_iterations": float(info[0]), "mean_reading_points": float(info[1]), "mean_patch_points": float(info[3]), "final_map_points": int(sts[-1][6]),
                       "stages_reference_kept": stages, "stage_sum_ms": float(med.sum()),
                       "reference_refresh_scan_ms_median": None if not len(refresh) else float(np.median(refresh.sum(1))), "gen_s": gen_s,
                       "first_pass": {"scans_per_s": float(len(ts_cold) / ts_cold.sum()), "scan_ms": B.spread(1000.0 * ts_cold),
                                      "note": "the pass before the timed one: the map, patch and scratch allocations grow with the map and the device's stream-ordered pool has to grow with them (a handful of multi-millisecond steps); afterwards the pool holds every block size"},
                       "value_vs_e2e": "the Mapper API takes host scans, so both arms include the upload: value = pageable numpy input, inline processing; e2e = pinned input, every next scan handed to Mapper.prefetch (uploaded and processed ahead on the worker)"},
            "roofline": {"bound": "hbm", "kernel": "normals stage (knnListKernel<32> self-kNN k=20 + normalsKernel) — the largest stage of a mapping step",
                         "achieved": stages["normals"]["frac"] * peak, "peak": peak, "unit": "GB/s", "frac": stages["normals"]["frac"], "traffic": None,
                         "peak_source": peak_src, "note": "3072 B/pt x mean merge-cloud size / median stage time (host wall around the synchronising call)"},
            "e2e": {"value": val_e, "unit": "scans/s", "h2d_bytes_per_step": int(raw_n * 24), "d2h_bytes_per_step": 128 + 144, "ms_per_step": 1000.0 / val_e},
            "gpu_launches": int(launches), "clocks": clocks}
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = _config3_cpu(args, poses, odom, scans, world, full_line=False, gpu_traj=traj)
    return line


def _config3_cpu(args, poses, odom, scans, world, full_line, gpu_traj=None):
    from oracle import oracle_py as O
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from test_gpu_mapper import OracleMapper
    O.lib()
    n_cpu = min(len(scans), 12 if not full_line else max(4, args.steps))
    om = OracleMapper(O, dict(kind=capi.CROP_MINMAX_RADIUS, rmin=2.0, rmax=80.0), dict(kind=capi.CROP_MAX_RADIUS, rmax=80.0), dict(C3_ICP),
                      scan_voxel=0.1, knn=20, radius=1.0, map_voxel=0.15, min_move=0.0, period=2.0)
    om.pose = poses[0].copy()
    tc, dmax = [], 0.0
    for i in range(n_cpu):
        t0 = time.perf_counter()
        ok, st = om.add(scans[i], 0.1 * i, odom[i])
        tc.append(time.perf_counter() - t0)
        assert ok
        if gpu_traj is not None:
            dmax = max(dmax, float(np.abs(om.pose[:3, 3] - gpu_traj[i][:3, 3]).max()))
    val = 1.0 / float(np.mean(tc[1:]))
    cb = {"value": val, "unit": "scans/s", "cores": 8, "kind": "port",
          "sample": f"first {n_cpu} scans of the trajectory through the oracle's composition of the same steps (8 threads in kNN / normals / ICP)",
          "gpu_vs_cpu_translation_diff_m": dmax}
    if not full_line:
        return cb
    return {"metric": "open3d_slam scan-to-map mapping loop (500-scan synthetic trajectory)", "value": val, "unit": "scans/s", "n_gpus": world,
            "steps": n_cpu, "warmup": 1, "ms_per_step": 1000.0 / val, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "impl": "reference", "config": {"workload": "config3: mapping loop", "step": "one scan"}, "cpu_baseline": cb,
            "e2e": {"value": val, "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# =====================================================================================================================
# config #5 — 50 M-point map stress
# =====================================================================================================================
def run_config5(args, rank, local_rank, world):
    if rank != 0:
        return None
    import scenes
    M, n_scans = args.big_map_points, (10 if args.scans == 500 else args.scans)
    side = float(np.sqrt(M / 50e6) * 400.0)
    t0 = time.time()
    scene = scenes.Scene(4001, size=(side, side, 15.0), n_boxes=int(600 * M / 50e6) + 20, n_cyl=int(400 * M / 50e6) + 10)
    map_xyz, map_nrm = scene.make_map(M, 4003, dedupe_voxel=0.03)
    ref_host, nrm_host = B.feat4(map_xyz), np.ascontiguousarray(map_nrm, dtype=np.float32)
    del map_xyz, map_nrm
    rng = np.random.default_rng(4002)
    scans, T_true, T_init = [], [], []
    for i in range(n_scans):
        c = np.array([side * rng.uniform(0.25, 0.75), side * rng.uniform(0.25, 0.75), 1.8])
        Tt = scenes.rigid(scenes.rot_axis_angle([0, 0, 1], rng.uniform(0, 2 * np.pi)), c)
        scans.append(B.feat4(scenes.lidar_scan(scene, Tt, 128, 1024, 0.01, 4100 + i)))
        T_true.append(Tt)
        T_init.append((Tt @ scenes.rigid(scenes.rot_axis_angle(rng.normal(size=3), np.deg2rad(1.0)), rng.uniform(-0.1, 0.1, 3))).astype(np.float32))
    gen_s = time.time() - t0
    kw = dict(knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=30, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
    if args.impl == "reference":
        return _config5_cpu(args, ref_host, nrm_host, scans, T_init, kw, world, full_line=True)
    import torch
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    from oracle import oracle_py as O   # checker: brute force over the whole map for a sample of queries
    ctx = capi.Context(local_rank)
    peak, peak_src = B.hbm_peak()
    t0 = time.perf_counter()
    ref = ctx.upload(ref_host, nrm_host)
    ctx.sync()
    upload_s = time.perf_counter() - t0
    icp = capi.ICP(ctx, **kw)
    build_ms = []
    for _ in range(4):
        ctx.flush_l2()
        ctx.timer_start()
        icp.set_reference(ref)
        build_ms.append(ctx.timer_stop())
    build_warm = float(np.median(build_ms[1:]))
    n = scans[0].shape[0]
    pinned = [torch.from_numpy(s).pin_memory().numpy() for s in scans]
    work = ctx.upload(scans[0])
    for _ in range(2):   # warm-up
        work.update(pinned[0])
        work.surface_normals(20, max_dist=1.0, orient_origin=np.zeros(3, np.float32))
        work.update(pinned[0])
        icp.register(work, T_init[0])
    sampler = B.ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    launches0 = ctx.launch_count
    per = []
    for arm in ("resident", "e2e"):
        for i in range(n_scans):
            cs = ctx.upload(scans[i]) if arm == "resident" else None
            ctx.flush_l2()
            ctx.sync()
            t0 = time.perf_counter()
            if arm == "e2e":
                work.update(pinned[i], wait=False)
                cs = work
            cs.surface_normals(20, max_dist=1.0, orient_origin=np.zeros(3, np.float32))      # (iii) SurfaceNormal k = 20 on the scan
            t1 = time.perf_counter()
            T, st = icp.register(cs, T_init[i])                                               # (ii) k = 1 matching inside the ICP
            t2 = time.perf_counter()
            E = np.linalg.inv(T_true[i]) @ T.astype(np.float64)
            per.append((arm, (t1 - t0) * 1e3, (t2 - t1) * 1e3, st.iterations, float(np.abs(E[:3, 3]).max()), _rot_err_deg(E), T))
    launches = ctx.launch_count - launches0
    clocks = sampler.stop()
    res = [p for p in per if p[0] == "resident"]
    e2e = [p for p in per if p[0] == "e2e"]
    assert all(np.array_equal(a[6], b[6]) for a, b in zip(res, e2e))
    assert max(p[4] for p in res) < 0.02 and max(p[5] for p in res) < 0.1, [(p[4], p[5]) for p in res]
    # (ii) standalone k = 1 queries at the converged poses, event-timed, L2 flushed
    mean = icp.reference_mean
    Tpre = np.eye(4, dtype=np.float32)
    Tpre[:3, 3] = -mean
    knn_ms, visits = [], 0
    for i in range(n_scans):
        cs = ctx.upload(scans[i])
        Tq = (Tpre.astype(np.float64) @ res[i][6].astype(np.float64)).astype(np.float32)
        ctx.knn(icp.index, cs, T=Tq, k=1, max_dist=1.0, want=False)
        ctx.flush_l2()
        ctx.timer_start()
        _, _, v = ctx.knn(icp.index, cs, T=Tq, k=1, max_dist=1.0, want=False)
        knn_ms.append(ctx.timer_stop())
        visits += v
    # parity at this size: a sample of queries against brute force over all M points; map points find themselves
    rs = np.random.default_rng(7)
    Tw = res[0][6].astype(np.float64)
    qs = B.feat4((scans[0][rs.choice(n, 128, replace=False), :3].astype(np.float64) @ Tw[:3, :3].T + Tw[:3, 3]))
    big = ctx.build_index(ref)
    ids_g, d2_g, _ = ctx.knn(big, ctx.upload(qs), k=1, max_dist=1.0)
    ids_b, d2_b = O.knn_brute(ref_host, qs, k=1, max_dist=1.0, nthreads=os.cpu_count() or 8)
    sample_ok = bool(np.array_equal(ids_g, ids_b) and np.array_equal(d2_g.view(np.uint32), d2_b.view(np.uint32)))
    own = rs.choice(M, 200_000, replace=False)
    ids_o, d2_o, _ = ctx.knn(big, ctx.upload(ref_host[own]), k=1, max_dist=0.5)
    self_ok = bool(np.array_equal(ids_o[:, 0], own.astype(np.int32)) and not d2_o.any())
    assert sample_ok and self_ok
    big.free()
    # (iii) SurfaceNormal k = 20 once over the whole map (stress)
    ctx.sync()
    t0 = time.perf_counter()
    info = ref.surface_normals(20, max_dist=1.0)        # first call: 12 GB of neighbour scratch (ids, d2, positions: 240 B/pt) come out of a growing pool
    map_normals_cold_s = time.perf_counter() - t0
    ctx.flush_l2()
    ctx.sync()
    t0 = time.perf_counter()
    info = ref.surface_normals(20, max_dist=1.0)        # the timed one: the pool holds the blocks
    map_normals_s = time.perf_counter() - t0
    knn_t = float(np.mean(knn_ms))
    scan_ms = float(np.mean([p[1] + p[2] for p in res]))
    scan_ms_e = float(np.mean([p[1] + p[2] for p in e2e]))
    line = {"metric": "large-map stress: scans/s (k=20 normals + k=1 ICP against a 50M-point map), kNN queries/s", "value": 1000.0 / scan_ms, "unit": "scans/s",
            "n_gpus": 1, "steps": n_scans, "warmup": 2, "ms_per_step": scan_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"config5: {M}-point map (scene 4001, {side:.0f} x {side:.0f} x 15 m, dedupe 0.03 m), {n_scans} scans of 131 072 points from random poses (seed 4002)",
                       "step": "one scan: SurfaceNormal k = 20 / 1 m on the scan, then ICP (k = 1, maxDist 1, TrimmedDist 0.8) against the map",
                       "map_upload_s": upload_s, "index_build_ms": build_warm, "index_build_ms_cold_first": build_ms[0], "index_build_frac": B.B_BUILD * M / (build_warm * 1e-3) / 1e9 / peak,
                       "cell_m": icp.index.cell, "normals_ms_per_scan": float(np.mean([p[1] for p in res])), "registration_ms_per_scan": float(np.mean([p[2] for p in res])),
                       "iterations": [p[3] for p in res], "worst_pose_error_m": max(p[4] for p in res), "worst_pose_error_deg": max(p[5] for p in res),
                       "map_normals_k20_s": map_normals_s, "map_normals_k20_first_call_s": map_normals_cold_s, "map_normals_frac": B.B_NORMALS * M / map_normals_s / 1e9 / peak, "map_normals_degenerate": info["degenerate"],
                       "parity": {"sampled_queries_vs_brute_force_over_the_whole_map": sample_ok, "map_points_find_themselves": self_ok}, "gen_s": gen_s},
            "knn_queries_per_s": n / (knn_t * 1e-3), "knn_visits_per_query": visits / (n * n_scans),
            "roofline": {"bound": "hbm", "kernel": "knn1GroupKernel<8> (k = 1 match at the converged pose, stand-alone, L2 flushed)", "achieved": B.B_KNN1 * n / (knn_t * 1e-3) / 1e9,
                         "peak": peak, "unit": "GB/s", "frac": B.B_KNN1 * n / (knn_t * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "avg_launch_ms": knn_t, "launch_ms": [float(x) for x in knn_ms]},
            "e2e": {"value": 1000.0 / scan_ms_e, "unit": "scans/s", "h2d_bytes_per_step": int(scans[0].nbytes + 3072), "d2h_bytes_per_step": 144 + 4, "ms_per_step": scan_ms_e},
            "gpu_launches": int(launches), "clocks": clocks}
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = _config5_cpu(args, ref_host, nrm_host, scans, T_init, kw, world, full_line=False)
    return line


def _config5_cpu(args, ref_host, nrm_host, scans, T_init, kw, world, full_line):
    """Bounded sample: the oracle's k-d tree over a 1/10 subsample of the map (building it over 50 M points takes minutes), one scan:
    normals k = 20 on the scan + ICP."""
    from oracle import oracle_py as O
    sub = max(1, ref_host.shape[0] // 5_000_000)
    cores = B.pick_threads(O, ref_host[::sub], scans[0])
    cfg = O.icp_config(precision=O.ORC_REF32, nthreads=cores, **kw)
    t0 = time.time()
    ref = O.IcpReference(cfg, np.ascontiguousarray(ref_host[::sub]), np.ascontiguousarray(nrm_host[::sub]))
    build_s = time.time() - t0
    t0 = time.time()
    O.surface_normals(scans[0], 20, max_dist=1.0, nthreads=cores)
    code, T, st = ref.compute(scans[0], None, T_init[0])
    dt = time.time() - t0
    cb = {"value": 1.0 / dt, "unit": "scans/s", "cores": cores, "kind": "port",
          "sample": f"one scan (normals k=20 + ICP, {st.iterations} iterations) against a 1/{sub} subsample of the map (kd-tree over {ref_host.shape[0] // sub} points built in {build_s:.1f} s)"}
    if not full_line:
        return cb
    return {"metric": "large-map stress: scans/s", "value": 1.0 / dt, "unit": "scans/s", "n_gpus": world, "steps": 1, "warmup": 0, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": "config5: 50M-point map stress", "step": "one scan"}, "cpu_baseline": cb,
            "e2e": {"value": 1.0 / dt, "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
