#!/usr/bin/env python3
"""bench.py — BASELINE.json metric: ICP registrations/sec (131k-point scan vs 5M-point voxel map) + kNN queries/sec,
% of the HBM roofline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config2|config3|config4|config5]

Default workload (the one the driver runs, at every N): BASELINE config #2 — synthetic 128x1024 ray-cast LiDAR scans against a
5 000 000-point map, KDTree(k1, maxDist 1.0) + TrimmedDist 0.8 + PointToPlane + Differential(1e-3,1e-2,3) + Counter 40 — through
libb200reg.so.  One "step" = one batch of 8 INDEPENDENT registrations (8 sensor poses in the scene, so the scans touch
different parts of the map) against the resident map index, kept in flight on the lanes of a b200reg_batch:
  value = registrations/s with scans and index resident in HBM, CUDA events around the batch, L2 flushed between steps;
  e2e   = the same through b200reg_batch_register_host with the scans in pinned HOST memory (uploads and the read-back of every
          result inside the timed region, host clock);
  config.single_stream = one registration at a time (round 1's definition of `value`: the latency a mapping loop sees).
N > 1: one process per GPU, every rank runs the SAME workload (replicas: a registration does not shard), one NCCL
all_gather of the results.  The other BASELINE configurations are separate workloads (--workload), each printing its own line
with roofline / cpu_baseline / e2e; raw outputs under profiles/r2/.
  --impl reference : the CPU restatement of the reference path (oracle, all host threads) on the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

B_KNN1 = 536          # algorithmic bytes per k=1 query   (SURVEY 8d / BASELINE.md §5)
B_ITER = 600          # algorithmic bytes per reading point per ICP iteration
B_BUILD = 156         # algorithmic bytes per reference point for the index build
B_KNN20 = 2736        # k = 20 query
B_NORMALS = 3072      # SurfaceNormal k = 20, per point
B_VOXEL = 180         # VoxelGrid / voxel down-sample, per point
ICP_KW = dict(knn=1, max_dist=1.0, trim_ratio=0.8, max_iter=40, min_diff_rot=1e-3, min_diff_trans=1e-2, smooth_length=3)
BATCH = 8             # registrations per step of the config #2 workload


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def ncu_traffic():
    """dram bytes per match launch from the committed ncu capture (profiles/r2/ncu_traffic.json names the commit it was taken at)."""
    p = os.path.join(ROOT, "profiles", "r2", "ncu_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("dram_bytes_per_launch"), d
    return None, None


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe); rank 0 only."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([s.strip() for s in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def feat4(xyz):
    out = np.ones((xyz.shape[0], 4), dtype=np.float32)
    out[:, :3] = xyz
    return out


def dist_setup(world, local_rank):
    import torch
    if world <= 1:
        return torch, None
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    return torch, dist


def gather_floats(torch, dist, local_rank, vals):
    """every rank's list of floats -> world x len array (rank order)"""
    if dist is None:
        return np.array([vals], dtype=np.float64)
    t = torch.tensor(vals, device=torch.device("cuda", local_rank), dtype=torch.float64)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return np.stack([o.cpu().numpy() for o in out])


def spread(a):
    a = np.asarray(a, dtype=np.float64)
    return {"min": float(a.min()), "median": float(np.median(a)), "p95": float(np.percentile(a, 95)), "max": float(a.max())}


# =====================================================================================================================
# config #2 — the driver's workload
# =====================================================================================================================
def make_config2(args):
    """Scene 1001 (the same on every rank, so the ranks do identical work), its 5 M-point map, and BATCH scans: scan 0 is round 1's
    (tests/scenes.py config2), the others are taken 30 m away from it at 45-degree steps of bearing and yaw.  Every scan starts from
    the same perturbation of its true pose (0.15 m, 1.5 deg)."""
    import scenes
    t = time.time()
    c = scenes.config2(m=args.map_points, n_beams=args.beams, n_az=args.azimuths, scene_seed=1001)
    sc = c["scene"]
    pert = scenes.rigid(scenes.rot_axis_angle([1, 2, 3], np.deg2rad(1.5)), [0.15, -0.10, 0.05])
    c0 = c["T_true"][:3, 3]
    scans, T_true, T_init = [feat4(c["scan"])], [c["T_true"]], [c["T_init"].astype(np.float32)]
    for j in range(1, BATCH):
        a = 2 * np.pi * j / 8 + 0.3
        Tt = scenes.rigid(scenes.rot_axis_angle([0, 0, 1], np.deg2rad(17.0 + 45.0 * j)), c0 + np.array([30 * np.cos(a), 30 * np.sin(a), 0.0]))
        scans.append(feat4(scenes.lidar_scan(sc, Tt, args.beams, args.azimuths, 0.01, 1002 + j)))
        T_true.append(Tt)
        T_init.append((Tt @ pert).astype(np.float32))
    return dict(ref=feat4(c["map_xyz"]), rn=c["map_nrm"].astype(np.float32), scans=scans, T_true=T_true, T_init=np.stack(T_init),
                gen_s=time.time() - t)


def config2_name(args):
    return (f"config2: synthetic {args.beams}x{args.azimuths} LiDAR scan ({args.beams * args.azimuths} pts) vs "
            f"{args.map_points}-pt voxelized map, point-to-plane ICP + TrimmedDist 0.8")


def pick_threads(O, ref, scan):
    """All the host threads the port can actually use: time one kNN sweep with 1 thread and with every core, keep the faster
    (on shared hosts the advertised cores are often not there)."""
    cores = os.cpu_count() or 1
    tree = O.KdTree(ref[:200_000])
    best, best_t = 1, None
    for nt in sorted({1, cores}):
        tree.knn(scan, k=1, max_dist=1.0, nthreads=nt)
        t0 = time.time()
        tree.knn(scan, k=1, max_dist=1.0, nthreads=nt)
        dt = time.time() - t0
        if best_t is None or dt < best_t:
            best, best_t = nt, dt
    return best


def reference_config2(args, world):
    """Reference arm: the CPU restatement (oracle/) of the reference's path on the box's host cores; a step = ONE registration
    (scan s mod 8 of the batch), value = registrations/s."""
    from oracle import oracle_py as O
    w = make_config2(args)
    cores = pick_threads(O, w["ref"], w["scans"][0])
    cfg = O.icp_config(precision=O.ORC_REF32, nthreads=cores, **ICP_KW)
    t0 = time.time()
    ref = O.IcpReference(cfg, w["ref"], w["rn"])
    build_s = time.time() - t0
    times, iters = [], []
    for s in range(args.warmup + args.steps):
        j = s % BATCH
        t0 = time.time()
        code, T, st = ref.compute(w["scans"][j], None, w["T_init"][j])
        dt = time.time() - t0
        assert code == 0, code
        if s >= args.warmup:
            times.append(dt)
            iters.append(st.iterations)
    ms = 1000.0 * float(np.mean(times))
    val = 1000.0 / ms
    return {"metric": "ICP registrations/sec (131k scan vs 5M map)", "value": val, "unit": "registrations/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": config2_name(args), "iterations": float(np.mean(iters)), "index_build_s": build_s,
                       "build": "amortised (kd-tree prebuilt)", "step": "one registration (scan s mod 8 of the 8-pose batch)"},
            "cpu_baseline": {"value": val, "unit": "registrations/s", "cores": cores, "kind": "port",
                             "sample": f"{args.steps} full registrations ({np.mean(iters):.1f} iterations each) of the same workload, kd-tree prebuilt ({build_s:.1f} s, 1 thread)"},
            "e2e": {"value": val, "unit": "registrations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_config2(args, rank, local_rank, world):
    torch, dist = dist_setup(world, local_rank)
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    ctx = capi.Context(local_rank)   # raises without a GPU / without the built library: no CPU fallback
    w = make_config2(args)
    n, m = w["scans"][0].shape[0], w["ref"].shape[0]
    peak, peak_src = hbm_peak()

    t0 = time.time()
    cref = ctx.upload(w["ref"], w["rn"])
    upload_ref_s = time.time() - t0
    icp = capi.ICP(ctx, cell=args.cell, **ICP_KW)
    build_ms = []
    for _ in range(5):   # the first build grows the stream-ordered pool (cold); later ones are what a running system pays
        ctx.flush_l2()
        ctx.timer_start()
        icp.set_reference(cref)
        build_ms.append(ctx.timer_stop())
    build_warm = float(np.median(build_ms[2:]))
    batch = capi.Batch(ctx, lanes=args.lanes, cell=args.cell, **ICP_KW)
    batch.set_reference(cref)
    cscans = [ctx.upload(s) for s in w["scans"]]
    pinned = [torch.from_numpy(s).pin_memory() for s in w["scans"]]
    host_scans = batch.host_clouds([(p.numpy(), None) for p in pinned])
    Ti = w["T_init"]

    def barrier():
        if dist is not None:
            dist.barrier()
        ctx.sync()
        torch.cuda.synchronize()

    # ---------------- resident-input arm: `value`
    for _ in range(args.warmup):
        T, stats, status = batch.register(None, cscans, Ti)
    assert not status.any()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    barrier()
    launches0 = ctx.launch_count
    step_ms = []
    wall0 = time.time()
    for _ in range(args.steps):
        ctx.flush_l2()                 # L2 flushed between timed steps (untimed)
        ctx.timer_start()
        T, stats, status = batch.register(None, cscans, Ti)
        step_ms.append(ctx.timer_stop())
    barrier()
    wall_s = time.time() - wall0
    launches = ctx.launch_count - launches0
    iters = [s.iterations for s in stats]
    total_ms = float(np.sum(step_ms))

    # ---------------- end-to-end arm: HOST buffers (pinned) through b200reg_batch_register_host: per step the upload of the 8 scans,
    # the registrations and the read-back of the 8 results; host wall clock around the call
    for _ in range(2):
        batch.register_host(None, host_scans, Ti)
    barrier()
    e2e_ms = []
    for _ in range(args.steps):
        ctx.flush_l2()
        ctx.sync()
        t0 = time.perf_counter()
        T_e, stats_e, status_e = batch.register_host(None, host_scans, Ti)
        e2e_ms.append(1000.0 * (time.perf_counter() - t0))
    barrier()
    assert np.array_equal(T_e, T) and not status_e.any(), "end-to-end arm and resident arm disagree"
    e2e_total_ms = float(np.sum(e2e_ms))
    clocks = sampler.stop() if sampler else None

    # ---------------- single stream: one registration at a time (round 1's `value` / `e2e`), scan 0
    for _ in range(3):
        T1, st1 = icp.register(cscans[0], Ti[0])
    single_ms, single_dev = [], []
    for _ in range(args.steps):
        ctx.flush_l2()
        ctx.timer_start()
        T1, st1 = icp.register(cscans[0], Ti[0])
        single_ms.append(ctx.timer_stop())
        single_dev.append(st1.device_ms)
    assert np.array_equal(T1, T[0]), "the batch lanes and the single registration disagree"
    cstream = ctx.upload(pinned[0].numpy())
    single_e2e = []
    for s in range(2 + args.steps):
        ctx.flush_l2()
        ctx.sync()
        t0 = time.perf_counter()
        cstream.update(pinned[0].numpy(), wait=False)
        icp.register(cstream, Ti[0])
        if s >= 2:
            single_e2e.append(1000.0 * (time.perf_counter() - t0))

    # ---------------- max over ranks (device-timed), whole-job aggregate; the path's only collective: ONE all_gather of
    # the step's results, timed on the device and charged to the K steps
    gather_ms = 0.0
    per_rank = gather_floats(torch, dist, local_rank, [total_ms, e2e_total_ms, float(np.mean(iters)), float(np.median(step_ms))])
    if dist is not None:
        dev = torch.device("cuda", local_rank)
        rec = torch.zeros((BATCH, 20), device=dev, dtype=torch.float32)
        rec[:, :16] = torch.from_numpy(T.reshape(BATCH, 16)).to(dev)
        out = [torch.empty_like(rec) for _ in range(world)]
        dist.all_gather(out, rec)     # warm-up of the communicator
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        dist.all_gather(out, rec)
        g1.record()
        torch.cuda.synchronize()
        gather_ms = g0.elapsed_time(g1)
        assert all(torch.equal(o, rec) for o in out), "ranks disagree on the same workload"
    total_ms = float(per_rank[:, 0].max()) + gather_ms
    e2e_total_ms = float(per_rank[:, 1].max()) + gather_ms
    ms_per_step = total_ms / args.steps
    value = world * BATCH * 1000.0 / ms_per_step
    e2e_value = world * BATCH * 1000.0 / (e2e_total_ms / args.steps)

    # ---------------- roofline of the dominant kernel (the k = 1 match), timed live with CUDA events on the launching stream:
    # profiling mode brackets EVERY match launch of a registration with its own event pair (kernel-only time, no host
    # work inside the bracket); registrations of scan 0, L2 flushed before each one.
    ctx.set_profiling(True)
    per_launch = []
    for r in range(2 + args.knn_reps):
        ctx.flush_l2()
        icp.register(cscans[0], Ti[0])
        ms = ctx.profile_read()
        if r >= 2:
            per_launch.append(ms)
    ctx.set_profiling(False)
    per_launch = np.array(per_launch)                   # reps x iterations
    knn_t = float(per_launch.mean())
    knn_conv = float(per_launch[:, -1].mean())
    mean = icp.reference_mean
    Tpre = np.eye(4, dtype=np.float32)
    Tpre[:3, 3] = -mean
    Tq = (Tpre.astype(np.float64) @ T[0].astype(np.float64)).astype(np.float32)    # converged pose in the refMean frame
    _, _, visits = ctx.knn(icp.index, cscans[0], T=Tq, k=1, max_dist=1.0, want=False)
    achieved = B_KNN1 * n / (knn_t * 1e-3) / 1e9
    iter_gbs = B_ITER * n * st1.iterations / (float(np.mean(single_dev)) * 1e-3) / 1e9
    batch_gbs = B_ITER * n * float(np.sum(iters)) / (float(np.mean(step_ms)) * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic()

    errs = [float(np.abs((np.linalg.inv(w["T_true"][j]) @ T[j].astype(np.float64))[:3, 3]).max()) for j in range(BATCH)]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return None
    line = {
        "metric": "ICP registrations/sec (131k scan vs 5M map)", "value": value, "unit": "registrations/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": config2_name(args), "step": f"{BATCH} independent registrations ({BATCH} sensor poses of the scene) in flight on {batch.lanes} lanes of a b200reg_batch against the resident map index",
                   "iterations": iters, "cell_m": icp.index.cell, "l2": "flushed between timed steps (256 MB memset)",
                   "build": "amortised (index resident, Mapper regime: one build per 20 scans)", "index_build_ms": build_warm,
                   "index_build_ms_cold_first": build_ms[0], "index_build_gbs": B_BUILD * m / (build_warm * 1e-3) / 1e9,
                   "index_build_frac": B_BUILD * m / (build_warm * 1e-3) / 1e9 / peak, "ref_upload_s": upload_ref_s,
                   "translation_error_vs_truth_m": max(errs), "step_ms": spread(step_ms), "e2e_step_ms": spread(e2e_ms),
                   "wall_s_timed_region": wall_s, "result_gather_ms": gather_ms,
                   "single_stream": {"registrations_per_s": 1000.0 / float(np.mean(single_ms)), "ms": spread(single_ms),
                                     "e2e_registrations_per_s": 1000.0 / float(np.mean(single_e2e)), "e2e_ms": spread(single_e2e),
                                     "device_ms_per_registration": float(np.mean(single_dev)), "iterations": st1.iterations,
                                     "note": "one registration at a time (scan 0): round 1's value / e2e definition"},
                   "per_rank": {"total_ms": spread(per_rank[:, 0]), "e2e_total_ms": spread(per_rank[:, 1]), "mean_iterations": spread(per_rank[:, 2]),
                                "median_step_ms": spread(per_rank[:, 3])},
                   "parallelism": f"{world} replicas of the workload, one per GPU (a registration does not shard), one NCCL all_gather of the results"},
        "knn_queries_per_s": n / (knn_t * 1e-3), "knn_visits_per_query": visits / n,
        "roofline": {"bound": "hbm", "kernel": "knn1GroupKernel<8 | 4> (k = 1 match, lanes per query chosen per iteration as in the graph loop, where knn1LoopKernel is the same body)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": B_KNN1 * n, "avg_launch_ms": knn_t,
                     "launch_ms_by_iteration": [float(x) for x in per_launch.mean(0)],
                     "converged_launch_ms": knn_conv, "converged_frac": B_KNN1 * n / (knn_conv * 1e-3) / 1e9 / peak,
                     "timing": "CUDA event pair around every match launch of %d single registrations (host-driven loop, same kernels), L2 flushed before each" % per_launch.shape[0],
                     "iteration_gbs_600B_per_pt": iter_gbs, "iteration_frac": iter_gbs / peak,
                     "batch_gbs_600B_per_pt_per_iteration": batch_gbs, "batch_frac": batch_gbs / peak},
        "e2e": {"value": e2e_value, "unit": "registrations/s", "h2d_bytes_per_step": int(BATCH * (w["scans"][0].nbytes + 3072)),
                "d2h_bytes_per_step": BATCH * 144, "ms_per_step": e2e_total_ms / args.steps},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if not args.no_cpu_baseline:
        from oracle import oracle_py as O
        cores = pick_threads(O, w["ref"], w["scans"][0])
        cfg = O.icp_config(precision=O.ORC_REF32, nthreads=cores, **ICP_KW)
        t0 = time.time()
        ref = O.IcpReference(cfg, w["ref"], w["rn"])
        cpu_build = time.time() - t0
        ts, dT = [], 0.0
        for j in (0, 1, 2, 3):
            t0 = time.time()
            code, T_o, st_o = ref.compute(w["scans"][j], None, w["T_init"][j])
            ts.append(time.time() - t0)
            dT = max(dT, float(np.abs(T_o[:3, 3] - T[j][:3, 3]).max()))
        line["cpu_baseline"] = {"value": 1.0 / float(np.mean(ts)), "unit": "registrations/s", "cores": cores, "kind": "port",
                                "sample": f"4 full registrations (scans 0-3 of the batch), kd-tree prebuilt in {cpu_build:.1f} s (1 thread); mean",
                                "gpu_vs_cpu_translation_diff_m": dT}
    if dist is not None:
        dist.destroy_process_group()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="config2", choices=["config2", "config3", "config4", "config5"])
    ap.add_argument("--map-points", type=int, default=5_000_000)
    ap.add_argument("--beams", type=int, default=128)
    ap.add_argument("--azimuths", type=int, default=1024)
    ap.add_argument("--cell", type=float, default=0.0)
    ap.add_argument("--lanes", type=int, default=0, help="lanes of the b200reg_batch (0: library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--knn-reps", type=int, default=20)
    ap.add_argument("--pairs", type=int, default=256, help="config4: submap pairs")
    ap.add_argument("--pair-points", type=int, default=200_000, help="config4: points per submap")
    ap.add_argument("--scans", type=int, default=500, help="config3: scans of the trajectory; config5: scans (default 10 there)")
    ap.add_argument("--big-map-points", type=int, default=50_000_000, help="config5: map points")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.workload != "config2":
        import bench_workloads   # the other BASELINE configurations (profiles/bench_workloads.py)
        line = bench_workloads.run(args, rank, local_rank, world)
    elif args.impl == "reference":
        line = reference_config2(args, world) if rank == 0 else None   # the other ranks exit 0 without work
    else:
        line = run_config2(args, rank, local_rank, world)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    sys.path.insert(0, os.path.join(ROOT, "profiles"))
    main()
