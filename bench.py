#!/usr/bin/env python
"""Benchmark of the dense RGB-D direct-alignment hot path (BASELINE.json metric: aligned frame pairs/sec at
640x480 RGB-D, 4 pyramid levels).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--frames F] [--impl ours|reference]

Workload (BASELINE.json configs[1]): a synthetic TUM-shape 640x480 odometry sequence, frame-to-frame
alignment of every consecutive pair in ONE batched call, LossTDistribution(ScalerTDistribution(nu=5)),
GaussNewton(1e-7, 100), minGradient 30, 4 levels.  A step = one pass over the whole sequence: pyramid +
gradients of every frame, then the alignment of all F-1 pairs.  `value` is timed with the raw frames already in
HBM; `e2e` is the same step through the C ABI with pinned HOST buffers (H2D of all frames and D2H of the poses
inside the timed region).  Under torchrun every rank runs its own sequence (weak scaling, no data-path
collective; poses are gathered once with all_gather).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, LEVELS = 640, 480, 4
ALGO_BYTES_PER_PIXEL_ITER = 10  # T u8 + dIx s16 + dIy s16 + depth f32 + gathered I u8


def load_traffic(frames):
    """dram__bytes_read.sum + dram__bytes_write.sum of the level-0 launch from the committed ncu --set full capture
    (profiles/traffic.json); only valid for the batch size it was captured on."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        t = json.load(open(p))
        if t.get("frames") == frames:
            return t.get("dram_bytes_per_launch")
    return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def make_jobs(capi, n_frames, id0=0):
    ident = (C.c_double * 7)(0, 0, 0, 1, 0, 0, 0)
    n = n_frames - 1
    jobs = (capi.Job * n)()
    ids = np.arange(id0, id0 + n_frames, dtype=np.uint64)
    poses = np.tile(np.array([0, 0, 0, 1, 0, 0, 0], np.float64), (n, 1))
    for i in range(n):
        jobs[i].to = id0 + i + 1
        jobs[i].to_pose = ident
        jobs[i].n_from = 1
        jobs[i].multi = 0
        jobs[i].from_ = C.cast(ids[i:].ctypes.data, C.POINTER(C.c_uint64))
        jobs[i].from_poses = C.cast(poses[i:].ctypes.data, C.POINTER(C.c_double))
    return jobs, (ids, poses)


def cpu_leg(n_pairs, threads, gray, depth, cam):
    """Times the CPU oracle (frame preparation + alignment) on the first n_pairs pairs."""
    from oracle import oracle as O
    O.build()
    cfg = O.make_config()  # tdist / tdist nu=5, GaussNewton(1e-7,100), minGradient 30, prior off
    t0 = time.perf_counter()
    poses, iters = O.baseline_sequence(cfg, gray[:n_pairs + 1], depth[:n_pairs + 1], *cam, LEVELS, threads)
    dt = time.perf_counter() - t0
    return n_pairs / dt, dt, poses, iters


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU aligner (oracle port; the reference itself cannot be built in this
    image) on the host cores, all threads, on a bounded sample of the same workload per step."""
    if rank != 0:
        return
    from vslam_b200 import synth
    threads = os.cpu_count() or 1
    n_pairs = max(4 * threads, 16)
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(n_pairs + 1, W, H, seed=1234, cam=cam)
    for _ in range(args.warmup):
        cpu_leg(min(n_pairs, threads), threads, gray, depth, cam)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_leg(n_pairs, threads, gray, depth, cam)
    dt = time.perf_counter() - t0
    v = n_pairs * args.steps / dt
    sample = f"first {n_pairs} pairs of the {args.frames}-frame sequence per step, pair-parallel OpenMP"
    print(json.dumps({
        "impl": "reference", "metric": "aligned frame pairs/sec at 640x480 RGB-D", "value": v, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"synthetic TUM-shape 640x480 sequence, frame-to-frame, t-distribution weights, "
                               f"{LEVELS} levels (sample of {n_pairs} pairs)"},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--cpu-pairs", type=int, default=32, help="pairs timed on the host cores for cpu_baseline")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--distinct-sequences", action="store_true",
                    help="every rank renders its own sequence (seed 1234 + rank) instead of the same one; the sequences "
                         "differ in Gauss-Newton iteration count by up to 30 %, so the per-GPU work is then not fixed")
    args = ap.parse_args()
    rank, world, local = dist_env()
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    from vslam_b200 import build, capi, synth
    if not os.path.exists(capi.LIB_PATH):
        build.build_all()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the alignment path has no CPU fallback")
    torch.cuda.set_device(local)
    use_dist = world > 1
    if use_dist:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    F = args.frames
    cam = synth.cam_for(W, H)
    # weak scaling = fixed work per GPU: every rank aligns the same sequence (the result gather still moves every rank's
    # poses).  Sequences rendered from different seeds need different numbers of Gauss-Newton iterations (measured: rank
    # 3's takes 27 % longer than rank 0's on identical GPUs), which would time the data instead of the scaling.
    gray, depth, gt = synth.make_sequence(F, W, H, seed=1234 + (rank if args.distinct_sequences else 0), cam=cam)
    h_gray = torch.from_numpy(gray).pin_memory()
    h_depth = torch.from_numpy(depth).pin_memory()
    d_gray = h_gray.cuda()
    d_depth = h_depth.cuda()
    ids = np.arange(F, dtype=np.uint64)
    cams = np.tile(np.array(cam, np.float64), (F, 1))
    ctx = capi.Context(capi.make_config(), device=local)
    stream = torch.cuda.Stream()
    capi.check(capi.lib().vslamb200_set_stream(ctx.h, C.c_void_p(stream.cuda_stream)))
    jobs, _keep = make_jobs(capi, F)
    results = (capi.Result * (F - 1))()

    def step_resident():
        ctx.upload_device(ids, d_gray.data_ptr(), d_depth.data_ptr(), F, W, H, cams, LEVELS)
        ctx.align_raw(jobs, F - 1, results)

    # End to end through the C ABI with pinned HOST buffers.  Successive steps (batches) alternate between two sets of
    # frame ids: the upload of the next step's frames (copy stream) is queued before this step's alignment is called, so
    # in steady state the PCIe transfer hides under the kernels.  Every step's upload, including the first one, and
    # every step's result read-back are inside the timed region.
    ids2 = [ids, ids + np.uint64(F)]
    jobs2, keep2 = [jobs], [_keep]
    j1, k1 = make_jobs(capi, F, id0=F)
    jobs2.append(j1)
    keep2.append(k1)

    def upload_set(q):
        ctx.upload_async(ids2[q], h_gray.data_ptr(), h_depth.data_ptr(), F, W, H, cams, LEVELS)

    def run_e2e(steps):
        upload_set(0)
        for st_ in range(steps):
            if st_ + 1 < steps:
                upload_set((st_ + 1) & 1)
            ctx.align_raw(jobs2[st_ & 1], F - 1, results)

    def timed(fn, steps, loop=True):
        if use_dist:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            if loop:
                for _ in range(steps):
                    fn()
            else:
                fn(steps)
            e1.record(stream)
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        timed.own_ms = float(ms.item())
        if use_dist:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.barrier()
        return float(ms.item())

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    ctx.reset_stats()
    ms = timed(step_resident, args.steps)
    own_ms = timed.own_ms
    st = ctx.stats()
    clocks = sampler.stop()
    pairs = (F - 1) * args.steps * world
    value = pairs / (ms * 1e-3)

    # end to end through the C ABI with host buffers
    resident_poses = np.array([list(r.pose) for r in results])
    run_e2e(2)
    ms_e2e = timed(run_e2e, args.steps, loop=False)
    e2e_diff = float(np.abs(np.array([list(r.pose) for r in results]) - resident_poses).max())
    e2e = pairs / (ms_e2e * 1e-3)
    h2d = F * W * H * 5
    d2h = (F - 1) * C.sizeof(capi.Result)

    # p50 / p95 latency of ONE alignment (BASELINE.json metric, second half): batch size 1, the first pair
    # of the sequence (frames resident), back-to-back calls, wall time from the host's submit to the result on the host
    lat = None
    if rank == 0:
        one = (capi.Result * 1)()
        for _ in range(50):
            ctx.align_raw(jobs, 1, one)
        ts = []
        for _ in range(300):
            t0 = time.perf_counter()
            ctx.align_raw(jobs, 1, one)
            ts.append((time.perf_counter() - t0) * 1e3)
        # a single job runs on larger clusters than a throughput batch (latency mode), so its sums are added in another
        # order: the same pose up to rounding, not bit for bit
        lat_diff = float(np.max(np.abs(np.array(list(one[0].pose)) - np.array(list(resident_poses[0])))))
        assert lat_diff < 1e-9, "single-pair call differs from the batched result"
        lat = {"p50_ms": float(np.percentile(ts, 50)), "p95_ms": float(np.percentile(ts, 95)), "calls": len(ts), "max_pose_diff_vs_batched": lat_diff,
               "what": "one 640x480 pair, 4 levels, frames resident, host submit -> pose on host"}

    # per-kernel timing pass (events around every residual / accumulate launch; separate from the timed runs)
    capi.check(capi.lib().vslamb200_set_timing(ctx.h, 1))
    ctx.reset_stats()
    step_resident()
    tst = ctx.stats()
    capi.check(capi.lib().vslamb200_set_timing(ctx.h, 0))
    peak, peak_kind = load_peaks()
    # dominant kernel: gn_level_kernel at level 0 (one launch per step); algorithmic bytes = 10 B per template pixel
    # and Gauss-Newton evaluation x the pixel-evaluations that launch processed
    l0_ms = tst["ms_gn_level"][0]
    l0_bytes = ALGO_BYTES_PER_PIXEL_ITER * tst["pixel_iterations_level"][0]
    all_ms = sum(tst["ms_gn_level"])
    achieved = l0_bytes / (l0_ms * 1e-3) / 1e9 if l0_ms > 0 else 0.0
    roofline = {
        "bound": "hbm", "kernel": "gn_level_kernel (level 0: all Gauss-Newton iterations of 640x480 for the batch)",
        "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak,
        "traffic": load_traffic(F),
        "algorithmic_bytes_per_launch": l0_bytes, "ms_per_launch": l0_ms,
        "all_levels": {"ms": tst["ms_gn_level"][:LEVELS], "pixel_iterations": tst["pixel_iterations_level"][:LEVELS],
                       "achieved": ALGO_BYTES_PER_PIXEL_ITER * tst["pixel_iterations"] / (all_ms * 1e-3) / 1e9 if all_ms > 0 else 0.0},
        "ms_lists": tst["ms_lists"],
        "point_iterations": tst["point_iterations"],
        "interest_density": tst["point_iterations"] / max(1, tst["pixel_iterations"]),
    }
    iters_per_pair = np.mean([sum(r.iters[:LEVELS]) for r in results])

    ranks = None
    if use_dist:
        # per-rank step time, clocks and throttle reasons (the reported time is the maximum over the ranks)
        ranks = [None] * world
        dist.all_gather_object(ranks, {"rank": rank, "ms_per_step": own_ms / args.steps, "sm_mhz": clocks.get("sm_mhz"),
                                       "reasons": clocks.get("reasons"),
                                       "ms_gn_level": [round(x, 3) for x in tst["ms_gn_level"][:LEVELS]],
                                       "ms_lists": round(tst["ms_lists"], 3), "ms_preprocess": round(tst["ms_preprocess"], 3)})
        # the only collective of the path: gather the poses (7 f64 per pair) of every rank
        mine = torch.tensor([list(r.pose) for r in results], device="cuda", dtype=torch.float64)
        allp = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allp, mine)

    cpu = None
    if rank == 0 and not args.no_cpu:
        v1, dt1, cposes, _ = cpu_leg(args.cpu_pairs, 1, gray, depth, cam)
        thr = os.cpu_count() or 1
        vN, dtN, _, _ = cpu_leg(max(args.cpu_pairs, thr), thr, gray, depth, cam)
        from oracle import oracle as O
        perr = max(np.linalg.norm(O.se3_log(O.se3_mul(np.array(results[i].pose), O.se3_inv(cposes[i]))))
                   for i in range(args.cpu_pairs))
        # spot checks across the whole batch (middle and last pair) against the oracle, and every job must have iterated
        for m in ((F - 1) // 2, F - 2):
            pm, _ = O.baseline_sequence(O.make_config(), gray[m:m + 2], depth[m:m + 2], *cam, LEVELS, 1)
            perr = max(perr, np.linalg.norm(O.se3_log(O.se3_mul(np.array(results[m].pose), O.se3_inv(pm[0])))))
        assert all(r.status == 0 and sum(r.iters[:LEVELS]) > 0 for r in results), "a job did not run"
        cpu = {"value": v1, "unit": "pairs/s", "cores": 1, "kind": "port",
               "sample": f"first {args.cpu_pairs} pairs of the sequence, frame preparation included ({dt1:.1f} s)",
               "all_cores": {"value": vN, "cores": thr, "pairs": max(args.cpu_pairs, thr)},
               "max_pose_diff_vs_gpu": perr}
    if rank == 0:
        print(json.dumps({
            "metric": "aligned frame pairs/sec at 640x480 RGB-D", "value": value, "unit": "pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"synthetic TUM-shape 640x480 sequence of {F} frames per GPU, frame-to-frame "
                                   f"alignment ({F - 1} pairs per batch), t-distribution weights, {LEVELS} levels, "
                                   "GaussNewton(1e-7,100), minGradient 30",
                       "sequences": ("one per rank, seeds 1234 + rank" if args.distinct_sequences else
                                     "the same sequence (seed 1234) on every rank: fixed work per GPU"),
                       "l2_policy": "inputs (1.5 GB of frames per step) are larger than the 126 MB L2",
                       "mean_gn_iterations_per_pair": float(iters_per_pair)},
            "e2e": {"value": e2e, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "max_pose_diff_vs_resident": e2e_diff,
                    "pipeline": "upload of step s+1 (copy stream) overlaps the alignment of step s; all uploads timed"},
            "gpu_launches": int(st["kernel_launches"]),
            "latency": lat, "clocks": clocks, "ranks": ranks, "roofline": roofline, "cpu_baseline": cpu,
        }))
    ctx.close()
    if use_dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
