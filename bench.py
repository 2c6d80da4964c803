#!/usr/bin/env python
"""Benchmark of the dense RGB-D direct-alignment hot path (BASELINE.json metric: aligned frame pairs/sec at
640x480 RGB-D, 4 pyramid levels).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload pairs|sequence] [--pairs P] [--impl ours|reference]

Default workload (BASELINE.json configs[2], SURVEY.md 8(d)'s throughput configuration): P = 4096 independent synthetic
640x480 frame pairs (seeds 1234 ... 1234 + P - 1), LossTDistribution(ScalerTDistribution(nu=5)), GaussNewton(1e-7, 100),
minGradient 30, 4 levels.  A step = one pass over the whole batch: pyramid + gradients of every frame, the alignment of
every pair in ONE batched call per GPU, and the gather of the poses.  Under torchrun the P pairs are dealt round-robin
over the ranks (vslam_b200/shard.py: STRONG scaling, no data-path collective); the only exchange is the all_gather of the
7-double poses over NCCL, inside the timed region.  `value` is timed with the raw frames already in HBM; `e2e` is the
same step through the C ABI with pinned HOST buffers (H2D of all frames and D2H of the results inside the timed region).
--workload sequence: BASELINE configs[1], a 1000-frame TUM-shape sequence per GPU, frame to frame (weak scaling, replicas).
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
DEPTH_DIVISOR = 5000.0
ALGO_BYTES_PER_PIXEL_ITER = 10  # T u8 + dIx s16 + dIy s16 + depth f32 + gathered I u8


def load_traffic(workload, pairs_this_gpu):
    """dram__bytes_read.sum + dram__bytes_write.sum of the level-0 launch from the committed ncu --set full capture
    (profiles/traffic.json); only valid for the workload and batch size it was captured on."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        t = json.load(open(p))
        if t.get("workload") == workload and t.get("pairs_this_gpu") == pairs_this_gpu:
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


IDENT7 = np.array([0, 0, 0, 1, 0, 0, 0], np.float64)
# Every pair starts from the same small motion guess instead of the exact identity.  At the identity a template pixel on
# the "1 < u'" visibility bound (InverseCompositional.cpp:98) re-projects to 1 +- 1 ulp, so whether it counts depends on
# the order of the floating-point operations in K^-1 and K: the oracle, the reference's sources against the stand-in
# Eigen and the real Eigen would each decide a few border pixels differently and end up to 3e-4 apart.  Off the identity
# all of them agree to 1e-15 (tests/test_oracle_ref_cpu.py), so the workload has one answer.
START_TWIST = np.array([1e-4, -1e-4, 1e-4, 1e-5, -1e-5, 1e-5])


def start_pose7():
    from vslam_b200 import synth
    R, t = synth.se3_exp(START_TWIST)
    return np.asarray(synth.pose7(R, t), np.float64)
WORKLOAD_TEXT = {
    "pairs": "BASELINE configs[2]: {P} independent synthetic 640x480 frame pairs (seeds 1234..{last}), dealt round-robin over "
             "the GPUs, t-distribution weights, 4 levels, GaussNewton(1e-7,100), minGradient 30, every pair starts 1e-4 off the identity, chi2 as the reference's Release build evaluates it",
    "sequence": "BASELINE configs[1]: synthetic TUM-shape 640x480 sequence of {F} frames per GPU, frame-to-frame alignment "
                "({P} pairs per batch), t-distribution weights, 4 levels, GaussNewton(1e-7,100), minGradient 30, chi2 as the reference's Release build evaluates it",
}


def render_pairs(synth, indices, cam):
    """Frames of the independent pairs `indices` (pair k: seed 1234 + k), stored ref, cur, ref, cur, ...; ground truth."""
    m = len(indices)
    gray = np.empty((2 * m, H, W), np.uint8)
    depth = np.empty((2 * m, H, W), np.float32)
    gt = np.empty((m, 7))
    for j, k in enumerate(indices):
        p = synth.make_pair(W, H, 1234 + int(k), cam=cam)
        gray[2 * j], gray[2 * j + 1] = p["gray0"], p["gray1"]
        depth[2 * j], depth[2 * j + 1] = p["depth0"], p["depth1"]
        gt[j] = p["pose1"]
    return gray, depth, gt


def quantise_depth(depth):
    """Sensor depth: 16-bit in units of 1 / 5000 m; returned as the f32 the aligners consume."""
    k16 = np.clip(np.rint(depth.astype(np.float64) * DEPTH_DIVISOR), 0, 65535).astype(np.uint16)
    return (k16.astype(np.float64) / DEPTH_DIVISOR).astype(np.float32)


def make_jobs(capi, n_jobs, stride, id0=0):
    """Job i aligns frame id0 + i * stride + 1 (initial guess: start_pose7()) against frame id0 + i * stride (identity)."""
    ident = (C.c_double * 7)(*start_pose7())
    jobs = (capi.Job * n_jobs)()
    ids = np.arange(id0, id0 + n_jobs * stride + 2, dtype=np.uint64)
    poses = np.tile(IDENT7, (n_jobs, 1))
    for i in range(n_jobs):
        jobs[i].to = id0 + i * stride + 1
        jobs[i].to_pose = ident
        jobs[i].n_from = 1
        jobs[i].multi = 0
        jobs[i].from_ = C.cast(ids[i * stride:].ctypes.data, C.POINTER(C.c_uint64))
        jobs[i].from_poses = C.cast(poses[i:].ctypes.data, C.POINTER(C.c_double))
    return jobs, (ids, poses)


def result_poses(capi, results, n):
    """(n, 7) poses out of a ctypes Result array, without a Python loop."""
    raw = np.frombuffer(results, dtype=np.uint8).reshape(len(results), C.sizeof(capi.Result))
    return raw[:n, :56].copy().view(np.float64).reshape(n, 7)


def odometry_leg(capi, synth, shard, cam, n_frames=150):
    """The shape of the reference's OWN caller (odometry/src/Odometry.cpp:21-62, NodeRgbdAlignment.cpp:121-150): frames
    arrive one at a time from host memory, each is aligned - motion prior on, starting from the constant-velocity
    prediction - against {last key frame, last frame} stacked (or the last frame alone when it is the key frame), key
    frame every second frame (KeyFrameSelectionIdx).  Sequential by construction, so this is a latency figure: frames/s
    of the loop and p50 / p95 from "frame in host memory" to "pose on the host" (upload, pyramids, gradients, alignment)."""
    import torch
    gray, depth, gt = synth.make_sequence(n_frames, W, H, seed=1234, cam=cam)
    k16 = np.clip(np.rint(depth.astype(np.float64) * DEPTH_DIVISOR), 0, 65535).astype(np.uint16)
    hg, hk = torch.from_numpy(gray).pin_memory(), torch.from_numpy(k16.view(np.int16)).pin_memory()
    ctx = capi.Context(capi.make_config(include_prior=True))
    cams = np.array(cam, np.float64).reshape(1, 4)
    npx = W * H

    def inv(p):
        q = p[:4] * np.array([-1.0, -1.0, -1.0, 1.0])
        return np.concatenate([q, -shard.quat_rot(q, p[4:])])

    poses, lat, steps = [IDENT7.copy()], [], 0
    last_kf, res1 = 0, (capi.Result * 1)()
    ctx.upload_u16_async(np.array([0], np.uint64), hg.data_ptr(), hk.data_ptr(), DEPTH_DIVISOR, 1, W, H, cams, LEVELS)
    ctx.synchronize()
    t_start = time.perf_counter()
    for t in range(1, n_frames):
        t0 = time.perf_counter()
        ctx.upload_u16_async(np.array([t], np.uint64), hg.data_ptr() + t * npx, hk.data_ptr() + 2 * t * npx, DEPTH_DIVISOR, 1, W, H,
                             cams, LEVELS)
        # constant velocity over equal time steps: exp(log(p_k p_{k-1}^-1)) p_k (MotionPrediction.cpp:38-55)
        guess = poses[-1] if t < 2 else shard.pose_mul(shard.pose_mul(poses[-1], inv(poses[-2])), poses[-1])
        frm = [last_kf, t - 1] if last_kf != t - 1 else [t - 1]
        job = (capi.Job * 1)()
        ids = np.array(frm, np.uint64)
        fp = np.ascontiguousarray(np.stack([poses[k] for k in frm]))
        job[0].to = t
        job[0].to_pose = (C.c_double * 7)(*guess)
        job[0].n_from = len(frm)
        job[0].multi = int(len(frm) > 1)
        job[0].from_ = ids.ctypes.data_as(C.POINTER(C.c_uint64))
        job[0].from_poses = fp.ctypes.data_as(C.POINTER(C.c_double))
        ctx.align_raw(job, 1, res1)
        poses.append(np.array(list(res1[0].pose)))
        steps += sum(res1[0].iters[:LEVELS])
        lat.append((time.perf_counter() - t0) * 1e3)
        if t % 2 == 0:
            last_kf = t
        if t >= 16:
            ctx.release(t - 16)  # out of the reference's rings of 7 frames / 7 key frames
    total = time.perf_counter() - t_start
    ctx.close()
    traj = np.array(poses)
    # frame-to-frame translation error against the rendered motion (P_frame = pose * P_world)
    rel = [shard.pose_mul(traj[k + 1], inv(traj[k])) for k in range(n_frames - 1)]
    rel_gt = [shard.pose_mul(gt[k + 1], inv(gt[k])) for k in range(n_frames - 1)]
    err = np.array([np.linalg.norm(shard.pose_mul(a, inv(b))[4:]) for a, b in zip(rel, rel_gt)])
    # frames_per_s: upload + alignment back to back (the per-frame times below); wall_frames_per_s also holds this loop's
    # housekeeping (releasing frames that left the rings, each a stream synchronisation; Python)
    return {"frames_per_s": (n_frames - 1) / (sum(lat) * 1e-3), "wall_frames_per_s": (n_frames - 1) / total,
            "p50_ms": float(np.percentile(lat, 50)), "p95_ms": float(np.percentile(lat, 95)),
            "frames": n_frames - 1, "mean_gn_iterations_per_frame": steps / (n_frames - 1),
            "rpe_translation_rmse_per_frame": float(np.sqrt(np.mean(err ** 2))),
            "what": "sequential odometry-shaped loop: one 640x480 frame (u8 + u16) from pinned host memory per step, stacked "
                    "alignment against {last key frame, last frame} with the motion prior, constant-velocity start, key frame "
                    "every 2nd frame; wall time per frame from host memory to the pose on the host"}


def cpu_leg(n_pairs, threads, gray, depth, cam, stride):
    """Times the CPU oracle (frame preparation + alignment) on the first n_pairs pairs of the frame array."""
    from oracle import oracle as O
    O.build()
    cfg = O.make_config()  # tdist / tdist nu=5, GaussNewton(1e-7,100), minGradient 30, prior off
    t0 = time.perf_counter()
    poses, iters = O.baseline_batch(cfg, gray, depth, *cam, LEVELS, threads, stride=stride, n_pairs=n_pairs,
                                    start=start_pose7())
    dt = time.perf_counter() - t0
    return n_pairs / dt, dt, poses, iters


REFERENCE_SAMPLE_PAIRS = 128  # the same sample on every box (the CPU arm's value is per pair)
REFERENCE_BUILD_PAIRS = 16    # pairs put through oracle/_ref (the reference's own sources, stand-in Eigen) beside the port


def reference_build_leg(gray, depth, cam, stride, threads, port_poses, port_iters):
    """oracle/_ref - the reference's own alignment sources compiled against the stand-in headers of oracle/ref_shim - on
    the first pairs of the sample, one pair per thread.  Reported beside the port, not as the baseline: the stand-in
    Eigen is a plain-loop, heap-allocating restatement of the surface the sources use, several times slower than the
    vectorised library the reference is really built with, so its time would flatter the GPU.  What it does show is that
    the port does the reference's work: step counts and poses of the two on the same pairs."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as O, ref as R
    if not R.available():
        return None
    n = min(REFERENCE_BUILD_PAIRS, len(port_poses))
    cfg = R.config_from(O.make_config())
    cam4 = np.array(cam, np.float64)

    def one(k):
        a, b = k * stride, k * stride + 1
        return R.align(cfg, gray[[a, b]], depth[[a, b]].astype(np.float64), cam4, np.stack([np.zeros(6), START_TWIST]), LEVELS,
                       want_trace=False)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=min(threads, n)) as ex:
        out = list(ex.map(one, range(n)))
    dt = time.perf_counter() - t0
    same = [k for k in range(n) if out[k]["pushed"].tolist() == [int(x) for x in port_iters[k]]]
    diff = max((float(np.abs(O.se3_log(port_poses[k]) - out[k]["twist"]).max()) for k in same), default=None)
    return {"value": n / dt, "unit": "pairs/s", "pairs": n, "threads": min(threads, n), "seconds": dt,
            "pairs_with_the_port's_step_counts": len(same), "max_pose_diff_vs_port_on_those": diff,
            "note": "stand-in Eigen (oracle/ref_shim): validates the port, not a fair timing of the reference"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU aligner on the host cores, all threads, on a bounded, FIXED sample of the
    same workload per step: the first 128 pairs.  The timed implementation is the oracle port (oracle/vslam_oracle.cpp,
    pinned against the reference's own sources by tests/test_oracle_ref_cpu.py); oracle/_ref, those sources compiled
    against a stand-in Eigen, runs beside it on 16 pairs (reference_build_leg says why it is not the timed one).  The
    1-core figure (the reference aligner is single-threaded) is measured once on 8 pairs."""
    if rank != 0:
        return
    from vslam_b200 import synth
    threads = os.cpu_count() or 1
    n_pairs = REFERENCE_SAMPLE_PAIRS
    cam = synth.cam_for(W, H)
    if args.workload == "pairs":
        gray, depth, _ = render_pairs(synth, range(n_pairs), cam)
        stride = 2
    else:
        gray, depth, _ = synth.make_sequence(n_pairs + 1, W, H, seed=1234, cam=cam)
        stride = 1
    depth = quantise_depth(depth)
    for _ in range(args.warmup):
        cpu_leg(min(n_pairs, max(threads, 8)), threads, gray, depth, cam, stride)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_leg(n_pairs, threads, gray, depth, cam, stride)
    dt = time.perf_counter() - t0
    v = n_pairs * args.steps / dt
    v1, dt1, _, _ = cpu_leg(8, 1, gray, depth, cam, stride)
    _, _, port_poses, port_iters = cpu_leg(REFERENCE_BUILD_PAIRS, threads, gray, depth, cam, stride)
    ref_build = reference_build_leg(gray, depth, cam, stride, threads, port_poses, port_iters)
    P = args.pairs if args.workload == "pairs" else args.frames - 1
    text = WORKLOAD_TEXT[args.workload].format(P=P, last=1234 + P - 1, F=args.frames)
    sample = f"the first {n_pairs} pairs of the workload per step, pair-parallel OpenMP over {threads} threads"
    print(json.dumps({
        "impl": "reference", "metric": "aligned frame pairs/sec at 640x480 RGB-D", "value": v, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong" if args.workload == "pairs" else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": text},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample,
                         "one_core": {"value": v1, "pairs": 8, "seconds": dt1}, "reference_build": ref_build},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", choices=["pairs", "sequence"], default="pairs")
    ap.add_argument("--pairs", type=int, default=4096, help="pairs workload: pairs of the whole job (all GPUs together)")
    ap.add_argument("--frames", type=int, default=1000, help="sequence workload: frames per GPU")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--cpu-pairs", type=int, default=32, help="pairs timed on ONE host core for cpu_baseline")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=4, help="sub-batches a step's frames are uploaded in (end-to-end leg)")
    ap.add_argument("--distinct-sequences", action="store_true",
                    help="sequence workload: every rank renders its own sequence (seed 1234 + rank) instead of the same one")
    args = ap.parse_args()
    rank, world, local = dist_env()
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    from vslam_b200 import build, capi, shard, synth
    if not os.path.exists(capi.LIB_PATH):
        build.build_all()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the alignment path has no CPU fallback")
    torch.cuda.set_device(local)
    use_dist = world > 1
    if use_dist:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    cam = synth.cam_for(W, H)
    if args.workload == "pairs":
        P = args.pairs
        mine = shard.deal_pairs(P, world, rank)           # strong scaling: the P pairs are dealt over the ranks
        gray, depth, gt = render_pairs(synth, mine, cam)
        n_jobs, stride, n_frames = len(mine), 2, 2 * len(mine)
        pairs_per_step, scaling, cap = P, "strong", (P + world - 1) // world
        text = WORKLOAD_TEXT["pairs"].format(P=P, last=1234 + P - 1)
    else:
        F = args.frames
        gray, depth, _ = synth.make_sequence(F, W, H, seed=1234 + (rank if args.distinct_sequences else 0), cam=cam)
        gt, mine = None, np.arange(F - 1)
        n_jobs, stride, n_frames = F - 1, 1, F
        pairs_per_step, scaling, cap = (F - 1) * world, "weak", F - 1
        text = WORKLOAD_TEXT["sequence"].format(P=F - 1, F=F)
    # The frames are what an RGB-D sensor delivers: 8-bit grey and 16-bit depth in units of 1 / 5000 m (the TUM convention,
    # test_alignment_se3.cpp:99).  The resident leg and the CPU oracle get the same values as f32 (k / 5000.0 in double,
    # narrowed), the end-to-end leg ships the 16-bit planes and converts on the device.
    k16 = np.clip(np.rint(depth.astype(np.float64) * DEPTH_DIVISOR), 0, 65535).astype(np.uint16)
    depth = (k16.astype(np.float64) / DEPTH_DIVISOR).astype(np.float32)
    h_gray = torch.from_numpy(gray).pin_memory()
    h_k16 = torch.from_numpy(k16.view(np.int16)).pin_memory()
    d_gray = h_gray.cuda()
    d_depth = torch.from_numpy(depth).cuda()
    ids = np.arange(n_frames, dtype=np.uint64)
    cams = np.tile(np.array(cam, np.float64), (n_frames, 1))
    ctx = capi.Context(capi.make_config(), device=local)
    stream = torch.cuda.Stream()
    capi.check(capi.lib().vslamb200_set_stream(ctx.h, C.c_void_p(stream.cuda_stream)))
    jobs, _keep = make_jobs(capi, n_jobs, stride)
    results = (capi.Result * n_jobs)()
    # the only collective of the path: every rank's poses (7 f64 per pair, shards padded to the largest one), all_gather
    # over NCCL on the alignment's stream, inside the timed region
    d_mine = torch.zeros((cap, 7), dtype=torch.float64, device="cuda")
    d_all = torch.zeros((world * cap, 7), dtype=torch.float64, device="cuda")

    def gather():
        if use_dist:
            d_mine[:n_jobs].copy_(torch.from_numpy(result_poses(capi, results, n_jobs)))
            dist.all_gather_into_tensor(d_all, d_mine)

    def step_resident():
        # the resident frames are used in place (vslamb200_frames_attach_device): no device-to-device copy of the inputs
        ctx.attach_device(ids, d_gray.data_ptr(), d_depth.data_ptr(), n_frames, W, H, cams, LEVELS)
        ctx.align_raw(jobs, n_jobs, results)
        gather()

    # End to end through the C ABI with pinned HOST buffers.  Successive steps (batches) alternate between two sets of
    # frame ids: the upload of the next step's frames (copy stream) is queued before this step's alignment is called, so
    # in steady state the PCIe transfer hides under the kernels.  Every step's upload, including the first one, every
    # step's result read-back and every step's gather are inside the timed region.
    ids2 = [ids, ids + np.uint64(n_frames)]
    j1, k1 = make_jobs(capi, n_jobs, stride, id0=n_frames)
    jobs2, _keep2 = [jobs, j1], [_keep, k1]

    # A step's frames travel in `chunks` sub-batches (independent pairs only): the alignment of sub-batch c runs under the
    # transfer of sub-batch c+1, so the first kernels start after 1/chunks of the step's bytes and the last step ends
    # one sub-batch's compute after its last byte - what vslamb200_frames_upload_u16_async is for (vslam_b200.h).
    chunks = max(1, min(args.e2e_chunks, n_jobs // 256)) if args.workload == "pairs" else 1
    cuts = [n_jobs * c_ // chunks for c_ in range(chunks + 1)]
    job_size, res_size = C.sizeof(capi.Job), C.sizeof(capi.Result)

    def upload_part(q, c_):
        a, b = stride * cuts[c_], stride * cuts[c_ + 1] + (0 if stride == 2 else 1)
        ctx.upload_u16_async(ids2[q][a:b], h_gray.data_ptr() + a * W * H, h_k16.data_ptr() + 2 * a * W * H, DEPTH_DIVISOR,
                             b - a, W, H, cams[a:b], LEVELS)

    def align_part(q, c_):
        a, b = cuts[c_], cuts[c_ + 1]
        capi.check(capi.lib().vslamb200_align(ctx.h, C.c_void_p(C.addressof(jobs2[q]) + a * job_size), b - a,
                                              C.c_void_p(C.addressof(results) + a * res_size)))

    def run_e2e(steps):
        work = [(st_, c_) for st_ in range(steps) for c_ in range(chunks)]
        upload_part(0, 0)
        for i, (st_, c_) in enumerate(work):
            if i + 1 < len(work):
                upload_part(work[i + 1][0] & 1, work[i + 1][1])
            align_part(st_ & 1, c_)
            if c_ == chunks - 1:
                gather()

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

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    ctx.reset_stats()
    ms = timed(step_resident, args.steps)
    own_ms = timed.own_ms
    st = ctx.stats()
    clocks = sampler.stop()
    value = pairs_per_step * args.steps / (ms * 1e-3)
    resident_poses = result_poses(capi, results, n_jobs)
    all_poses = d_all.cpu().numpy() if use_dist else None

    # end to end through the C ABI with host buffers (the raw device copies of the resident leg are no longer needed)
    ctx.synchronize()
    ctx.release_all()  # the attached frames point into these buffers
    del d_gray, d_depth
    torch.cuda.empty_cache()
    with torch.cuda.stream(stream):
        run_e2e(2)
    ms_e2e = timed(run_e2e, args.steps, loop=False)
    e2e_diff = float(np.abs(result_poses(capi, results, n_jobs) - resident_poses).max())
    e2e = pairs_per_step * args.steps / (ms_e2e * 1e-3)
    h2d = n_frames * W * H * 3  # u8 grey + u16 depth
    d2h = n_jobs * C.sizeof(capi.Result)

    # p50 / p95 latency of ONE alignment (BASELINE.json metric, second half): batch size 1, this rank's first
    # pair (frames resident), back-to-back calls, wall time from the host's submit to the result on the host
    lat = None
    if rank == 0:
        one = (capi.Result * 1)()
        for _ in range(50):
            ctx.align_raw(jobs2[(args.steps - 1) & 1], 1, one)
        ts = []
        for _ in range(300):
            t0 = time.perf_counter()
            ctx.align_raw(jobs2[(args.steps - 1) & 1], 1, one)
            ts.append((time.perf_counter() - t0) * 1e3)
        # a single job runs on larger clusters than a throughput batch (latency mode), so its sums are added in another
        # order: the same pose up to rounding, not bit for bit
        lat_diff = float(np.max(np.abs(np.array(list(one[0].pose)) - resident_poses[0])))
        assert lat_diff < 1e-9, "single-pair call differs from the batched result"
        lat = {"p50_ms": float(np.percentile(ts, 50)), "p95_ms": float(np.percentile(ts, 95)), "calls": len(ts),
               "max_pose_diff_vs_batched": lat_diff,
               "what": "one 640x480 pair, 4 levels, frames resident, host submit -> pose on host"}

    # per-kernel timing pass (events around every level launch; separate from the timed runs, one stream)
    capi.check(capi.lib().vslamb200_set_timing(ctx.h, 1))
    ctx.reset_stats()
    ctx.align_raw(jobs2[(args.steps - 1) & 1], n_jobs, results)
    tst = ctx.stats()
    capi.check(capi.lib().vslamb200_set_timing(ctx.h, 0))
    peak, peak_kind = load_peaks()
    # dominant kernel: gn_level_kernel at level 0 (one launch per step and GPU); algorithmic bytes = 10 B per template
    # pixel and Gauss-Newton evaluation x the pixel-evaluations that launch processed
    l0_ms = tst["ms_gn_level"][0]
    l0_bytes = ALGO_BYTES_PER_PIXEL_ITER * tst["pixel_iterations_level"][0]
    all_ms = sum(tst["ms_gn_level"])
    achieved = l0_bytes / (l0_ms * 1e-3) / 1e9 if l0_ms > 0 else 0.0
    traffic = load_traffic(args.workload, n_jobs)
    roofline = {
        "bound": "hbm", "kernel": "gn_level_kernel (level 0: all Gauss-Newton iterations of 640x480 for this GPU's share of the batch)",
        "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic, "traffic_over_algorithmic": (traffic / l0_bytes) if traffic and l0_bytes else None,
        "algorithmic_bytes_per_launch": l0_bytes, "ms_per_launch": l0_ms,
        "all_levels": {"ms": tst["ms_gn_level"][:LEVELS], "pixel_iterations": tst["pixel_iterations_level"][:LEVELS],
                       "achieved": ALGO_BYTES_PER_PIXEL_ITER * tst["pixel_iterations"] / (all_ms * 1e-3) / 1e9 if all_ms > 0 else 0.0},
        "ms_lists": tst["ms_lists"], "point_iterations": tst["point_iterations"],
        "interest_density": tst["point_iterations"] / max(1, tst["pixel_iterations"]),
        "level_kernel_launches": {"dense": int(tst["dense_launches"]), "moments": int(tst["moment_launches"])},
    }
    iters_per_pair = float(np.mean([sum(r.iters[:LEVELS]) for r in results]))
    assert all(r.status == 0 and sum(r.iters[:LEVELS]) > 0 for r in results), "a job did not run"

    ranks = None
    if use_dist:
        # per-rank step time, work, clocks and throttle reasons (the reported time is the maximum over the ranks)
        ranks = [None] * world
        dist.all_gather_object(ranks, {"rank": rank, "pairs": n_jobs, "ms_per_step": own_ms / args.steps,
                                       "gn_iterations": int(st["gn_iterations"] // max(1, args.steps)),
                                       "sm_mhz": clocks.get("sm_mhz"), "reasons": clocks.get("reasons"),
                                       "ms_gn_level": [round(x, 3) for x in tst["ms_gn_level"][:LEVELS]],
                                       "ms_lists": round(tst["ms_lists"], 3)})

    checks = {}
    if rank == 0 and args.workload == "pairs":
        # every pair of the whole batch (all ranks: the gathered array) recovers its ground-truth motion to the accuracy
        # the reference's own alignment tests ask for (test_alignment_se3.cpp:119: 0.01)
        from oracle import oracle as O
        allp = resident_poses if not use_dist else np.stack(
            [all_poses[(k % world) * cap + k // world] for k in range(args.pairs)])
        errs = np.empty(args.pairs)
        for k in range(args.pairs):
            xi = synth.pair_twist(1234 + k)
            Rk, tk = synth.se3_exp(xi)
            d = O.se3_log(O.se3_mul(allp[k], O.se3_inv(synth.pose7(Rk, tk))))
            errs[k] = max(float(np.linalg.norm(d[:3])), float(np.linalg.norm(d[3:])))
        # (a Gauss-Newton run can stop in a local minimum; whether the GPU's answer is the reference algorithm's answer is
        # what the comparison with the CPU oracle below and the parity tests check)
        checks["error_vs_ground_truth"] = {"max": float(errs.max()), "median": float(np.median(errs)),
                                           "pairs_within_0.01": int((errs < 0.01).sum()), "pairs": int(args.pairs),
                                           "worst_pair": int(errs.argmax())}
        # With the chi2 the reference's Release build computes (VSLAMB200_CHI2_REFERENCE: w[0] * sum r^2, vslam_b200.h) the
        # levels stop earlier than with the weighted sum: about nine pairs in ten reach 0.01 (the CPU oracle and oracle/_ref
        # give the same poses); every pair is within 0.02.
        assert np.median(errs) < 0.01 and (errs < 0.02).mean() > 0.99, checks

    odo = odometry_leg(capi, synth, shard, cam) if rank == 0 and args.workload == "pairs" else None
    cpu = None
    if rank == 0 and not args.no_cpu:
        from oracle import oracle as O
        v1, dt1, _, _ = cpu_leg(args.cpu_pairs, 1, gray, depth, cam, stride)
        thr = os.cpu_count() or 1
        nN = min(n_jobs, REFERENCE_SAMPLE_PAIRS)
        vN, dtN, cposes, citers = cpu_leg(nN, thr, gray, depth, cam, stride)
        # The GPU's results for the CPU sample: the same iteration counts per level and the same pose.  A pair may differ
        # where GaussNewton.cpp:70's test "chi2 increased" meets an exact tie (unchanged residual histogram: the GPU's
        # histogram-based chi2 is then bit-identical, a point-order sum differs in the last bit and stops one step
        # earlier) - such pairs are counted, not hidden
        same = [list(results[i].iters[:LEVELS]) == list(citers[i]) for i in range(nN)]
        perr = max([np.linalg.norm(O.se3_log(O.se3_mul(resident_poses[i], O.se3_inv(cposes[i])))) for i in range(nN) if same[i]] + [0.0])
        ties = [int(mine[i]) for i in range(nN) if not same[i]]
        assert perr < 1e-4 and len(ties) <= max(1, nN // 50), (perr, ties)
        cpu = {"value": v1, "unit": "pairs/s", "cores": 1, "kind": "port",
               "sample": f"this GPU's first {args.cpu_pairs} pairs, frame preparation included ({dt1:.1f} s)",
               "all_cores": {"value": vN, "cores": thr, "pairs": nN, "seconds": dtN},
               "max_pose_diff_vs_gpu": float(perr), "pairs_checked": nN, "pairs_with_equal_iteration_counts": int(sum(same)),
               "chi2_tie_pairs": ties}
        # and against oracle/_ref, the reference's own alignment sources (stand-in Eigen), on the first pairs of the sample
        gpu_iters = [[int(x) for x in results[i].iters[:LEVELS]] for i in range(min(nN, REFERENCE_BUILD_PAIRS))]
        rb = reference_build_leg(gray, depth, cam, stride, thr, resident_poses, gpu_iters)
        if rb is not None:
            rb["compared_with"] = "the GPU's poses and step counts"
            rb["pairs_with_the_gpu's_step_counts"] = rb.pop("pairs_with_the_port's_step_counts")
            rb["max_pose_diff_vs_gpu_on_those"] = rb.pop("max_pose_diff_vs_port_on_those")
            assert rb["pairs_with_the_gpu's_step_counts"] >= rb["pairs"] - 1 and (rb["max_pose_diff_vs_gpu_on_those"] or 0) < 1e-6, rb
        cpu["reference_build"] = rb
    if rank == 0:
        print(json.dumps({
            "metric": "aligned frame pairs/sec at 640x480 RGB-D", "value": value, "unit": "pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": text, "pairs_per_step": pairs_per_step, "pairs_this_gpu": n_jobs,
                       "collective": ("all_gather of the poses (7 f64 per pair) over NCCL, inside the timed region"
                                      if use_dist else "none (one GPU)"),
                       "l2_policy": f"inputs ({n_frames * W * H * 5 / 1e9:.1f} GB of level-0 planes per step and GPU) are larger than the 126 MB L2",
                       "mean_gn_iterations_per_pair": iters_per_pair},
            "e2e": {"value": e2e, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "max_pose_diff_vs_resident": e2e_diff,
                    "pipeline": "host frames are u8 grey + u16 depth (3 B per pixel, vslamb200_frames_upload_u16_async); "
                                f"a step travels in {chunks} sub-batches, the transfer of a sub-batch (copy stream) overlaps the "
                                "alignment of the one before it; all uploads timed; bytes are per GPU"},
            "gpu_launches": int(st["kernel_launches"]),
            "latency": lat, "odometry_shaped": odo, "clocks": clocks, "ranks": ranks, "roofline": roofline, "cpu_baseline": cpu, "checks": checks,
        }))
    ctx.close()
    if use_dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
