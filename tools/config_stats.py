"""Throughput of the BASELINE configurations that bench.py does not time (they are parity-test cases): KITTI-shape pairs with
5 levels (GN and the LM extension), 1 frame against 8 keyframes at 1280x720, and the bilinear / depth-residual extension
at 640x480.  Frames resident, CUDA-synchronised wall time of vslamb200_align over `reps` repetitions (best)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi as A, synth  # noqa: E402

IDENT = np.array([0, 0, 0, 1, 0, 0, 0.0])


def timed(ctx, jobs, reps=5):
    ctx.align(jobs)
    best = 1e9
    for _ in range(reps):
        ctx.synchronize()
        t0 = time.perf_counter()
        res = ctx.align(jobs)
        best = min(best, time.perf_counter() - t0)
    return best, res


def pairs_case(name, W, H, L, n_frames, cfg):
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(n_frames, W, H, seed=77, cam=cam)
    ctx = A.Context(cfg)
    ctx.upload(np.arange(n_frames), gray, depth, cam, L)
    jobs = [dict(to=i + 1, to_pose=IDENT, frm=[i], from_poses=IDENT, multi=False) for i in range(n_frames - 1)]
    t, res = timed(ctx, jobs)
    t1, _ = timed(ctx, jobs[:1], reps=20)
    it = np.mean([sum(r["iters"][:L]) for r in res])
    print(f"{name}: {len(jobs)} pairs in {t * 1e3:.2f} ms = {len(jobs) / t:.0f} pairs/s, one pair {t1 * 1e3:.2f} ms, "
          f"mean iterations per pair {it:.1f}")
    ctx.close()


def keyframes_case(W=1280, H=720, L=4, K=8, n_targets=16, tag="", **cfg_kw):
    cam = synth.cam_for(W, H)
    xis = []
    for k in range(K):
        a = 2 * np.pi * k / K
        xis.append(np.array([0.03 * np.cos(a), 0.03 * np.sin(a), 0.0, 0.015 * np.sin(a), 0.015 * np.cos(a), 0.0]))
    rng = np.random.default_rng(5)
    for _ in range(n_targets):
        xis.append(np.concatenate([rng.uniform(-0.01, 0.01, 3), rng.uniform(-0.005, 0.005, 3)]))
    gray, depth, poses = [], [], []
    for xi in xis:
        R, t = synth.se3_exp(xi)
        g, d = synth.render(W, H, cam, R, t, scene_seed=505)
        gray.append(g); depth.append(d); poses.append(synth.pose7(R, t))
    ctx = A.Context(A.make_config(max_iterations=60, **cfg_kw))
    ctx.upload(np.arange(len(xis)), np.stack(gray), np.stack(depth), cam, L)
    jobs = [dict(to=K + j, to_pose=poses[0], frm=list(range(K)), from_poses=np.stack(poses[:K]), multi=True)
            for j in range(n_targets)]
    t, res = timed(ctx, jobs)
    t1, _ = timed(ctx, jobs[:1], reps=10)
    it = np.mean([sum(r["iters"][:L]) for r in res])
    print(f"config 5 (1 frame vs {K} keyframes, {W}x{H}, {L} levels{tag}): {n_targets} alignments in {t * 1e3:.2f} ms = "
          f"{n_targets / t:.0f} alignments/s, one alignment {t1 * 1e3:.2f} ms, mean iterations {it:.1f}")
    ctx.close()


if __name__ == "__main__":
    only = sys.argv[1] if len(sys.argv) > 1 else ""
    if only == "keyframes":
        keyframes_case()
        keyframes_case(tag=", dense kernel", level_kernel=A.KERNEL_DENSE)
        keyframes_case(tag=", 64 targets", n_targets=64)
        keyframes_case(tag=", 64 targets, dense kernel", n_targets=64, level_kernel=A.KERNEL_DENSE)
        sys.exit(0)
    # chi2 as the reference's Release build evaluates it (the default) and the weighted sum (vslam_b200.h)
    pairs_case("config 4 (KITTI-shape 1232x368, 5 levels, GN)", 1232, 368, 5, 129, A.make_config(max_iterations=60))
    pairs_case("config 4 (KITTI-shape 1232x368, 5 levels, GN, weighted chi2)", 1232, 368, 5, 129,
               A.make_config(max_iterations=60, chi2_mode=A.CHI2_WEIGHTED))
    pairs_case("config 4 (KITTI-shape 1232x368, 5 levels, LM extension, weighted chi2)", 1232, 368, 5, 129,
               A.make_config(max_iterations=60, solver=A.SOLVER_LM, lm_lambda0=1e-2, lm_up=8.0, lm_down=0.25,
                             chi2_mode=A.CHI2_WEIGHTED))
    keyframes_case()
    keyframes_case(tag=", dense kernel", level_kernel=A.KERNEL_DENSE)
    keyframes_case(tag=", weighted chi2", chi2_mode=A.CHI2_WEIGHTED)
    keyframes_case(tag=", weighted chi2, dense kernel", chi2_mode=A.CHI2_WEIGHTED, level_kernel=A.KERNEL_DENSE)
    pairs_case("extension (bilinear sampling, 640x480, 4 levels, GN)", 640, 480, 4, 65,
               A.make_config(max_iterations=30, sampling=A.SAMPLING_BILINEAR))
    pairs_case("extension (bilinear + depth residual, 640x480, 4 levels, GN)", 640, 480, 4, 65,
               A.make_config(max_iterations=30, sampling=A.SAMPLING_BILINEAR, depth_weight=1e6))
