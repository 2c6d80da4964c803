"""Profiling harness: F-1 frame pairs, ONE pyramid level (640x480), a fixed small number of Gauss-Newton
iterations, so that every residual / accumulate launch ncu sees is a full-batch level-0 launch."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=129)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--levels", type=int, default=1)
args = ap.parse_args()
W, H = 640, 480
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(args.frames, W, H, seed=1234, cam=cam)
ctx = capi.Context(capi.make_config(max_iterations=args.iters))
ctx.upload(np.arange(args.frames), gray, depth, cam, args.levels)
ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(args.frames - 1)]
for _ in range(2):
    res = ctx.align(jobs)
st = ctx.stats()
print("iters", res[0]["iters"][:args.levels])
names = ["phaseA", "bar_hist", "scale", "phaseC", "blockred", "bar_sums", "solve", "sc_scaler", "sc_publish_bar", "sc_table"]
for l in range(args.levels):
    pc = st["phase_clocks"][l]
    tot = sum(pc) or 1
    ev = st["pixel_iterations_level"][l] / (W * H / 4 ** l)   # job-evaluations at this level (single keyframe)
    print("level", l, "evals", int(ev), "clocks/job-eval", {n: round(c / max(1, ev), 1) for n, c in zip(names, pc)},
          "share", {n: round(c / tot, 3) for n, c in zip(names, pc)})
print({k: v for k, v in st.items() if k not in ("phase_clocks",)})
ctx.close()
