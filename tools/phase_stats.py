"""Per-level timing and per-phase SM clocks of a full alignment batch (F-1 pairs of a 640x480 sequence, 4 levels).
Phase slots of the dense kernel: phaseA, bar_hist, scale, phaseC, blockred, bar_sums, solve (7-9: scale sub-laps); of the
moment kernel: 0 warp+sample+moment moves (evaluations after the first), 1 histogram barrier, 2 scale+weights, 3 A/b from
the moments, 4 sums barrier, 5 cluster sums, 6 solve, 7 first evaluation's warp+sample, 8 moment build, 9 moved points (a
count).  VSLAMB200_KERNEL=dense|moments, VSLAMB200_CLUSTER, VSLAMB200_PARTS select the execution (capi.make_config)."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=257)
ap.add_argument("--levels", type=int, default=4)
ap.add_argument("--timing", type=int, default=1)
args = ap.parse_args()
W, H = 640, 480
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(args.frames, W, H, seed=1234, cam=cam)
ctx = capi.Context(capi.make_config())
ctx.upload(np.arange(args.frames), gray, depth, cam, args.levels)
ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(args.frames - 1)]
ctx.align(jobs)
capi.check(capi.lib().vslamb200_set_timing(ctx.h, args.timing))
ctx.reset_stats()
res = ctx.align(jobs)
st = ctx.stats()
print("KERNEL", os.environ.get("VSLAMB200_KERNEL"), "moment launches", st["moment_launches"],
      "dense launches", st["dense_launches"], "ms_gn_level",
      [round(x, 3) for x in st["ms_gn_level"][:args.levels]], "ms_lists", round(st["ms_lists"], 3))
for l in range(args.levels):
    pc = st["phase_clocks"][l]
    tot = sum(pc[:9] if st["moment_launches"] else pc[:7]) or 1
    ev = st["pixel_iterations_level"][l] / (W * H / 4 ** l)
    nsl = 9 if st["moment_launches"] else 7
    print(" level", l, "evals/job", round(ev / len(jobs), 2), "kclocks/job", [round(c / len(jobs) / 1e3, 1) for c in pc[:nsl]],
          "share", [round(c / tot, 3) for c in pc[:nsl]])
    if st["moment_launches"]:
        pts = st["point_iterations"] / max(1, st["gn_iterations"]) if l == 0 else 0
        print("   moved points per job and evaluation", round(pc[9] / len(jobs) / max(ev / len(jobs), 1), 1))
    else:
        print("   slots 7-9 (scale sub-laps)", pc[7:10])
ctx.close()
