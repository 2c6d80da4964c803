"""Time of the interest-list construction (CUDA events inside the library, all pyramid levels) for a batch of 640x480
pairs, against its algorithmic bytes: per pixel 2 + 2 + 4 + 1 read (dIx, dIy, depth, intensity) and per retained point
8 (warp-stream tile entry) + 16 (list record) written."""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi, synth
ap = argparse.ArgumentParser(); ap.add_argument("--pairs", type=int, default=1024); ap.add_argument("--levels", type=int, default=4)
a = ap.parse_args()
W, H = 640, 480
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(9, W, H, seed=1234, cam=cam)
F = a.pairs + 1
gray = np.tile(gray, (F // 9 + 1, 1, 1))[:F]; depth = np.tile(depth, (F // 9 + 1, 1, 1))[:F]
ctx = capi.Context(capi.make_config(max_iterations=2))
ctx.upload(np.arange(F), gray, depth, cam, a.levels)
capi.check(capi.lib().vslamb200_set_timing(ctx.h, 1))
ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(a.pairs)]
best = 1e9
for rep in range(4):
    ctx.reset_stats()
    ctx.align(jobs)
    st = ctx.stats()
    best = min(best, st["ms_lists"])
n_all = sum((W >> l) * (H >> l) for l in range(a.levels))
points = st["point_iterations"] / max(1, st["gn_iterations"] / a.levels / a.pairs) / a.pairs  # retained points per pair, all levels
bytes_ = a.pairs * (9 * n_all + 24 * points)
p = os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")
peak = json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6547.2
print(f"lists {a.pairs} pairs: {best:.3f} ms ({best / a.pairs * 1e3:.2f} us per pair), {points / n_all:.3f} of the pixels retained, "
      f"algorithmic {bytes_ / 1e9:.3f} GB -> {bytes_ / best / 1e6:.0f} GB/s = {bytes_ / best / 1e6 / peak:.3f} of {peak} GB/s")
ctx.close()
