"""Time of the frame preprocessing (pyramids + gradients, CUDA events inside the library) and of the interest lists for a
batch of 640x480 frames, with the algorithmic bytes of SURVEY 8(d) (5 N0 + 9 N_all per frame) against the HBM peak."""
import argparse, os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi, synth
ap = argparse.ArgumentParser(); ap.add_argument("--frames", type=int, default=500); ap.add_argument("--levels", type=int, default=4)
a = ap.parse_args()
W, H = 640, 480
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(8, W, H, seed=1234, cam=cam)
gray = np.tile(gray, (a.frames // 8 + 1, 1, 1))[:a.frames]; depth = np.tile(depth, (a.frames // 8 + 1, 1, 1))[:a.frames]
ctx = capi.Context(capi.make_config())
ctx.upload(np.arange(a.frames), gray, depth, cam, a.levels)
capi.check(capi.lib().vslamb200_set_timing(ctx.h, 1))
best = 1e9
for rep in range(5):
    ctx.reset_stats()
    ctx.upload(np.arange(a.frames), gray, depth, cam, a.levels)
    best = min(best, ctx.stats()["ms_preprocess"])
n_all = sum((W >> l) * (H >> l) for l in range(a.levels))
bytes_ = a.frames * (5 * W * H + 9 * n_all)
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6547.2
print(f"preprocess {a.frames} frames: {best:.3f} ms, algorithmic {bytes_/1e9:.3f} GB -> {bytes_/best/1e6:.0f} GB/s = {bytes_/best/1e6/peak:.3f} of {peak} GB/s")
ctx.close()
