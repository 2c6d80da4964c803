"""Dense against moment kernel on the same batch: per-job iteration counts, stop reasons and poses (and the moment
kernel against itself, run to run: it must be bit-identical)."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi, synth  # noqa: E402
ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=257)
ap.add_argument("--levels", type=int, default=4)
ap.add_argument("--cluster", type=int, default=0)
ap.add_argument("--width", type=int, default=640)
ap.add_argument("--height", type=int, default=480)
args = ap.parse_args()
W, H = args.width, args.height
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(args.frames, W, H, seed=1234, cam=cam)
ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(args.frames - 1)]
out = {}
for name, k in (("dense", capi.KERNEL_DENSE), ("moments", capi.KERNEL_MOMENTS)):
    ctx = capi.Context(capi.make_config(level_kernel=k, cluster_size=args.cluster if name == "moments" else 0))
    ctx.upload(np.arange(args.frames), gray, depth, cam, args.levels)
    out[name] = ctx.align(jobs)
    out[name + "2"] = ctx.align(jobs)
    ctx.close()
bad = 0
for j, (a, b, b2) in enumerate(zip(out["dense"], out["moments"], out["moments2"])):
    same = list(a["iters"]) == list(b["iters"]) and list(a["stop"]) == list(b["stop"])
    dp = np.abs(a["pose"] - b["pose"]).max()
    rep = np.abs(b["pose"] - b2["pose"]).max()
    if not same or dp > 1e-9 or rep > 0:
        bad += 1
        if bad <= 6:
            print("job", j, "dense", [int(x) for x in a["iters"][:4]], "moments", [int(x) for x in b["iters"][:4]], "pose diff", dp, "run-to-run", rep)
print("jobs that differ:", bad, "of", len(jobs))
