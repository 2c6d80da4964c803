import ctypes as C, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vslam_b200 import capi, synth
W, H, L, F = 640, 480, 4, 401
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(F, W, H, seed=1234, cam=cam)
out = {}
for parts in (1, 8):
    os.environ["VSLAMB200_PARTS"] = str(parts)
    ctx = capi.Context(capi.make_config())
    ctx.upload(np.arange(F, dtype=np.uint64), gray, depth, cam, L)
    jobs, keep = bench.make_jobs(capi, F)
    results = (capi.Result * (F - 1))()
    ctx.reset_stats()
    ctx.align_raw(jobs, F - 1, results)
    st = ctx.stats()
    out[parts] = (np.array([list(r.pose) for r in results]), np.array([list(r.iters[:L]) for r in results]))
    print(parts, "gn_iterations", st["gn_iterations"], "pixel_iterations_level", st["pixel_iterations_level"][:4], "sum iters", out[parts][1].sum(0))
    ctx.close()
print("max pose diff", np.abs(out[1][0] - out[8][0]).max(), "iters equal", np.array_equal(out[1][1], out[8][1]))
