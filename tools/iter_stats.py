import ctypes as C, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vslam_b200 import capi, synth
W, H, L, F = 640, 480, 4, 1000
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(F, W, H, seed=1234, cam=cam)
ctx = capi.Context(capi.make_config())
ids = np.arange(F, dtype=np.uint64); cams = np.tile(np.array(cam, np.float64), (F, 1))
ctx.upload(ids, gray, depth, cam, L)
jobs, keep = bench.make_jobs(capi, F)
results = (capi.Result * (F - 1))()
ctx.align_raw(jobs, F - 1, results)
it = np.array([list(r.iters[:L]) for r in results])
st = np.array([list(r.stop[:L]) for r in results])
for l in range(L):
    x = it[:, l]
    print(f"level {l}: mean {x.mean():.1f} p50 {np.percentile(x,50):.0f} p90 {np.percentile(x,90):.0f} p99 {np.percentile(x,99):.0f} max {x.max()}  stop reasons {np.bincount(st[:, l], minlength=6)}")
print("corr(L0, L1) =", np.corrcoef(it[:, 0], it[:, 1])[0, 1])
# greedy list-scheduling bound for L0 on 74 cluster slots in index order vs longest-first
def makespan(d, slots=74):
    import heapq
    h = [0.0] * slots
    for x in d:
        t = heapq.heappop(h); heapq.heappush(h, t + x)
    return max(h)
d = it[:, 0].astype(float) + 1
print("L0 evals: sum/74 =", d.sum() / 74, " in-order makespan =", makespan(d), " longest-first =", makespan(sorted(d, reverse=True)))
d1 = it[:, 1].astype(float) + 1
print("L1 evals: sum/148 =", d1.sum() / 148, " in-order =", makespan(d1, 148), " longest-first =", makespan(sorted(d1, reverse=True), 148))
ctx.close()
