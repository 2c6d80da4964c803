import ctypes as C, sys, time, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vslam_b200 import capi, synth
W, H, L = 640, 480, 4
F = 1000
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(F, W, H, seed=1234, cam=cam)
hg, hd = torch.from_numpy(gray).pin_memory(), torch.from_numpy(depth).pin_memory()
ids = np.arange(F, dtype=np.uint64); cams = np.tile(np.array(cam, np.float64), (F, 1))
ctx = capi.Context(capi.make_config())
jobs = [bench.make_jobs(capi, F, 0), bench.make_jobs(capi, F, F)]
results = (capi.Result * (F - 1))()
idset = [ids, ids + np.uint64(F)]
def up(q): ctx.upload_async(idset[q], hg.data_ptr(), hd.data_ptr(), F, W, H, cams, L)
def al(q): ctx.align_raw(jobs[q][0], F - 1, results)
now = time.perf_counter
up(0); up(1); al(0); al(1); ctx.synchronize()
for rep in range(2):
    t0 = now(); up(0); t1 = now(); ctx.synchronize(); t2 = now()
    print(f"upload_async call {1e3*(t1-t0):.2f} ms, until done {1e3*(t2-t0):.2f} ms")
    t0 = now(); al(0); t1 = now()
    print(f"align alone {1e3*(t1-t0):.2f} ms")
    t0 = now(); up(1); t1 = now(); al(0); t2 = now(); ctx.synchronize(); t3 = now()
    print(f"upload_async(1) call {1e3*(t1-t0):.2f}; + align(0) {1e3*(t2-t0):.2f}; + sync {1e3*(t3-t0):.2f}")
ctx.close()
