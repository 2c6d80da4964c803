"""Times variants of the end-to-end step (1000-frame sequence): synchronous upload + one align; pipelined sub-batches;
sub-batched aligns on resident frames (no PCIe)."""
import ctypes as C, sys, time, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vslam_b200 import capi, synth
W, H, L = 640, 480, 4
F = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
cam = synth.cam_for(W, H)
gray, depth, _ = synth.make_sequence(F, W, H, seed=1234, cam=cam)
hg, hd = torch.from_numpy(gray).pin_memory(), torch.from_numpy(depth).pin_memory()
dg, dd = hg.cuda(), hd.cuda()
ids = np.arange(F, dtype=np.uint64); cams = np.tile(np.array(cam, np.float64), (F, 1))
ctx = capi.Context(capi.make_config())
jobs, keep = bench.make_jobs(capi, F)
results = (capi.Result * (F - 1))()
jp, rp = C.cast(jobs, C.c_void_p).value, C.cast(results, C.c_void_p).value
def align(p0, p1):
    capi.check(capi.lib().vslamb200_align(ctx.h, C.c_void_p(jp + p0 * C.sizeof(capi.Job)), p1 - p0, C.c_void_p(rp + p0 * C.sizeof(capi.Result))))
def bounds(n): return [0] + [1 + (F - 1) * (k + 1) // n for k in range(n)]
def sync_all():
    capi.check(capi.lib().vslamb200_frames_upload(ctx.h, F, capi.ptr(ids), W, H, C.c_void_p(hg.data_ptr()), 0, C.c_void_p(hd.data_ptr()), capi.DEPTH_F32, 0, capi.ROW_MAJOR, capi.ptr(cams), L))
    align(0, F - 1)
def piped(n):
    b = bounds(n)
    def up(k): ctx.upload_async(ids[b[k]:b[k+1]], hg.data_ptr() + b[k] * W * H, hd.data_ptr() + b[k] * W * H * 4, b[k+1] - b[k], W, H, cams[b[k]:b[k+1]], L)
    up(0)
    for k in range(n):
        if k + 1 < n: up(k + 1)
        align(max(b[k] - 1, 0), b[k + 1] - 1)
def resident_sub(n):
    b = bounds(n)
    ctx.upload_device(ids, dg.data_ptr(), dd.data_ptr(), F, W, H, cams, L)
    for k in range(n): align(max(b[k] - 1, 0), b[k + 1] - 1)
def upload_only():
    capi.check(capi.lib().vslamb200_frames_upload(ctx.h, F, capi.ptr(ids), W, H, C.c_void_p(hg.data_ptr()), 0, C.c_void_p(hd.data_ptr()), capi.DEPTH_F32, 0, capi.ROW_MAJOR, capi.ptr(cams), L))
def upload_async_only(n=4):
    b = bounds(n)
    for k in range(n): ctx.upload_async(ids[b[k]:b[k+1]], hg.data_ptr() + b[k] * W * H, hd.data_ptr() + b[k] * W * H * 4, b[k+1] - b[k], W, H, cams[b[k]:b[k+1]], L)
    ctx.synchronize()
def t(name, fn, reps=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    print(f"{name:28s} {1e3 * (time.perf_counter() - t0) / reps:8.2f} ms", flush=True)
t("sync upload + align", sync_all)
t("upload only (sync)", upload_only)
t("upload only (async x4)", upload_async_only)
for n in (1, 2, 4, 8): t(f"resident, {n} sub-batches", lambda n=n: resident_sub(n))
for n in (2, 4, 8, 16): t(f"pipelined, {n} sub-batches", lambda n=n: piped(n))
ctx.close()
