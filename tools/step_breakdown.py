"""Where a bench step goes: wall time of vslamb200_align for P resident pairs against the CUDA-event times of its kernels
(lists + level kernels, measured in a second, timing-enabled call), and of the frame preparation alone."""
import argparse, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi, synth
import bench
ap = argparse.ArgumentParser(); ap.add_argument("--pairs", type=int, default=2048)
a = ap.parse_args()
W, H, L = 640, 480, 4
cam = synth.cam_for(W, H)
g8, d8, _ = bench.render_pairs(synth, range(8), cam)
F = 2 * a.pairs
gray = np.tile(g8, (F // 16 + 1, 1, 1))[:F]; depth = np.tile(d8, (F // 16 + 1, 1, 1))[:F]
ctx = capi.Context(capi.make_config())
t0 = time.perf_counter(); ctx.upload(np.arange(F), gray, depth, cam, L); ctx.synchronize(); t_up = time.perf_counter() - t0
jobs, keep = bench.make_jobs(capi, a.pairs, 2)
res = (capi.Result * a.pairs)()
for _ in range(2):
    ctx.align_raw(jobs, a.pairs, res)
ts = []
for _ in range(5):
    ctx.synchronize(); t0 = time.perf_counter(); ctx.align_raw(jobs, a.pairs, res); ts.append((time.perf_counter() - t0) * 1e3)
capi.check(capi.lib().vslamb200_set_timing(ctx.h, 1))
ctx.reset_stats(); ctx.align_raw(jobs, a.pairs, res); st = ctx.stats()
k = st["ms_lists"] + sum(st["ms_gn_level"][:L])
print(f"{a.pairs} pairs: align wall {min(ts):.2f} ms (median {np.median(ts):.2f}); kernels with timing on (one stream, serialised): "
      f"lists {st['ms_lists']:.2f} + levels {[round(x, 2) for x in st['ms_gn_level'][:L]]} = {k:.2f} ms; upload+prepare of {F} frames {t_up * 1e3:.1f} ms (first call)")
ctx.close()
