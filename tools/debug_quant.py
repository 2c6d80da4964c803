"""GPU against oracle on sensor-quantised depth: which pairs differ, and where (level, iteration, residual plane)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vslam_b200 import capi as A, synth
from oracle import oracle as O
O.build()
W, H, L, N = 640, 480, 4, int(sys.argv[1]) if len(sys.argv) > 1 else 128
cam = synth.cam_for(W, H)
gray = np.empty((2 * N, H, W), np.uint8); depth = np.empty((2 * N, H, W), np.float32)
for k in range(N):
    p = synth.make_pair(W, H, 1234 + k, cam=cam)
    gray[2 * k], gray[2 * k + 1], depth[2 * k], depth[2 * k + 1] = p["gray0"], p["gray1"], p["depth0"], p["depth1"]
k16 = np.clip(np.rint(depth.astype(np.float64) * 5000.0), 0, 65535).astype(np.uint16)
depth = (k16.astype(np.float64) / 5000.0).astype(np.float32)
cposes, citers = O.baseline_batch(O.make_config(), gray, depth, *cam, L, os.cpu_count(), stride=2)
for kern in (A.KERNEL_DENSE, A.KERNEL_MOMENTS):
    ctx = A.Context(A.make_config(level_kernel=kern))
    ctx.upload(np.arange(2 * N), gray, depth, cam, L)
    jobs = [dict(to=2 * k + 1, to_pose=O.IDENTITY, frm=[2 * k], from_poses=O.IDENTITY, multi=False) for k in range(N)]
    ctx.set_trace(True)
    got = ctx.align(jobs)
    bad = [k for k in range(N) if list(got[k]["iters"][:L]) != list(citers[k])]
    print("kernel", kern, "pairs with different iteration counts:", bad[:20], len(bad))
    if bad:
        k = bad[0]
        print(" pair", k, "gpu", list(got[k]["iters"][:L]), "cpu", list(citers[k]), "stop", list(got[k]["stop"][:L]))
        f0 = O.Frame(gray[2 * k], depth[2 * k].astype(np.float64), *cam, L)
        f1 = O.Frame(gray[2 * k + 1], depth[2 * k + 1].astype(np.float64), *cam, L)
        ref = O.align(O.make_config(), f0, O.IDENTITY, f1, O.IDENTITY, want_trace=True)
        tr = ctx.get_trace(k, L)
        done = False
        for l in range(L - 1, -1, -1):
            for i in range(min(got[k]["iters"][l], ref["iters"][l])):
                a, b = tr[l, i], ref["trace"][l, i]
                dA = np.abs(a[18:54] - b[18:54]).max() / np.abs(b[18:54]).max()
                if dA > 1e-9 or a[1] != b[1]:
                    print("  first difference at level", l, "iteration", i, "rel dA", dA, "n", a[1], b[1], "chi2", a[2], b[2], "scale", a[4], b[4])
                    # the iterate both used at this point: x before the step = previous x (or the level's start)
                    done = True
                    break
            if done:
                break
        # residual planes at identity on every level
        for l in range(L):
            e_g = ctx.evaluate(2 * k, 2 * k + 1, l, O.IDENTITY, False)
            e_c = O.evaluate(O.make_config(), f0, f1, l, O.IDENTITY, False)
            nd = int((e_g["rcode"] != e_c["rcode"]).sum())
            print("  level", l, "residual-plane differences at identity:", nd, "n", e_g["n"], e_c["n"])
            if nd:
                ys, xs = np.nonzero(e_g["rcode"] != e_c["rcode"])
                for y, x in list(zip(ys, xs))[:5]:
                    print("    (v,u)", y, x, "gpu", e_g["rcode"][y, x], "cpu", e_c["rcode"][y, x], "depth", depth[2 * k][y, x] if l == 0 else None)
    ctx.close()
# chi2 sequences of the level whose iteration count differs (last kernel's trace)
k = 28
if N > k:
    ctx = A.Context(A.make_config(level_kernel=A.KERNEL_DENSE))
    ctx.upload(np.arange(2 * N), gray, depth, cam, L)
    ctx.set_trace(True)
    got = ctx.align([dict(to=2 * k + 1, to_pose=O.IDENTITY, frm=[2 * k], from_poses=O.IDENTITY, multi=False)])
    tr = ctx.get_trace(0, L)
    f0 = O.Frame(gray[2 * k], depth[2 * k].astype(np.float64), *cam, L)
    f1 = O.Frame(gray[2 * k + 1], depth[2 * k + 1].astype(np.float64), *cam, L)
    ref = O.align(O.make_config(), f0, O.IDENTITY, f1, O.IDENTITY, want_trace=True)
    for l in (3, 2):
        print("level", l, "gpu chi2", [repr(tr[l, i][2]) for i in range(got[0]["iters"][l])], "stop", got[0]["stop"][l])
        print("level", l, "cpu chi2", [repr(ref["trace"][l, i][2]) for i in range(ref["iters"][l])], "stop", ref["stop"][l])
        print("   gpu step sizes", [tr[l, i][3] for i in range(got[0]["iters"][l])])
    # the evaluation at which the CPU stopped: its iterate is the GPU's x after its 2nd step at level 2
    se3 = O.IDENTITY
    for l in (3, 2):
        for i in range(got[0]["iters"][l] if l == 3 else 2):
            se3 = O.se3_mul(se3, O.se3_exp(-tr[l, i][54:60]))
    e_c = O.evaluate(O.make_config(), f0, f1, 2, se3, False)
    e_g = ctx.evaluate(2 * k, 2 * k + 1, 2, se3, False)
    print("third evaluation at level 2: cpu chi2", repr(e_c["chi2"]), "gpu chi2", repr(e_g["chi2"]), "previous", repr(tr[2, 1][2]))
    ctx.close()
