"""GPU parity tests: CUDA path (through the C ABI) against the CPU oracle on identical synthetic inputs.
Tolerances are BASELINE.json's: masks / indices bit-exact, pyramid and gradient images <= 1e-5 relative
(integer planes are required to be exact), per-iteration normal equations <= 1e-4 relative Frobenius,
final pose <= 1e-4 m and 1e-4 rad."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import ctypes as _C


def C_void(p):
    return _C.c_void_p(p)


NE_TOL = 1e-4
POSE_TOL = 1e-4


def _mk_cfgs(O, A, **kw):
    okw = dict(kw)
    return O.make_config(**okw), A.make_config(**kw)


def _pair(synth, W=640, H=480, seed=1234, holes=0):
    return synth.make_pair(W, H, seed, holes_permille=holes)


def _frames(O, p, levels):
    cam = p["cam"]
    f0 = O.Frame(p["gray0"], p["depth0"].astype(np.float64), *cam, levels)
    f1 = O.Frame(p["gray1"], p["depth1"].astype(np.float64), *cam, levels)
    return f0, f1


def _upload(ctx, p, levels, ids=(1, 2)):
    ctx.upload(np.array(ids), np.stack([p["gray0"], p["gray1"]]), np.stack([p["depth0"], p["depth1"]]), p["cam"], levels)


@pytest.fixture(params=["moments", "dense"])
def level_kernel(request, monkeypatch):
    """The two Gauss-Newton level kernels: the moment-based one (gn_mom.cu; the default wherever it applies: scalers
    without offset) and the dense cluster kernel (gn_level.cu; every configuration).  capi.make_config() turns the
    variable into vslamb200_config.level_kernel; the `ctx` fixture is created after this one (tests/conftest.py)."""
    monkeypatch.setenv("VSLAMB200_KERNEL", request.param)
    return request.param


def test_level_kernels_agree_and_repeat_on_a_multi_wave_batch(capi, synth, chi2_mode):
    """512 small jobs (more CTAs than the device holds at once, so SMs are shared between jobs in different phases): the
    moment kernel and the dense kernel give the same iteration counts and poses, and the moment kernel is bit-identical
    run to run.  (This is the case that exposed a stage refill overtaking the loads of the stage: see start_chunk.)"""
    F, W, H = 513, 160, 120
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=1234, cam=cam)
    ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
    jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(F - 1)]
    out = {}
    for name, k, cl in (("dense", capi.KERNEL_DENSE, 0), ("moments", capi.KERNEL_MOMENTS, 1)):
        c = capi.Context(capi.make_config(level_kernel=k, cluster_size=cl))
        try:
            c.upload(np.arange(F), gray, depth, cam, 1)
            out[name] = [c.align(jobs) for _ in range(2)]
        finally:
            c.close()
    for a, b, b2 in zip(out["dense"][0], out["moments"][0], out["moments"][1]):
        assert list(a["iters"]) == list(b["iters"]) and list(a["stop"]) == list(b["stop"])
        assert np.abs(a["pose"] - b["pose"]).max() < 1e-9
        assert np.array_equal(b["pose"], b2["pose"])


def _kernel_was_used(ctx, level_kernel, applicable=True):
    st = ctx.stats()
    m, d = st["moment_launches"], st["dense_launches"]
    if level_kernel == "moments" and applicable:
        return m > 0 and d == 0
    return d > 0 and m == 0


def rel_fro(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(np.asarray(b)), 1e-300)


def pose_err(O, a, b):
    d = O.se3_log(O.se3_mul(a, O.se3_inv(b)))
    return np.linalg.norm(d[:3]), np.linalg.norm(d[3:])


@pytest.mark.parametrize("W,H,levels,holes", [(640, 480, 4, 0), (640, 480, 4, 50), (160, 96, 3, 80), (1232, 368, 5, 0)])
def test_pyramid_and_gradients_exact(ctx, oracle, synth, W, H, levels, holes):
    p = _pair(synth, W, H, 7, holes)
    f0, _ = _frames(oracle, p, levels)
    _upload(ctx, p, levels)
    for l in range(levels):
        ref = f0.level(l)
        got = ctx.download(1, l)
        assert np.array_equal(got["I"], ref["I"]), f"intensity level {l}"
        assert np.array_equal(got["dIx"], ref["dIx"]) and np.array_equal(got["dIy"], ref["dIy"]), f"gradients level {l}"
        assert np.array_equal(got["Z"].astype(np.float64), ref["Z"]), f"depth level {l}"


@pytest.mark.parametrize("W,H,levels", [(8, 4, 1), (16, 8, 2), (40, 24, 2), (48, 40, 3), (36, 20, 2), (132, 36, 2),
                                        (1280, 720, 4)])
def test_pyramid_and_gradients_exact_on_random_images(ctx, oracle, W, H, levels):
    """The register-resident pyrDown / blur+Sobel kernels (widths that are multiples of 8 / 4) and the tiled ones (other
    widths) against the oracle on random and saturated images: smallest sizes, widths that switch kernels between
    levels, borders on every side."""
    rng = np.random.default_rng(W * 1000 + H)
    gray = rng.integers(0, 256, (2, H, W), dtype=np.uint8)
    gray[1] = np.where(rng.random((H, W)) < 0.5, 0, 255).astype(np.uint8)     # saturated checker noise
    depth = (rng.uniform(0.5, 4.0, (2, H, W)) * (rng.random((2, H, W)) > 0.2)).astype(np.float32)
    cam = (100.0, 100.0, W / 2 - 0.5, H / 2 - 0.5)
    ctx.upload(np.array([1, 2]), gray, depth, cam, levels)
    for k in range(2):
        f = oracle.Frame(gray[k], depth[k].astype(np.float64), *cam, levels)
        for l in range(levels):
            ref, got = f.level(l), ctx.download(k + 1, l)
            assert np.array_equal(got["I"], ref["I"]), (k, l)
            assert np.array_equal(got["dIx"], ref["dIx"]) and np.array_equal(got["dIy"], ref["dIy"]), (k, l)
            assert np.array_equal(got["Z"].astype(np.float64), ref["Z"]), (k, l)


def test_real_tum_frame_pair_matches_oracle(ctx, capi, oracle, level_kernel, chi2_mode):
    """The one real RGB-D frame the reference ships (odometry/test/resource; crops in tests/golden/tum_frame.npz, made by
    tests/golden/make_tum_frame.py): real texture and 40 % invalid depth.  Every pyramid level bit-exact, then the
    alignment of the two crops: same iteration counts, stop reasons and pose as the oracle, with every level kernel."""
    O, A = oracle, capi
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tum_frame.npz"))
    cam, L = (525.0 / 2, 525.0 / 2, 159.5, 119.5), 3
    gray = np.stack([g["gray0"], g["gray1"]])
    depth = np.stack([g["depth0"], g["depth1"]]).astype(np.float64) / 5000.0  # test_alignment_se3.cpp:99
    fr = [O.Frame(gray[k], depth[k], *cam, L) for k in range(2)]
    ctx.upload(np.array([1, 2]), gray, depth.astype(np.float32), cam, L)
    for k in range(2):
        for l in range(L):
            ref, got = fr[k].level(l), ctx.download(k + 1, l)
            assert np.array_equal(got["I"], ref["I"]), (k, l)
            assert np.array_equal(got["dIx"], ref["dIx"]) and np.array_equal(got["dIy"], ref["dIy"]), (k, l)
            assert np.array_equal(got["Z"].astype(np.float64), ref["Z"]), (k, l)
    kw = dict(max_iterations=50)
    ctx.set_config(A.make_config(**kw))
    ctx.reset_stats()
    ref = O.align(O.make_config(**kw), fr[0], O.IDENTITY, fr[1], O.IDENTITY)
    got = ctx.align([dict(to=2, to_pose=O.IDENTITY, frm=[1], from_poses=O.IDENTITY, multi=False)])[0]
    assert _kernel_was_used(ctx, level_kernel)
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:L]) == list(ref["iters"]), (got["iters"], ref["iters"])
    assert list(got["stop"][:L]) == list(ref["stop"])
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    assert rel_fro(got["cov"], ref["cov"]) < 1e-3


def test_single_frame_uploads(ctx, oracle, synth):
    """One frame per call (what the odometry loop does), f32 row-major: the direct copy path with a one-row pitch."""
    p = _pair(synth, 320, 240, 3)
    f0, _ = _frames(oracle, p, 3)
    ctx.upload(np.array([9]), p["gray0"][None], p["depth0"][None].astype(np.float32), p["cam"], 3)
    for l in range(3):
        ref, got = f0.level(l), ctx.download(9, l)
        assert np.array_equal(got["I"], ref["I"]) and np.array_equal(got["Z"].astype(np.float64), ref["Z"]), l


@pytest.mark.parametrize("depth_kind,bgr", [("u16", False), ("f32", True), ("f64", False)])
def test_rgbd_ingestion_matches_the_node(ctx, oracle, synth, depth_kind, bgr):
    """RGB -> grey, depth factor and NaN -> 0 fused into the upload (NodeRgbdAlignment.cpp:177-199),
    then the usual pyramids: every level equals the oracle fed with the oracle's own ingestion."""
    O = oracle
    W, H, L = 320, 240, 3
    cam = synth.cam_for(W, H)
    rng = np.random.default_rng(77)
    p = synth.make_pair(W, H, 9, cam=cam)
    # a colour image whose grey version is structured: the rendered grey plus per-channel offsets
    g = p["gray0"].astype(np.int32)
    rgb = np.clip(np.stack([g + rng.integers(-20, 21, g.shape), g + rng.integers(-20, 21, g.shape),
                            g + rng.integers(-20, 21, g.shape)], -1), 0, 255).astype(np.uint8)
    if depth_kind == "u16":
        depth = np.round(p["depth0"].astype(np.float64) * 5000.0).astype(np.uint16)
        depth[10:20, 30:50] = 0                                   # sensor holes
        divisor = 5000.0
        ref_depth = O.depth_ingest(depth.astype(np.float64), divisor)
    else:
        depth = p["depth0"].astype(np.float32 if depth_kind == "f32" else np.float64).copy()
        depth[40:60, 100:140] = np.nan                            # ROS depth images mark invalid pixels with NaN
        divisor = 1.0
        ref_depth = O.depth_ingest(depth.astype(np.float64), divisor)
    ctx.ingest_rgbd(np.array([5]), rgb[None], depth[None], cam, L, bgr=bgr, depth_divisor=divisor)
    ref = O.Frame(O.rgb_to_gray(rgb, bgr=bgr), ref_depth, *cam, L)   # f32_storage: the device keeps depth as f32
    for l in range(L):
        r, got = ref.level(l), ctx.download(5, l)
        assert np.array_equal(got["I"], r["I"]), f"intensity level {l}"
        assert np.array_equal(got["dIx"], r["dIx"]) and np.array_equal(got["dIy"], r["dIy"]), f"gradients level {l}"
        assert np.array_equal(got["Z"].astype(np.float64), r["Z"]), f"depth level {l}"
    assert not np.isnan(ctx.download(5, 0)["Z"]).any()


@pytest.mark.parametrize("loss,scaler", [("TDIST", "TDIST"), ("NONE", "IDENTITY"), ("HUBER", "MEAN"), ("TUKEY", "MEDIAN"),
                                         ("HUBER", "MEDIAN"), ("QUADRATIC", "IDENTITY")])
@pytest.mark.parametrize("holes", [0, 50])
def test_single_iteration_masks_and_normal_equations(ctx, capi, oracle, synth, loss, scaler, holes, chi2_mode):
    O, A = oracle, capi
    p = _pair(synth, 640, 480, 1234, holes)
    f0, f1 = _frames(O, p, 4)
    _upload(ctx, p, 4)
    ocfg = O.make_config(loss=getattr(O, "LOSS_" + loss), scaler=getattr(O, "SCALER_" + scaler), huber_c=10.0)
    ctx.set_config(A.make_config(loss=getattr(A, "LOSS_" + loss), scaler=getattr(A, "SCALER_" + scaler), huber_c=10.0))
    xi = 0.6 * p["xi"]
    se3 = O.se3_exp(xi)
    for level in range(4):
        for multi in (False, True):
            ref = O.evaluate(ocfg, f0, f1, level, se3, multi)
            got = ctx.evaluate(1, 2, level, se3, multi)
            assert np.array_equal(got["rcode"], ref["rcode"]), f"mask/residual plane differs at level {level}"
            assert got["n"] == ref["n"]
            assert abs(got["scale"] - ref["scale"]) <= 1e-9 * abs(ref["scale"])
            assert abs(got["offset"] - ref["offset"]) <= 1e-9 * max(1.0, abs(ref["offset"]))
            assert rel_fro(got["A"], ref["A"]) < NE_TOL
            assert rel_fro(got["b"], ref["b"]) < NE_TOL
            assert abs(got["chi2"] - ref["chi2"]) <= NE_TOL * abs(ref["chi2"])


def _teacher_forced_check(O, ocfg, frames_from, from_poses, f1, to_pose, multi, tr, iters, prior, n_levels):
    """Per-iteration parity of the normal equations: the nearest-neighbour sampling makes the chained iteration
    chaotic, so each GPU iteration is compared with the oracle evaluated at the GPU's own iterate."""
    K = len(frames_from)
    se3 = [O.se3_mul(to_pose, O.se3_inv(fp)) for fp in from_poses]
    x_pred = [O.se3_log(s) for s in se3]
    for l in range(n_levels - 1, -1, -1):
        # every level re-creates the warps from the current pose of keyframe 0's problem (SE3Alignment.cpp:113-121)
        pose = O.se3_mul(se3[0], from_poses[0])
        se3 = [O.se3_mul(pose, O.se3_inv(fp)) for fp in from_poses]
        for i in range(iters[l]):
            A = np.zeros((6, 6))
            b = np.zeros(6)
            chi2, n = 0.0, 0
            for k in range(K):
                e = O.evaluate(ocfg, frames_from[k], f1, l, se3[k], multi)
                Ak, bk = e["A"].copy(), e["b"].copy()
                if prior:
                    Ak = Ak / 255.0 ** 2 + np.eye(6)
                    bk = bk / 255.0 ** 2 + (x_pred[k] - O.se3_log(se3[k]))
                A += Ak
                b += bk
                chi2 += e["chi2"]
                n += e["n"]
            a = tr[l, i]
            assert a[0] == 1 and a[1] == n, (l, i)
            assert rel_fro(a[18:54].reshape(6, 6), A) < NE_TOL, (l, i)
            assert rel_fro(a[12:18], b) < NE_TOL, (l, i, a[12:18], b)
            assert abs(a[2] - chi2) <= NE_TOL * abs(chi2), (l, i)
            inc = O.se3_exp(-a[54:60])
            se3 = [O.se3_mul(s, inc) for s in se3]
            assert np.allclose(O.se3_log(se3[0]), a[6:12], rtol=0, atol=1e-9), (l, i)


@pytest.mark.parametrize("loss,scaler,prior", [("TDIST", "TDIST", False), ("TDIST", "TDIST", True), ("NONE", "IDENTITY", False),
                                               ("HUBER", "MEAN", False), ("TUKEY", "MEDIAN", True)])
def test_align_matches_oracle_per_iteration(ctx, capi, oracle, synth, loss, scaler, prior, level_kernel, chi2_mode):
    O, A = oracle, capi
    p = _pair(synth, 640, 480, 1234)
    f0, f1 = _frames(O, p, 4)
    _upload(ctx, p, 4)
    kw = dict(huber_c=10.0, max_iterations=50, include_prior=prior)
    ocfg = O.make_config(loss=getattr(O, "LOSS_" + loss), scaler=getattr(O, "SCALER_" + scaler), **kw)
    ctx.set_config(A.make_config(loss=getattr(A, "LOSS_" + loss), scaler=getattr(A, "SCALER_" + scaler), **kw))
    ctx.set_trace(True)
    ctx.reset_stats()
    ref = O.align(ocfg, f0, O.IDENTITY, f1, O.IDENTITY, want_trace=True)
    got = ctx.align([dict(to=2, to_pose=O.IDENTITY, frm=[1], from_poses=O.IDENTITY, multi=False)])[0]
    assert _kernel_was_used(ctx, level_kernel, applicable=scaler in ("TDIST", "IDENTITY"))
    tr = ctx.get_trace(0, 4)
    _teacher_forced_check(O, ocfg, [f0], [O.IDENTITY], f1, O.IDENTITY, False, tr, got["iters"], prior, 4)
    # chained run: same number of steps, same stop reasons, same pose
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:4]) == list(ref["iters"]), (got["iters"], ref["iters"])
    assert list(got["stop"][:4]) == list(ref["stop"])
    for l in range(4):
        for i in range(ref["iters"][l]):
            a, b = tr[l, i], ref["trace"][l, i]
            assert rel_fro(a[18:54], b[18:54]) < NE_TOL, (l, i)
            assert rel_fro(a[12:18], b[12:18]) < NE_TOL, (l, i)
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    assert rel_fro(got["cov"], ref["cov"]) < 1e-3


def test_batch_of_pairs_matches_oracle(ctx, capi, oracle, synth, level_kernel, chi2_mode):
    """BASELINE configs[2] in small: independent pairs in ONE batched call, each against the oracle."""
    O, A = oracle, capi
    n, L = 6, 4
    pairs = [_pair(synth, 640, 480, 1234 + k, holes=(30 if k % 2 else 0)) for k in range(n)]
    gray = np.stack([x for p in pairs for x in (p["gray0"], p["gray1"])])
    depth = np.stack([x for p in pairs for x in (p["depth0"], p["depth1"])])
    ctx.upload(np.arange(2 * n), gray, depth, pairs[0]["cam"], L)
    ctx.reset_stats()
    got = ctx.align([dict(to=2 * k + 1, to_pose=O.IDENTITY, frm=[2 * k], from_poses=O.IDENTITY, multi=False)
                     for k in range(n)])
    assert _kernel_was_used(ctx, level_kernel)
    ocfg = O.make_config()
    for k, p in enumerate(pairs):
        f0, f1 = _frames(O, p, L)
        ref = O.align(ocfg, f0, O.IDENTITY, f1, O.IDENTITY)
        assert list(got[k]["iters"][:L]) == list(ref["iters"]), k
        assert list(got[k]["stop"][:L]) == list(ref["stop"]), k
        dt, dr = pose_err(O, got[k]["pose"], ref["pose"])
        assert dt < POSE_TOL and dr < POSE_TOL, k
        # and the alignment recovers the ground-truth motion to the accuracy the reference's own tests ask for
        dt, dr = pose_err(O, got[k]["pose"], p["pose1"])
        assert dt < 0.01 and dr < 0.01, k


@pytest.mark.parametrize("cluster", [0, 1, 2, 4])
def test_moment_kernel_stays_exact_under_many_changes(capi, oracle, synth, cluster, chi2_mode):
    """The moment-based kernel moves the moments of every point whose residual changed from the old to the new cell;
    the fixed-point cells must stay exactly what a fresh build would give.  Pairs with a deliberately bad initial guess
    (large per-level motion: many points change per iteration), depth holes, every pyramid level: every iteration's normal
    equations against the oracle evaluated at the device's own iterate to 1e-9 (measured ~1e-14), and the count of moved
    points reported by the kernel is non-zero.  cluster: CTAs per job (0 = the automatic choice, which is the 16-CTA
    latency mode for this small batch; 1, 2, 4 are the slice sizes of the throughput batches)."""
    O, A = oracle, capi
    n, L = 4, 4
    pairs = [_pair(synth, 640, 480, 4321 + k, holes=(40 if k % 2 else 0)) for k in range(n)]
    gray = np.stack([x for p in pairs for x in (p["gray0"], p["gray1"])])
    depth = np.stack([x for p in pairs for x in (p["depth0"], p["depth1"])])
    kw = dict(max_iterations=40)
    ctx = A.Context(A.make_config(level_kernel=A.KERNEL_MOMENTS, cluster_size=cluster, **kw))
    try:
        ctx.upload(np.arange(2 * n), gray, depth, pairs[0]["cam"], L)
        bad = O.se3_exp(np.array([0.01, -0.008, 0.012, 0.004, -0.003, 0.005]))
        jobs = [dict(to=2 * k + 1, to_pose=(bad if k >= n // 2 else O.IDENTITY), frm=[2 * k], from_poses=O.IDENTITY,
                     multi=False) for k in range(n)]
        ctx.set_trace(True)
        got = ctx.align(jobs)
        st = ctx.stats()
        traces = [ctx.get_trace(k, L) for k in range(n)]
    finally:
        ctx.close()
    assert st["moment_launches"] == L and st["dense_launches"] == 0
    assert sum(st["phase_clocks"][l][9] for l in range(L)) > 0  # points whose moments moved
    ocfg = O.make_config(**kw)
    worst = 0.0
    for k, p in enumerate(pairs):
        f0, f1 = _frames(O, p, L)
        se3 = jobs[k]["to_pose"]
        for l in range(L - 1, -1, -1):
            for i in range(got[k]["iters"][l]):
                e = O.evaluate(ocfg, f0, f1, l, se3, False)
                a = traces[k][l, i]
                assert a[0] == 1 and a[1] == e["n"], (k, l, i)
                worst = max(worst, rel_fro(a[18:54].reshape(6, 6), e["A"]), rel_fro(a[12:18], e["b"]),
                            abs(a[2] - e["chi2"]) / max(abs(e["chi2"]), 1e-300))
                se3 = O.se3_mul(se3, O.se3_exp(-a[54:60]))
    assert worst < 1e-9, worst
    ocfg = O.make_config()
    for k, p in enumerate(pairs):
        f0, f1 = _frames(O, p, L)
        ref = O.align(ocfg, f0, O.IDENTITY, f1, jobs[k]["to_pose"])
        assert list(got[k]["iters"][:L]) == list(ref["iters"]), k
        dt, dr = pose_err(O, got[k]["pose"], ref["pose"])
        assert dt < POSE_TOL and dr < POSE_TOL, k


@pytest.mark.parametrize("K,prior", [(2, True), (3, False)])
def test_multi_keyframe_matches_oracle(ctx, capi, oracle, synth, K, prior, chi2_mode):
    """SE3Alignment::align(vector<from>, to): stacked normal equations over K keyframes (320x240, 3 levels)."""
    O, A = oracle, capi
    W, H, L = 320, 240, 3
    cam = synth.cam_for(W, H)
    rng = np.random.default_rng(5)
    xis = [np.concatenate([rng.uniform(-0.02, 0.02, 3), rng.uniform(-0.01, 0.01, 3)]) for _ in range(K + 1)]
    frames, poses, raw = [], [], []
    for xi in xis:
        R, t = synth.se3_exp(xi)
        g, d = synth.render(W, H, cam, R, t, scene_seed=77)
        raw.append((g, d))
        frames.append(O.Frame(g, d.astype(np.float64), *cam, L))
        poses.append(synth.pose7(R, t))
    ctx.upload(np.arange(K + 1), np.stack([g for g, _ in raw]), np.stack([d for _, d in raw]), cam, L)
    kw = dict(max_iterations=40, include_prior=prior)
    ocfg = O.make_config(**kw)
    ctx.set_config(A.make_config(**kw))
    ctx.set_trace(True)
    # keyframes at their true poses, target starts at keyframe 0's pose
    ref = O.align(ocfg, frames[:K], poses[:K], frames[K], poses[0], multi=True, want_trace=True)
    got = ctx.align([dict(to=K, to_pose=poses[0], frm=list(range(K)), from_poses=np.stack(poses[:K]), multi=True)])[0]
    tr = ctx.get_trace(0, L)
    _teacher_forced_check(O, ocfg, frames[:K], poses[:K], frames[K], poses[0], True, tr, got["iters"], prior, L)
    assert list(got["iters"][:L]) == list(ref["iters"])
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    if not prior:  # the stacked motion prior (+I per keyframe against A/255^2) pins the pose to the initial guess
        dt, dr = pose_err(O, got["pose"], poses[K])
        assert dt < 0.01 and dr < 0.01


def test_config4_kitti_shape_five_levels(ctx, capi, oracle, synth, level_kernel):
    """BASELINE configs[3] in the form the reference is well defined for: KITTI-shape 1232x368
    intensity + depth pair, 5 pyramid levels, Gauss-Newton (the reference has no LM solver on this path)."""
    O, A = oracle, capi
    W, H, L = 1232, 368, 5
    p = _pair(synth, W, H, 4004)
    f0, f1 = _frames(O, p, L)
    _upload(ctx, p, L)
    kw = dict(max_iterations=60)
    ctx.set_config(A.make_config(**kw))
    ctx.reset_stats()
    ref = O.align(O.make_config(**kw), f0, O.IDENTITY, f1, O.IDENTITY)
    got = ctx.align([dict(to=2, to_pose=O.IDENTITY, frm=[1], from_poses=O.IDENTITY, multi=False)])[0]
    assert _kernel_was_used(ctx, level_kernel)
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:L]) == list(ref["iters"]), (got["iters"], ref["iters"])
    assert list(got["stop"][:L]) == list(ref["stop"])
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    dt, dr = pose_err(O, got["pose"], p["pose1"])
    assert dt < 0.01 and dr < 0.01
    # the un-cropped KITTI size is ill-defined in the reference (intensity and depth pyramids disagree): refused
    q = _pair(synth, 1241, 376, 4004)
    ctx.upload(np.array([11, 12]), np.stack([q["gray0"], q["gray1"]]), np.stack([q["depth0"], q["depth1"]]), q["cam"], L)
    with pytest.raises(A.VslamB200Error):
        ctx.align([dict(to=12, to_pose=O.IDENTITY, frm=[11], from_poses=O.IDENTITY, multi=False)])


def test_config5_one_frame_against_eight_keyframes_720p(ctx, capi, oracle, synth, chi2_mode):
    """BASELINE configs[4]: frame-to-multi-keyframe alignment, 1 frame against 8 keyframes jointly at 1280x720,
    keyframes on a circle of radius 0.03 m / 0.015 rad around the target, stacked normal equations
    (SE3Alignment.cpp:106-146, InverseCompositionalStacked.cpp:25-48)."""
    O, A = oracle, capi
    W, H, L, K = 1280, 720, 4, 8
    cam = synth.cam_for(W, H)
    xis = []
    for k in range(K):
        a = 2 * np.pi * k / K
        xis.append(np.array([0.03 * np.cos(a), 0.03 * np.sin(a), 0.0, 0.015 * np.sin(a), 0.015 * np.cos(a), 0.0]))
    xis.append(np.zeros(6))  # the target frame sits in the middle
    frames, poses, raw = [], [], []
    for xi in xis:
        R, t = synth.se3_exp(xi)
        g, d = synth.render(W, H, cam, R, t, scene_seed=505)
        raw.append((g, d))
        frames.append(O.Frame(g, d.astype(np.float64), *cam, L))
        poses.append(synth.pose7(R, t))
    ctx.upload(np.arange(K + 1), np.stack([g for g, _ in raw]), np.stack([d for _, d in raw]), cam, L)
    kw = dict(max_iterations=30)
    ctx.set_config(A.make_config(**kw))
    # initial guess: keyframe 0's pose (3 cm / 0.9 deg away from the truth)
    ref = O.align(O.make_config(**kw), frames[:K], poses[:K], frames[K], poses[0], multi=True)
    got = ctx.align([dict(to=K, to_pose=poses[0], frm=list(range(K)), from_poses=np.stack(poses[:K]), multi=True)])[0]
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:L]) == list(ref["iters"]), (got["iters"], ref["iters"])
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    dt, dr = pose_err(O, got["pose"], poses[K])
    assert dt < 0.01 and dr < 0.01


@pytest.mark.gpu
def test_cpp_host_mirror_matches_oracle(oracle, synth, capi):
    """include/vslam_b200.hpp (the C++ mirror of AlignmentSE3 / SE3Alignment / Frame) driven like the reference's
    alignment test, with column-major u8 / f64 frames: same pose as the oracle on the same rendered pair."""
    import subprocess
    from vslam_b200 import build
    O = oracle
    exe = build.build_host_mirror()
    W, H, seed, L = 320, 240, 77, 3
    r = subprocess.run([exe, str(W), str(H), str(seed), str(L)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = dict((ln.split()[0], ln.split()[1:]) for ln in r.stdout.strip().splitlines())
    assert "ok" in lines
    got = np.array([float(x) for x in lines["pose"]])
    gotm = np.array([float(x) for x in lines["multi"]])
    p = synth.make_pair(W, H, seed)
    f0, f1 = _frames(O, p, L)
    cfg = O.make_config(max_iterations=30)
    ref = O.align(cfg, f0, O.IDENTITY, f1, O.IDENTITY)
    dt, dr = pose_err(O, got, ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    refm = O.align(cfg, [f0], [O.IDENTITY], f1, O.IDENTITY, multi=True)
    dt, dr = pose_err(O, gotm, refm["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    # solver->solve(problem) (least_squares::Problem / Solver::Results, Solver.h:22-63) reproduces align's iterates on one
    # level, without and with the motion prior: same number of steps, x / chi2 / step size / A per iteration, final pose
    for key in ("solve0", "solve1"):
        n_solve, n_trace, worst, pose_diff, stop = lines[key]
        assert int(n_solve) == int(n_trace) > 1, lines[key]
        assert float(worst) < 1e-9 and float(pose_diff) < 1e-9, lines[key]
    # Frame::intensity(l) / dIx(l) / dIy(l) / depth(l) / camera(l) / pcl(l) of a resident frame against the oracle's frame
    lv = [float(x) for x in lines["levels"]]
    for l in range(L):
        Wl, Hl, sI, sx, sy, sZ, fx, px, py, pz = lv[10 * l:10 * l + 10]
        ref_l = f1.level(l)
        assert (Wl, Hl) == (ref_l["I"].shape[1], ref_l["I"].shape[0])
        assert sI == float(ref_l["I"].astype(np.float64).sum())
        assert sx == float(np.abs(ref_l["dIx"].astype(np.float64)).sum()) and sy == float(np.abs(ref_l["dIy"].astype(np.float64)).sum())
        assert abs(sZ - float(ref_l["Z"].sum())) <= 1e-9 * abs(sZ)
        assert fx == p["cam"][0] * 0.5 ** l
        idx = (int(Wl) * int(Hl)) // 2 + int(Wl) // 3
        v, u = divmod(idx, int(Wl))
        d = float(ref_l["Z"][v, u])
        cam_l = [c * 0.5 ** l for c in p["cam"]]
        assert np.allclose([px, py, pz], [(u - cam_l[2]) / cam_l[0] * d, (v - cam_l[3]) / cam_l[1] * d, d], rtol=1e-12)


@pytest.mark.gpu
def test_async_upload_pipeline_is_identical_to_the_synchronous_path(capi, oracle, synth):
    """vslamb200_frames_upload_async + per-sub-batch align (the end-to-end pipeline of bench.py) returns bit-identical
    poses to one synchronous upload + one batched align, including when frame ids are re-uploaded."""
    import torch
    F, W, H, L = 9, 320, 240, 3
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=99, cam=cam)
    ident = oracle.IDENTITY
    jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(F - 1)]
    ctx = capi.Context(capi.make_config(max_iterations=30))
    try:
        ctx.upload(np.arange(F), gray, depth, cam, L)
        want = ctx.align(jobs)
        hg = torch.from_numpy(gray).pin_memory()
        hd = torch.from_numpy(depth).pin_memory()
        for rep in range(2):  # the second round replaces every frame while the arenas are recycled
            got = []
            bounds = [0, 4, 7, F]
            ids = np.arange(F, dtype=np.uint64)
            cams = np.tile(np.array(cam, np.float64), (F, 1))

            def up(k):
                a, b = bounds[k], bounds[k + 1]
                ctx.upload_async(ids[a:b], hg.data_ptr() + a * W * H, hd.data_ptr() + a * W * H * 4, b - a, W, H, cams[a:b], L)

            up(0)
            for k in range(3):
                if k + 1 < 3:
                    up(k + 1)
                p0, p1 = max(bounds[k] - 1, 0), bounds[k + 1] - 1
                got += ctx.align(jobs[p0:p1])
            assert len(got) == F - 1
            for a, b in zip(got, want):
                assert np.array_equal(a["pose"], b["pose"]) and list(a["iters"]) == list(b["iters"])
        # download after an asynchronous upload waits for it
        ctx.upload_async(ids[:2], hg.data_ptr(), hd.data_ptr(), 2, W, H, cams[:2], L)
        lv = ctx.download(1, 1)
        f1 = oracle.Frame(gray[1], depth[1].astype(np.float64), *cam, L)
        assert np.array_equal(lv["I"], f1.level(1)["I"])
        ctx.synchronize()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_attached_device_frames_are_identical_to_copied_ones(capi, oracle, synth):
    """vslamb200_frames_attach_device (level-0 planes used in place in the caller's device memory) gives the same
    pyramids, gradients and poses as vslamb200_frames_upload_device, which copies them; a misaligned buffer is refused."""
    import torch
    F, W, H, L = 6, 320, 240, 3
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=321, cam=cam)
    d_gray, d_depth = torch.from_numpy(gray).cuda(), torch.from_numpy(depth).cuda()
    ident = oracle.IDENTITY
    jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(F - 1)]
    jobs.append(dict(to=F - 1, to_pose=ident, frm=[0, 2], from_poses=np.stack([ident, ident]), multi=True))
    out = {}
    for how in ("copy", "attach"):
        ctx = capi.Context(capi.make_config(max_iterations=30))
        try:
            up = ctx.upload_device if how == "copy" else ctx.attach_device
            up(np.arange(F), d_gray.data_ptr(), d_depth.data_ptr(), F, W, H, cam, L)
            out[how] = (ctx.align(jobs), [ctx.download(2, l) for l in range(L)])
            if how == "attach":
                with pytest.raises(capi.VslamB200Error, match="aligned"):
                    ctx.attach_device(np.arange(F - 1) + 100, d_gray.data_ptr() + 1, d_depth.data_ptr(), F - 1, W, H, cam, L)
        finally:
            ctx.close()
    for a, b in zip(out["copy"][0], out["attach"][0]):
        assert np.array_equal(a["pose"], b["pose"]) and list(a["iters"]) == list(b["iters"])
    for a, b in zip(out["copy"][1], out["attach"][1]):
        for k in a:
            assert np.array_equal(a[k], b[k]), k


@pytest.mark.gpu
def test_reupload_of_an_id_whose_asynchronous_upload_is_still_running(capi, oracle, synth):
    """A frame id that is uploaded asynchronously and then, at once, synchronously with other content ends up holding the
    second content (make_frames orders the frame-table update behind the pending upload); the same id twice in one call
    is refused."""
    import torch
    F, W, H, L = 24, 640, 480, 4
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(2, W, H, seed=5, cam=cam)
    first_g = torch.from_numpy(np.repeat(gray[:1], F, 0)).pin_memory()
    first_d = torch.from_numpy(np.repeat(depth[:1], F, 0)).pin_memory()
    ids = np.arange(F, dtype=np.uint64)
    cams = np.tile(np.array(cam, np.float64), (F, 1))
    want = oracle.Frame(gray[1], depth[1].astype(np.float64), *cam, L)
    ctx = capi.Context(capi.make_config())
    try:
        for rep in range(3):
            ctx.upload_async(ids, first_g.data_ptr(), first_d.data_ptr(), F, W, H, cams, L)
            ctx.upload(ids[5:7], np.repeat(gray[1:], 2, 0), np.repeat(depth[1:], 2, 0), cam, L)   # other arena, same ids
            for fid in (5, 6):
                for lvl in (0, L - 1):
                    got, ref = ctx.download(fid, lvl), want.level(lvl)
                    assert np.array_equal(got["I"], ref["I"]) and np.array_equal(got["dIx"], ref["dIx"])
                    assert np.array_equal(got["Z"], ref["Z"].astype(np.float32))
            ctx.synchronize()
        with pytest.raises(capi.VslamB200Error, match="twice"):
            ctx.upload(np.array([3, 4, 3]), np.repeat(gray[:1], 3, 0), np.repeat(depth[:1], 3, 0), cam, L)
    finally:
        ctx.close()


@pytest.mark.gpu
def test_u16_depth_async_upload_matches_the_reference_ingestion(capi, oracle, synth):
    """vslamb200_frames_upload_u16_async: 16-bit sensor depth / 5000 (test_alignment_se3.cpp:99) converted on the device;
    every pyramid level equals the oracle's frame built from oracle.depth_ingest of the same values, the alignment is
    bit-identical to uploading the converted f32 depth, and a second upload re-uses the staging buffer."""
    import torch
    F, W, H, L = 5, 320, 240, 3
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=21, cam=cam)
    k16 = np.clip(np.rint(depth.astype(np.float64) * 5000.0), 0, 65535).astype(np.uint16)
    k16[:, 50:60, 70:90] = 0  # no measurement
    d32 = np.stack([oracle.depth_ingest(k16[i], 5000.0) for i in range(F)]).astype(np.float32)
    assert np.array_equal(d32, (k16.astype(np.float64) / 5000.0).astype(np.float32))
    ident = oracle.IDENTITY
    jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(F - 1)]
    ctx = capi.Context(capi.make_config(max_iterations=30))
    try:
        ctx.upload(np.arange(F), gray, d32, cam, L)
        want = ctx.align(jobs)
        hg = torch.from_numpy(gray).pin_memory()
        hk = torch.from_numpy(k16.view(np.int16)).pin_memory()
        for rep in range(2):
            ctx.upload_u16_async(np.arange(F), hg.data_ptr(), hk.data_ptr(), 5000.0, F, W, H, cam, L)
            got = ctx.align(jobs)
            for a, b in zip(got, want):
                assert np.array_equal(a["pose"], b["pose"]) and list(a["iters"]) == list(b["iters"])
        for i in (0, F - 1):
            fr = oracle.Frame(gray[i], d32[i].astype(np.float64), *cam, L)
            for l in range(L):
                lv, ref = ctx.download(i, l), fr.level(l)
                assert np.array_equal(lv["Z"].astype(np.float64), ref["Z"]) and np.array_equal(lv["I"], ref["I"])
        assert capi.lib().vslamb200_frames_upload_u16_async(ctx.h, 1, capi.ptr(np.arange(1, dtype=np.uint64)), W, H,
                                                            C_void(hg.data_ptr()), C_void(hk.data_ptr()), 0.0,
                                                            capi.ptr(np.array(cam, np.float64)), L) == -1
    finally:
        ctx.close()


@pytest.mark.gpu
def test_large_batch_two_stream_path_matches_oracle(capi, oracle, synth, level_kernel):
    """Batches of >= 192 jobs are cut into two halves that run on two streams (vslamb200_align); every pair of a
    200-frame sequence must still match the oracle (pose, and bit-identical to the single-stream run)."""
    import os
    F, W, H, L = 200, 160, 120, 3
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=5, cam=cam)
    ident = oracle.IDENTITY
    jobs = [dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(F - 1)]
    ref_poses, _ = oracle.baseline_sequence(oracle.make_config(), gray, depth, *cam, L, os.cpu_count() or 1)
    got = {}
    for parts in ("1", "2", "5"):
        ctx = capi.Context(capi.make_config(stream_parts=int(parts)))
        ctx.upload(np.arange(F), gray, depth, cam, L)
        got[parts] = ctx.align(jobs)
        ctx.close()
    ctx = capi.Context(capi.make_config())  # default policy: two halves for this batch size
    ctx.upload(np.arange(F), gray, depth, cam, L)
    got["default"] = ctx.align(jobs)
    ctx.close()
    for k in range(F - 1):
        assert got["1"][k]["status"] == 0 and sum(got["1"][k]["iters"][:L]) > 0, k
        dt, dr = pose_err(oracle, got["1"][k]["pose"], ref_poses[k])
        assert dt < POSE_TOL and dr < POSE_TOL, k
        for key in ("2", "5", "default"):
            assert np.array_equal(got[key][k]["pose"], got["1"][k]["pose"]), (key, k)
            assert list(got[key][k]["iters"]) == list(got["1"][k]["iters"]), (key, k)


@pytest.mark.gpu
@pytest.mark.parametrize("model,period", [("ConstantMotion", 3), ("NoMotion", 2)])
def test_cpp_odometry_mirror_matches_the_oracle_loop(oracle, synth, capi, tmp_path, model, period):
    """OdometryRgbd + Map rings + KeyFrameSelectionIdx + MotionPrediction of
    include/vslam_b200_odometry.hpp, driven like NodeRgbdAlignment::processFrame over a rendered sequence; every
    pose and speed equals oracle/odometry.py's on the same frames (single and stacked alignments, prior on)."""
    import subprocess
    from vslam_b200 import build
    from oracle import odometry as OD
    O = oracle
    exe = build.build_odometry_mirror()
    W, H, F, L = 320, 240, 9, 3
    traj = tmp_path / "trajectory.txt"
    r = subprocess.run([exe, str(W), str(H), str(F), str(L), str(period), model, str(traj)], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stderr
    rows = [ln.split() for ln in r.stdout.strip().splitlines()]
    assert rows[-1] == ["ok"]
    got_pose = {int(x[1]): np.array([float(v) for v in x[2:]]) for x in rows if x[0] == "pose"}
    got_speed = {int(x[1]): np.array([float(v) for v in x[2:]]) for x in rows if x[0] == "speed"}
    got_kind = {int(x[1]): x[2] for x in rows if x[0] == "kind"}
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=1234, cam=cam)
    frames = [O.Frame(gray[k], depth[k].astype(np.float64), *cam, L) for k in range(F)]
    stamps = [1000000000 + k * 33333333 for k in range(F)]
    cfg = O.make_config(max_iterations=50, include_prior=True)
    poses, speeds, kinds = OD.run(cfg, frames, stamps, prediction=model, kf_period=period)
    assert [got_kind[k] for k in range(F)] == kinds and "stacked" in kinds and "single" in kinds
    for k in range(F):
        dt, dr = pose_err(O, got_pose[k], poses[k])
        assert dt < POSE_TOL and dr < POSE_TOL, k
        dt, dr = pose_err(O, got_speed[k], O.se3_exp(speeds[k]))
        assert dt < 30 * POSE_TOL and dr < 30 * POSE_TOL, k       # speed = log(relative pose) / 0.0333 s
    assert len(traj.read_text().splitlines()) == 3 + F


@pytest.mark.gpu
def test_sequence_trajectory_rpe_and_ate_against_ground_truth(capi, oracle, synth, tmp_path, chi2_mode):
    """BASELINE configs[1] in small, end to end on both sides of the path: a rendered 40-frame sequence, frame-to-frame
    alignment of all pairs in one batched call, relative poses chained into a trajectory (vslam_b200/shard.py),
    written as TUM files and scored with the RPE / ATE of the reference's evaluation scripts
    (vslam_b200/evaluation.py).  The bounds are the CPU oracle's figures on the same sequence plus head-room: with the
    weighted chi2 RPE 4.6 mm / 1.9 mrad per frame at 320x240 and 3 levels, ATE 25 mm over 40 frames; with the chi2 of the
    reference's Release build (levels stop earlier) 7.1 mm / 3.3 mrad and 75 mm.  The reference's own tests ask < 0.01
    per pair."""
    from vslam_b200 import evaluation as E, shard
    F, W, H, L = 40, 320, 240, 3
    cam = synth.cam_for(W, H)
    gray, depth, gt = synth.make_sequence(F, W, H, seed=1234, cam=cam)
    ident = oracle.IDENTITY
    ctx = capi.Context(capi.make_config())
    try:
        ctx.upload(np.arange(F), gray, depth, cam, L)
        res = ctx.align([dict(to=i + 1, to_pose=ident, frm=[i], from_poses=ident, multi=False) for i in range(F - 1)])
    finally:
        ctx.close()
    traj = shard.chain_trajectory(np.stack([r["pose"] for r in res]))
    stamps = 1305031100.0 + np.arange(F) / 30.0
    E.write_trajectory(str(tmp_path / "est.txt"), stamps, traj)
    E.write_trajectory(str(tmp_path / "gt.txt"), stamps, gt)
    tg, te = E.read_trajectory(str(tmp_path / "gt.txt")), E.read_trajectory(str(tmp_path / "est.txt"))
    rpe = E.relative_pose_error(tg, te, 1, "f")
    assert len(rpe) == F - 2                                         # the script leaves out pairs ending at the last pose
    t_rmse, r_rmse, t_max, ate_rmse = (0.006, 0.003, 0.01, 0.035) if chi2_mode == capi.CHI2_WEIGHTED else (0.009, 0.0045, 0.013, 0.1)
    assert E.stats(rpe[:, 4])["rmse"] < t_rmse and E.stats(rpe[:, 5])["rmse"] < r_rmse
    assert E.stats(rpe[:, 4])["max"] < t_max
    ate = E.absolute_trajectory_error(tg, te)
    assert len(ate["matches"]) == F and ate["rmse"] < ate_rmse


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,L,K,prior", [(640, 480, 4, 1, False), (1232, 368, 5, 1, False), (320, 240, 3, 2, True)])
def test_levenberg_marquardt_extension_matches_oracle(ctx, capi, oracle, synth, W, H, L, K, prior):
    """SURVEY 8(f)-3 / BASELINE configs[3] ("LM solver"): Levenberg-Marquardt damping on the reference's normal
    equations.  The reference has no LM solver: oracle/vslam_oracle.cpp (levenberg_marquardt) is the definition, parity
    is GPU <-> that restatement - every evaluation's accepted A, b, lambda and step, the stop reasons and the pose."""
    O, A = oracle, capi
    cam = synth.cam_for(W, H)
    rng = np.random.default_rng(11)
    xis = [np.zeros(6)] + [np.concatenate([rng.uniform(-0.02, 0.02, 3), rng.uniform(-0.01, 0.01, 3)]) for _ in range(K)]
    frames, poses, raw = [], [], []
    for xi in xis:
        R, t = synth.se3_exp(xi)
        g, d = synth.render(W, H, cam, R, t, scene_seed=4004)
        raw.append((g, d))
        frames.append(O.Frame(g, d.astype(np.float64), *cam, L))
        poses.append(synth.pose7(R, t))
    ctx.upload(np.arange(K + 1), np.stack([g for g, _ in raw]), np.stack([d for _, d in raw]), cam, L)
    kw = dict(max_iterations=40, include_prior=prior, lm_lambda0=1e-2, lm_up=8.0, lm_down=0.25,
              chi2_mode=O.CHI2_WEIGHTED)  # accept / reject needs the weighted sum the steps minimise
    ocfg = O.make_config(solver=O.SOLVER_LM, **kw)
    ctx.set_config(A.make_config(solver=A.SOLVER_LM, **kw))
    ctx.set_trace(True)
    multi = K > 1
    # keyframes 0..K-1 at their true poses, the last frame is the target and starts at keyframe 0's pose
    ref = O.align(ocfg, frames[:K] if multi else frames[0], poses[:K] if multi else poses[0], frames[K], poses[0],
                  multi=multi, want_trace=True)
    got = ctx.align([dict(to=K, to_pose=poses[0], frm=list(range(K)), from_poses=np.stack(poses[:K]), multi=multi)])[0]
    tr = ctx.get_trace(0, L)
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:L]) == list(ref["iters"]), (got["iters"], ref["iters"])
    assert list(got["stop"][:L]) == list(ref["stop"])
    rejected = 0
    for l in range(L):
        for i in range(ref["iters"][l]):
            a, b = tr[l, i], ref["trace"][l, i]
            assert a[0] == b[0] == 1
            assert rel_fro(a[18:54], b[18:54]) < NE_TOL and rel_fro(a[12:18], b[12:18]) < NE_TOL, (l, i)
            assert abs(a[60] - b[60]) <= 1e-12 * b[60], (l, i)                      # lambda
            assert np.allclose(a[54:60], b[54:60], rtol=1e-6, atol=1e-12), (l, i)     # the damped step
            if i > 0 and b[60] > ref["trace"][l, i - 1][60]:
                rejected += 1
    assert rejected > 0, "the case never exercised a rejected step"
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    assert rel_fro(got["cov"], ref["cov"]) < 1e-3
    if not prior:
        dt, dr = pose_err(O, got["pose"], poses[K])
        assert dt < 0.01 and dr < 0.01


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,L,K,prior,dw,solver,loss", [
    (640, 480, 4, 1, False, 0.0, "GN", "TDIST"),
    (640, 480, 4, 1, False, 1e6, "GN", "TDIST"),
    (320, 240, 3, 2, True, 1e6, "GN", "TDIST"),
    (320, 240, 3, 1, False, 1e6, "LM", "TDIST"),
    (320, 240, 3, 1, False, 50.0, "GN", "NONE"),
    (1232, 368, 5, 1, False, 0.0, "GN", "TDIST"),
])
def test_bilinear_depth_extension_matches_oracle(ctx, capi, oracle, synth, W, H, L, K, prior, dw, solver, loss):
    """bilinear sampling with per-iteration Jacobians and the optional depth residual (ext.cu).  The
    reference's shipped aligner is nearest-pixel inverse compositional; this mode is defined by
    oracle/vslam_oracle.cpp (view_normal_equations_ext), parity is GPU <-> that restatement: every evaluation's A, b,
    chi2, scale and step, the iteration counts, stop reasons, pose and covariance."""
    O, A = oracle, capi
    cam = synth.cam_for(W, H)
    rng = np.random.default_rng(23)
    xis = [np.zeros(6)] + [np.concatenate([rng.uniform(-0.02, 0.02, 3), rng.uniform(-0.01, 0.01, 3)]) for _ in range(K)]
    frames, poses, raw = [], [], []
    for xi in xis:
        R, t = synth.se3_exp(xi)
        g, d = synth.render(W, H, cam, R, t, scene_seed=5005)
        raw.append((g, d))
        frames.append(O.Frame(g, d.astype(np.float64), *cam, L))
        poses.append(synth.pose7(R, t))
    ctx.upload(np.arange(K + 1), np.stack([g for g, _ in raw]), np.stack([d for _, d in raw]), cam, L)
    kw = dict(max_iterations=30, include_prior=prior, depth_weight=dw)
    if loss == "NONE":
        okw, akw = dict(loss=O.LOSS_NONE, scaler=O.SCALER_IDENTITY), dict(loss=A.LOSS_NONE, scaler=A.SCALER_IDENTITY)
    else:
        okw, akw = {}, {}
    ocfg = O.make_config(sampling=O.SAMPLING_BILINEAR, solver=O.SOLVER_LM if solver == "LM" else O.SOLVER_GN, **kw, **okw)
    ctx.set_config(A.make_config(sampling=A.SAMPLING_BILINEAR, solver=A.SOLVER_LM if solver == "LM" else A.SOLVER_GN,
                                 **kw, **akw))
    ctx.set_trace(True)
    multi = K > 1
    ref = O.align(ocfg, frames[:K] if multi else frames[0], poses[:K] if multi else poses[0], frames[K], poses[0],
                  multi=multi, want_trace=True)
    got = ctx.align([dict(to=K, to_pose=poses[0], frm=list(range(K)), from_poses=np.stack(poses[:K]), multi=multi)])[0]
    tr = ctx.get_trace(0, L)
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:L]) == list(ref["iters"]), (got["iters"], ref["iters"])
    assert list(got["stop"][:L]) == list(ref["stop"])
    for l in range(L):
        for i in range(ref["iters"][l]):
            a, b = tr[l, i], ref["trace"][l, i]
            assert a[0] == b[0] == 1 and a[1] == b[1], (l, i)
            assert rel_fro(a[18:54], b[18:54]) < NE_TOL and rel_fro(a[12:18], b[12:18]) < NE_TOL, (l, i)
            assert abs(a[2] - b[2]) <= NE_TOL * abs(b[2]), (l, i)                       # chi2
            assert abs(a[4] - b[4]) <= 1e-6 * abs(b[4]) and a[5] == b[5] == 0, (l, i)  # scale of the photometric residuals
            assert np.allclose(a[54:60], b[54:60], rtol=1e-4, atol=1e-9), (l, i)       # the step
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    assert rel_fro(got["cov"], ref["cov"]) < 1e-3
    if not prior:
        dt, dr = pose_err(O, got["pose"], poses[K])
        assert dt < 0.01 and dr < 0.01
    # the nearest-pixel entry points refuse the mode instead of silently answering for another one
    with pytest.raises(Exception):
        ctx.evaluate(0, K, 0, poses[0])
    ctx.set_config(A.make_config())


@pytest.mark.gpu
def test_bilinear_extension_batch_is_independent_of_batch_order(ctx, capi, synth):
    """The extension's sums are combined in block order by the last block to finish: a job's result does not depend on
    what else is in the batch (bit-identical poses alone, in a batch of 6 and in the reversed batch)."""
    A = capi
    W, H, L = 320, 240, 3
    cam = synth.cam_for(W, H)
    rng = np.random.default_rng(5)
    xis = [np.zeros(6)] + [np.concatenate([rng.uniform(-0.02, 0.02, 3), rng.uniform(-0.01, 0.01, 3)]) for _ in range(6)]
    raw = []
    for xi in xis:
        R, t = synth.se3_exp(xi)
        raw.append(synth.render(W, H, cam, R, t, scene_seed=6006))
    ctx.upload(np.arange(7), np.stack([g for g, _ in raw]), np.stack([d for _, d in raw]), cam, L)
    ctx.set_config(A.make_config(max_iterations=30, sampling=A.SAMPLING_BILINEAR, depth_weight=1e6))
    ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
    jobs = [dict(to=k, to_pose=ident, frm=[0], from_poses=ident[None], multi=False) for k in range(1, 7)]
    alone = [ctx.align([j])[0] for j in jobs]
    batch = ctx.align(jobs)
    rev = ctx.align(jobs[::-1])[::-1]
    for a, b, c in zip(alone, batch, rev):
        assert a["status"] == b["status"] == c["status"] == 0
        assert np.array_equal(a["pose"], b["pose"]) and np.array_equal(a["pose"], c["pose"])
        assert list(a["iters"]) == list(b["iters"]) == list(c["iters"])
    ctx.set_config(A.make_config())
