"""CPU suite, part 5: the oracle (oracle/vslam_oracle.cpp) pinned against the REFERENCE'S OWN CODE.

oracle/_ref/libvslam_ref.so is the reference's alignment path - SE3Alignment.cpp, InverseCompositional(Stacked).cpp,
Warp.cpp, GaussNewton.cpp, NormalEquations.cpp, Loss.cpp, Scaler.cpp, Frame.cpp, Camera.cpp, PoseWithCovariance.cpp -
compiled UNCHANGED from /root/reference by oracle/build_ref.py against the stand-in headers of oracle/ref_shim (the Eigen /
Sophus / OpenCV / logging surface those files use; the real libraries are not in this image).  What is pinned: the
reference's control flow and arithmetic.  What the stand-ins decide: the summation order inside matrix products, the LU /
LDL^T pivoting details and Sophus' closed forms (restated from their published definitions), and the three OpenCV image
operators, which are the oracle's own cv2-pinned functions.

Every alignment starts from a small non-zero motion guess, as a motion model gives: at the exact identity a point on the
"1 < u'" bound (InverseCompositional.cpp:98) re-projects to 1 +- 1 ulp and its visibility is decided by rounding order.
"""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
START = np.array([0.004, -0.003, 0.002, 0.001, -0.002, 0.0015])


@pytest.fixture(scope="module")
def ref(oracle):
    from oracle import ref as R
    if not R.available():
        pytest.skip("no /root/reference and no prebuilt oracle/_ref/libvslam_ref.so")
    return R


def _tum():
    d = np.load(os.path.join(HERE, "golden", "tum_frame.npz"))
    cam = np.array([525.0, 525.0, 319.5 - 160.0, 239.5 - 120.0])  # a 320 x 240 crop around the principal point
    cam[2], cam[3] = 159.5, 119.5
    gray = np.stack([d["gray0"], d["gray1"]])
    depth = np.stack([d["depth0"], d["depth1"]]).astype(np.float64) / 5000.0
    return gray, depth, cam


def _run_both(O, R, gray, depth, cam, L, K=1, multi=False, start=START, **kw):
    """Keyframes gray[:K] at the identity, the last frame starts at `start`; the same job through the oracle and the
    reference build."""
    cfg = O.make_config(f32_storage=False, **kw)
    assert cfg.chi2_mode == O.CHI2_REFERENCE_RELEASE
    tw = np.zeros((K + 1, 6))
    tw[K] = start
    fr = [O.Frame(gray[i], depth[i], *cam, L, f32_storage=False) for i in range(K + 1)]
    poses = [O.se3_exp(t) for t in tw]
    if K == 1 and not multi:
        o = O.align(cfg, fr[0], poses[0], fr[1], poses[1], want_trace=True)
    else:
        o = O.align(cfg, fr[:K], poses[:K], fr[K], poses[K], multi=True, want_trace=True)
    r = R.align(R.config_from(cfg), gray, depth, cam, tw, L, multi=multi or K > 1)
    return o, r


def _assert_same_run(O, o, r, L):
    assert o["iters"].tolist() == r["pushed"].tolist()                      # iterations pushed per level
    assert [i - 1 for i in o["iters"]] == r["last_iteration"].tolist()      # Solver::Results::iteration
    for l in range(L):
        for i in range(o["iters"][l]):
            a, b = o["trace"][l, i], r["trace"][l, i]
            assert a[1] == b[1]                                                              # nConstraints
            assert np.allclose(a[18:54], b[18:54], rtol=0, atol=1e-12 * np.abs(b[18:54]).max())  # A
            assert np.allclose(a[12:18], b[12:18], rtol=0, atol=1e-12 * np.abs(b[12:18]).max())  # b
            assert abs(a[2] - b[2]) <= 1e-12 * abs(b[2])                                     # chi2
            assert np.allclose(a[6:12], b[6:12], rtol=0, atol=1e-12)                         # x after the step
            assert abs(a[3] - b[3]) <= 1e-12                                                 # step size
    assert np.allclose(O.quat_to_R(o["pose"][:4]), r["R"], rtol=0, atol=1e-12)
    assert np.allclose(o["pose"][4:], r["t"], rtol=0, atol=1e-12)
    assert np.allclose(O.se3_log(o["pose"]), r["twist"], rtol=0, atol=1e-12)
    assert np.allclose(o["cov"], r["cov"], rtol=0, atol=1e-9 * np.abs(r["cov"]).max())


def _punch(depth):
    """Rectangular holes (sensor shadow / range dropout) whose 3 x 3 windows never hold exactly two valid depths at any
    pyramid level: that case is undefined in the reference (see test_frame_pyramid_matches_the_reference_frame)."""
    d = depth.copy()
    H, W = d.shape[-2:]
    d[..., H // 4: H // 4 + 16, W // 8: W // 8 + 24] = 0.0
    d[..., H // 2: H // 2 + 24, W // 2 + 8: W // 2 + 24] = 0.0
    d[..., : 8, :] = 0.0
    return d


def _two_valid_sites(Z):
    """Centres (of the decimated grid) whose 3 x 3 window in Z holds exactly two valid depths."""
    H, W = Z.shape
    valid = (Z > 0).astype(np.int32)
    cnt = sum(np.roll(np.roll(valid, dy, 0), dx, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1))
    return (cnt == 2)[: (H // 2) * 2: 2, : (W // 2) * 2: 2]


def test_frame_pyramid_matches_the_reference_frame(oracle, ref, synth):
    """Frame.cpp:131-259 compiled unchanged: intensity pyramid, Sobel planes, median-filtered / resized depth, camera
    per level and point cloud against oracle.Frame, on a rendered frame (with holes) and on the TUM crop (u16 / 5000).

    One site class is left out: a 3 x 3 window with exactly TWO valid depths.  algorithm::median (algorithm.cpp:109-120)
    then reads v[n / 2 + 1] = v[2] past the end of its vector - undefined; this build returns stale heap contents - so
    there is nothing to pin.  The oracle (and the device) return the smaller of the two values and count those sites."""
    p = synth.make_pair(160, 120, 7)
    L = 3
    gray, depth, cam = p["gray0"], _punch(p["depth0"].astype(np.float64)), np.array(p["cam"], float)
    f = oracle.Frame(gray, depth, *cam, L, f32_storage=False)
    assert f.n2_hits() == 0
    for lvl in range(L):
        a, b = f.level(lvl), ref.frame_level(gray, depth, cam, L, lvl)
        assert np.array_equal(a["I"], b["I"])
        assert np.array_equal(a["dIx"].astype(np.float64), b["dIx"]) and np.array_equal(a["dIy"].astype(np.float64), b["dIy"])
        assert np.array_equal(a["Z"], b["Z"]) and (lvl == 0 or np.any(a["Z"] == 0))
        assert np.array_equal(a["cam"], b["cam"])
        assert np.allclose(a["pcl"], b["pcl"], rtol=0, atol=2e-15 * max(1.0, np.abs(b["pcl"]).max()))
    g, d, cam = _tum()
    f = oracle.Frame(g[0], d[0], *cam, 2, f32_storage=False)
    a, b = f.level(1), ref.frame_level(g[0], d[0], cam, 2, 1)
    assert np.array_equal(a["I"], b["I"]) and np.array_equal(a["dIx"].astype(np.float64), b["dIx"])
    undefined = _two_valid_sites(d[0])
    assert 0 < undefined.sum() < 200 and f.n2_hits() == undefined[1:, 1:].sum()   # row / column 0 are the blur's border
    assert np.array_equal(a["Z"][~undefined], b["Z"][~undefined])


def test_odd_sizes_split_intensity_and_depth_levels(oracle, ref):
    """cv::buildPyramid rounds (W + 1) / 2 up, algorithm::resize truncates W * 0.5."""
    rng = np.random.default_rng(3)
    gray = rng.integers(0, 255, (61, 81), dtype=np.uint8)
    depth = rng.uniform(0.5, 3.0, (61, 81))
    cam = np.array([80.0, 80.0, 40.0, 30.0])
    assert ref.frame_dims(gray, depth, cam, 2, 1) == (41, 31, 40, 30)
    f = oracle.Frame(gray, depth, *cam, 2, f32_storage=False)
    assert f.dims(1) == (41, 31, 40, 30)


@pytest.mark.parametrize("name,kw", [
    ("none", dict(loss="NONE", scaler="IDENTITY")), ("quadratic", dict(loss="QUADRATIC", scaler="IDENTITY")),
    ("tukey-median", dict(loss="TUKEY", scaler="MEDIAN")), ("huber-median", dict(loss="HUBER", scaler="MEDIAN")),
    ("huber-mean", dict(loss="HUBER", scaler="MEAN")), ("tdist-tdist", dict(loss="TDIST", scaler="TDIST")),
    ("tdist-prior", dict(loss="TDIST", scaler="TDIST", include_prior=True)),
])
def test_single_keyframe_alignment_follows_the_reference(oracle, ref, synth, name, kw):
    """SE3Alignment::align(from, to) (SE3Alignment.cpp:63-105) with every loss / scaler family: the same number of
    Gauss-Newton steps per level, the same normal equations, chi2, iterates, pose and covariance."""
    O = oracle
    kw = dict(kw, loss=getattr(O, "LOSS_" + kw["loss"]), scaler=getattr(O, "SCALER_" + kw["scaler"]))
    for W, H, seed, L in ((160, 120, 7, 3), (320, 240, 1234, 3)):
        p = synth.make_pair(W, H, seed)
        gray = np.stack([p["gray0"], p["gray1"]])
        depth = np.stack([p["depth0"], p["depth1"]]).astype(np.float64)
        if W == 320:
            depth = _punch(depth)
        o, r = _run_both(O, ref, gray, depth, np.array(p["cam"], float), L, max_iterations=30, **kw)
        _assert_same_run(O, o, r, L)


def test_first_point_weight_drives_the_chi2_stops(oracle, ref, synth):
    """NormalEquations.cpp:58 in a Release build: chi2 = w[0] * sum r^2.  With the first interest point outside the
    image chi2 is 0 and every level ends 'converged' at its second step; with it inside, chi2 follows the unweighted sum
    of squares and levels run on until it rises - both as the reference build does."""
    O = oracle
    p = synth.make_pair(160, 120, 7)
    gray = np.stack([p["gray0"], p["gray1"]])
    depth = np.stack([p["depth0"], p["depth1"]]).astype(np.float64)
    o, r = _run_both(O, ref, gray, depth, np.array(p["cam"], float), 3)
    assert r["pushed"].tolist() == [2, 2, 2] and np.all(r["trace"][:, :2, 2] == 0.0)
    assert o["stop"].tolist() == [O.STOP_NAMES.index("converged")] * 3
    _assert_same_run(O, o, r, 3)
    # a frame whose border carries no gradient: the first interest point lies well inside and stays visible
    g = gray.copy()
    g[:, :12] = g[:, -12:] = 90
    g[:, :, :12] = g[:, :, -12:] = 90
    o, r = _run_both(O, ref, g, depth, np.array(p["cam"], float), 2, max_iterations=40)
    assert np.all(r["trace"][1, : r["pushed"][1], 2] > 0.0) and max(r["pushed"]) > 2
    _assert_same_run(O, o, r, 2)


@pytest.mark.parametrize("K,prior", [(1, False), (2, False), (3, True)])
def test_stacked_keyframes_follow_the_reference(oracle, ref, synth, K, prior):
    """The vector overload (SE3Alignment.cpp:107-150, InverseCompositionalStacked.cpp:25-48): K keyframes rendered from
    different poses, stacked normal equations."""
    O = oracle
    W, H, L = 160, 120, 3
    cam = synth.cam_for(W, H)
    rng = np.random.default_rng(21)
    xis = [np.concatenate([rng.uniform(-0.02, 0.02, 3), rng.uniform(-0.01, 0.01, 3)]) for _ in range(K + 1)]
    gray, depth, tw = [], [], []
    for xi in xis:
        Rm, t = synth.se3_exp(xi)
        g, d = synth.render(W, H, cam, Rm, t, scene_seed=4004)
        gray.append(g)
        depth.append(d.astype(np.float64))
        tw.append(O.se3_log(synth.pose7(Rm, t)))
    tw = np.array(tw)
    tw[K] = tw[0] + START * 0.5                                             # the target starts near keyframe 0
    cfg = O.make_config(f32_storage=False, max_iterations=30, include_prior=prior)
    fr = [O.Frame(gray[i], depth[i], *cam, L, f32_storage=False) for i in range(K + 1)]
    poses = [O.se3_exp(t) for t in tw]
    o = O.align(cfg, fr[:K], poses[:K], fr[K], poses[K], multi=True, want_trace=True)
    r = ref.align(ref.config_from(cfg), np.stack(gray), np.stack(depth), np.array(cam, float), tw, L, multi=True)
    _assert_same_run(O, o, r, L)


def test_f32_depth_storage_against_the_reference(oracle, ref, synth):
    """The device keeps depth pyramids in f32 (the oracle's f32_storage=True models it); the reference keeps f64
    (types.h DepthMap).  Measured against the reference build on sensor-style depth (u16 / 5000, holes): every level
    differs by at most one f32 ulp (relative 1.2e-7), no pixel changes validity, the interest-point counts of every
    evaluation are the same, the step counts are the same and the aligned pose moves by less than 1e-6."""
    p = synth.make_pair(320, 240, 1234)
    L = 3
    cam = np.array(p["cam"], float)
    gray = np.stack([p["gray0"], p["gray1"]])
    depth = _punch(np.round(np.stack([p["depth0"], p["depth1"]]).astype(np.float64) * 5000.0) / 5000.0)
    f32 = oracle.Frame(gray[0], depth[0], *cam, L, f32_storage=True)
    for lvl in range(L):
        a, b = f32.level(lvl), ref.frame_level(gray[0], depth[0], cam, L, lvl)
        assert np.array_equal(a["Z"] > 0, b["Z"] > 0)
        ok = b["Z"] > 0
        assert np.all(np.abs(a["Z"][ok] - b["Z"][ok]) <= 1.2e-7 * b["Z"][ok])
        assert lvl == 0 or np.any(a["Z"][ok] != b["Z"][ok])
    cfg = oracle.make_config(f32_storage=True, max_iterations=30)
    tw = np.zeros((2, 6))
    tw[1] = START
    fr = [oracle.Frame(gray[i], depth[i], *cam, L, f32_storage=True) for i in range(2)]
    o = oracle.align(cfg, fr[0], oracle.se3_exp(tw[0]), fr[1], oracle.se3_exp(tw[1]), want_trace=True)
    r = ref.align(ref.config_from(cfg), gray, depth, cam, tw, L)
    assert o["iters"].tolist() == r["pushed"].tolist()
    for l in range(L):
        assert np.array_equal(o["trace"][l, : o["iters"][l], 1], r["trace"][l, : r["pushed"][l], 1])   # same point counts
    assert np.abs(oracle.se3_log(o["pose"]) - r["twist"]).max() < 1e-6


def test_bench_workload_pairs_follow_the_reference(oracle, ref, synth):
    """Two pairs of bench.py's default workload (BASELINE configs[2]: 640x480, 4 levels, 16-bit depth / 5000, the bench's
    start guess, t-distribution weights, chi2 of the Release build) through the oracle's batch helper - what bench.py times
    on the CPU and checks the GPU against - and through the reference build."""
    import bench
    cam = synth.cam_for(bench.W, bench.H)
    gray, depth, _ = bench.render_pairs(synth, range(2), cam)
    depth = bench.quantise_depth(depth)
    cfg = oracle.make_config()
    poses, iters = oracle.baseline_batch(cfg, gray, depth, *cam, bench.LEVELS, 2, stride=2, n_pairs=2, start=bench.start_pose7())
    for k in range(2):
        r = ref.align(ref.config_from(cfg), gray[2 * k: 2 * k + 2], depth[2 * k: 2 * k + 2].astype(np.float64), np.array(cam, float),
                      np.stack([np.zeros(6), bench.START_TWIST]), bench.LEVELS, want_trace=False)
        assert r["pushed"].tolist() == iters[k].tolist()
        assert np.allclose(oracle.se3_log(poses[k]), r["twist"], rtol=0, atol=1e-12)


def test_kitti_shape_five_levels_follows_the_reference(oracle, ref, synth):
    """BASELINE configs[3]'s shape: 1232 x 368, five pyramid levels (77 x 23 at the coarsest)."""
    p = synth.make_pair(1232, 368, 77)
    gray = np.stack([p["gray0"], p["gray1"]])
    depth = _punch(np.stack([p["depth0"], p["depth1"]]).astype(np.float64))
    o, r = _run_both(oracle, ref, gray, depth, np.array(p["cam"], float), 5, max_iterations=30)
    _assert_same_run(oracle, o, r, 5)


@pytest.mark.parametrize("model,period", [("ConstantMotion", 2), ("NoMotion", 3)])
def test_odometry_loop_follows_the_reference(oracle, ref, synth, model, period):
    """SURVEY.md 8(f)-2, the aligner's caller: the reference's own OdometryRgbd (Odometry.cpp:21-62: prior always on, stacked
    alignment against {last key frame, last frame} when they differ), MotionPrediction (MotionPrediction.cpp:22-55),
    KeyFrameSelectionIdx and Map, driven like the node's processFrame, against oracle/odometry.py - the restatement the
    GPU host mirror (include/vslam_b200_odometry.hpp) is tested against.  The first alignment of the loop starts at the exact
    identity, where the visibility of border pixels is a matter of rounding order (see the header of this file), so the
    frames carry a blank border that keeps every interest point away from it."""
    from oracle import odometry as OD
    F, W, H, L = 7, 320, 240, 3
    cam = synth.cam_for(W, H)
    gray, depth, _ = synth.make_sequence(F, W, H, seed=4242, cam=cam)
    gray = gray.copy()
    gray[:, :40] = gray[:, -40:] = 90
    gray[:, :, :40] = gray[:, :, -40:] = 90
    depth = depth.astype(np.float64)
    stamps = (1305031100.0e9 + np.arange(F) * 1e9 / 30.0).astype(np.uint64)
    cfg = oracle.make_config(f32_storage=False, max_iterations=30, include_prior=True)
    frames = [oracle.Frame(gray[i], depth[i], *cam, L, f32_storage=False) for i in range(F)]
    poses, speeds, kinds = OD.run(cfg, frames, stamps, prediction=model, kf_period=period)
    assert "stacked" in kinds and "single" in kinds
    rposes, rspeeds = ref.odometry(ref.config_from(cfg), gray, depth, np.array(cam, float), stamps, L, model, period)
    for i in range(F):
        assert np.allclose(oracle.se3_log(poses[i]), rposes[i], rtol=0, atol=1e-11), i
        assert np.allclose(speeds[i], rspeeds[i], rtol=0, atol=1e-9), i
    assert np.abs(rposes[-1]).max() > 1e-3                                 # the sequence moves
