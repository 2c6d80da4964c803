"""CPU suite, part 3: the solver side of the oracle against independent restatements (published closed forms, the
OpenCV Jacobian the reference's own test compares with, scipy's matrix exponential) and against the ground truth
of the synthetic scenes.  The reference holds no golden vectors for this part (its alignment tests are DISABLED_
and need a dataset that is not shipped): parity of the solver path stays "unpinned", these tests bound it."""
import numpy as np
import pytest


def _small_pair(synth, W=160, H=120, seed=7):
    return synth.make_pair(W, H, seed)


def _frames(oracle, p, levels):
    f0 = oracle.Frame(p["gray0"], p["depth0"].astype(np.float64), *p["cam"], levels)
    f1 = oracle.Frame(p["gray1"], p["depth1"].astype(np.float64), *p["cam"], levels)
    return f0, f1


def test_se3_exp_matches_scipy_expm(oracle):
    """Sophus SE3d::exp with the tangent ordered (translation, rotation); pose layout qx qy qz qw tx ty tz."""
    from scipy.linalg import expm
    rng = np.random.default_rng(11)
    for scale in (1e-9, 1e-3, 0.3, 2.0):
        xi = rng.normal(0, scale, 6)
        wx = np.array([[0, -xi[5], xi[4]], [xi[5], 0, -xi[3]], [-xi[4], xi[3], 0]])
        M = np.zeros((4, 4))
        M[:3, :3] = wx
        M[:3, 3] = xi[:3]
        T = expm(M)
        pose = oracle.se3_exp(xi)
        assert np.allclose(oracle.quat_to_R(pose[:4]), T[:3, :3], atol=1e-12)
        assert np.allclose(pose[4:], T[:3, 3], atol=1e-12)
        if np.linalg.norm(xi[3:]) < 3.0:  # inside the principal branch the logarithm returns the same tangent
            assert np.allclose(oracle.se3_log(pose), xi, atol=1e-9 * max(1.0, scale))
        back = oracle.se3_exp(oracle.se3_log(pose))
        assert np.allclose(oracle.quat_to_R(back[:4]), T[:3, :3], atol=1e-12)  # q and -q are the same rotation
        assert np.allclose(back[4:], T[:3, 3], atol=1e-12)


def test_se3_group_laws(oracle):
    rng = np.random.default_rng(12)
    a, b = oracle.se3_exp(rng.normal(0, 0.4, 6)), oracle.se3_exp(rng.normal(0, 0.4, 6))
    ident = oracle.se3_mul(a, oracle.se3_inv(a))
    assert np.allclose(ident, oracle.IDENTITY, atol=1e-14)
    ab = oracle.se3_mul(a, b)
    Ra, Rb = oracle.quat_to_R(a[:4]), oracle.quat_to_R(b[:4])
    assert np.allclose(oracle.quat_to_R(ab[:4]), Ra @ Rb, atol=1e-14)
    assert np.allclose(ab[4:], Ra @ b[4:] + a[4:], atol=1e-14)


def test_jacobian_matches_the_opencv_form(oracle, synth):
    """lukas_kanade/test/test_warp.cpp:29-41,67-83: the warp Jacobian row times the image gradient equals OpenCV's
    calcRgbdEquationCoeffs (rgbd odometry) for the same point."""
    p = _small_pair(synth)
    f0, f1 = _frames(oracle, p, 1)
    cfg = oracle.make_config(min_gradient=30.0)
    ev = oracle.evaluate(cfg, f0, f1, 0, want_J=True)
    lv = f0.level(0)
    mask = ev["rcode"] != oracle.R_NOT_INTEREST
    vs, us = np.nonzero(mask)  # row-major == interest point order
    assert len(vs) == len(ev["J"]) > 500
    fx, fy = lv["cam"][0], lv["cam"][1]
    for k in np.linspace(0, len(vs) - 1, 200).astype(int):
        v, u = vs[k], us[k]
        dIdx, dIdy = float(lv["dIx"][v, u]), float(lv["dIy"][v, u])
        X, Y, Z = lv["pcl"][v, u]
        invz = 1.0 / Z
        v0, v1 = dIdx * fx * invz, dIdy * fy * invz
        v2 = -(v0 * X + v1 * Y) * invz
        Cc = np.array([v0, v1, v2, -Z * v1 + Y * v2, Z * v0 - X * v2, -Y * v0 + X * v1])
        assert np.allclose(ev["J"][k], Cc, rtol=1e-11, atol=1e-9), (k, ev["J"][k], Cc)


def test_interest_point_selection(oracle, synth):
    """SE3Alignment.cpp:54,87-94: |grad|^2 >= minGradient^2 and a valid depth; raising the threshold shrinks the set."""
    p = _small_pair(synth)
    f0, f1 = _frames(oracle, p, 1)
    lv = f0.level(0)
    g2 = lv["dIx"].astype(np.int64) ** 2 + lv["dIy"].astype(np.int64) ** 2
    prev = None
    for mg in (10.0, 30.0, 80.0):
        ev = oracle.evaluate(oracle.make_config(min_gradient=mg), f0, f1, 0)
        mask = ev["rcode"] != oracle.R_NOT_INTEREST
        want = (g2 >= mg * mg) & (lv["Z"] > 0)
        assert np.array_equal(mask, want)
        if prev is not None:
            assert mask.sum() < prev
        prev = mask.sum()


def _residual_vector(oracle, ev):
    """The residual vector the reference hands to the loss: one entry per interest point, in row-major order, 0 for
    points that leave the image (InverseCompositional.cpp:100-121 leaves r = 0, w = 0 there); visible mask."""
    rc = ev["rcode"]
    order = rc[rc != oracle.R_NOT_INTEREST]
    vis = order != oracle.R_INVISIBLE
    return np.where(vis, order, 0).astype(np.float64), vis


def _tdist_scale_numpy(r, v):
    """ScalerTDistribution::compute (least_squares/src/Scaler.cpp:88-107) restated independently with numpy."""
    sigma, step, it = 1.0, np.inf, 0
    while it < 30 and step > 1e-5:
        s = np.sqrt(np.sum(r * r * (v + 1.0) / (v + (r / sigma) ** 2)) / len(r))
        step, sigma, it = abs(sigma - s), s, it + 1
    return sigma


def _normal_equations_numpy(J, r, w, vis):
    """NormalEquations(J, r, w) / n (NormalEquations.cpp:53-76, InverseCompositional.cpp:139-146); n counts every
    interest point, weights of invisible points are 0."""
    w = np.where(vis, w, 0.0)
    n = len(r)
    return (J * w[:, None]).T @ J / n, (J * w[:, None]).T @ r / n, np.sum(w * r * r) / n


def _check_chi2(oracle, cfg, f0, f1, r, w, vis, weighted, tol):
    """The configuration's default is CHI2_REFERENCE_RELEASE: w[0] * sum r^2, what NormalEquations.cpp:58
    `(r * w).transpose() * r` gives in a Release build (w[0] = 0 when the first interest point is not visible);
    CHI2_WEIGHTED is the documented r^T W r."""
    cfg.chi2_mode = oracle.CHI2_WEIGHTED
    assert abs(oracle.evaluate(cfg, f0, f1, 0)["chi2"] - weighted) <= tol * weighted
    cfg.chi2_mode = oracle.CHI2_REFERENCE_RELEASE
    expect = np.where(vis, w, 0.0)[0] * np.sum(r * r) / len(r)
    assert abs(oracle.evaluate(cfg, f0, f1, 0)["chi2"] - expect) <= tol * max(expect, 1e-300)


def test_tdistribution_scale_and_weights(oracle, synth):
    p = _small_pair(synth)
    f0, f1 = _frames(oracle, p, 1)
    v = 5.0
    cfg = oracle.make_config(v=v)
    ev = oracle.evaluate(cfg, f0, f1, 0, want_J=True)
    r, vis = _residual_vector(oracle, ev)
    assert ev["n"] == len(r) > 1000 and 0 < vis.sum() < len(r)
    sigma = _tdist_scale_numpy(r, v)
    assert ev["offset"] == 0.0
    assert abs(ev["scale"] - sigma) < 1e-9 * sigma
    # LossTDistribution::computeWeight (Loss.cpp:92): w = (v + 1) / (v + (r / sigma)^2)
    A, b, chi2 = _normal_equations_numpy(ev["J"], r, (v + 1.0) / (v + (r / sigma) ** 2), vis)
    assert np.allclose(ev["A"], A, rtol=1e-10, atol=1e-10 * np.abs(A).max())
    assert np.allclose(ev["b"], b, rtol=1e-10, atol=1e-10 * np.abs(b).max())
    _check_chi2(oracle, cfg, f0, f1, r, (v + 1.0) / (v + (r / sigma) ** 2), vis, chi2, 1e-10)


@pytest.mark.parametrize("loss,scaler", [("QUADRATIC", "IDENTITY"), ("HUBER", "MEDIAN"), ("TUKEY", "MEDIAN"),
                                         ("QUADRATIC", "MEAN"), ("HUBER", "MEAN")])
def test_other_losses_and_scalers(oracle, synth, loss, scaler):
    """Loss.cpp:51-72 weights of the residual standardised by Scaler.cpp:25-60 (MAD-like spread: sqrt of the mean
    absolute deviation over n - 1, the reference's own definition)."""
    p = _small_pair(synth)
    f0, f1 = _frames(oracle, p, 1)
    cfg = oracle.make_config(loss=getattr(oracle, "LOSS_" + loss), scaler=getattr(oracle, "SCALER_" + scaler))
    ev = oracle.evaluate(cfg, f0, f1, 0, want_J=True)
    r, vis = _residual_vector(oracle, ev)
    if scaler == "IDENTITY":
        off, sc = 0.0, 1.0
    else:
        if scaler == "MEAN":
            off = r.mean()
        else:
            srt, m = np.sort(r), len(r)
            off = srt[m // 2] if m % 2 else (srt[m // 2 - 1] + srt[m // 2 + 1]) / 2  # algorithm.cpp:109-120
        sc = np.sqrt(np.sum(np.abs(r - off)) / (len(r) - 1))
    assert abs(ev["offset"] - off) <= 1e-9 * max(1.0, abs(off)) and abs(ev["scale"] - sc) <= 1e-9 * sc
    x = (r - off) / sc
    if loss == "QUADRATIC":
        w = np.ones_like(x)
    elif loss == "HUBER":
        c = cfg.huber_c
        with np.errstate(divide="ignore"):
            w = np.where(np.abs(x) < c, 1.0, np.where(c * x > 0, 1.0, -1.0) / x)  # Loss.cpp:65-72: 1 / |x| outside
    else:
        c = 4.6851  # TukeyLoss::C
        w = np.where(np.abs(x) < c, (1 - (x / c) ** 2) ** 2, 0.0)
    A, b, chi2 = _normal_equations_numpy(ev["J"], r, w, vis)
    assert np.allclose(ev["A"], A, rtol=1e-9, atol=1e-10 * np.abs(A).max())
    assert np.allclose(ev["b"], b, rtol=1e-9, atol=1e-10 * np.abs(b).max())
    _check_chi2(oracle, cfg, f0, f1, r, w, vis, chi2, 1e-9)


def test_align_recovers_the_synthetic_motion(oracle, synth):
    """End to end on a rendered pair: the estimated relative pose is the one the scene was rendered with."""
    p = synth.make_pair(320, 240, 1234)
    f0, f1 = _frames(oracle, p, 3)
    res = oracle.align(oracle.make_config(max_iterations=50, chi2_mode=oracle.CHI2_WEIGHTED), f0, oracle.IDENTITY, f1,
                       oracle.IDENTITY)
    d = oracle.se3_log(oracle.se3_mul(res["pose"], oracle.se3_inv(p["pose1"])))
    assert np.linalg.norm(d[:3]) < 5e-3 and np.linalg.norm(d[3:]) < 5e-3, d
    assert res["status"] == 0
    # The reference's Release-build chi2 (w[0] * sum r^2, vslam_b200.h) follows the unweighted sum of squares scaled by
    # the first interest point's weight: it stops the levels earlier, a little further from the rendered motion.
    ref = oracle.align(oracle.make_config(max_iterations=50), f0, oracle.IDENTITY, f1, oracle.IDENTITY)
    assert ref["status"] == 0 and sum(ref["iters"]) < sum(res["iters"])
    dr = oracle.se3_log(oracle.se3_mul(ref["pose"], oracle.se3_inv(p["pose1"])))
    assert np.linalg.norm(dr[:3]) < 1.5e-2 and np.linalg.norm(dr[3:]) < 1e-2, dr
    assert all(1 <= i <= 50 for i in res["iters"])
    # covariance is the inverse of a symmetric positive definite matrix
    assert np.allclose(res["cov"], res["cov"].T, rtol=1e-8, atol=1e-14)
    assert np.all(np.linalg.eigvalsh(0.5 * (res["cov"] + res["cov"].T)) > 0)


def test_gauss_newton_stop_rules(oracle, synth):
    """GaussNewton.cpp:53-98: iteration cap; fewer than 6 constraints ends the level before any step."""
    p = synth.make_pair(160, 120, 5)
    f0, f1 = _frames(oracle, p, 2)
    res = oracle.align(oracle.make_config(max_iterations=2), f0, oracle.IDENTITY, f1, oracle.IDENTITY)
    assert all(i <= 2 for i in res["iters"])
    flat = np.full((120, 160), 90, np.uint8)  # no gradient anywhere: no interest points at all
    g0 = oracle.Frame(flat, p["depth0"].astype(np.float64), *p["cam"], 2)
    g1 = oracle.Frame(flat, p["depth1"].astype(np.float64), *p["cam"], 2)
    res = oracle.align(oracle.make_config(), g0, oracle.IDENTITY, g1, oracle.IDENTITY)
    assert res["status"] == 1 and list(res["iters"]) == [0, 0]
    assert np.array_equal(res["pose"], oracle.IDENTITY)
    assert np.array_equal(res["cov"], np.eye(6))


def test_trace_is_consistent_with_evaluate(oracle, synth):
    """The oracle's own per-iteration trace replays through evaluate(): the same teacher-forced check the GPU parity
    tests apply to the device trace (normal equations at every iterate, the pose update x <- x * exp(-dx))."""
    from test_gpu_parity import _teacher_forced_check
    p = _small_pair(synth)
    L = 3
    f0, f1 = _frames(oracle, p, L)
    for prior in (False, True):
        cfg = oracle.make_config(max_iterations=8, include_prior=prior)
        res = oracle.align(cfg, f0, oracle.IDENTITY, f1, oracle.IDENTITY, want_trace=True)
        assert sum(res["iters"]) >= 4
        _teacher_forced_check(oracle, cfg, [f0], [oracle.IDENTITY], f1, oracle.IDENTITY, False, res["trace"],
                              res["iters"], prior, L)


def test_levenberg_marquardt_extension(oracle, synth):
    """The LM extension's own contract (oracle/vslam_oracle.cpp, levenberg_marquardt): accepted chi2 never increases
    within a level, lambda grows by `up` exactly when a step is rejected and shrinks by `down` otherwise, the damped
    step solves (A + lambda diag A) dx = b, and the alignment still recovers the rendered motion."""
    O = oracle
    p = synth.make_pair(320, 240, 1234)
    L = 3
    f0 = O.Frame(p["gray0"], p["depth0"].astype(np.float64), *p["cam"], L)
    f1 = O.Frame(p["gray1"], p["depth1"].astype(np.float64), *p["cam"], L)
    cfg = O.make_config(solver=O.SOLVER_LM, max_iterations=40, lm_lambda0=1e-2, lm_up=8.0, lm_down=0.25,
                        chi2_mode=O.CHI2_WEIGHTED)  # accept / reject needs the weighted sum the steps minimise
    r = O.align(cfg, f0, O.IDENTITY, f1, O.IDENTITY, want_trace=True)
    assert r["status"] == 0
    ups = downs = 0
    for l in range(L):
        n = r["iters"][l]
        assert n > 0
        chi2, lam = r["trace"][l, :n, 2], r["trace"][l, :n, 60]
        assert np.all(np.diff(chi2) <= 0)                       # chi2 of the accepted state
        assert lam[0] == 1e-2
        for i in range(1, n):
            if chi2[i] == chi2[i - 1] and lam[i] > lam[i - 1]:
                assert np.isclose(lam[i], lam[i - 1] * 8.0)
                ups += 1
            else:
                assert np.isclose(lam[i], max(lam[i - 1] * 0.25, 1e-12))
                downs += 1
        for i in range(n):
            t = r["trace"][l, i]
            A, b, dx = t[18:54].reshape(6, 6), t[12:18], t[54:60]
            assert np.allclose((A + t[60] * np.diag(np.diag(A))) @ dx, b, rtol=1e-9, atol=1e-12)
    assert ups > 0 and downs > 0
    d = O.se3_log(O.se3_mul(r["pose"], O.se3_inv(p["pose1"])))
    assert np.linalg.norm(d[:3]) < 0.01 and np.linalg.norm(d[3:]) < 0.01
    gn = O.align(O.make_config(max_iterations=40, chi2_mode=O.CHI2_WEIGHTED), f0, O.IDENTITY, f1, O.IDENTITY)
    d = O.se3_log(O.se3_mul(r["pose"], O.se3_inv(gn["pose"])))
    assert np.linalg.norm(d) < 5e-3                             # both solvers end in the same basin


def test_bilinear_depth_extension(oracle, synth):
    """The extended mode's own contract (oracle/vslam_oracle.cpp, view_normal_equations_ext): the first evaluation's b
    is the derivative of chi2 with respect to a right perturbation of the warp (checked by central differences on the
    quadratic loss, where chi2 is smooth), the mode reaches the rendered motion in fewer iterations than nearest-pixel
    sampling, and the depth residual keeps the solution in the same place."""
    O = oracle
    p = synth.make_pair(320, 240, 1234)
    L = 3
    f0 = O.Frame(p["gray0"], p["depth0"].astype(np.float64), *p["cam"], L)
    f1 = O.Frame(p["gray1"], p["depth1"].astype(np.float64), *p["cam"], L)
    near = O.align(O.make_config(max_iterations=40, chi2_mode=O.CHI2_WEIGHTED), f0, O.IDENTITY, f1, O.IDENTITY)
    res = {}
    for dw in (0.0, 1e6):
        cfg = O.make_config(max_iterations=40, sampling=O.SAMPLING_BILINEAR, depth_weight=dw)
        r = O.align(cfg, f0, O.IDENTITY, f1, O.IDENTITY, want_trace=True)
        assert r["status"] == 0
        d = O.se3_log(O.se3_mul(r["pose"], O.se3_inv(p["pose1"])))
        assert np.linalg.norm(d[:3]) < 0.005 and np.linalg.norm(d[3:]) < 0.005
        res[dw] = r
    assert sum(res[0.0]["iters"]) < sum(near["iters"])
    assert not np.allclose(res[0.0]["trace"][L - 1, 0, 18:54], res[1e6]["trace"][L - 1, 0, 18:54], rtol=1e-3)
    # gradient check: one evaluation on a one-level pyramid with a quadratic loss (weights 1, scale 1): b is half the
    # derivative of chi2 along a right perturbation of the warp.  The images are smoothed first: the gradient planes are
    # Sobel derivatives of a blurred image, on sharp edges they are not the slope of the bilinear interpolant.
    from scipy.ndimage import gaussian_filter
    q = _small_pair(synth)
    sm = lambda g: np.clip(np.round(gaussian_filter(g.astype(float), 2.5)), 0, 255).astype(np.uint8)
    g0 = O.Frame(sm(q["gray0"]), q["depth0"].astype(np.float64), *q["cam"], 1)
    g1 = O.Frame(sm(q["gray1"]), q["depth1"].astype(np.float64), *q["cam"], 1)
    start = O.se3_exp(np.array([0.004, -0.003, 0.002, 0.002, -0.001, 0.0015]))
    grads = {}
    for dw in (0.0, 5000.0):
        c1 = O.make_config(min_gradient=5.0, max_iterations=1, loss=O.LOSS_QUADRATIC, scaler=O.SCALER_IDENTITY,
                           sampling=O.SAMPLING_BILINEAR, depth_weight=dw)

        def chi2_at(xi):
            t = O.align(c1, g0, O.IDENTITY, g1, O.se3_mul(start, O.se3_exp(xi)), want_trace=True)["trace"][0, 0]
            return t[2], t[12:18].copy()
        _, b = chi2_at(np.zeros(6))
        num = np.zeros(6)
        for k in range(6):
            e = np.zeros(6)
            e[k] = 2e-4
            num[k] = (chi2_at(e)[0] - chi2_at(-e)[0]) / 4e-4
        grads[dw] = (num, 2 * b)
        assert np.allclose(num, 2 * b, rtol=0.08, atol=0.01 * np.abs(b).max()), (dw, num, 2 * b)
    # the depth term alone (difference of the two runs): its Jacobian is the exact derivative of the bilinear depth sample
    dn, da = grads[5000.0][0] - grads[0.0][0], grads[5000.0][1] - grads[0.0][1]
    assert np.abs(da).max() > 10 and np.allclose(dn, da, rtol=0.05, atol=0.02 * np.abs(da).max()), (dn, da)
