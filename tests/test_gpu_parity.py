"""GPU parity tests: CUDA path (through the C ABI) against the CPU oracle on identical synthetic inputs.
Tolerances are BASELINE.json's: masks / indices bit-exact, pyramid and gradient images <= 1e-5 relative
(integer planes are required to be exact), per-iteration normal equations <= 1e-4 relative Frobenius,
final pose <= 1e-4 m and 1e-4 rad."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

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


@pytest.mark.parametrize("loss,scaler", [("TDIST", "TDIST"), ("NONE", "IDENTITY"), ("HUBER", "MEAN"), ("TUKEY", "MEDIAN"),
                                         ("HUBER", "MEDIAN"), ("QUADRATIC", "IDENTITY")])
@pytest.mark.parametrize("holes", [0, 50])
def test_single_iteration_masks_and_normal_equations(ctx, capi, oracle, synth, loss, scaler, holes):
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


@pytest.mark.parametrize("loss,scaler,prior", [("TDIST", "TDIST", False), ("TDIST", "TDIST", True), ("NONE", "IDENTITY", False),
                                               ("HUBER", "MEAN", False)])
def test_align_matches_oracle_per_iteration(ctx, capi, oracle, synth, loss, scaler, prior):
    O, A = oracle, capi
    p = _pair(synth, 640, 480, 1234)
    f0, f1 = _frames(O, p, 4)
    _upload(ctx, p, 4)
    kw = dict(huber_c=10.0, max_iterations=50, include_prior=prior)
    ocfg = O.make_config(loss=getattr(O, "LOSS_" + loss), scaler=getattr(O, "SCALER_" + scaler), **kw)
    ctx.set_config(A.make_config(loss=getattr(A, "LOSS_" + loss), scaler=getattr(A, "SCALER_" + scaler), **kw))
    ctx.set_trace(True)
    ref = O.align(ocfg, f0, O.IDENTITY, f1, O.IDENTITY, want_trace=True)
    got = ctx.align([dict(to=2, to_pose=O.IDENTITY, frm=[1], from_poses=O.IDENTITY, multi=False)])[0]
    tr = ctx.get_trace(0, 4)
    assert got["status"] == ref["status"] == 0
    assert list(got["iters"][:4]) == list(ref["iters"]), (got["iters"], ref["iters"])
    assert list(got["stop"][:4]) == list(ref["stop"])
    for l in range(4):
        for i in range(ref["iters"][l]):
            a, b = tr[l, i], ref["trace"][l, i]
            assert a[0] == 1 and a[1] == b[1]
            assert rel_fro(a[18:54], b[18:54]) < NE_TOL, (l, i)
            assert rel_fro(a[12:18], b[12:18]) < NE_TOL, (l, i)
            assert abs(a[2] - b[2]) <= NE_TOL * abs(b[2])
    dt, dr = pose_err(O, got["pose"], ref["pose"])
    assert dt < POSE_TOL and dr < POSE_TOL
    assert rel_fro(got["cov"], ref["cov"]) < 1e-3
