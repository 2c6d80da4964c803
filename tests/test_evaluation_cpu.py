"""Trajectory evaluation: vslam_b200/evaluation.py against the outputs of the reference's own
scripts (script/vslam_evaluation/tum/evaluate_rpe.py, evaluate_ate.py, associate.py) stored in
tests/golden/tum_eval.npz by tests/golden/make_tum_golden.py."""
import os

import numpy as np
import pytest

from vslam_b200 import evaluation as E

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tum_eval.npz")


@pytest.fixture(scope="module")
def case(tmp_path_factory):
    g = np.load(GOLDEN)
    d = tmp_path_factory.mktemp("tum")
    (d / "gt.txt").write_text(str(g["gt_text"]))
    (d / "est.txt").write_text(str(g["est_text"]))
    return g, E.read_trajectory(str(d / "gt.txt")), E.read_trajectory(str(d / "est.txt"))


def test_read_trajectory_drops_nan_and_zero_quaternion_lines(case):
    g, gt, est = case
    assert len(gt) == int(g["n_gt"]) and len(est) == int(g["n_est"])
    assert 1305031104.5 not in est and 1305031104.6 not in est


@pytest.mark.parametrize("unit,delta", [("s", 1.0), ("f", 5), ("m", 0.1), ("deg", 2.0)])
def test_relative_pose_error_matches_the_reference_script(case, unit, delta):
    g, gt, est = case
    got, want = E.relative_pose_error(gt, est, delta, unit), g["rpe_" + unit]
    assert got.shape == want.shape
    assert np.array_equal(got[:, :4], want[:, :4])                 # the same pose pairs, ties included
    assert np.allclose(got[:, 4:], want[:, 4:], rtol=0, atol=1e-12)


def test_relative_pose_error_with_time_offset_and_scale(case):
    g, gt, est = case
    got = E.relative_pose_error(gt, est, 1.0, "s", offset=0.01, scale=1.02)
    assert np.allclose(got, g["rpe_s_offset_scale"], rtol=0, atol=1e-12)


def test_absolute_trajectory_error_matches_the_reference_script(case):
    g, gt, est = case
    a = E.absolute_trajectory_error(gt, est)
    assert np.array_equal(np.array(a["matches"]), g["ate_matches"])
    assert np.allclose(a["rot"], g["ate_rot"], atol=1e-12) and np.allclose(a["trans"], g["ate_trans"], atol=1e-12)
    assert np.allclose(a["errors"], g["ate_err"], atol=1e-12)
    assert abs(a["rmse"] - np.sqrt(np.mean(g["ate_err"] ** 2))) < 1e-15


def test_errors_on_disjoint_trajectories(case):
    _, gt, _ = case
    far = {t + 1e6: T for t, T in list(gt.items())[:10]}
    with pytest.raises(ValueError):
        E.absolute_trajectory_error(gt, far)
    with pytest.raises(ValueError):
        E.relative_pose_error(gt, far)


def test_write_then_read_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    q = rng.normal(size=(5, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    poses7 = np.concatenate([q, rng.normal(size=(5, 3))], 1)          # C-ABI order: qx qy qz qw tx ty tz
    stamps = 100.0 + np.arange(5) / 30.0
    E.write_trajectory(str(tmp_path / "t.txt"), stamps, poses7)
    back = E.read_trajectory(str(tmp_path / "t.txt"))
    assert len(back) == 5
    for s, p in zip(sorted(back), poses7):
        assert np.allclose(back[s], E.pose_matrix(p[4:], p[:4]), atol=1e-8)
