#!/usr/bin/env python
"""Generates tests/golden/tum_eval.npz: outputs of the reference's OWN evaluation scripts
(/root/reference/script/vslam_evaluation/tum/evaluate_rpe.py, evaluate_ate.py, associate.py) on a small synthetic
ground-truth / estimate trajectory pair, for vslam_b200/evaluation.py to be checked against.  Run in the build
container (the reference tree is not on the GPU box; the .npz travels):

    python tests/golden/make_tum_golden.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/script/vslam_evaluation/tum")
import associate  # noqa: E402
import evaluate_ate  # noqa: E402
import evaluate_rpe  # noqa: E402


class ListKeysDict(dict):
    """The reference's "m" / "deg" / "rad" units call traj.keys().sort() (Python 2); handing it a dict whose keys() is a
    list runs that code unmodified under Python 3."""

    def keys(self):
        return list(dict.keys(self))


def quat(axis, angle):
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    return np.concatenate([axis * np.sin(angle / 2), [np.cos(angle / 2)]])


def trajectories():
    """Ground truth at 100 Hz with 4-decimal stamps (TUM mocap style), estimate at 30 Hz with drift, noise, a gap, a
    line with a NaN and a line with an all-zero quaternion."""
    rng = np.random.default_rng(2022)
    gt, est = [], []
    for k in range(400):
        t = 1305031100.0 + round(k * 0.01, 4)
        p = [np.sin(0.4 * k * 0.01), 0.3 * k * 0.01, 0.1 * np.cos(0.7 * k * 0.01)]
        gt.append([t] + p + list(quat([0.2, 1.0, 0.1], 0.25 * k * 0.01)))
    for k in range(110):
        if 40 <= k < 46:
            continue                                                   # tracking loss
        t = 1305031100.0037 + k / 30.0
        s = t - 1305031100.0
        p = np.array([np.sin(0.4 * s), 0.3 * s, 0.1 * np.cos(0.7 * s)]) * 1.01 + 0.002 * s + rng.normal(0, 0.003, 3)
        est.append([t] + list(p + [0.5, -0.2, 0.1]) + list(quat([0.2, 1.0, 0.1], 0.25 * s + 0.004 * s + rng.normal(0, 0.002))))
    return np.array(gt), np.array(est)


def write(path, rows, extra=()):
    with open(path, "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for r in rows:
            f.write("%.4f %.6f %.6f %.6f %.8f %.8f %.8f %.8f\n" % tuple(r))
        for line in extra:
            f.write(line + "\n")


def main():
    gt, est = trajectories()
    d = tempfile.mkdtemp()
    fg, fe = os.path.join(d, "gt.txt"), os.path.join(d, "est.txt")
    write(fg, gt)
    write(fe, est, extra=["1305031104.5000 nan 0 0 0 0 0 1", "1305031104.6000 1 2 3 0 0 0 0"])
    data = {"gt_text": np.array(open(fg).read()), "est_text": np.array(open(fe).read())}
    tg, te = ListKeysDict(evaluate_rpe.read_trajectory(fg)), ListKeysDict(evaluate_rpe.read_trajectory(fe))
    data["n_gt"], data["n_est"] = np.array(len(tg)), np.array(len(te))
    for unit, delta in (("s", 1.0), ("f", 5), ("m", 0.1), ("deg", 2.0)):
        res = evaluate_rpe.evaluate_trajectory(tg, te, 0, True, delta, unit, 0.0, 1.0)
        data["rpe_%s" % unit] = np.array(res, np.float64)
    res = evaluate_rpe.evaluate_trajectory(tg, te, 0, True, 1.0, "s", 0.01, 1.02)
    data["rpe_s_offset_scale"] = np.array(res, np.float64)
    lg, le = associate.read_file_list(fg), associate.read_file_list(fe)
    # the estimate file's NaN line makes associate's reader keep strings; ATE in the reference reads the same lists
    m = associate.associate(lg, le, 0.0, 0.02)
    data["ate_matches"] = np.array(m, np.float64)
    a = np.matrix([[float(v) for v in lg[x][0:3]] for x, _ in m]).transpose()
    b = np.matrix([[float(v) for v in le[y][0:3]] for _, y in m]).transpose()
    rot, trans, err = evaluate_ate.align(b, a)
    data["ate_rot"], data["ate_trans"], data["ate_err"] = np.asarray(rot), np.asarray(trans), np.asarray(err).ravel()
    np.savez_compressed(os.path.join(HERE, "tum_eval.npz"), **data)
    print("wrote tum_eval.npz", {k: np.shape(v) for k, v in data.items()})


if __name__ == "__main__":
    main()
