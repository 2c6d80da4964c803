"""Trajectory evaluation either side of the odometry loop: TUM trajectory files, relative pose
error (RPE) and absolute trajectory error (ATE) with the semantics of the reference's evaluation scripts
(/root/reference/script/vslam_evaluation/tum/evaluate_rpe.py, evaluate_ate.py, associate.py - the TUM RGB-D benchmark
tools the reference's README quotes its 0.036 m RPE with).  Host-side numpy, written for this repository; checked
against the reference's own scripts through tests/golden/tum_eval.npz (tests/golden/make_tum_golden.py).

Trajectory: dict stamp (float seconds) -> 4x4 homogeneous matrix, file lines "stamp tx ty tz qx qy qz qw".
"""
import numpy as np

_EPS = np.finfo(float).eps * 4.0


def pose_matrix(t, q):
    """4x4 matrix of translation t and quaternion q = (x, y, z, w); a (near) zero quaternion means no rotation
    (evaluate_rpe.py:46-74)."""
    M = np.eye(4)
    M[:3, 3] = t
    q = np.asarray(q, np.float64)
    n = float(q @ q)
    if n >= _EPS:
        x, y, z, w = q * np.sqrt(2.0 / n)
        M[:3, :3] = [[1 - y * y - z * z, x * y - z * w, x * z + y * w],
                     [x * y + z * w, 1 - x * x - z * z, y * z - x * w],
                     [x * z - y * w, y * z + x * w, 1 - x * x - y * y]]
    return M


def read_trajectory(path):
    """evaluate_rpe.py:76-108: '#' comments, blanks / commas / tabs as separators; lines with an all-zero quaternion
    or a NaN are dropped."""
    traj = {}
    with open(path) as f:
        for line in f:
            line = line.replace(",", " ").replace("\t", " ").strip()
            if not line or line[0] == "#":
                continue
            v = [float(x) for x in line.split() if x.strip() != ""]
            if len(v) < 8 or v[4:8] == [0, 0, 0, 0] or any(np.isnan(v)):
                continue
            traj[v[0]] = pose_matrix(v[1:4], v[4:8])
    return traj


def write_trajectory(path, stamps, poses7):
    """Poses as the C ABI returns them (qx qy qz qw tx ty tz) -> TUM order (tx ty tz qx qy qz qw)."""
    with open(path, "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for s, p in zip(stamps, poses7):
            f.write("%.6f %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n" % (s, p[4], p[5], p[6], p[0], p[1], p[2], p[3]))


def closest_index(sorted_values, t):
    """Index of the value closest to t; of two equally close values the one the reference's bisection keeps
    (evaluate_rpe.py:110-136) - the search is restated step for step because ties matter for parity."""
    lo, hi, best = 0, len(sorted_values), 0
    diff = abs(sorted_values[0] - t)
    while lo < hi:
        mid = (lo + hi) // 2
        d = abs(sorted_values[mid] - t)
        if d < diff:
            diff, best = d, mid
        if t == sorted_values[mid]:
            return mid
        if sorted_values[mid] > t:
            hi = mid
        else:
            lo = mid + 1
    return best


def _rotation_angle(R):
    return float(np.arccos(min(1.0, max(-1.0, (np.trace(R) - 1.0) / 2.0))))


def _index_along(traj, stamps, unit):
    if unit == "s":
        return list(stamps)
    if unit == "f":
        return list(range(len(stamps)))
    if unit == "m":                                    # evaluate_rpe.py:175-187
        out, acc = [0.0], 0.0
        for a, b in zip(stamps[:-1], stamps[1:]):
            acc += float(np.linalg.norm((np.linalg.inv(traj[b]) @ traj[a])[:3, 3]))
            out.append(acc)
        return out
    if unit in ("rad", "deg"):                         # evaluate_rpe.py:189-202
        k = 1.0 if unit == "rad" else 180.0 / np.pi
        out, acc = [0.0], 0.0
        for a, b in zip(stamps[:-1], stamps[1:]):
            acc += _rotation_angle((np.linalg.inv(traj[b]) @ traj[a])[:3, :3]) * k
            out.append(acc)
        return out
    raise ValueError("Unknown unit for delta: '%s'" % unit)


def relative_pose_error(traj_gt, traj_est, delta=1.0, unit="s", offset=0.0, scale=1.0):
    """evaluate_rpe.py:204-297 with --fixed_delta (deterministic: every pair (i, closest(index_i + delta)), no
    sub-sampling).  -> array of rows [stamp_est_0, stamp_est_1, stamp_gt_0, stamp_gt_1, trans (m), rot (rad)]."""
    s_gt, s_est = sorted(traj_gt), sorted(traj_est)
    overlap = {s_est[closest_index(s_est, s_gt[closest_index(s_gt, t + offset)] - offset)] for t in s_est}
    if len(overlap) < 2:
        raise ValueError("Number of overlap in the timestamps is too small.")
    index = _index_along(traj_est, s_est, unit)
    max_dt = 2.0 * float(np.median(np.diff(s_gt)))
    rows = []
    for i in range(len(s_est)):
        j = closest_index(index, index[i] + delta)
        if j == len(s_est) - 1:
            continue
        e0, e1 = s_est[i], s_est[j]
        g0, g1 = s_gt[closest_index(s_gt, e0 + offset)], s_gt[closest_index(s_gt, e1 + offset)]
        if abs(g0 - (e0 + offset)) > max_dt or abs(g1 - (e1 + offset)) > max_dt:
            continue
        d_est = np.linalg.inv(traj_est[e1]) @ traj_est[e0]
        d_est[:3, 3] *= scale
        d_gt = np.linalg.inv(traj_gt[g1]) @ traj_gt[g0]
        err = np.linalg.inv(d_est) @ d_gt
        rows.append([e0, e1, g0, g1, float(np.linalg.norm(err[:3, 3])), _rotation_angle(err[:3, :3])])
    if len(rows) < 2:
        raise ValueError("Couldn't find matching timestamp pairs between groundtruth and estimated trajectory!")
    return np.array(rows)


def associate(first, second, offset=0.0, max_difference=0.02):
    """associate.py:71-101: greedy closest-stamp matching, every stamp used at most once.  -> sorted [(a, b)]."""
    a_keys, b_keys = list(first), list(second)
    cand = sorted((abs(a - (b + offset)), a, b) for a in a_keys for b in b_keys if abs(a - (b + offset)) < max_difference)
    used_a, used_b, out = set(), set(), []
    for _, a, b in cand:
        if a not in used_a and b not in used_b:
            used_a.add(a)
            used_b.add(b)
            out.append((a, b))
    return sorted(out)


def horn_alignment(model, data):
    """evaluate_ate.py:47-79: rotation / translation that map `model` (3xn) onto `data` (3xn) in the least-squares
    sense (Horn's closed form through an SVD), and the remaining distance per point."""
    mc, dc = model - model.mean(1, keepdims=True), data - data.mean(1, keepdims=True)
    U, _, Vh = np.linalg.svd((mc @ dc.T).T)
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vh) < 0:
        S[2, 2] = -1
    rot = U @ S @ Vh
    trans = data.mean(1, keepdims=True) - rot @ model.mean(1, keepdims=True)
    err = rot @ model + trans - data
    return rot, trans, np.sqrt((err * err).sum(0))


def absolute_trajectory_error(traj_gt, traj_est, offset=0.0, scale=1.0, max_difference=0.02):
    """evaluate_ate.py:139-160: associate by stamp, align the estimate to the ground truth, distances per pose."""
    m = associate(traj_gt, traj_est, offset, max_difference)
    if len(m) < 2:
        raise ValueError("Couldn't find matching timestamp pairs between groundtruth and estimated trajectory!")
    gt = np.array([traj_gt[a][:3, 3] for a, _ in m]).T
    est = np.array([traj_est[b][:3, 3] * scale for _, b in m]).T
    rot, trans, err = horn_alignment(est, gt)
    return dict(matches=m, rot=rot, trans=trans, errors=err, **stats(err))


def stats(e):
    e = np.asarray(e, np.float64)
    return dict(rmse=float(np.sqrt(e @ e / len(e))), mean=float(e.mean()), median=float(np.median(e)), std=float(e.std()),
                min=float(e.min()), max=float(e.max()))
