"""Deterministic synthetic RGB-D scenes; ctypes front-end of synth/synth.c."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "synth", "synth.c")
_SO = os.path.join(_HERE, "lib", "libvslam_synth.so")

CAM_640 = (525.0, 525.0, 319.5, 239.5)          # odometry/test/test_alignment_se3.cpp:69
CAM_1280 = (1050.0, 1050.0, 639.5, 359.5)


def cam_for(W, H):
    """Camera of SURVEY §8(d): TUM intrinsics scaled with the width, centred principal point."""
    if (W, H) == (640, 480):
        return CAM_640
    if (W, H) == (1280, 720):
        return CAM_1280
    f = 525.0 * W / 640.0
    return (f, f, W / 2 - 0.5, H / 2 - 0.5)


def build(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["/usr/bin/gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", _SO, _SRC, "-lm"])
    return _SO


_lib = None


def _L():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.vsynth_render.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                    C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p]
        L.vsynth_se3_exp.argtypes = [C.c_void_p] * 3
        L.vsynth_pair_twist.argtypes = [C.c_uint64, C.c_void_p]
        L.vsynth_sequence_pose.argtypes = [C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def se3_exp(xi):
    xi = np.ascontiguousarray(xi, np.float64)
    R = np.empty(9)
    t = np.empty(3)
    _L().vsynth_se3_exp(_p(xi), _p(R), _p(t))
    return R.reshape(3, 3), t


def R_to_quat(R):
    """Rotation matrix -> quaternion (x, y, z, w)."""
    R = np.asarray(R, np.float64)
    tr = np.trace(R)
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = [(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s]
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0) * 2
        q = [0, 0, 0, (R[k, j] - R[j, k]) / s]
        q[i] = 0.25 * s
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
    q = np.array(q)
    if q[3] < 0:
        q = -q
    return q / np.linalg.norm(q)


def pose7(R, t):
    return np.concatenate([R_to_quat(R), np.asarray(t, np.float64)])


def render(W, H, cam, R=None, t=None, scene_seed=1234, holes_permille=0, hole_seed=99):
    """-> (gray u8 HxW, depth f32 HxW) of the textured plane seen from pose (R, t)."""
    R = np.eye(3) if R is None else R
    t = np.zeros(3) if t is None else t
    R = np.ascontiguousarray(R, np.float64)
    t = np.ascontiguousarray(t, np.float64)
    gray = np.empty((H, W), np.uint8)
    depth = np.empty((H, W), np.float32)
    _L().vsynth_render(W, H, cam[0], cam[1], cam[2], cam[3], _p(R), _p(t), scene_seed, holes_permille,
                       hole_seed, _p(gray), _p(depth))
    return gray, depth


def pair_twist(seed):
    xi = np.empty(6)
    _L().vsynth_pair_twist(seed, _p(xi))
    return xi


def make_pair(W=640, H=480, seed=1234, cam=None, holes_permille=0):
    """Config 1/3 of SURVEY §8(d): reference frame at identity, second frame at exp(xi(seed)).
    -> dict(gray0, depth0, gray1, depth1, xi, pose1 (7,), cam)"""
    cam = cam or cam_for(W, H)
    xi = pair_twist(seed)
    R, t = se3_exp(xi)
    g0, d0 = render(W, H, cam, scene_seed=seed, holes_permille=holes_permille, hole_seed=seed * 3 + 1)
    g1, d1 = render(W, H, cam, R, t, scene_seed=seed, holes_permille=holes_permille, hole_seed=seed * 3 + 2)
    return dict(gray0=g0, depth0=d0, gray1=g1, depth1=d1, xi=xi, pose1=pose7(R, t), cam=cam)


def sequence_pose(seed, k):
    R = np.empty(9)
    t = np.empty(3)
    _L().vsynth_sequence_pose(seed, k, _p(R), _p(t))
    return R.reshape(3, 3), t


def make_sequence(n_frames, W=640, H=480, seed=1234, cam=None):
    """Config 2: smooth trajectory; -> gray (F,H,W) u8, depth (F,H,W) f32, poses (F,7)."""
    cam = cam or cam_for(W, H)
    gray = np.empty((n_frames, H, W), np.uint8)
    depth = np.empty((n_frames, H, W), np.float32)
    poses = np.empty((n_frames, 7))
    # incremental pose composition (vsynth_sequence_pose is O(k) per call)
    for k in range(n_frames):
        R, t = sequence_pose(seed, k)
        poses[k] = pose7(R, t)
        _L().vsynth_render(W, H, cam[0], cam[1], cam[2], cam[3], _p(np.ascontiguousarray(R)),
                           _p(np.ascontiguousarray(t)), seed, 0, 0, _p(gray[k]), _p(depth[k]))
    return gray, depth, poses
