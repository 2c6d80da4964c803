"""ctypes front-end of oracle/_ref/libvslam_ref.so: the reference's OWN aligner (SE3Alignment, GaussNewton, losses, scalers,
warp, Frame), built from /root/reference by oracle/build_ref.py against the stand-in headers of oracle/ref_shim/.

TEST INFRASTRUCTURE ONLY (tests/): never imported by the product package.
"""
import ctypes as C

import numpy as np

from . import build_ref

TRACE_STRIDE = 64


class Config(C.Structure):
    _fields_ = [
        ("min_gradient", C.c_double),
        ("loss_kind", C.c_int),
        ("huber_c", C.c_double),
        ("loss_v", C.c_double),
        ("scaler_kind", C.c_int),
        ("scaler_v", C.c_double),
        ("min_step_size", C.c_double),
        ("max_iterations", C.c_int),
        ("include_prior", C.c_int),
    ]


def config_from(oracle_cfg):
    """The reference-side subset of an oracle Config (oracle.make_config)."""
    c = oracle_cfg
    return Config(c.min_gradient, c.loss_kind, c.huber_c, c.loss_v, c.scaler_kind, c.scaler_v, c.min_step_size,
                  c.max_iterations, c.include_prior)


_lib = None


def available():
    return build_ref.build() is not None


def lib():
    global _lib
    if _lib is None:
        so = build_ref.build()
        if so is None:
            raise RuntimeError("oracle/_ref/libvslam_ref.so: no reference tree and no prebuilt library")
        _lib = C.CDLL(so)
        _lib.ref_align.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_int] + \
            [C.c_void_p] * 6 + [C.c_char_p, C.c_int]
        _lib.ref_frame_level.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_int] + [C.c_void_p] * 7
        _lib.ref_odometry.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int] * 3 + \
            [C.c_void_p] * 2 + [C.c_char_p, C.c_int]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def align(cfg, gray, depth, cam4, twists, n_levels, multi=False, want_trace=True):
    """gray (F, H, W) u8 and depth (F, H, W) f64: the keyframes, then the frame to align; twists (F, 6) their poses.
    -> dict(twist, R, t, cov, pushed[level], last_iteration[level], trace[level, iteration, 64])."""
    gray = np.ascontiguousarray(gray, np.uint8)
    depth = np.ascontiguousarray(depth, np.float64)
    twists = np.ascontiguousarray(twists, np.float64)
    cam4 = np.ascontiguousarray(cam4, np.float64)
    F, H, W = gray.shape
    assert depth.shape == gray.shape and twists.shape == (F, 6)
    twist, Rt, cov = np.empty(6), np.empty(12), np.empty(36)
    pushed = np.zeros(n_levels, np.int32)
    last = np.full(n_levels, -1, np.int32)
    trace = np.full((n_levels, cfg.max_iterations, TRACE_STRIDE), np.nan) if want_trace else None
    err = C.create_string_buffer(512)
    st = lib().ref_align(C.byref(cfg), F - 1, W, H, _p(gray), _p(depth), _p(cam4), _p(twists), n_levels, int(multi),
                         _p(twist), _p(Rt), _p(cov), _p(pushed), _p(last), _p(trace), err, 512)
    if st != 0:
        raise RuntimeError("reference align: " + err.value.decode())
    return dict(twist=twist, R=Rt[:9].reshape(3, 3), t=Rt[9:], cov=cov.reshape(6, 6), pushed=pushed, last_iteration=last,
                trace=trace)


def frame_dims(gray, depth, cam4, n_levels, level):
    """(W, H) of the intensity level and (W, H) of the depth level of the reference's Frame."""
    gray = np.ascontiguousarray(gray, np.uint8)
    depth = np.ascontiguousarray(depth, np.float64)
    cam4 = np.ascontiguousarray(cam4, np.float64)
    H, W = gray.shape
    dims = np.zeros(4, np.int32)
    if lib().ref_frame_level(W, H, _p(gray), _p(depth), _p(cam4), n_levels, level, _p(dims), *[None] * 6) != 0:
        raise RuntimeError("reference frame: construction failed")
    return tuple(int(x) for x in dims)


def frame_level(gray, depth, cam4, n_levels, level):
    """One level of the reference's Frame after computePyramid / computeDerivatives / computePcl."""
    gray = np.ascontiguousarray(gray, np.uint8)
    depth = np.ascontiguousarray(depth, np.float64)
    cam4 = np.ascontiguousarray(cam4, np.float64)
    H, W = gray.shape
    w, h, wd, hd = frame_dims(gray, depth, cam4, n_levels, level)
    if (w, h) != (wd, hd):
        raise RuntimeError("reference frame: level %d: intensity %dx%d and depth %dx%d sizes disagree" % (level, w, h, wd, hd))
    dims = np.zeros(4, np.int32)
    I = np.empty((h, w), np.uint8)
    Z, dIx, dIy = (np.empty((h, w)) for _ in range(3))
    pcl = np.empty((h, w, 3))
    cam = np.empty(4)
    st = lib().ref_frame_level(W, H, _p(gray), _p(depth), _p(cam4), n_levels, level, _p(dims), _p(I), _p(Z), _p(dIx),
                               _p(dIy), _p(pcl), _p(cam))
    if st != 0:
        raise RuntimeError("reference frame: failed")
    return dict(I=I, Z=Z, dIx=dIx, dIy=dIy, pcl=pcl, cam=cam)


def odometry(cfg, gray, depth, cam4, stamps_ns, n_levels, motion_model="ConstantMotion", kf_period=2):
    """The reference's OdometryRgbd + MotionPrediction + KeyFrameSelectionIdx + Map over a frame sequence, in the order of
    the ROS node's processFrame.  gray (F, H, W) u8, depth (F, H, W) f64, stamps in nanoseconds.
    -> poses (F, 6) and speeds (F, 6) as twists (speed per second)."""
    gray = np.ascontiguousarray(gray, np.uint8)
    depth = np.ascontiguousarray(depth, np.float64)
    cam4 = np.ascontiguousarray(cam4, np.float64)
    stamps = np.ascontiguousarray(stamps_ns, np.uint64)
    F, H, W = gray.shape
    poses, speeds = np.empty((F, 6)), np.empty((F, 6))
    err = C.create_string_buffer(512)
    st = lib().ref_odometry(C.byref(cfg), F, W, H, _p(gray), _p(depth), _p(cam4), _p(stamps), n_levels,
                            int(motion_model == "ConstantMotion"), kf_period, _p(poses), _p(speeds), err, 512)
    if st != 0:
        raise RuntimeError("reference odometry: " + err.value.decode())
    return poses, speeds
