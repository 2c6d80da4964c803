"""ctypes front-end of the CPU oracle (oracle/vslam_oracle.cpp).

TEST INFRASTRUCTURE ONLY: import from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never from the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libvslam_oracle.so")

LOSS_NONE, LOSS_QUADRATIC, LOSS_TUKEY, LOSS_HUBER, LOSS_TDIST = range(5)
SCALER_IDENTITY, SCALER_MEDIAN, SCALER_MEAN, SCALER_TDIST = range(4)
SOLVER_GN, SOLVER_LM = range(2)
SAMPLING_NEAREST, SAMPLING_BILINEAR = range(2)
# chi2 of a view: w[0] * sum r^2 as the reference's Release build evaluates NormalEquations.cpp:58, or sum w r^2
CHI2_REFERENCE_RELEASE, CHI2_WEIGHTED = range(2)
DEFAULT_CHI2_MODE = CHI2_REFERENCE_RELEASE  # what make_config() uses when not told; tests/conftest.py switches it
R_NOT_INTEREST, R_INVISIBLE = -32768, -32767
TRACE_STRIDE = 64
STOP_NAMES = ["maxiter", "constraints", "hessian", "chi2", "converged", "nan"]


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
        ("f32_storage", C.c_int),
        ("solver_kind", C.c_int),
        ("lm_lambda0", C.c_double),
        ("lm_up", C.c_double),
        ("lm_down", C.c_double),
        ("sampling", C.c_int),
        ("depth_weight", C.c_double),
        ("chi2_mode", C.c_int),
    ]


def make_config(min_gradient=30.0, loss=LOSS_TDIST, scaler=SCALER_TDIST, huber_c=1.345, v=5.0,
                min_step_size=1e-7, max_iterations=100, include_prior=False, f32_storage=True, solver=SOLVER_GN,
                lm_lambda0=1e-3, lm_up=10.0, lm_down=0.1, sampling=SAMPLING_NEAREST, depth_weight=0.0,
                chi2_mode=None):
    if chi2_mode is None:
        chi2_mode = DEFAULT_CHI2_MODE
    return Config(min_gradient, loss, huber_c, v, scaler, v, min_step_size, max_iterations,
                  int(include_prior), int(f32_storage), solver, lm_lambda0, lm_up, lm_down, sampling, depth_weight,
                  chi2_mode)


def build(force=False):
    src = os.path.join(_HERE, "vslam_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.vo_frame_create.restype = C.c_void_p
        L.vo_frame_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                      C.c_double, C.c_double, C.c_int, C.c_int]
        L.vo_frame_destroy.argtypes = [C.c_void_p]
        L.vo_frame_n2_hits.restype = C.c_long
        L.vo_frame_n2_hits.argtypes = [C.c_void_p]
        L.vo_frame_dims.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.vo_frame_get_level.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 6
        L.vo_evaluate.restype = C.c_long
        L.vo_evaluate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p]
        L.vo_align.restype = C.c_int
        L.vo_align.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.vo_depth_down.restype = C.c_long
        L.vo_depth_down.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.vo_bilinear_f64.restype = C.c_double
        L.vo_bilinear_f64.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double]
        L.vo_bilinear_u8.restype = C.c_double
        L.vo_bilinear_u8.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double]
        L.vo_median.restype = C.c_double
        L.vo_median.argtypes = [C.c_void_p, C.c_int]
        L.vo_camera2image.restype = C.c_int
        L.vo_camera2image.argtypes = [C.c_double] * 4 + [C.c_void_p, C.c_void_p]
        L.vo_image2camera.argtypes = [C.c_double] * 4 + [C.c_int, C.c_int, C.c_double, C.c_void_p]
        L.vo_baseline_sequence.restype = C.c_long
        L.vo_baseline_sequence.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                           C.c_void_p]
        L.vo_baseline_batch.restype = C.c_long
        L.vo_baseline_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                        C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        for n in ("vo_pyr_down_u8", "vo_blur3_u8"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.vo_rgb_to_gray_u8.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_void_p]
        L.vo_depth_ingest.argtypes = [C.c_void_p, C.c_long, C.c_double, C.c_void_p]
        L.vo_sobel3_s16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        for n in ("vo_se3_exp", "vo_se3_log", "vo_se3_inv", "vo_quat_to_R"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_void_p]
        L.vo_se3_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


IDENTITY = np.array([0, 0, 0, 1, 0, 0, 0], dtype=np.float64)


def rgb_to_gray(rgb, bgr=False):
    """cv::cvtColor(COLOR_RGB2GRAY / COLOR_BGR2GRAY) of an (H, W, 3) u8 image (NodeRgbdAlignment.cpp:182)."""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    out = np.empty(rgb.shape[:-1], np.uint8)
    lib().vo_rgb_to_gray_u8(_p(rgb), C.c_long(out.size), int(bgr), _p(out))
    return out


def depth_ingest(depth, divisor=1.0):
    """NaN -> 0 and the dataset's depth factor (NodeRgbdAlignment.cpp:190, test_alignment_se3.cpp:99); -> f64."""
    d = np.ascontiguousarray(depth, np.float64)
    out = np.empty_like(d)
    lib().vo_depth_ingest(_p(d), C.c_long(d.size), C.c_double(divisor), _p(out))
    return out


def pyr_down_u8(img):
    H, W = img.shape
    out = np.empty(((H + 1) // 2, (W + 1) // 2), np.uint8)
    lib().vo_pyr_down_u8(_p(np.ascontiguousarray(img)), W, H, _p(out))
    return out


def blur3_u8(img):
    H, W = img.shape
    out = np.empty((H, W), np.uint8)
    lib().vo_blur3_u8(_p(np.ascontiguousarray(img)), W, H, _p(out))
    return out


def sobel3_s16(img):
    H, W = img.shape
    dx = np.empty((H, W), np.int16)
    dy = np.empty((H, W), np.int16)
    lib().vo_sobel3_s16(_p(np.ascontiguousarray(img)), W, H, _p(dx), _p(dy))
    return dx, dy


def depth_down(depth, f32_storage=False):
    depth = np.ascontiguousarray(depth, np.float64)
    H, W = depth.shape
    out = np.empty((int(H * 0.5), int(W * 0.5)), np.float64)
    n2 = lib().vo_depth_down(_p(depth), W, H, _p(out), int(f32_storage))
    return out, n2


def bilinear(m, x, y):
    m = np.ascontiguousarray(m)
    if m.dtype == np.uint8:
        return lib().vo_bilinear_u8(_p(m), m.shape[1], x, y)
    m = m.astype(np.float64)
    return lib().vo_bilinear_f64(_p(m), m.shape[1], x, y)


def median(v):
    v = np.ascontiguousarray(v, np.float64)
    return lib().vo_median(_p(v), len(v))


def se3_exp(xi):
    xi = np.ascontiguousarray(xi, np.float64)
    o = np.empty(7)
    lib().vo_se3_exp(_p(xi), _p(o))
    return o


def se3_log(p):
    p = np.ascontiguousarray(p, np.float64)
    o = np.empty(6)
    lib().vo_se3_log(_p(p), _p(o))
    return o


def se3_mul(a, b):
    a = np.ascontiguousarray(a, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    o = np.empty(7)
    lib().vo_se3_mul(_p(a), _p(b), _p(o))
    return o


def se3_inv(a):
    a = np.ascontiguousarray(a, np.float64)
    o = np.empty(7)
    lib().vo_se3_inv(_p(a), _p(o))
    return o


def quat_to_R(q):
    q = np.ascontiguousarray(q, np.float64)
    o = np.empty(9)
    lib().vo_quat_to_R(_p(q), _p(o))
    return o.reshape(3, 3)


class Frame:
    """Frame ctor + computePyramid + computeDerivatives + computePcl (node order)."""

    def __init__(self, gray, depth, fx, fy, cx, cy, n_levels, f32_storage=True):
        gray = np.ascontiguousarray(gray, np.uint8)
        depth = np.ascontiguousarray(depth, np.float64)
        H, W = gray.shape
        self.n_levels = n_levels
        self.h = lib().vo_frame_create(W, H, _p(gray), _p(depth), fx, fy, cx, cy, n_levels, int(f32_storage))
        if not self.h:
            raise RuntimeError("Inconsistent camera parameters / image / depth dimensions detected.")

    def __del__(self):
        if getattr(self, "h", None):
            lib().vo_frame_destroy(self.h)
            self.h = None

    def dims(self, level):
        d = np.zeros(4, np.int32)
        lib().vo_frame_dims(self.h, level, _p(d))
        return tuple(int(x) for x in d)

    def n2_hits(self):
        return lib().vo_frame_n2_hits(self.h)

    def level(self, level):
        W, H, Wd, Hd = self.dims(level)
        I = np.empty((H, W), np.uint8)
        Z = np.empty((Hd, Wd), np.float64)
        dx = np.empty((H, W), np.int16)
        dy = np.empty((H, W), np.int16)
        pcl = np.empty((Hd, Wd, 3), np.float64)
        cam = np.empty(4, np.float64)
        lib().vo_frame_get_level(self.h, level, _p(I), _p(Z), _p(dx), _p(dy), _p(pcl), _p(cam))
        return dict(I=I, Z=Z, dIx=dx, dIy=dy, pcl=pcl, cam=cam)


def evaluate(cfg, frm, to, level, se3=IDENTITY, multi_mask=False, want_J=False):
    W, H, _, _ = frm.dims(level)
    rcode = np.empty((H, W), np.int16)
    ne = np.zeros(64)
    J = np.empty((W * H, 6)) if want_J else None
    se3 = np.ascontiguousarray(se3, np.float64)
    n = lib().vo_evaluate(C.byref(cfg), frm.h, to.h, level, _p(se3), int(multi_mask), _p(rcode), _p(ne), _p(J))
    out = dict(n=int(ne[0]), chi2=ne[1], scale=ne[2], offset=ne[3], b=ne[4:10].copy(),
               A=ne[10:46].reshape(6, 6).copy(), rcode=rcode)
    if want_J:
        out["J"] = J[:n].copy()
    return out


def align(cfg, frm, from_poses, to, to_pose, multi=None, want_trace=False):
    if isinstance(frm, Frame):
        frm = [frm]
        from_poses = [from_poses]
        if multi is None:
            multi = False
    elif multi is None:
        multi = True
    n = len(frm)
    handles = (C.c_void_p * n)(*[f.h for f in frm])
    fp = np.ascontiguousarray(np.stack(from_poses), np.float64)
    tp = np.ascontiguousarray(to_pose, np.float64)
    nl = frm[0].n_levels
    pose = np.empty(7)
    cov = np.empty(36)
    iters = np.zeros(nl, np.int32)
    stop = np.zeros(nl, np.int32)
    trace = np.zeros((nl, cfg.max_iterations, TRACE_STRIDE)) if want_trace else None
    st = lib().vo_align(C.byref(cfg), n, handles, _p(fp), to.h, _p(tp), int(multi), _p(pose), _p(cov),
                        _p(iters), _p(stop), _p(trace))
    if st < 0:
        raise RuntimeError("oracle align: pyramid level sizes disagree (odd dimensions, SURVEY 9.9-3)")
    return dict(pose=pose, cov=cov.reshape(6, 6), iters=iters, stop=stop, status=st, trace=trace)


def baseline_sequence(cfg, gray, depth, fx, fy, cx, cy, n_levels, n_threads=1):
    """gray: (F,H,W) u8, depth: (F,H,W) f32 -> (F-1,7) relative poses, total iterations."""
    gray = np.ascontiguousarray(gray, np.uint8)
    depth = np.ascontiguousarray(depth, np.float32)
    F, H, W = gray.shape
    poses = np.zeros((F - 1, 7))
    it = lib().vo_baseline_sequence(C.byref(cfg), F, W, H, _p(gray), _p(depth), fx, fy, cx, cy, n_levels,
                                    n_threads, _p(poses))
    return poses, it


def baseline_batch(cfg, gray, depth, fx, fy, cx, cy, n_levels, n_threads=1, stride=2, n_pairs=None, start=None):
    """Independent pairs out of one frame array: pair i aligns frame i * stride + 1 against frame i * stride (stride 2:
    frames stored ref0, cur0, ref1, cur1, ...; stride 1: the consecutive pairs of a sequence).
    start: the initial guess (pose7) of every pair's second frame, identity when None.
    -> (n_pairs, 7) poses, (n_pairs, n_levels) completed Gauss-Newton steps per level."""
    gray = np.ascontiguousarray(gray, np.uint8)
    depth = np.ascontiguousarray(depth, np.float32)
    F, H, W = gray.shape
    if n_pairs is None:
        n_pairs = (F - 2) // stride + 1 if F >= 2 else 0
    assert (n_pairs - 1) * stride + 1 < F
    poses = np.zeros((n_pairs, 7))
    iters = np.zeros((n_pairs, n_levels), np.int32)
    start = None if start is None else np.ascontiguousarray(start, np.float64)
    lib().vo_baseline_batch(C.byref(cfg), n_pairs, stride, W, H, _p(gray), _p(depth), fx, fy, cx, cy, n_levels, n_threads,
                            _p(poses), _p(iters), _p(start))
    return poses, iters
