"""ctypes binding of include/vslam_b200.h.  Fails loudly when the CUDA library is missing: there is no
CPU fallback."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# VSLAMB200_LIB: another build of the same library (A/B kernel experiments, tools/ab.sh)
LIB_PATH = os.environ.get("VSLAMB200_LIB") or os.path.join(_HERE, "lib", "libvslam_b200.so")

MAX_LEVELS = 8
TRACE_STRIDE = 64
LOSS_NONE, LOSS_QUADRATIC, LOSS_TUKEY, LOSS_HUBER, LOSS_TDIST = range(5)
SCALER_IDENTITY, SCALER_MEDIAN, SCALER_MEAN, SCALER_TDIST = range(4)
ROW_MAJOR, COL_MAJOR = 0, 1
DEPTH_F32, DEPTH_F64, DEPTH_U16 = 0, 1, 2
RGB, BGR = 0, 1
SOLVER_GN, SOLVER_LM = 0, 1
SAMPLING_NEAREST, SAMPLING_BILINEAR = 0, 1
KERNEL_AUTO, KERNEL_DENSE, KERNEL_MOMENTS = 0, 1, 2
# chi2 of a view's normal equations: as the reference's Release build evaluates NormalEquations.cpp:58 (w_0 * sum r^2),
# or the documented r^T W r (vslam_b200.h)
CHI2_REFERENCE, CHI2_WEIGHTED = 0, 1
DEFAULT_CHI2_MODE = CHI2_REFERENCE  # make_config()'s default (a zeroed C config means the same)
R_NOT_INTEREST, R_INVISIBLE = -32768, -32767
STOP_NAMES = ["maxiter", "constraints", "hessian", "chi2", "converged", "nan"]


class Config(C.Structure):
    _fields_ = [("min_gradient", C.c_double), ("loss_kind", C.c_int), ("huber_c", C.c_double),
                ("loss_v", C.c_double), ("scaler_kind", C.c_int), ("scaler_v", C.c_double),
                ("min_step_size", C.c_double), ("max_iterations", C.c_int), ("include_prior", C.c_int),
                ("solver_kind", C.c_int), ("lm_lambda0", C.c_double), ("lm_up", C.c_double), ("lm_down", C.c_double),
                ("sampling", C.c_int), ("depth_weight", C.c_double), ("chi2_mode", C.c_int), ("level_kernel", C.c_int),
                ("cluster_size", C.c_int), ("stream_parts", C.c_int)]


class Job(C.Structure):
    _fields_ = [("to", C.c_uint64), ("to_pose", C.c_double * 7), ("n_from", C.c_int), ("multi", C.c_int),
                ("from_", C.POINTER(C.c_uint64)), ("from_poses", C.POINTER(C.c_double))]


class Result(C.Structure):
    _fields_ = [("pose", C.c_double * 7), ("cov", C.c_double * 36), ("status", C.c_int),
                ("iters", C.c_int * MAX_LEVELS), ("stop", C.c_int * MAX_LEVELS)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("pixel_iterations", C.c_uint64),
                ("point_iterations", C.c_uint64), ("gn_iterations", C.c_uint64), ("frames_uploaded", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("ms_lists", C.c_double),
                ("ms_preprocess", C.c_double), ("ms_gn_level", C.c_double * MAX_LEVELS),
                ("pixel_iterations_level", C.c_uint64 * MAX_LEVELS), ("phase_clocks", (C.c_uint64 * 10) * MAX_LEVELS),
                ("moment_launches", C.c_uint64), ("dense_launches", C.c_uint64)]


EXPORTS = [
    "vslamb200_last_error", "vslamb200_version", "vslamb200_create", "vslamb200_destroy", "vslamb200_set_config",
    "vslamb200_set_stream", "vslamb200_frames_upload", "vslamb200_frames_upload_device", "vslamb200_frames_attach_device",
    "vslamb200_frames_ingest_rgbd",
    "vslamb200_frame_release",
    "vslamb200_frame_release_all", "vslamb200_frame_dims", "vslamb200_frame_download", "vslamb200_align",
    "vslamb200_set_trace", "vslamb200_get_trace", "vslamb200_evaluate", "vslamb200_get_stats",
    "vslamb200_reset_stats", "vslamb200_set_timing", "vslamb200_synchronize", "vslamb200_frames_upload_async",
    "vslamb200_interest_points_noalloc", "vslamb200_frames_upload_u16_async",
]
NOALLOC_MAX_PIXELS = 4096 * 4096

_lib = None


class VslamB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vslam_b200 error {code}: {msg}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m vslam_b200.build` "
                               "(vslam_b200 has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, i, u64, sz = C.c_void_p, C.c_int, C.c_uint64, C.c_size_t
        L.vslamb200_last_error.restype = C.c_char_p
        L.vslamb200_create.argtypes = [vp, i, vp]
        L.vslamb200_destroy.argtypes = [vp]
        L.vslamb200_destroy.restype = None
        L.vslamb200_set_config.argtypes = [vp, vp]
        L.vslamb200_set_stream.argtypes = [vp, vp]
        L.vslamb200_frames_upload.argtypes = [vp, i, vp, i, i, vp, sz, vp, i, sz, i, vp, i]
        L.vslamb200_frames_upload_async.argtypes = [vp, i, vp, i, i, vp, sz, vp, i, sz, i, vp, i]
        L.vslamb200_frames_upload_u16_async.argtypes = [vp, i, vp, i, i, vp, vp, C.c_double, vp, i]
        L.vslamb200_frames_upload_device.argtypes = [vp, i, vp, i, i, vp, vp, vp, i]
        L.vslamb200_frames_attach_device.argtypes = [vp, i, vp, i, i, vp, vp, vp, i]
        L.vslamb200_frames_ingest_rgbd.argtypes = [vp, i, vp, i, i, vp, i, vp, i, C.c_double, vp, i]
        L.vslamb200_frame_release.argtypes = [vp, u64]
        L.vslamb200_frame_release_all.argtypes = [vp]
        L.vslamb200_frame_dims.argtypes = [vp, u64, i, vp]
        L.vslamb200_frame_download.argtypes = [vp, u64, i, vp, vp, vp, vp, vp]
        L.vslamb200_align.argtypes = [vp, vp, i, vp]
        L.vslamb200_set_trace.argtypes = [vp, i]
        L.vslamb200_get_trace.argtypes = [vp, i, vp]
        L.vslamb200_evaluate.argtypes = [vp, u64, u64, i, vp, i, vp, vp]
        L.vslamb200_get_stats.argtypes = [vp, vp]
        L.vslamb200_reset_stats.argtypes = [vp]
        L.vslamb200_set_timing.argtypes = [vp, i]
        L.vslamb200_synchronize.argtypes = [vp]
        L.vslamb200_interest_points_noalloc.argtypes = [vp, vp, vp, i, i, C.c_double, vp, vp, vp]
        _lib = L
    return _lib


def check(code):
    if code != 0:
        raise VslamB200Error(code, lib().vslamb200_last_error().decode())


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def make_config(min_gradient=30.0, loss=LOSS_TDIST, scaler=SCALER_TDIST, huber_c=1.345, v=5.0,
                min_step_size=1e-7, max_iterations=100, include_prior=False, solver=SOLVER_GN, lm_lambda0=1e-3,
                lm_up=10.0, lm_down=0.1, sampling=SAMPLING_NEAREST, depth_weight=0.0, level_kernel=None,
                cluster_size=None, stream_parts=None, chi2_mode=None):
    """level_kernel / cluster_size / stream_parts: execution choices (0 = automatic).  When not given they come from the
    environment (VSLAMB200_KERNEL = dense | moments, VSLAMB200_CLUSTER, VSLAMB200_PARTS): the switches of the tuning
    tools live in this binding, the library itself reads no environment variable."""
    if level_kernel is None:
        level_kernel = {"dense": KERNEL_DENSE, "moments": KERNEL_MOMENTS}.get(os.environ.get("VSLAMB200_KERNEL", ""), KERNEL_AUTO)
    if cluster_size is None:
        cluster_size = int(os.environ.get("VSLAMB200_CLUSTER", "0"))
    if chi2_mode is None:
        chi2_mode = DEFAULT_CHI2_MODE
    if stream_parts is None:
        stream_parts = int(os.environ.get("VSLAMB200_PARTS", "0"))
    return Config(min_gradient, loss, huber_c, v, scaler, v, min_step_size, max_iterations, int(include_prior), solver,
                  lm_lambda0, lm_up, lm_down, sampling, depth_weight, chi2_mode, level_kernel, cluster_size, stream_parts)


class Context:
    """Thin object wrapper over the C ABI (one context = one device + one stream)."""

    def __init__(self, cfg=None, device=0):
        self.cfg = cfg or make_config()
        h = C.c_void_p()
        check(lib().vslamb200_create(C.byref(self.cfg), device, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().vslamb200_destroy(self.h)
            self.h = None

    __del__ = close

    def set_config(self, cfg):
        self.cfg = cfg
        check(lib().vslamb200_set_config(self.h, C.byref(cfg)))

    def upload(self, ids, gray, depth, cams, n_levels, layout=ROW_MAJOR):
        """gray (n,H,W) u8, depth (n,H,W) f32|f64 (or (n,W,H) memory when layout == COL_MAJOR: pass W,H via shape
        of the row-major interpretation), cams (n,4) or (4,)."""
        gray = np.ascontiguousarray(gray)
        depth = np.ascontiguousarray(depth)
        if gray.ndim == 2:
            gray, depth = gray[None], depth[None]
        n = gray.shape[0]
        if layout == ROW_MAJOR:
            H, W = gray.shape[1:]
        else:
            W, H = gray.shape[1:]
        ids = np.ascontiguousarray(np.atleast_1d(ids), np.uint64)
        cams = np.ascontiguousarray(np.broadcast_to(np.asarray(cams, np.float64).reshape(-1, 4), (n, 4)))
        assert gray.dtype == np.uint8 and depth.dtype in (np.float32, np.float64) and len(ids) == n
        dt = DEPTH_F32 if depth.dtype == np.float32 else DEPTH_F64
        check(lib().vslamb200_frames_upload(self.h, n, ptr(ids), W, H, ptr(gray), 0, ptr(depth), dt, 0, layout,
                                            ptr(cams), n_levels))

    def upload_async(self, ids, gray_ptr, depth_ptr, n, W, H, cams, n_levels):
        """Queues the upload of n row-major u8 / f32 frames from (pinned) host memory at the given addresses and returns;
        the buffers must stay untouched until the next align() / synchronize()."""
        ids = np.ascontiguousarray(np.atleast_1d(ids), np.uint64)
        cams = np.ascontiguousarray(np.broadcast_to(np.asarray(cams, np.float64).reshape(-1, 4), (n, 4)))
        check(lib().vslamb200_frames_upload_async(self.h, n, ptr(ids), W, H, C.c_void_p(gray_ptr), 0,
                                                  C.c_void_p(depth_ptr), DEPTH_F32, 0, ROW_MAJOR, ptr(cams), n_levels))

    def upload_u16_async(self, ids, gray_ptr, depth_u16_ptr, depth_divisor, n, W, H, cams, n_levels):
        """Queues the upload of n row-major u8 grey / u16 depth frames (depth = value / depth_divisor) from (pinned) host
        memory and returns; the buffers must stay untouched until the next align() / synchronize()."""
        ids = np.ascontiguousarray(np.atleast_1d(ids), np.uint64)
        cams = np.ascontiguousarray(np.broadcast_to(np.asarray(cams, np.float64).reshape(-1, 4), (n, 4)))
        check(lib().vslamb200_frames_upload_u16_async(self.h, n, ptr(ids), W, H, C.c_void_p(gray_ptr), C.c_void_p(depth_u16_ptr),
                                                      float(depth_divisor), ptr(cams), n_levels))

    def ingest_rgbd(self, ids, rgb, depth, cams, n_levels, bgr=False, depth_divisor=1.0):
        """Node-side ingestion (NodeRgbdAlignment.cpp:177-199): rgb (n,H,W,3) u8, depth (n,H,W) f32 / f64 / u16."""
        ids = np.ascontiguousarray(np.atleast_1d(ids), np.uint64)
        rgb = np.ascontiguousarray(rgb, np.uint8)
        if rgb.ndim == 3:
            rgb = rgb[None]
        n, H, W, _ = rgb.shape
        depth = np.ascontiguousarray(depth)
        dt = {np.dtype(np.float32): DEPTH_F32, np.dtype(np.float64): DEPTH_F64, np.dtype(np.uint16): DEPTH_U16}[depth.dtype]
        cams = np.ascontiguousarray(np.broadcast_to(np.asarray(cams, np.float64).reshape(-1, 4), (n, 4)))
        check(lib().vslamb200_frames_ingest_rgbd(self.h, n, ptr(ids), W, H, ptr(rgb), BGR if bgr else RGB, ptr(depth), dt,
                                                 float(depth_divisor), ptr(cams), n_levels))

    def upload_device(self, ids, d_gray_ptr, d_depth_ptr, n, W, H, cams, n_levels):
        ids = np.ascontiguousarray(np.atleast_1d(ids), np.uint64)
        cams = np.ascontiguousarray(np.broadcast_to(np.asarray(cams, np.float64).reshape(-1, 4), (n, 4)))
        check(lib().vslamb200_frames_upload_device(self.h, n, ptr(ids), W, H, C.c_void_p(d_gray_ptr),
                                                   C.c_void_p(d_depth_ptr), ptr(cams), n_levels))

    def attach_device(self, ids, d_gray_ptr, d_depth_ptr, n, W, H, cams, n_levels):
        """Zero-copy: the level-0 planes are the caller's device buffers (tightly packed u8 / f32, 16-byte aligned), which
        must stay valid and unchanged while the frames are resident."""
        ids = np.ascontiguousarray(np.atleast_1d(ids), np.uint64)
        cams = np.ascontiguousarray(np.broadcast_to(np.asarray(cams, np.float64).reshape(-1, 4), (n, 4)))
        check(lib().vslamb200_frames_attach_device(self.h, n, ptr(ids), W, H, C.c_void_p(d_gray_ptr),
                                                   C.c_void_p(d_depth_ptr), ptr(cams), n_levels))

    def release(self, fid):
        check(lib().vslamb200_frame_release(self.h, int(fid)))

    def release_all(self):
        check(lib().vslamb200_frame_release_all(self.h))

    def dims(self, fid, level):
        d = np.zeros(4, np.int32)
        check(lib().vslamb200_frame_dims(self.h, int(fid), level, ptr(d)))
        return tuple(int(x) for x in d)

    def download(self, fid, level):
        W, H, Wd, Hd = self.dims(fid, level)
        out = dict(I=np.empty((H, W), np.uint8), Z=np.empty((Hd, Wd), np.float32),
                   Zsf=np.empty((Hd, Wd), np.float32), dIx=np.empty((H, W), np.int16),
                   dIy=np.empty((H, W), np.int16))
        check(lib().vslamb200_frame_download(self.h, int(fid), level, ptr(out["I"]), ptr(out["Z"]), ptr(out["Zsf"]),
                                             ptr(out["dIx"]), ptr(out["dIy"])))
        return out

    def align(self, jobs):
        """jobs: list of dict(to=id, to_pose=(7,), frm=[ids], from_poses=(k,7), multi=bool) -> list of dict"""
        n = len(jobs)
        arr = (Job * n)()
        keep = []
        for k, j in enumerate(jobs):
            frm = np.ascontiguousarray(np.atleast_1d(j["frm"]), np.uint64)
            fp = np.ascontiguousarray(np.asarray(j["from_poses"], np.float64).reshape(-1, 7))
            keep += [frm, fp]
            arr[k].to = int(j["to"])
            arr[k].to_pose = (C.c_double * 7)(*np.asarray(j["to_pose"], np.float64))
            arr[k].n_from = len(frm)
            arr[k].multi = int(bool(j.get("multi", len(frm) > 1)))
            arr[k].from_ = frm.ctypes.data_as(C.POINTER(C.c_uint64))
            arr[k].from_poses = fp.ctypes.data_as(C.POINTER(C.c_double))
        res = (Result * n)()
        check(lib().vslamb200_align(self.h, arr, n, res))
        out = []
        for r in res:
            out.append(dict(pose=np.array(r.pose), cov=np.array(r.cov).reshape(6, 6), status=r.status,
                            iters=np.array(r.iters), stop=np.array(r.stop)))
        return out

    def align_raw(self, job_array, n, result_array):
        check(lib().vslamb200_align(self.h, job_array, n, result_array))

    def set_trace(self, on):
        check(lib().vslamb200_set_trace(self.h, int(on)))

    def get_trace(self, job, n_levels):
        t = np.zeros((n_levels, self.cfg.max_iterations, TRACE_STRIDE))
        check(lib().vslamb200_get_trace(self.h, job, ptr(t)))
        return t

    def evaluate(self, frm, to, level, se3, multi_mask=False):
        W, H, _, _ = self.dims(frm, level)
        rcode = np.empty((H, W), np.int16)
        ne = np.zeros(64)
        se3 = np.ascontiguousarray(se3, np.float64)
        check(lib().vslamb200_evaluate(self.h, int(frm), int(to), level, ptr(se3), int(multi_mask), ptr(rcode), ptr(ne)))
        return dict(n=int(ne[0]), chi2=ne[1], scale=ne[2], offset=ne[3], b=ne[4:10].copy(),
                    A=ne[10:46].reshape(6, 6).copy(), rcode=rcode)

    def stats(self):
        s = Stats()
        check(lib().vslamb200_get_stats(self.h, C.byref(s)))
        out = {f[0]: getattr(s, f[0]) for f in Stats._fields_}
        out["ms_gn_level"] = list(out["ms_gn_level"])
        out["pixel_iterations_level"] = list(out["pixel_iterations_level"])
        out["phase_clocks"] = [list(r) for r in out["phase_clocks"]]
        return out

    def reset_stats(self):
        check(lib().vslamb200_reset_stats(self.h))

    def synchronize(self):
        check(lib().vslamb200_synchronize(self.h))
