"""CPU restatement of the direct caller of the alignment path - TEST INFRASTRUCTURE, like the rest
of oracle/: only tests/ may import it.

Follows, relative to /root/reference/src/:
  ros/nodes/NodeRgbdAlignment.cpp:121-150   processFrame: predict -> odometry update -> prediction update ->
                                            key-frame selection -> map insert
  vslam/src/odometry/src/Odometry.cpp:21-62 OdometryRgbd: aligner with the motion prior always on (:24); stacked
                                            alignment against {lastKf, lastFrame} when they differ (:33-35), else
                                            lastFrame (:41); speed = log(pose * lastPose^-1) / dT (:43-50)
  vslam/src/odometry/src/mapping/Map.cpp:19-33  rings of at most 7 frames / 7 key frames, newest first
  vslam/src/odometry/src/KeyFrameSelection.h:36-51  KeyFrameSelectionIdx: counter, key frame iff ctr % period == 0
  vslam/src/odometry/src/MotionPrediction.cpp:38-55, MotionPrediction.h:41-76  NoMotion / ConstantMotion
The alignment itself is oracle.align (vslam_oracle.cpp).  Parity: pinned - tests/test_oracle_ref_cpu.py runs this loop
against the reference's own Odometry.cpp / MotionPrediction.cpp / KeyFrameSelection.cpp / Map.cpp (oracle/_ref, compiled
unchanged against the stand-in headers); the GPU host mirror is compared with this restatement.
"""
import numpy as np

from . import oracle as O


class Map:
    def __init__(self, max_frames=7, max_key_frames=7):
        self.frames, self.key_frames = [], []
        self.max_frames, self.max_key_frames = max_frames, max_key_frames

    def insert(self, frame, is_key_frame):
        if len(self.frames) >= self.max_frames:
            self.frames.pop()
        self.frames.insert(0, frame)
        if is_key_frame:
            if len(self.key_frames) >= self.max_key_frames:
                self.key_frames.pop()
            self.key_frames.insert(0, frame)

    def last_kf(self):
        return self.key_frames[0] if self.key_frames else None

    def last_frame(self):
        return self.frames[0] if self.frames else None


class KeyFrameSelectionIdx:
    def __init__(self, period=2):
        self.period, self.ctr = period, 0

    def update(self, _frame):
        self.ctr += 1

    def is_key_frame(self):
        return self.ctr == 0 or self.ctr % self.period == 0


class MotionPredictionNoMotion:
    def __init__(self):
        self.last_pose = O.IDENTITY.copy()

    def update(self, pose, _t):
        self.last_pose = pose

    def predict(self, _t):
        return self.last_pose.copy()


class MotionPredictionConstant:
    def __init__(self):
        self.speed = np.zeros(6)
        self.last_pose = O.IDENTITY.copy()
        self.last_t = 0

    def update(self, pose, t):
        if t < self.last_t:
            raise RuntimeError("New timestamp is older than last one!")
        dT = (float(t) - float(self.last_t)) / 1e9
        self.speed = O.se3_log(O.se3_mul(pose, O.se3_inv(self.last_pose))) / dT
        self.last_pose = pose
        self.last_t = t

    def predict(self, t):
        dT = (float(t) - float(self.last_t)) / 1e9
        return O.se3_mul(O.se3_exp(self.speed * dT), self.last_pose)


class _F:
    def __init__(self, frame, t):
        self.frame, self.t, self.pose = frame, t, O.IDENTITY.copy()


def run(cfg, frames, stamps, prediction="ConstantMotion", kf_period=2):
    """frames: oracle.Frame list, stamps: ns.  cfg must have include_prior=True (Odometry.cpp:24).
    -> poses (F,7), speeds (F,6), which alignment ran per frame ('first' | 'single' | 'stacked')."""
    m, sel = Map(), KeyFrameSelectionIdx(kf_period)
    pred = MotionPredictionConstant() if prediction == "ConstantMotion" else MotionPredictionNoMotion()
    poses, speeds, kinds = [], [], []
    for fr, t in zip(frames, stamps):
        f = _F(fr, int(t))
        f.pose = pred.predict(f.t)                       # frame->set(*_prediction->predict(frame->t()))
        last = m.last_frame()
        if last is not None:
            kf = m.last_kf()
            if kf is not None and kf is not last:        # _includeKeyFrame (Odometry.cpp:33-35)
                res = O.align(cfg, [kf.frame, last.frame], [kf.pose, last.pose], f.frame, f.pose, multi=True)
                kinds.append("stacked")
            else:                                        # Odometry.cpp:40-41
                res = O.align(cfg, last.frame, last.pose, f.frame, f.pose)
                kinds.append("single")
            pose = res["pose"]
            dT = (f.t - last.t) / 1e9
            speed = O.se3_log(O.se3_mul(pose, O.se3_inv(last.pose))) / dT
        else:
            pose, speed = f.pose.copy(), np.zeros(6)
            kinds.append("first")
        f.pose = pose                                    # frame->set(*_odometry->pose())
        pred.update(pose, f.t)
        sel.update(f)
        m.insert(f, sel.is_key_frame())
        poses.append(pose)
        speeds.append(speed)
    return np.array(poses), np.array(speeds), kinds
