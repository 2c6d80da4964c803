"""Builds oracle/_ref/libvslam_ref.so: the REFERENCE's own alignment path, compiled from the sources where they lie under
/root/reference against the small stand-in headers of oracle/ref_shim/ (Eigen / Sophus / OpenCV / logging surface; the
real libraries are not in this image).  Test infrastructure: tests/test_oracle_ref_cpu.py pins oracle/vslam_oracle.cpp
against it.

Compiled UNCHANGED, straight from the reference tree:
  least_squares/src  Scaler.cpp Loss.cpp NormalEquations.cpp GaussNewton.cpp
  core/src           Camera.cpp PoseWithCovariance.cpp Frame.cpp Feature2D.cpp Point3D.cpp
  lukas_kanade/src   Warp.cpp InverseCompositional.cpp InverseCompositionalStacked.cpp
  odometry/src/direct_image_alignment  SE3Alignment.cpp
  odometry/src       Odometry.cpp MotionPrediction.cpp KeyFrameSelection.cpp mapping/Map.cpp   (the aligner's caller)
core/src/algorithm.cpp also holds eigenvalue / quaternion helpers that need Eigen modules the stand-in does not have, so
only the five functions the path calls (median x2, resize x2, computeRelativeTransform) are cut out of it, by signature, into a temporary file
that is compiled and deleted: no reference source is ever written into the repository.

The GPU box has no /root/reference: there the prebuilt .so (git-ignored, not gpurun-ignored) is used as it is.
"""
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("VSLAM_REFERENCE", "/root/reference/src/vslam/src")
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "libvslam_ref.so")
ORACLE_SO = os.path.join(HERE, "_build", "libvslam_oracle.so")
UNCHANGED = [
    "least_squares/src/Scaler.cpp", "least_squares/src/Loss.cpp", "least_squares/src/NormalEquations.cpp",
    "least_squares/src/GaussNewton.cpp", "core/src/Camera.cpp", "core/src/PoseWithCovariance.cpp", "core/src/Frame.cpp",
    "core/src/Feature2D.cpp", "core/src/Point3D.cpp", "lukas_kanade/src/Warp.cpp",
    "lukas_kanade/src/InverseCompositional.cpp", "lukas_kanade/src/InverseCompositionalStacked.cpp",
    "odometry/src/direct_image_alignment/SE3Alignment.cpp",
    # the aligner's direct caller: odometry update, motion prediction, key-frame selection, map rings
    "odometry/src/Odometry.cpp", "odometry/src/MotionPrediction.cpp", "odometry/src/KeyFrameSelection.cpp",
    "odometry/src/mapping/Map.cpp",
]
ALGORITHM_FUNCTIONS = [  # signatures inside namespace pd::vslam::algorithm of core/src/algorithm.cpp
    r"Image resize\(const Image & mat, double scale\)",
    r"Eigen::MatrixXd resize\(const Eigen::MatrixXd & mat, double scale\)",
    r"double median\(const Eigen::VectorXd & d, bool isSorted\)",
    r"double median\(std::vector<double> & v, bool isSorted\)",
    r"Sophus::SE3d computeRelativeTransform\(const Sophus::SE3d & t0, const Sophus::SE3d & t1\)",
]
INCLUDES = ["core/src", "core/include", "least_squares/src", "least_squares/include", "lukas_kanade/src",
            "lukas_kanade/include", "odometry/src/direct_image_alignment", "odometry/src", "kalman/include", "kalman/src"]
# -DNDEBUG: the reference's Dockerfile builds Release; -ffp-contract=off: no fused multiply-adds, like the oracle
CXXFLAGS = ["-std=c++17", "-O2", "-DNDEBUG", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-w"]


def available():
    return os.path.isdir(REF)


def _cut_function(text, signature):
    m = re.search(signature, text)
    if not m:
        raise RuntimeError("build_ref: %s not found in algorithm.cpp" % signature)
    i = text.index("{", m.end())
    depth, j = 0, i
    while True:
        depth += {"{": 1, "}": -1}.get(text[j], 0)
        j += 1
        if depth == 0:
            return text[m.start():j]


def _algorithm_subset():
    text = open(os.path.join(REF, "core/src/algorithm.cpp")).read()
    body = "\n".join(_cut_function(text, s) for s in ALGORITHM_FUNCTIONS)
    return '#include "algorithm.h"\nnamespace pd::vslam::algorithm\n{\n%s\n}\n' % body


def sources_newer_than(t):
    mine = [os.path.join(HERE, "ref_harness.cpp"), os.path.abspath(__file__), ORACLE_SO]
    for root, _, files in os.walk(os.path.join(HERE, "ref_shim")):
        mine += [os.path.join(root, f) for f in files]
    return any(os.path.getmtime(p) > t for p in mine if os.path.exists(p))


def build(force=False):
    """-> path of the library, or None when neither the reference tree nor a prebuilt library is there."""
    if not available():
        return OUT if os.path.exists(OUT) else None
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", HERE], stdout=subprocess.DEVNULL)
    if not force and os.path.exists(OUT) and not sources_newer_than(os.path.getmtime(OUT)):
        return OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    inc = ["-I", os.path.join(HERE, "ref_shim")]
    for d in INCLUDES:
        inc += ["-I", os.path.join(REF, d)]
    with tempfile.TemporaryDirectory() as tmp:
        subset = os.path.join(tmp, "algorithm_subset.cpp")
        with open(subset, "w") as f:
            f.write(_algorithm_subset())
        objs = []
        jobs = [(os.path.join(REF, s), s.replace("/", "_")) for s in UNCHANGED]
        jobs += [(subset, "algorithm_subset.cpp"), (os.path.join(HERE, "ref_harness.cpp"), "ref_harness.cpp")]
        procs = []
        for src, name in jobs:
            obj = os.path.join(tmp, name + ".o")
            objs.append(obj)
            procs.append((src, subprocess.Popen(["g++"] + CXXFLAGS + inc + ["-c", src, "-o", obj])))
        for src, p in procs:
            if p.wait() != 0:
                raise RuntimeError("build_ref: compiling %s failed" % src)
        subprocess.check_call(["g++", "-shared", "-o", OUT] + objs + [
            "-L", os.path.dirname(ORACLE_SO), "-lvslam_oracle", "-Wl,-rpath,$ORIGIN/../_build", "-Wl,--no-undefined"])
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
