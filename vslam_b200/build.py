"""In-tree build of the CUDA library (sm_100a only) and the synthetic-scene helper."""
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIBDIR = os.path.join(_HERE, "lib")
LIB = os.path.join(LIBDIR, "libvslam_b200.so")
SOURCES = ["preprocess.cu", "lists.cu", "level_misc.cu", "gn_level.cu", "gn_mom.cu", "ext.cu", "noalloc.cu", "api.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++",
]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_cuda(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, h) for h in os.listdir(CSRC) if h.endswith((".cuh", ".h"))]
    headers.append(os.path.join(_HERE, "..", "include", "vslam_b200.h"))
    objs, jobs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            jobs.append([NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj])
    if jobs:
        with ThreadPoolExecutor(max_workers=len(jobs)) as ex:
            for r in ex.map(lambda c: subprocess.run(c, capture_output=True, text=True), jobs):
                if verbose and r.stderr:
                    print(r.stderr)
                if r.returncode != 0:
                    raise RuntimeError("nvcc failed:\n" + " ".join(r.args) + "\n" + r.stderr)
    if jobs or force or _stale(LIB, objs):
        subprocess.check_call([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-ccbin", "/usr/bin/g++"])
    return LIB


HOST_MIRROR = os.path.join(LIBDIR, "host_mirror")


def build_host_mirror(force=False):
    """g++ build of tests/cpp/host_mirror.cpp: the C++ mirror of the reference interface (include/vslam_b200.hpp)
    linked against the C-ABI library and the synthetic-scene helper."""
    root = os.path.join(_HERE, "..")
    src = os.path.join(root, "tests", "cpp", "host_mirror.cpp")
    inc = os.path.join(root, "include")
    deps = [src, os.path.join(inc, "vslam_b200.hpp"), os.path.join(inc, "vslam_b200.h")]
    if force or _stale(HOST_MIRROR, deps):
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I", inc, src, "-o", HOST_MIRROR,
                               "-L", LIBDIR, "-lvslam_b200", "-lvslam_synth", "-Wl,-rpath,$ORIGIN", "-lpthread"])
    return HOST_MIRROR


ODOMETRY_MIRROR = os.path.join(LIBDIR, "odometry_mirror")


def build_odometry_mirror(force=False):
    """g++ build of tests/cpp/odometry_mirror.cpp: the C++ mirror of the odometry loop (include/vslam_b200_odometry.hpp)."""
    root = os.path.join(_HERE, "..")
    src = os.path.join(root, "tests", "cpp", "odometry_mirror.cpp")
    inc = os.path.join(root, "include")
    deps = [src] + [os.path.join(inc, h) for h in ("vslam_b200_odometry.hpp", "vslam_b200.hpp", "vslam_b200.h")]
    if force or _stale(ODOMETRY_MIRROR, deps):
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I", inc, src, "-o",
                               ODOMETRY_MIRROR, "-L", LIBDIR, "-lvslam_b200", "-lvslam_synth", "-Wl,-rpath,$ORIGIN",
                               "-lpthread"])
    return ODOMETRY_MIRROR


HOST_ALGEBRA = os.path.join(LIBDIR, "host_algebra")


def build_host_algebra(force=False):
    """g++ build of tests/cpp/host_algebra.cpp (host-only parts of include/vslam_b200_odometry.hpp; no GPU needed)."""
    root = os.path.join(_HERE, "..")
    src = os.path.join(root, "tests", "cpp", "host_algebra.cpp")
    inc = os.path.join(root, "include")
    deps = [src] + [os.path.join(inc, h) for h in ("vslam_b200_odometry.hpp", "vslam_b200.hpp", "vslam_b200.h")]
    if force or _stale(HOST_ALGEBRA, deps):
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I", inc, src, "-o",
                               HOST_ALGEBRA, "-L", LIBDIR, "-lvslam_b200", "-Wl,-rpath,$ORIGIN", "-lpthread"])
    return HOST_ALGEBRA


HOST_LEAST_SQUARES = os.path.join(LIBDIR, "host_least_squares")


def build_host_least_squares(force=False):
    """g++ build of tests/cpp/host_least_squares.cpp (least-squares half of include/vslam_b200.hpp; no GPU needed)."""
    root = os.path.join(_HERE, "..")
    src = os.path.join(root, "tests", "cpp", "host_least_squares.cpp")
    inc = os.path.join(root, "include")
    deps = [src] + [os.path.join(inc, h) for h in ("vslam_b200.hpp", "vslam_b200.h")]
    if force or _stale(HOST_LEAST_SQUARES, deps):
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I", inc, src, "-o",
                               HOST_LEAST_SQUARES, "-L", LIBDIR, "-lvslam_b200", "-Wl,-rpath,$ORIGIN", "-lpthread"])
    return HOST_LEAST_SQUARES


def build_all(force=False):
    from . import synth
    synth.build(force)
    lib = build_cuda(force)
    build_host_mirror(force)
    build_odometry_mirror(force)
    build_host_algebra(force)
    build_host_least_squares(force)
    return lib


if __name__ == "__main__":
    import sys
    print(build_all(force="--force" in sys.argv))
