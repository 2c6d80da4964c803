"""CPU suite, part 2: the C-ABI library (loads, exports every declared symbol, fails loudly without a GPU), the
ctypes mirror of its structs, and the host-side sharding logic incl. a world_size-2 gloo run.  No compute calls."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vslam_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vslamb200_[a-z_0-9]+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    syms = _declared_symbols()
    for s in ("vslamb200_create", "vslamb200_destroy", "vslamb200_frames_upload", "vslamb200_frames_upload_device",
              "vslamb200_align", "vslamb200_evaluate", "vslamb200_get_stats", "vslamb200_last_error"):
        assert s in syms
    assert len(syms) >= 20


def test_library_loads_and_exports_every_declared_symbol(capi):
    assert os.path.exists(capi.LIB_PATH), "libvslam_b200.so missing: run `python -c 'import __graft_entry__ as g; g.build()'`"
    lib = C.CDLL(capi.LIB_PATH)
    missing = [s for s in _declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    lib.vslamb200_version.restype = C.c_int
    assert lib.vslamb200_version() >= 100


def test_library_is_built_for_sm_100a(capi):
    out = subprocess.run(["cuobjdump", "-lelf", capi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout, out.stdout[:400]


def test_ctypes_structs_match_the_header_layout(capi, tmp_path):
    """The header compiles as plain C, and the ctypes mirrors have the sizes / key offsets the C compiler gives."""
    src = tmp_path / "sizes.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "vslam_b200.h"\n'
        'int main(void) { printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(vslamb200_config), sizeof(vslamb200_job), '
        'sizeof(vslamb200_result), sizeof(vslamb200_stats), offsetof(vslamb200_result, iters), '
        'offsetof(vslamb200_job, from_poses), offsetof(vslamb200_stats, phase_clocks)); return 0; }\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(capi.Config), C.sizeof(capi.Job), C.sizeof(capi.Result), C.sizeof(capi.Stats),
            capi.Result.iters.offset, capi.Job.from_poses.offset, capi.Stats.phase_clocks.offset]
    assert got == want


def test_create_fails_loudly_without_a_gpu(capi):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.VslamB200Error) as e:
        capi.Context(capi.make_config())
    assert "no CPU fallback" in str(e.value)


def test_bad_arguments_are_rejected_before_any_device_work(capi):
    lib = capi.lib()
    out = C.c_void_p()
    assert lib.vslamb200_create(None, 0, C.byref(out)) == -1
    cfg = capi.make_config()
    cfg.max_iterations = 0
    assert lib.vslamb200_create(C.byref(cfg), 0, C.byref(out)) == -1
    cfg = capi.make_config()
    cfg.loss_kind = 99
    assert lib.vslamb200_create(C.byref(cfg), 0, C.byref(out)) == -1
    assert lib.vslamb200_align(None, None, 0, None) < 0
    assert lib.vslamb200_get_stats(None, None) < 0
    lib.vslamb200_destroy(None)  # no-op


# ---------------------------------------------------------------------------------------------------------
# sharding of a sequence over ranks
# ---------------------------------------------------------------------------------------------------------
def test_pair_ranges_partition_the_sequence():
    from vslam_b200 import shard
    for n in (0, 1, 7, 8, 999, 1000):
        for world in (1, 2, 3, 8):
            got = [shard.pair_range(n, world, r) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == n
            for a, b in zip(got, got[1:]):
                assert a[1] == b[0]
            sizes = [b - a for a, b in got]
            assert max(sizes) - min(sizes) <= 1
            for r in range(world):
                lo, hi = shard.frame_range(n, world, r)
                assert hi - lo == (sizes[r] + 1 if sizes[r] else 0)
    with pytest.raises(ValueError):
        shard.pair_range(10, 2, 2)


def test_dealt_pairs_partition_the_batch():
    from vslam_b200 import shard
    for n in (0, 1, 7, 8, 4096):
        for world in (1, 2, 3, 8):
            got = [shard.deal_pairs(n, world, r) for r in range(world)]
            assert sorted(np.concatenate(got).tolist()) == list(range(n))
            assert max(len(g) for g in got) - min(len(g) for g in got) <= 1
    with pytest.raises(ValueError):
        shard.deal_pairs(10, 2, -1)


def test_oracle_batch_of_independent_pairs_matches_single_alignments(oracle, synth):
    """vo_baseline_batch (the CPU arm of bench.py) against oracle.align pair by pair, stride 2 and stride 1."""
    O = oracle
    W, H, L = 96, 64, 2
    pairs = [synth.make_pair(W, H, 1234 + k) for k in range(3)]
    gray = np.stack([x for p in pairs for x in (p["gray0"], p["gray1"])])
    depth = np.stack([x for p in pairs for x in (p["depth0"], p["depth1"])])
    cam = pairs[0]["cam"]
    cfg = O.make_config(max_iterations=20)
    poses, iters = O.baseline_batch(cfg, gray, depth, *cam, L, n_threads=2, stride=2)
    assert poses.shape == (3, 7) and iters.shape == (3, L)
    for k, p in enumerate(pairs):
        f0 = O.Frame(p["gray0"], p["depth0"].astype(np.float64), *cam, L)
        f1 = O.Frame(p["gray1"], p["depth1"].astype(np.float64), *cam, L)
        ref = O.align(cfg, f0, O.IDENTITY, f1, O.IDENTITY)
        assert np.array_equal(poses[k], ref["pose"]) and list(iters[k]) == list(ref["iters"])
    seq, _ = O.baseline_sequence(cfg, gray, depth, *cam, L, 1)
    p1, _ = O.baseline_batch(cfg, gray, depth, *cam, L, n_threads=1, stride=1)
    assert np.array_equal(seq, p1)


def test_chain_trajectory_composes_relative_motions(oracle):
    from vslam_b200 import shard
    rng = np.random.default_rng(3)
    rel = np.stack([oracle.se3_exp(rng.normal(0, 0.05, 6)) for _ in range(6)])
    traj = shard.chain_trajectory(rel)
    cur = oracle.IDENTITY
    for i, r in enumerate(rel):
        cur = oracle.se3_mul(r, cur)
        d = oracle.se3_log(oracle.se3_mul(traj[i + 1], oracle.se3_inv(cur)))
        assert np.linalg.norm(d) < 1e-12


_GLOO_WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from vslam_b200 import shard
rank, world, n_pairs = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(sys.argv[2])
dist.init_process_group("gloo", rank=rank, world_size=world)
def fake_pose(i):  # stands in for the aligned pose of pair i
    return np.array([0.001 * i, -0.002 * i, 0.0005 * i, 1.0, 0.01 * i, 0.02 * i, -0.03 * i])
lo, hi = shard.pair_range(n_pairs, world, rank)
local = np.stack([fake_pose(i) for i in range(lo, hi)]) if hi > lo else np.zeros((0, 7))
allp = shard.gather_poses(local, n_pairs, dist)
want = np.stack([fake_pose(i) for i in range(n_pairs)])
assert allp.shape == (n_pairs, 7) and np.array_equal(allp, want), rank
# independent pairs dealt round-robin (BASELINE configs[2])
mine = shard.deal_pairs(n_pairs, world, rank)
local = np.stack([fake_pose(i) for i in mine]) if len(mine) else np.zeros((0, 7))
allp = shard.gather_dealt(local, n_pairs, dist)
assert allp.shape == (n_pairs, 7) and np.array_equal(allp, want), rank
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok", lo, hi)
"""


@pytest.mark.parametrize("n_pairs", [9, 1])
def test_gather_poses_world_size_2_gloo(tmp_path, n_pairs):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    port = 29500 + (os.getpid() % 2000)
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT, str(n_pairs)], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=180)
        except subprocess.TimeoutExpired:
            p.kill()
            o, _ = p.communicate()
        outs.append(o)
    for rank, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o[-2000:]
        assert f"rank {rank} ok" in o


# ---------------------------------------------------------------------------------------------------------
# C++ mirror of the reference interface (include/vslam_b200.hpp)
# ---------------------------------------------------------------------------------------------------------
def test_cpp_host_mirror_builds_and_fails_loudly_without_a_gpu(capi):
    """-Wall -Wextra -Werror build of the header-only mirror + link against the C ABI; without a device the aligner
    constructor throws std::runtime_error (what OdometryRgbd catches, odometry/src/Odometry.cpp:52-56)."""
    import torch
    from vslam_b200 import build
    exe = build.build_host_mirror()
    assert os.path.exists(exe)
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: covered by the gpu test")
    r = subprocess.run([exe, "160", "120", "5", "2"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3, (r.returncode, r.stderr)
    assert "no CPU fallback" in r.stderr


def test_odometry_host_mirror_algebra_rings_prediction_and_trajectory_files(oracle, tmp_path):
    """include/vslam_b200_odometry.hpp without a GPU (tests/cpp/host_algebra.cpp): SE3 closed forms against the
    oracle's, Map rings (Map.cpp:19-33), KeyFrameSelectionIdx (KeyFrameSelection.h:36-51), constant-motion prediction
    (MotionPrediction.cpp:38-55) against oracle/odometry.py, TUM trajectory files (utils.cpp:76-117,134-163)."""
    import subprocess
    from vslam_b200 import build
    from oracle import odometry as OD
    O = oracle
    exe = build.build_host_algebra()
    gt = tmp_path / "groundtruth.txt"
    gt.write_text("# ground truth trajectory\n# timestamp tx ty tz qx qy qz qw\n"
                  "1305031102.1758 1.3405 0.6266 1.6575 0.6574 0.6126 -0.2949 -0.3248\n"
                  "1305031102.2758 1.5 -0.25 0.125 0 0 0 1\n")
    out = tmp_path / "algo.txt"
    x = np.array([0.01, -0.02, 0.03, 0.1, -0.2, 0.05])
    y = np.array([-0.02, 0.01, 0.0, 0.0, 0.02, -0.01])
    r = subprocess.run([exe, str(gt), str(out)] + [repr(float(v)) for v in np.concatenate([x, y])],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    rows = [ln.split() for ln in r.stdout.strip().splitlines()]
    val = {row[0]: np.array([float(v) for v in row[1:]]) for row in rows if row[0] != "stamp"}
    X, Y = O.se3_exp(x), O.se3_exp(y)
    assert np.allclose(val["expx"], X, atol=1e-14)
    assert np.allclose(val["mul"], O.se3_mul(X, Y), atol=1e-14)
    assert np.allclose(val["inv"], O.se3_inv(X), atol=1e-14)
    assert np.allclose(val["rel"], O.se3_mul(Y, O.se3_inv(X)), atol=1e-14)     # algorithm.cpp:82-85: t1 * t0^-1
    assert np.allclose(val["log"], O.se3_log(O.se3_mul(X, Y)), atol=1e-13)
    assert list(val["kf"]) == [0, 0, 1] * 4                                      # period 3, counter starts at 1
    assert list(val["rings"]) == [7, 4, 11, 11]                                 # 7 + 7 rings, newest first
    pred = OD.MotionPredictionConstant()
    pred.update(X, 1000000000)
    pred.update(O.se3_mul(Y, X), 1500000000)
    assert np.allclose(val["predict"], pred.predict(2500000000), atol=1e-13)
    assert val["older"][0] == 1                                                  # "New timestamp is older than last one!"
    stamps = [row for row in rows if row[0] == "stamp"]
    assert [int(row[1]) for row in stamps] == [int(1305031102 * 1e9 + 1758 * 100000), int(1305031102 * 1e9 + 2758 * 100000)]
    q = np.array([0.6574, 0.6126, -0.2949, -0.3248])
    assert np.allclose([float(v) for v in stamps[0][2:6]], q / np.linalg.norm(q), atol=1e-15)
    assert [float(v) for v in stamps[1][6:9]] == [1.5, -0.25, 0.125]
    text = out.read_text().splitlines()
    assert text[0] == "# Algorithm Trajectory" and text[2] == "# timestamp tx ty tz qx qy qz qw"
    assert text[4].split()[1:] == ["1.5", "-0.25", "0.125", "0", "0", "0", "1"]


def test_least_squares_mirror_scalers_losses_normal_equations_and_gauss_newton(oracle):
    """The least-squares half of include/vslam_b200.hpp on the host (tests/cpp/host_least_squares.cpp): scalers
    (Scaler.cpp:21-67) and loss weights (Loss.cpp:23-100) against independent numpy restatements, the reference's
    constructor signatures and defaults (Loss.h:70,87,101), NormalEquations (NormalEquations.cpp:53-76) against J^T W J,
    the 6x6 helpers against numpy.linalg, and GaussNewton::solve (GaussNewton.cpp:33-102) on a linear problem."""
    import subprocess
    from vslam_b200 import build, capi as A
    exe = build.build_host_least_squares()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-500:] + r.stderr[-500:]
    rows = [ln.split() for ln in r.stdout.strip().splitlines()]
    first = {}
    for row in rows:
        first.setdefault(row[0], row)
    res = np.array([((k * 37) % 101 - 50) * 0.37 for k in range(200)])
    sc = np.array([float(v) for v in first["scale"][1:]])
    srt = np.sort(res)
    med = (srt[99] + srt[101]) / 2            # the reference's even-count rule: v[n/2-1], v[n/2+1]
    assert list(sc[:2]) == [0.0, 1.0]
    assert np.allclose(sc[2:4], [med, np.sqrt(np.abs(res - med).sum() / 199)], rtol=1e-14)
    assert np.allclose(sc[4:6], [res.mean(), np.sqrt(np.abs(res - res.mean()).sum() / 199)], rtol=1e-13, atol=1e-15)
    sigma = 1.0
    for _ in range(30):
        nxt = np.sqrt(np.mean(res * res * 6.0 / (5.0 + (res / sigma) ** 2)))
        step, sigma = abs(sigma - nxt), nxt
        if not step > 1e-5:
            break
    assert sc[6] == 0.0 and abs(sc[7] - sigma) < 1e-12 * sigma
    assert [float(first["median4"][i]) for i in (1, 3, 5)] == [3.0, 5.0, 3.0]  # (1+3)/2... v[1], v[3] of 1 2 3 4 -> 3
    kinds = [int(v) for v in first["kinds"][1:8]]
    scalers = [int(v) for v in first["kinds"][9:16]]
    assert kinds == [A.LOSS_TUKEY, A.LOSS_HUBER, A.LOSS_HUBER, A.LOSS_HUBER, A.LOSS_TDIST, A.LOSS_TDIST, A.LOSS_QUADRATIC]
    # defaults of the reference: Tukey / Huber / LossTDistribution default to the MedianScaler (Loss.h:70,87,101)
    assert scalers == [A.SCALER_MEDIAN, A.SCALER_MEDIAN, A.SCALER_MEAN, A.SCALER_MEDIAN, A.SCALER_MEDIAN, A.SCALER_TDIST,
                       A.SCALER_IDENTITY]
    assert float(first["kinds"][17]) == float(np.float32(1.345)) and float(first["kinds"][18]) == 10.0
    for row in (rw for rw in rows if rw[0] == "w"):
        x, wt, wh, wd, wq, rt, rh, rq = [float(v) for v in row[1:] if v != "rho"]
        C = 4.6851
        assert abs(wt - ((1 - (x / C) ** 2) ** 2 if abs(x) < C else 0.0)) < 1e-15
        assert abs(wh - (1.0 if abs(x) < 10 else 1.0 / abs(x))) < 1e-15       # 1/|r|, not c/|r| (Loss.cpp:65-72)
        assert abs(wd - 6.0 / (5.0 + x * x)) < 1e-15 and wq == 1.0 and rq == 0.5 * x * x
        assert abs(rt - (C / 6 * (1 - (1 - (x / C) ** 2) ** 3) if abs(x) < C else C / 6)) < 1e-15
        assert abs(rh - (0.5 * x * x if abs(x) < 10 else 10 * abs(x) - 0.5 * x * x)) < 1e-12
    assert [float(v) for v in first["lossscale"][1:]] == [sc[6], sc[7]]
    k = np.arange(40)
    J = np.sin(0.7 * k[:, None] + 1.3 * np.arange(6)[None, :]) + 1.5 * (np.arange(6)[None, :] == (k % 6)[:, None])
    y = np.cos(0.31 * k) * 2.0
    w = 0.5 + (k % 5) * 0.25
    ne = np.array([float(v) for v in first["ne"][1:]])
    # NormalEquations.cpp:58 as the reference's Release build evaluates it: w[0] * sum r^2 (vslam_b200.h, CHI2_REFERENCE)
    assert ne[0] == 40 and abs(ne[1] - w[0] * (y * y).sum()) < 1e-12
    assert np.allclose(ne[2:8], J.T @ (w * y), rtol=1e-13, atol=1e-13)
    Aw = J.T @ (w[:, None] * J)
    assert np.allclose(ne[8:44].reshape(6, 6), Aw, rtol=1e-13, atol=1e-13)
    cb = [float(v) for v in first["combine"][1:]]
    # the second part was built with addConstraint, which accumulates r^T W r (NormalEquations.cpp:66)
    assert cb[0] == 80 and abs(cb[1] - (ne[1] + (w * y * y).sum())) < 1e-12 and abs(cb[2] - cb[3]) < 1e-12
    assert abs(float(first["det"][1]) - np.linalg.det(Aw)) < 1e-9 * abs(np.linalg.det(Aw))
    assert np.allclose([float(v) for v in first["solve"][1:]], np.linalg.solve(Aw, J.T @ (w * y)), rtol=1e-11)
    gn = [float(v) for v in first["gn"][1:]]
    assert gn[0] == 2 and gn[1] == 1 and gn[2] == A.STOP_NAMES.index("converged")
    assert abs(gn[3] - (y * y).sum()) < 1e-12                                  # chi2 at x = 0
    assert np.allclose(gn[4:10], np.linalg.lstsq(J, y, rcond=None)[0], rtol=1e-10, atol=1e-12)
    assert first["throws"][1:] == ["1", "1"]
