"""CPU suite, part 1: pins the oracle (oracle/vslam_oracle.cpp) against the golden vectors and the known-answer
structures the reference's own tests hold for the path.  No GPU, no /root/reference at run time.

The OpenCV goldens (tests/golden/cv_ops.npz) come from tests/golden/make_golden.py."""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv_ops.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def _names(g):
    return sorted({k.split("/")[0] for k in g.files if "/" in k and not k.startswith("rgb_")})


# ---------------------------------------------------------------------------------------------------------
# image operations the reference delegates to OpenCV (core/src/Frame.cpp:225-228, 264): bit-exact
# ---------------------------------------------------------------------------------------------------------
def test_pyr_down_matches_opencv_golden(oracle, golden):
    checked = 0
    for n in _names(golden):
        cur, l = golden[n + "/img"], 1
        while f"{n}/pyr{l}" in golden.files:
            cur = oracle.pyr_down_u8(cur)
            want = golden[f"{n}/pyr{l}"]
            assert cur.shape == want.shape, (n, l)
            assert np.array_equal(cur, want), (n, l)
            l += 1
            checked += 1
    assert checked >= 12


def test_blur_and_sobel_match_opencv_golden(oracle, golden):
    for n in _names(golden):
        blur = oracle.blur3_u8(golden[n + "/img"])
        assert np.array_equal(blur, golden[n + "/blur"]), n
        dx, dy = oracle.sobel3_s16(blur)
        assert dx.dtype == np.int16
        assert np.array_equal(dx, golden[n + "/dx"]), n
        assert np.array_equal(dy, golden[n + "/dy"]), n


def test_rgb_to_gray_matches_opencv_golden(oracle, golden):
    """cv::cvtColor(COLOR_RGB2GRAY) of the node's ingestion (NodeRgbdAlignment.cpp:182), both channel orders."""
    for name in ("rgb_noise_56x40", "rgb_grid"):
        img = golden[name + "/img"]
        assert np.array_equal(oracle.rgb_to_gray(img), golden[name + "/gray_rgb"]), name
        assert np.array_equal(oracle.rgb_to_gray(img, bgr=True), golden[name + "/gray_bgr"]), name


def test_depth_ingest_zeroes_nan_and_applies_the_dataset_factor(oracle):
    """depth.array().isNaN().select(0, depth) (NodeRgbdAlignment.cpp:190); loadDepth / 5000.0 (test_alignment_se3.cpp:99)."""
    d = np.array([[5000.0, np.nan, 0.0], [12345.0, 65535.0, np.inf]])
    out = oracle.depth_ingest(d, 5000.0)
    assert out[0, 0] == 1.0 and out[0, 1] == 0.0 and out[0, 2] == 0.0
    assert out[1, 0] == 12345.0 / 5000.0 and out[1, 1] == 65535.0 / 5000.0 and np.isinf(out[1, 2])
    assert np.array_equal(oracle.depth_ingest(np.array([1.5, np.nan])), np.array([1.5, 0.0]))


def test_frame_pyramid_uses_the_same_ops(oracle, golden):
    """Frame (ctor + computePyramid + computeDerivatives) must chain exactly those operations per level."""
    img = golden["noise_64x48/img"]
    depth = np.full(img.shape, 2.0)
    f = oracle.Frame(img, depth, 50.0, 50.0, 32.0, 24.0, 3)
    cur = img
    for l in range(3):
        if l:
            cur = oracle.pyr_down_u8(cur)
        lv = f.level(l)
        assert np.array_equal(lv["I"], cur)
        dx, dy = oracle.sobel3_s16(oracle.blur3_u8(cur))
        assert np.array_equal(lv["dIx"], dx) and np.array_equal(lv["dIy"], dy)
        assert np.allclose(lv["cam"], np.array([50.0, 50.0, 32.0, 24.0]) * 0.5 ** l, rtol=0, atol=0)


# ---------------------------------------------------------------------------------------------------------
# known-answer tests of the reference (core/test/test_algorithm.cpp, core/test/test_camera.cpp)
# ---------------------------------------------------------------------------------------------------------
def test_bilinear_known_answer(oracle):
    # core/test/test_algorithm.cpp:27-33 : (255 + 128) / 2 == 191 after the uint8 cast
    m = np.array([[128, 128, 128], [255, 255, 255], [255, 255, 255]], np.uint8)
    assert int(np.uint8(oracle.bilinear(m, 0.5, 0.5))) == (255 + 128) // 2
    assert oracle.bilinear(m, 0.0, 0.0) == 128 and oracle.bilinear(m, 1.0, 1.0) == 255


def test_resize_known_answer(oracle):
    # core/test/test_algorithm.cpp:57-68: algorithm::resize(m, 0.5) of a 4x4 picks m(2i, 2j) through the bilinear
    # sampler at integer coordinates: (0,0) == 128 and (1,1) == 255
    m = np.array([[128, 128, 128, 128], [128, 128, 255, 255], [255, 128, 255, 255], [255, 255, 255, 255]], np.uint8)
    res = np.array([[oracle.bilinear(m, 2.0 * j, 2.0 * i) for j in range(2)] for i in range(2)])
    assert res.shape == (2, 2)
    assert res[0, 0] == 128 and res[1, 1] == 255
    assert np.array_equal(res, m[::2, ::2])


def test_masked_median_quirks(oracle):
    # core/src/algorithm.cpp:109-120 median(): odd count -> middle; even count -> (v[n/2-1] + v[n/2+1]) / 2, i.e. NOT
    # the two middle values (reference quirk, kept); n == 2 reads past the end there
    assert oracle.median([3.0, 1.0, 2.0]) == 2.0
    assert oracle.median([4.0, 1.0, 3.0, 2.0]) == (2.0 + 4.0) / 2
    assert oracle.median([6.0, 5.0, 4.0, 3.0, 2.0, 1.0]) == (3.0 + 5.0) / 2
    assert oracle.median([5.0]) == 5.0
    assert oracle.median([7.0, 3.0]) == 3.0  # documented rule for the reference's out-of-bounds n == 2 read


def test_depth_down_ignores_invalid(oracle):
    # Frame.cpp:270-274: medianBlur(3x3, v <= 0 masked) then resize(0.5).  medianBlur never writes its one-pixel
    # border (algorithm.h:157-184), so row 0 / column 0 of every coarser depth level are 0 == invalid.
    d = np.full((12, 12), 2.0)
    d[5:10, 5:10] = 0.0  # a hole: fully masked 3x3 windows stay 0, partially masked ones take the valid median
    out, _ = oracle.depth_down(d)
    want = np.full((6, 6), 2.0)
    want[0, :] = 0.0
    want[:, 0] = 0.0
    want[3:5, 3:5] = 0.0  # blur(6,6), (6,8), (8,6), (8,8): windows entirely inside the hole
    assert np.array_equal(out, want)
    out2, _ = oracle.depth_down(np.zeros((6, 6)))
    assert np.all(out2 == 0.0)
    # odd sizes: int(H * 0.5) x int(W * 0.5) (algorithm.h:93-94), unlike the (W + 1) / 2 of cv::pyrDown
    out3, _ = oracle.depth_down(np.full((9, 11), 1.5))
    assert out3.shape == (4, 5)


def test_camera_round_trip(oracle):
    # core/test/test_camera.cpp:59-92 structure: project, compare with the closed form, back-project
    import ctypes as C
    lib = oracle.lib()
    lib.vo_camera2image.restype = C.c_int
    f, cx, cy = 10.0, 320.0, 240.0
    rng = np.random.default_rng(5)
    for _ in range(50):
        p = np.array([rng.uniform(-3, 3), rng.uniform(-3, 3), rng.uniform(0.5, 9)])
        uv = np.zeros(2)
        ok = lib.vo_camera2image(C.c_double(f), C.c_double(f), C.c_double(cx), C.c_double(cy), oracle._p(p), oracle._p(uv))
        assert ok
        assert abs(uv[0] - (f * p[0] + cx * p[2]) / p[2]) < 1e-9
        assert abs(uv[1] - (f * p[1] + cy * p[2]) / p[2]) < 1e-9
    # integer pixel back-projection round trip (image2camera is what computePcl uses, Frame.cpp:241-252)
    for (u, v, d) in [(0, 0, 1.0), (319, 17, 2.5), (639, 479, 7.25)]:
        p = np.zeros(3)
        lib.vo_image2camera(C.c_double(f), C.c_double(f), C.c_double(cx), C.c_double(cy), u, v, C.c_double(d), oracle._p(p))
        assert np.allclose(p, [(u - cx) / f * d, (v - cy) / f * d, d], rtol=1e-15)
        uv = np.zeros(2)
        assert lib.vo_camera2image(C.c_double(f), C.c_double(f), C.c_double(cx), C.c_double(cy), oracle._p(p), oracle._p(uv))
        assert np.allclose(uv, [u, v], atol=1e-9)


def test_camera_projecting_invalid_is_nan(oracle):
    # core/test/test_camera.cpp:94-111: z <= 0 projects to NaN
    import ctypes as C
    lib = oracle.lib()
    for z in (0.0, -1.0):
        uv = np.zeros(2)
        p = np.array([1.0, 1.0, z])
        ok = lib.vo_camera2image(C.c_double(10), C.c_double(10), C.c_double(320), C.c_double(240), oracle._p(p), oracle._p(uv))
        assert not ok
        assert np.isnan(uv).all()


# ---------------------------------------------------------------------------------------------------------
# the one real RGB-D frame the reference ships (odometry/test/resource/rgb.png, depth.png; crops of it in
# tests/golden/tum_frame.npz, made by tests/golden/make_tum_frame.py): real texture, 40 % invalid depth
# ---------------------------------------------------------------------------------------------------------
TUM_FRAME = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tum_frame.npz")
TUM_CAM = (525.0 / 2, 525.0 / 2, 159.5, 119.5)  # test_alignment_se3.cpp:69 scaled to the 320x240 crop


def test_real_frame_pyramid_and_gradients_match_opencv(oracle):
    g = np.load(TUM_FRAME)
    depth = g["depth0"].astype(np.float64) / 5000.0  # test_alignment_se3.cpp:99
    f = oracle.Frame(g["gray0"], depth, *TUM_CAM, 3)
    for l in range(3):
        lv = f.level(l)
        if l:
            assert np.array_equal(lv["I"], g[f"pyr{l}"]), l
        assert np.array_equal(lv["dIx"], g[f"dx{l}"]) and np.array_equal(lv["dIy"], g[f"dy{l}"]), l
        # the depth pyramid keeps holes as zeros and never invents depth where the 3x3 window holds none
        Z = lv["Z"]
        assert np.isfinite(Z).all() and (Z >= 0).all()
    Z0, Z1 = f.level(0)["Z"], f.level(1)["Z"]
    assert 0.3 < (Z0 == 0).mean() < 0.5 and (Z1 == 0).mean() > 0.25
    # a level-1 pixel is the masked median of the 3x3 level-0 window around (2i, 2j): it lies within the window's range
    for (i, j) in [(10, 10), (40, 100), (100, 30), (77, 141)]:
        w = Z0[2 * i - 1:2 * i + 2, 2 * j - 1:2 * j + 2]
        valid = w[w > 0]
        if valid.size == 0:
            assert Z1[i, j] == 0
        else:
            assert valid.min() - 1e-12 <= Z1[i, j] <= valid.max() + 1e-12


def test_real_frame_pair_aligns_on_the_oracle(oracle):
    """Two crops of the real frame offset by (3, 2) pixels: the aligner must run to a regular stop on real imagery with
    depth holes (this is the pair the GPU parity test repeats)."""
    O = oracle
    g = np.load(TUM_FRAME)
    fr = [O.Frame(g[f"gray{k}"], g[f"depth{k}"].astype(np.float64) / 5000.0, *TUM_CAM, 3) for k in range(2)]
    ref = O.align(O.make_config(max_iterations=50), fr[0], O.IDENTITY, fr[1], O.IDENTITY)
    assert ref["status"] == 0 and sum(ref["iters"]) >= 3
    xi = O.se3_log(ref["pose"])
    assert np.all(np.isfinite(xi)) and 1e-4 < np.linalg.norm(xi) < 0.2  # the crops differ: the pose moved, sanely
