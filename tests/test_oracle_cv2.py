"""Oracle primitives vs the real OpenCV (cv2 4.13.0, the only OpenCV in the image): bit-exact.

Pins the restated OpenCV arithmetic the reference relies on: resize INTER_LINEAR 8U,
copyMakeBorder REFLECT_101, FAST-9/16 + NMS + score, GaussianBlur 7x7 sigma 2, fastAtan2, cvRound.
"""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
from oracle import pyoracle as O

cv2.setNumThreads(1)


def _cvfast(img, t):
    f = cv2.FastFeatureDetector_create(threshold=t, nonmaxSuppression=True,
                                       type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    return np.array([[int(p.pt[0]), int(p.pt[1]), int(p.response)] for p in f.detect(img, None)],
                    np.int32).reshape(-1, 3)


@pytest.mark.parametrize("dst", [(627, 400), (522, 333), (376, 240), (700, 401), (751, 479)])
def test_resize_fixture(fixtures, dst):
    im = fixtures["EuRoC_l"]
    assert np.array_equal(O.resize_linear(im, *dst), cv2.resize(im, dst, interpolation=cv2.INTER_LINEAR))


def test_resize_random_cascade():
    rng = np.random.default_rng(1)
    im = rng.integers(0, 256, (480, 640), dtype=np.uint8)
    for (w, h) in [(533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]:
        a = O.resize_linear(im, w, h)
        assert np.array_equal(a, cv2.resize(im, (w, h), interpolation=cv2.INTER_LINEAR))
        im = a


def test_resize_exact_half_is_area():
    rng = np.random.default_rng(2)
    im = rng.integers(0, 256, (96, 128), dtype=np.uint8)
    assert np.array_equal(O.resize_linear(im, 64, 48), cv2.resize(im, (64, 48), interpolation=cv2.INTER_LINEAR))


def test_border(fixtures):
    im = fixtures["TUM_l"]
    assert np.array_equal(O.border_reflect101(im, 19), cv2.copyMakeBorder(im, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))


@pytest.mark.parametrize("shape", [(480, 752), (134, 210), (31, 45), (9, 9)])
def test_blur(shape):
    rng = np.random.default_rng(shape[0])
    for im in (rng.integers(0, 256, shape, dtype=np.uint8), np.full(shape, 255, np.uint8)):
        ref = cv2.GaussianBlur(im, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(O.gaussian_blur(im), ref)


def test_blur_fixture(fixtures):
    for n in ("EuRoC_l", "TUM_r"):
        im = fixtures[n]
        assert np.array_equal(O.gaussian_blur(im), cv2.GaussianBlur(im, (7, 7), 2, sigmaY=2,
                                                                     borderType=cv2.BORDER_REFLECT_101))


def test_fast_atan2():
    rng = np.random.default_rng(3)
    ys = rng.integers(-3_000_000, 3_000_000, 20000).astype(np.float32)
    xs = rng.integers(-3_000_000, 3_000_000, 20000).astype(np.float32)
    ys[:50] = 0; xs[50:100] = 0; xs[100:150] = ys[100:150]
    for y, x in zip(ys, xs):
        assert O.fast_atan2(y, x) == np.float32(cv2.fastAtan2(float(y), float(x)))


def test_cv_round_half_even():
    for v, e in [(0.5, 0), (1.5, 2), (2.5, 2), (-0.5, 0), (-1.5, -2), (626.6667, 627), (3.4999, 3)]:
        assert O.cv_round(v) == e
    assert O.cv_round_f(np.float32(752) * (np.float32(1) / np.float32(1.2))) == 627


@pytest.mark.parametrize("t", [20, 7])
def test_fast_fixture_and_cells(fixtures, t):
    im = fixtures["EuRoC_l"]
    assert np.array_equal(O.fast9(im, t), _cvfast(im, t))
    rng = np.random.default_rng(t)
    for _ in range(40):  # cell-sized crops incl. the degenerate narrow/short ones the cell loop can produce
        w, h = int(rng.integers(4, 44)), int(rng.integers(4, 44))
        x, y = int(rng.integers(0, 752 - w)), int(rng.integers(0, 480 - h))
        cell = np.ascontiguousarray(im[y:y + h, x:x + w])
        assert np.array_equal(O.fast9(cell, t), _cvfast(cell, t))


@pytest.mark.parametrize("t", [20, 7])
def test_fast_random_noise(t):
    rng = np.random.default_rng(10 + t)
    for lo, hi in [(0, 256), (100, 140), (0, 2)]:
        im = rng.integers(lo, hi, (40, 38), dtype=np.uint8)
        assert np.array_equal(O.fast9(im, t), _cvfast(im, t))
    # equal-score plateaus: strict NMS must drop both neighbours
    im = np.full((20, 20), 100, np.uint8); im[8:12, 8:12] = 200
    assert np.array_equal(O.fast9(im, t), _cvfast(im, t))


def test_match_bruteforce_oracle_vs_numpy():
    """pins the oracle's restatement of the matcher loop (src/ORBmatcher.cc:562-576) against an independent numpy
    formulation: nearest = argmin (first index), second = second smallest of the multiset, then both tests"""
    rng = np.random.default_rng(11)
    for n, m, ratio, th in [(300, 400, 0.6, 50), (50, 1, 0.9, 100), (64, 200, 0.75, 80)]:
        a = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        b = a[rng.integers(0, n, m)] ^ np.packbits(rng.random((m, 256)) < 0.1, axis=1)
        r = O.match_bruteforce(a, b, ratio, th, nthreads=3)
        d = np.unpackbits(a[:, None, :] ^ b[None, :, :], axis=2).sum(2).astype(np.int64)
        best = d.min(1); idx = d.argmin(1)
        second = np.sort(d, 1)[:, 1] if m > 1 else np.full(n, np.iinfo(np.int32).max)
        ok = (best <= th) & (best.astype(np.float32) < second.astype(np.float32) * np.float32(ratio))
        assert np.array_equal(r["best_idx"], idx) and np.array_equal(r["best_dist"], best)
        assert np.array_equal(r["second_dist"], second) and np.array_equal(r["match"], np.where(ok, idx, -1))


# ---- pre-processing (oracle/preprocess.cpp) against the genuine cv2.remap / cv2.cvtColor ----
def _maps_cases(w, h):
    from synth import rectify_maps
    rng = np.random.default_rng(5)
    yield rectify_maps(w, h)
    yield rectify_maps(w, h, k1=0.35, k2=-0.1, tilt=-0.05, shift=(-40.0, 25.5), scale=1.3)      # large out-of-source areas
    mx = (rng.random((h, w), dtype=np.float32) * (w + 40) - 20).astype(np.float32)               # random gather, all fractions
    my = (rng.random((h, w), dtype=np.float32) * (h + 40) - 20).astype(np.float32)
    mx[:4, :64] = (np.arange(64, dtype=np.float32) / 32 - 1.5); my[:4, :64] = np.float32(-0.5)  # 1/32 steps around the border
    mx[4:8, :64] = np.float32(w) - np.arange(64, dtype=np.float32) / 32; my[4:8, :64] = np.float32(h) - np.arange(64, dtype=np.float32) / 64
    yield mx, my


def test_preprocess_remap_gray8_vs_cv2(fixtures):
    im = fixtures["EuRoC_l"]
    h, w = im.shape
    for mx, my in _maps_cases(w, h):
        assert np.array_equal(O.preprocess(im, O.PIX_GRAY8, mx, my), cv2.remap(im, mx, my, cv2.INTER_LINEAR))
    mx, my = next(_maps_cases(320, 200))                       # destination smaller than the source
    assert np.array_equal(O.preprocess(im, O.PIX_GRAY8, mx, my), cv2.remap(im, mx, my, cv2.INTER_LINEAR))


def test_preprocess_colour_and_16bit_vs_cv2():
    rng = np.random.default_rng(6)
    h, w = 120, 168
    mx, my = next(_maps_cases(w, h))
    for fmt, code, ch in ((O.PIX_RGB8, cv2.COLOR_RGB2GRAY, 3), (O.PIX_BGR8, cv2.COLOR_BGR2GRAY, 3),
                          (O.PIX_RGBA8, cv2.COLOR_RGBA2GRAY, 4), (O.PIX_BGRA8, cv2.COLOR_BGRA2GRAY, 4)):
        im = rng.integers(0, 256, (h, w, ch), dtype=np.uint8)
        assert np.array_equal(O.preprocess(im, fmt), cv2.cvtColor(im, code))
        # reference order: remap in the caller (per channel), cvtColor inside Tracking
        assert np.array_equal(O.preprocess(im, fmt, mx, my), cv2.cvtColor(cv2.remap(im, mx, my, cv2.INTER_LINEAR), code))
    im16 = rng.integers(0, 65536, (h, w), dtype=np.uint16)
    hi = (im16 >> 8).astype(np.uint8)
    assert np.array_equal(O.preprocess(im16, O.PIX_GRAY16), hi)
    assert np.array_equal(O.preprocess(im16, O.PIX_GRAY16, mx, my), cv2.remap(hi, mx, my, cv2.INTER_LINEAR))
    # every (R, G, B) on a coarse lattice + the extremes
    lat = np.stack(np.meshgrid(np.arange(0, 256, 5), np.arange(0, 256, 3), np.arange(0, 256, 7), indexing="ij"), -1)
    lat = np.ascontiguousarray(lat.reshape(1, -1, 3).astype(np.uint8))
    assert np.array_equal(O.preprocess(lat, O.PIX_RGB8), cv2.cvtColor(lat, cv2.COLOR_RGB2GRAY))


def test_undistort_points_vs_cv2():
    """oracle/undistort.cpp == cv2.undistortPoints, bit for bit, on 100k points per camera form"""
    from test_undistort import CAMERAS
    rng = np.random.default_rng(8)
    pts = np.stack([rng.random(100000) * 1100 - 150, rng.random(100000) * 800 - 150], 1).astype(np.float32)
    for name, (K, D, R, P) in CAMERAS.items():
        Pn = K if P is None else P
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K.astype(np.float32), D.astype(np.float32), R=R,
                                  P=Pn.astype(np.float32) if P is None else Pn).reshape(-1, 2)
        assert np.array_equal(O.undistort_points(pts, K, D, R, Pn), ref), name
    # points far outside the image, and a much stronger distortion whose iteration diverges out there (OpenCV's
    # icdist < 0 escape, regression 14583, is taken): still bit for bit
    K, D4 = CAMERAS["mono4"][0], CAMERAS["mono4"][1]
    far = np.stack([rng.random(20000) * 4000 - 1500, rng.random(20000) * 3000 - 1200], 1).astype(np.float32)
    D = np.array([-0.6, 0.1, 0.001, -0.002, 0.0], np.float32).astype(np.float64)        # float-representable, as mDistCoef is CV_32F
    for dist in (D4, D):
        ref = cv2.undistortPoints(far.reshape(-1, 1, 2), K.astype(np.float32), dist.astype(np.float32), P=K.astype(np.float32)).reshape(-1, 2)
        assert np.array_equal(O.undistort_points(far, K, dist, None, K), ref, equal_nan=True)
