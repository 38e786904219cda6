"""Frame::UndistortKeyPoints / UndistortKeyPointsStereo / ComputeImageBounds / ComputeImageBoundsStereo
(src/Frame.cc:670-887) run from THE REFERENCE'S OWN TEXT (oracle/_ref/libgforb_ref.so, lifted verbatim by
oracle/ref_extract.py; cv::undistortPoints = the cv2-pinned primitive of oracle/undistort.cpp behind the shim).

CPU: the restatement (O.undistort_points + the key copy / floor-ceil rule written out below) == the reference's text.
GPU (-m gpu): the CUDA path through the C ABI == the reference's text.
The reference keeps K, distCoef, R and P as CV_32F (src/Tracking.cc:144-168), so every camera here is float32-valued."""
import ctypes as C
import math

import numpy as np
import pytest

from oracle import pyoracle as O
from oracle import pyref
from synth import KEYPOINT
from test_undistort import CAMERAS, points

pytestmark = pytest.mark.skipif(not pyref.available("ref"), reason="oracle/_ref is not built and /root/reference is absent")

F32 = np.float32
vp, i32 = C.c_void_p, C.c_int


@pytest.fixture(scope="module")
def L():
    lib = pyref.load("ref").lib()
    lib.ref_undistort_keypoints.argtypes = [vp, vp, i32, vp, i32, vp]; lib.ref_undistort_keypoints.restype = None
    lib.ref_undistort_keypoints_stereo.argtypes = [vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, i32, vp, vp]
    lib.ref_undistort_keypoints_stereo.restype = None
    lib.ref_compute_image_bounds.argtypes = [i32, vp, vp, i32, vp, vp, i32, i32, vp]; lib.ref_compute_image_bounds.restype = None
    return lib


def P(a):
    return None if a is None else a.ctypes.data_as(vp)


def cam32(name):
    """the camera as the reference holds it: float32 K (3x3), dist, R (3x3), P (3x4); None where the mono form has none"""
    K, D, R, Pm = CAMERAS[name]
    f = lambda a: None if a is None else np.ascontiguousarray(np.asarray(a, np.float64).astype(F32))
    return f(K), f(D).ravel(), f(R), f(Pm)


def keys_of(pts, seed):
    rng = np.random.default_rng(seed)
    k = np.zeros(len(pts), KEYPOINT)
    k["x"], k["y"] = pts[:, 0], pts[:, 1]
    k["size"] = 31; k["angle"] = rng.random(len(pts)).astype(F32) * 360; k["response"] = rng.integers(7, 200, len(pts))
    k["octave"] = rng.integers(0, 8, len(pts)); k["class_id"] = -1
    return k


def restated_keys(keys, K, D, R, Pm):
    """src/Frame.cc:724-758 / 670-722: distCoef[0] == 0 -> copy; else only pt changes"""
    out = keys.copy()
    if D[0] == 0.0 or len(keys) == 0:
        return out
    d = lambda a: None if a is None else np.asarray(a, np.float64)
    xy = O.undistort_points(np.stack([keys["x"], keys["y"]], 1), d(K), d(D), d(R), d(K) if Pm is None else d(Pm))
    out["x"], out["y"] = xy[:, 0], xy[:, 1]
    return out


def restated_bounds(K, D, R, Pm, cols, rows):
    """src/Frame.cc:760-828 / 830-887: eight border points, floor / ceil applied only when a point moves the bound"""
    if D[0] == 0.0:
        return (0.0, float(cols), 0.0, float(rows))
    h2, w2 = F32(rows) / F32(2.0), F32(cols) / F32(2.0)
    b = np.array([[0, 0], [cols, 0], [0, rows], [cols, rows], [0, h2], [w2, 0], [w2, rows], [cols, h2]], F32)
    d = lambda a: None if a is None else np.asarray(a, np.float64)
    u = O.undistort_points(b, d(K), d(D), d(R), d(K) if Pm is None else d(Pm))
    mnx, mxx, mny, mxy = F32(2 ** 31 - 1), F32(-2 ** 31), F32(2 ** 31 - 1), F32(-2 ** 31)     # INT_MAX / INT_MIN stored in floats
    for x, y in u:
        if mnx > x: mnx = F32(math.floor(x))
        if mny > y: mny = F32(math.floor(y))
        if mxx < x: mxx = F32(math.ceil(x))
        if mxy < y: mxy = F32(math.ceil(y))
    return (float(mnx), float(mxx), float(mny), float(mxy))


def ref_mono(L, K, D, keys):
    out = np.zeros(max(len(keys), 1), KEYPOINT)
    L.ref_undistort_keypoints(P(K), P(D), len(D), P(keys), len(keys), P(out))
    return out[:len(keys)]


def ref_bounds(L, stereo, K, D, R, Pm, cols, rows):
    out = np.zeros(4, F32)
    L.ref_compute_image_bounds(int(stereo), P(K), P(D), len(D), P(R), P(Pm), cols, rows, P(out))
    return tuple(float(v) for v in out)


def test_restatement_matches_reference_text_mono(L):
    for name in ("mono4", "mono5"):
        K, D, _, _ = cam32(name)
        keys = keys_of(points(name), 5)
        got = ref_mono(L, K, D, keys)
        assert got.tobytes() == restated_keys(keys, K, D, None, None).tobytes(), name
        assert not np.array_equal(got["x"], keys["x"])
        for cols, rows in ((752, 480), (640, 481), (1241, 376)):
            assert ref_bounds(L, 0, K, D, None, None, cols, rows) == restated_bounds(K, D, None, None, cols, rows), (name, cols, rows)
    # distCoef[0] == 0: keys copied, bounds = the image rectangle (:672-677, :726-730, :821-827)
    K, D, _, _ = cam32("mono4")
    D0 = D.copy(); D0[0] = 0
    keys = keys_of(points("mono4"), 6)
    assert ref_mono(L, K, D0, keys).tobytes() == keys.tobytes()
    assert ref_bounds(L, 0, K, D0, None, None, 752, 480) == (0.0, 752.0, 0.0, 480.0)
    assert len(ref_mono(L, K, D, keys[:0])) == 0


def test_restatement_matches_reference_text_stereo(L):
    K, D, R, Pm = cam32("stereo")
    # a right camera of its own: different intrinsics, distortion, rectification
    KR = (K.astype(np.float64) * np.array([[1.002, 1, 0.99], [1, 0.998, 1.01], [1, 1, 1]])).astype(F32)
    DR = (D.astype(np.float64) * np.array([0.97, 1.1, -1.0, 2.0])).astype(F32)
    RR = np.ascontiguousarray(R.T); PR = Pm.copy(); PR[0, 3] = F32(-47.90639384423901)
    kl = keys_of(points("stereo"), 7); kr = keys_of(points("mono4")[:2500], 8)
    outL = np.zeros(len(kl), KEYPOINT); outR = np.zeros(len(kr), KEYPOINT)
    L.ref_undistort_keypoints_stereo(P(K), P(D), len(D), P(R), P(Pm), P(KR), P(DR), len(DR), P(RR), P(PR),
                                     P(kl), len(kl), P(kr), len(kr), P(outL), P(outR))
    assert outL.tobytes() == restated_keys(kl, K, D, R, Pm).tobytes()
    assert outR.tobytes() == restated_keys(kr, KR, DR, RR, PR).tobytes()
    for cols, rows in ((752, 480), (1280, 720)):
        assert ref_bounds(L, 1, K, D, R, Pm, cols, rows) == restated_bounds(K, D, R, Pm, cols, rows)
    # :672-677: only the LEFT distCoef[0] is looked at -- both key sets are copied when it is 0
    D0 = D.copy(); D0[0] = 0
    L.ref_undistort_keypoints_stereo(P(K), P(D0), len(D0), P(R), P(Pm), P(KR), P(DR), len(DR), P(RR), P(PR),
                                     P(kl), len(kl), P(kr), len(kr), P(outL), P(outR))
    assert outL.tobytes() == kl.tobytes() and outR.tobytes() == kr.tobytes()


@pytest.mark.gpu
def test_gpu_matches_reference_text(L):
    import importlib
    gf = importlib.import_module("gf-orb-slam2_b200")
    m = gf.ORBmatcher()
    d = lambda a: None if a is None else np.asarray(a, np.float64)
    for name in ("mono4", "mono5", "stereo"):
        K, D, R, Pm = cam32(name)
        cam = gf.Camera.make(d(K), d(D), d(R), d(Pm))
        keys = keys_of(points(name), 9).astype(gf.KP_DTYPE)
        un = gf.UndistortKeyPoints(m, cam, keys)
        if name == "stereo":
            outL = np.zeros(len(keys), KEYPOINT); outR = np.zeros(len(keys), KEYPOINT)
            L.ref_undistort_keypoints_stereo(P(K), P(D), len(D), P(R), P(Pm), P(K), P(D), len(D), P(R), P(Pm),
                                             P(keys), len(keys), P(keys), len(keys), P(outL), P(outR))
            want = outL
        else:
            want = ref_mono(L, K, D, keys)
        assert un.tobytes() == want.tobytes(), name
        for cols, rows in ((752, 480), (1241, 376)):
            assert gf.ComputeImageBounds(m, cam, cols, rows) == ref_bounds(L, name == "stereo", K, D, R, Pm, cols, rows), (name, cols, rows)
