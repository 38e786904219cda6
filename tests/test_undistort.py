"""Keypoint undistortion + image bounds (src/Frame.cc:670-887).  CPU: the oracle reproduces the
cv2-made golden vectors (tools/gen_golden_undistort.py) bit for bit.  GPU (-m gpu): the CUDA path through the C ABI ==
oracle == golden, single and batched, and the undistorted keys feed the device grid / SearchForInitialization."""
import importlib
import math
from pathlib import Path

import numpy as np
import pytest

from oracle import pyoracle as O

GOLD = Path(__file__).resolve().parent / "golden" / "undistort.npz"
_K = np.array([[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]], np.float32).astype(np.float64)
_D4 = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32).astype(np.float64)
_D5 = np.array([-0.28, 0.07, 0.0002, 1.8e-5, 0.011], np.float32).astype(np.float64)


def _rodrigues(rv):
    th = float(np.linalg.norm(rv)); k = np.asarray(rv, np.float64) / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + math.sin(th) * Kx + (1 - math.cos(th)) * (Kx @ Kx)


_R = np.round(_rodrigues([0.01, -0.02, 0.005]), 12)                       # rounded: the same doubles on every machine
_P = np.array([[435.2046959714599, 0, 367.4517211914062, 0], [0, 435.2046959714599, 252.2008514404297, 0], [0, 0, 1, 0]])
CAMERAS = {"mono4": (_K, _D4, None, None), "mono5": (_K, _D5, None, None), "stereo": (_K, _D4, _R, _P)}


def points(name):
    rng = np.random.default_rng({"mono4": 1, "mono5": 2, "stereo": 3}[name])
    pts = np.stack([rng.random(4000) * 752, rng.random(4000) * 480], 1).astype(np.float32)
    pts[:200] = np.round(pts[:200])                                       # integer pixel positions (level-0 keys)
    pts[200:260] = np.stack([rng.random(60) * 3000 - 1000, rng.random(60) * 2000 - 800], 1).astype(np.float32)   # far outside
    return pts


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def _ref_bounds(b):
    mnx, mxx, mny, mxy = float(2 ** 31 - 1), float(-2 ** 31), float(2 ** 31 - 1), float(-2 ** 31)
    for x, y in b:
        if mnx > x: mnx = math.floor(x)
        if mny > y: mny = math.floor(y)
        if mxx < x: mxx = math.ceil(x)
        if mxy < y: mxy = math.ceil(y)
    return (mnx, mxx, mny, mxy)


def test_oracle_reproduces_cv2_golden(gold):
    for name, (K, D, R, P) in CAMERAS.items():
        pts = points(name)
        assert np.array_equal(pts, gold[name + "_in"])
        got = O.undistort_points(pts, K, D, R, K if P is None else P)
        assert np.array_equal(got, gold[name + "_out"]), name


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


def _keys(gf, pts, rng):
    k = np.zeros(len(pts), gf.KP_DTYPE)
    k["x"], k["y"] = pts[:, 0], pts[:, 1]
    k["size"] = 31; k["angle"] = rng.random(len(pts)).astype(np.float32) * 360; k["response"] = rng.integers(7, 200, len(pts))
    k["octave"] = rng.integers(0, 8, len(pts)); k["class_id"] = -1
    return k


@pytest.mark.gpu
def test_gpu_undistort_matches_golden_and_oracle(gf, gold):
    m = gf.ORBmatcher()
    rng = np.random.default_rng(4)
    for name, (K, D, R, P) in CAMERAS.items():
        cam = gf.Camera.make(K, D, R, P)
        keys = _keys(gf, points(name), rng)
        un = gf.UndistortKeyPoints(m, cam, keys)
        assert np.array_equal(np.stack([un["x"], un["y"]], 1), gold[name + "_out"]), name
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(un[f], keys[f])
        assert gf.ComputeImageBounds(m, cam, 752, 480) == _ref_bounds(gold[name + "_border"]), name
    # distCoef[0] == 0: the reference copies the keys and uses the image rectangle
    cam0 = gf.Camera.make(_K, np.zeros(4))
    keys = _keys(gf, points("mono4"), rng)
    assert gf.UndistortKeyPoints(m, cam0, keys).tobytes() == keys.tobytes()
    assert gf.ComputeImageBounds(m, cam0, 752, 480) == (0.0, 752.0, 0.0, 480.0)
    assert len(gf.UndistortKeyPoints(m, cam0, keys[:0])) == 0


@pytest.mark.gpu
def test_gpu_undistort_batch_then_grid_and_search(gf):
    """Device-resident chain of the default build (src/Frame.cc:61-147): keys -> UndistortKeyPointsStereo -> grid /
    SearchForInitialization on the undistorted keys with the undistorted image bounds; nothing returns to the host
    in between.  Checked against the oracle run step by step."""
    from test_oracle_grid import _pair
    rng = np.random.default_rng(21)
    K, D, R, P = CAMERAS["stereo"]
    cam = gf.Camera.make(K, D, R, P)
    m = gf.ORBmatcher(0.9, True)
    bounds = gf.ComputeImageBounds(m, cam, 752, 480)
    npair, S = 5, 900
    k1 = np.zeros((npair, S), gf.KP_DTYPE); k2 = np.zeros((npair, S), gf.KP_DTYPE)
    d1 = np.zeros((npair, S, 32), np.uint8); d2 = np.zeros((npair, S, 32), np.uint8)
    n1 = np.array([900, 640, 0, 333, 900], np.int32); n2 = np.array([900, 700, 50, 400, 1], np.int32)
    for p in range(npair):
        if n1[p] and n2[p]:
            a, da, b, db = _pair(rng, int(n1[p]), int(n2[p]))
            k1[p, :n1[p]], d1[p, :n1[p]], k2[p, :n2[p]], d2[p, :n2[p]] = a, da, b, db
    Dv = gf.DeviceArray.from_host
    dk1, dk2, dd1, dd2, dn1, dn2 = Dv(k1), Dv(k2), Dv(d1), Dv(d2), Dv(n1), Dv(n2)
    ptr = lambda x: x.ptr
    gf._check(gf.lib.gforb_undistort_keypoints_batch_device(m._h, gf.C.byref(cam), ptr(dk1), ptr(dn1), S, npair, ptr(dk1), None))
    gf._check(gf.lib.gforb_undistort_keypoints_batch_device(m._h, gf.C.byref(cam), ptr(dk2), ptr(dn2), S, npair, ptr(dk2), None))
    u1, u2 = dk1.to_host(), dk2.to_host()
    prev = np.ascontiguousarray(np.stack([u1["x"], u1["y"]], 2).astype(np.float32))
    dprev = Dv(prev); dm12 = gf.DeviceArray((npair, S), np.int32); dnm = gf.DeviceArray((npair,), np.int32)
    m.SearchForInitializationBatchDevice(bounds, npair, dk1, dd1, dn1, S, dk2, dd2, dn2, S, dprev, 100, dm12, dnm)
    m.sync()
    m12, nm = dm12.to_host(), dnm.to_host()
    for p in range(npair):
        a, b = int(n1[p]), int(n2[p])
        o1 = k1[p, :a].copy(); o2 = k2[p, :b].copy()
        for o in (o1, o2):
            if len(o):
                xy = O.undistort_points(np.stack([o["x"], o["y"]], 1), K, D, R, P)
                o["x"], o["y"] = xy[:, 0], xy[:, 1]
        assert u1[p, :a].tobytes() == o1.tobytes() and u2[p, :b].tobytes() == o2.tobytes()
        on, om12, _ = O.search_for_initialization(o1, d1[p, :a], o2, d2[p, :b], bounds,
                                                  np.stack([o1["x"], o1["y"]], 1).astype(np.float32), 100, 0.9, True)
        assert nm[p] == on and np.array_equal(m12[p, :a], om12), p
