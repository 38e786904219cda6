"""GPU parity for the Frame feature grid and ORBmatcher::SearchForInitialization, through the
C ABI, against the sequential CPU oracle (oracle/frame_grid.cpp).  Everything here is index / integer work or float
compares on identical float sequences: the bar is bit-exact, order of candidate lists included."""
import importlib

import numpy as np
import pytest

from test_oracle_grid import BOUNDS, F32, _pair, make_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.fixture(scope="module")
def O():
    from oracle import pyoracle
    return pyoracle


def test_grid_and_area_queries_bit_exact(gf, O):
    rng = np.random.default_rng(3)
    for n, spread in [(1000, 0.0), (2100, 25.0), (1, 0.0), (0, 0.0), (8192, 5.0)]:
        keys = make_keys(rng, n, spread=spread)
        if n >= 1000:                                    # exact cell-boundary and half-way coordinates
            keys["x"][:64] = (np.arange(64) * 11.75).astype(F32); keys["y"][:64] = (np.arange(64) % 48 * 10.0 + 5.0).astype(F32)
        g = gf.FrameGrid(keys, BOUNDS)
        start, idx = O.assign_features_to_grid(keys, BOUNDS)
        assert np.array_equal(g.cell_start, start) and np.array_equal(g.cell_idx, idx), n
        qs = []
        for q in range(300):
            lo, hi = [(-1, -1), (0, 0), (2, -1), (1, 3), (0, 7)][q % 5]
            qs.append((float(F32(rng.random() * 900 - 70)), float(F32(rng.random() * 600 - 60)),
                       float(rng.choice([5, 15.5, 40, 100, 2000])), lo, hi))
        got = g.GetFeaturesInArea(qs)
        for q, t in enumerate(qs):
            assert np.array_equal(got[q], O.features_in_area(keys, BOUNDS, *t)), (n, q, t)
    with pytest.raises(gf.GforbError):
        gf.FrameGrid(make_keys(rng, 8193), BOUNDS)
    g = gf.FrameGrid(make_keys(rng, 500), BOUNDS)
    with pytest.raises(gf.GforbError):                   # capacity is checked, nothing is truncated
        g.GetFeaturesInArea([(376.0, 240.0, 2000.0)] * 4, cap=100)


def _check_search(gf, O, k1, d1, k2, d2, bounds, win, ratio, ori, prev=None):
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(F32) if prev is None else prev
    on, om12, oprev = O.search_for_initialization(k1, d1, k2, d2, bounds, prev, win, ratio, ori)
    m = gf.ORBmatcher(ratio, ori)
    p = prev.copy()
    n, m12 = m.SearchForInitialization(k1, d1, k2, d2, p, win, bounds=bounds)
    assert n == on and np.array_equal(m12, om12) and np.array_equal(p, oprev)
    return n


def test_search_for_initialization_synthetic(gf, O):
    rng = np.random.default_rng(11)
    total = 0
    for n1, n2, win, ratio, ori in [(1000, 1000, 100, 0.9, True), (2100, 2000, 100, 0.9, True), (700, 900, 30, 0.6, True),
                                    (900, 800, 100, 0.9, False), (600, 600, 1000, 0.95, True), (333, 4000, 50, 0.8, True),
                                    (5000, 5000, 20, 0.9, True)]:
        k1, d1, k2, d2 = _pair(rng, n1, n2)
        total += _check_search(gf, O, k1, d1, k2, d2, BOUNDS, win, ratio, ori)
    assert total > 1000
    # claim conflicts on purpose: many queries share one descriptor and one window
    k1, d1, k2, d2 = _pair(rng, 400, 400, p0=1.0)
    d1[:200] = d1[0]; k1["x"][:200] = k1["x"][0]; k1["y"][:200] = k1["y"][0]
    d1[:200] ^= np.packbits(rng.random((200, 256)) < 0.03, axis=1)
    _check_search(gf, O, k1, d1, k2, d2, BOUNDS, 100, 0.99, True)
    _check_search(gf, O, k1, d1, k2, d2, BOUNDS, 100, 0.99, False)


def test_search_for_initialization_edges(gf, O):
    rng = np.random.default_rng(12)
    k1, d1, k2, d2 = _pair(rng, 300, 300)
    e = np.zeros(0, k1.dtype); ed = np.zeros((0, 32), np.uint8)
    assert _check_search(gf, O, e, ed, k2, d2, BOUNDS, 100, 0.9, True) == 0
    assert _check_search(gf, O, k1, d1, e, ed, BOUNDS, 100, 0.9, True) == 0
    hi = k1.copy(); hi["octave"] = 3                                           # no level-0 query at all
    assert _check_search(gf, O, hi, d1, k2, d2, BOUNDS, 100, 0.9, True) == 0
    far = np.full((300, 2), -5000.0, F32)                                       # every window misses the grid
    assert _check_search(gf, O, k1, d1, k2, d2, BOUNDS, 100, 0.9, True, prev=far) == 0
    k2o = k2.copy(); k2o["x"] += 400                                            # half of frame 2 outside the bounds
    _check_search(gf, O, k1, d1, k2o, d2, BOUNDS, 500, 0.9, True)
    _check_search(gf, O, k1, d1, k2, d2, (-20.5, 770.25, -11.0, 495.5), 100, 0.9, True)   # undistorted bounds
    same = _check_search(gf, O, k1, d1, k1, d1, BOUNDS, 100, 0.9, True)         # identical frames: distance 0 ties
    assert same > 100


def test_search_for_initialization_on_extracted_frames(gf, O, fixtures):
    """Frames as the reference builds them: ORBextractor output of two views (the stereo fixtures serve as
    consecutive frames), window 100 (src/Tracking.cc: SearchForInitialization(..., 100))."""
    for a, b in (("EuRoC_l", "EuRoC_r"), ("TUM_l", "TUM_r")):
        ex = gf.ORBextractor(2000)
        k1, d1 = ex(fixtures[a]); k2, d2 = ex(fixtures[b])
        h, w = fixtures[a].shape
        n = _check_search(gf, O, k1, d1, k2, d2, (0.0, float(w), 0.0, float(h)), 100, 0.9, True)
        assert n > 30, (a, n)


def test_batched_device_form_equals_single(gf, O):
    rng = np.random.default_rng(13)
    P, S1, S2 = 9, 1100, 1024
    k1 = np.zeros((P, S1), gf.KP_DTYPE); k2 = np.zeros((P, S2), gf.KP_DTYPE)
    d1 = np.zeros((P, S1, 32), np.uint8); d2 = np.zeros((P, S2, 32), np.uint8)
    n1 = np.zeros(P, np.int32); n2 = np.zeros(P, np.int32)
    prev = np.zeros((P, S1, 2), F32)
    ref = []
    for p in range(P):
        a, b = (0, 0) if p == 4 else (int(rng.integers(200, S1 + 1)), int(rng.integers(200, S2 + 1)))
        if p == 7:
            a, b = S1, S2
        n1[p], n2[p] = a, b
        if a:
            ka, da, kb, db = _pair(rng, a, b)
            k1[p, :a], d1[p, :a], k2[p, :b], d2[p, :b] = ka, da, kb, db
            prev[p, :a, 0], prev[p, :a, 1] = ka["x"], ka["y"]
        ref.append(O.search_for_initialization(k1[p, :a], d1[p, :a], k2[p, :b], d2[p, :b], BOUNDS, prev[p, :a], 100, 0.9, True))
    D = gf.DeviceArray.from_host
    dk1, dd1, dn1, dk2, dd2, dn2, dprev = D(k1), D(d1), D(n1), D(k2), D(d2), D(n2), D(prev)
    dm12 = gf.DeviceArray((P, S1), np.int32); dnm = gf.DeviceArray((P,), np.int32)
    m = gf.ORBmatcher(0.9, True)
    m.SearchForInitializationBatchDevice(BOUNDS, P, dk1, dd1, dn1, S1, dk2, dd2, dn2, S2, dprev, 100, dm12, dnm)
    assert m.sync() > 0
    m12, nm, pv = dm12.to_host(), dnm.to_host(), dprev.to_host()
    for p in range(P):
        on, om12, oprev = ref[p]
        a = n1[p]
        assert nm[p] == on and np.array_equal(m12[p, :a], om12) and np.array_equal(pv[p, :a], oprev), p
