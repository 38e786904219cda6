"""GPU parity for ORBmatcher::SearchByBoW (src/ORBmatcher.cc:270-404, 635-768) through the C ABI
against the sequential CPU oracle (oracle/frame_grid.cpp: orc_search_by_bow, itself checked against literal per-form
Python restatements in tests/test_oracle_bow.py).  Integer work and float compares on identical floats: bit-exact."""
import importlib

import numpy as np
import pytest

from synth import bow_case

pytestmark = pytest.mark.gpu
TH_LOW = 50


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.fixture(scope="module")
def O():
    from oracle import pyoracle
    return pyoracle


def _check(gf, O, case, key_frame_form, ratio, ori):
    k1, d1, v1, fv1, k2, d2, v2, fv2 = case
    v2 = v2 if key_frame_form else None
    th = TH_LOW - 1 if key_frame_form else TH_LOW
    want = O.search_by_bow(k1, d1, v1, O.flatten_feature_vector(fv1), k2, d2, v2, O.flatten_feature_vector(fv2), th, ratio, ori)
    got = gf.ORBmatcher(ratio, ori).SearchByBoW(k1, d1, v1, fv1, k2, d2, v2, fv2, th)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])
    return want


@pytest.mark.parametrize("key_frame_form", [False, True])
def test_forms_bit_exact(gf, O, key_frame_form):
    for seed, n1, n2, nodes, ratio, ori in [(0, 1000, 1000, 300, 0.7, True), (1, 2000, 1500, 800, 0.75, True), (2, 1500, 2000, 40, 0.9, True),
                                            (3, 1000, 1000, 300, 0.7, False), (4, 4000, 4000, 1, 0.8, True), (5, 8000, 6000, 2000, 0.7, True)]:
        nm, m12 = _check(gf, O, bow_case(seed, n1, n2, nodes), key_frame_form, ratio, ori)
        if nodes > 1:
            assert nm > min(n1, n2) // 20
        taken = m12[m12 >= 0]
        assert len(np.unique(taken)) == len(taken)            # a candidate is matched at most once


def test_edges(gf, O):
    _check(gf, O, bow_case(10, 300, 0, 20), False, 0.7, True)           # empty frame 2
    _check(gf, O, bow_case(11, 0, 300, 20), True, 0.7, True)            # empty frame 1
    k1, d1, v1, fv1, k2, d2, v2, fv2 = bow_case(12, 400, 400, 50)
    nm, _ = _check(gf, O, (k1, d1, np.zeros_like(v1), fv1, k2, d2, v2, fv2), False, 0.7, True)
    assert nm == 0                                                      # no query has a MapPoint
    nm, _ = _check(gf, O, (k1, d1, None, fv1, k2, d2, np.zeros_like(v2), fv2), True, 0.7, True)
    assert nm == 0                                                      # no candidate has one
    _check(gf, O, (k1, d1, None, fv1, k2, d2, None, fv2), True, 0.7, True)
    disjoint = {n + 1: v for n, v in fv2.items()}                       # no common node id
    nm, _ = _check(gf, O, (k1, d1, v1, fv1, k2, d2, v2, disjoint), False, 0.7, True)
    assert nm == 0
    with pytest.raises(gf.GforbError):
        gf.ORBmatcher(0.7, True).SearchByBoW(k1, d1, v1, {3: [0, 400]}, k2, d2, None, fv2)        # index out of range


def test_on_extracted_frames(gf, O, fixtures):
    """the stereo fixtures as key frame and frame; vocabulary nodes stood in for by the first descriptor byte with two
    bits masked (a re-observed feature mostly keeps its node)"""
    for a, b in (("EuRoC_l", "EuRoC_r"), ("TUM_l", "TUM_r")):
        ex = gf.ORBextractor(2000)
        k1, d1 = ex(fixtures[a]); k2, d2 = ex(fixtures[b])

        def fv(d):
            out = {}
            for i, row in enumerate(d):
                out.setdefault(int(row[0] & 0xF0), []).append(i)
            return out
        want = O.search_by_bow(k1, d1, None, O.flatten_feature_vector(fv(d1)), k2, d2, None, O.flatten_feature_vector(fv(d2)), TH_LOW, 0.7, True)
        got = gf.ORBmatcher(0.7, True).SearchByBoW(k1, d1, None, fv(d1), k2, d2, None, fv(d2), TH_LOW)
        assert got[0] == want[0] > 20 and np.array_equal(got[1], want[1])


def test_search_for_triangulation_bit_exact(gf, O):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:770-935): float epipolar tests, last-of-equals rule"""
    from synth import triangulation_case
    for seed, n1, n2, nodes, ori in [(0, 1000, 1000, 300, True), (1, 2000, 1500, 100, True), (2, 1500, 1500, 600, False),
                                     (3, 3000, 3000, 1, True), (4, 50, 0, 10, True), (5, 6000, 8000, 1500, True)]:
        c = triangulation_case(seed, n1, n2, nodes)
        k1, d1, v1, st1, fv1, k2, d2, v2, st2, fv2, F12, ex, ey, sf, sig2 = c
        want = O.search_for_triangulation(k1, d1, v1, st1, O.flatten_feature_vector(fv1), k2, d2, v2, st2, O.flatten_feature_vector(fv2),
                                          F12, ex, ey, sf, sig2, TH_LOW, ori)
        got = gf.ORBmatcher(0.6, ori).SearchForTriangulation(k1, d1, v1, st1, fv1, k2, d2, v2, st2, fv2, F12, ex, ey, sf, sig2, TH_LOW)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]), seed
        if n1 >= 1000 and n2 >= 1000 and nodes > 1:
            assert want[0] > 50
    # monocular key frames (no stereo flags), every feature without a MapPoint
    got = gf.ORBmatcher(0.6, True).SearchForTriangulation(k1, d1, None, None, fv1, k2, d2, None, None, fv2, F12, ex, ey, sf, sig2)
    want = O.search_for_triangulation(k1, d1, None, None, O.flatten_feature_vector(fv1), k2, d2, None, None, O.flatten_feature_vector(fv2),
                                      F12, ex, ey, sf, sig2, TH_LOW, True)
    assert got[0] == want[0] > 0 and np.array_equal(got[1], want[1])
