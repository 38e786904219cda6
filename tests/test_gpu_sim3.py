"""ORBmatcher::SearchBySim3 on the device (gforb_search_by_sim3, src/ORBmatcher.cc:1214-1438) == the oracle, bit for
bit: both vnMatch arrays, the agreed pairs and the return value."""
import importlib

import numpy as np
import pytest

from oracle import pyoracle as O
from synth import sim3_case

pytestmark = pytest.mark.gpu
BOUNDS = (0.0, 752.0, 0.0, 480.0)


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.mark.parametrize("seed,n1,n2,th", [(0, 300, 280, 7.5), (1, 1000, 1100, 7.5), (2, 64, 2000, 10.0), (3, 3000, 40, 7.5),
                                            (4, 1, 1, 7.5), (5, 2000, 2000, 3.0), (6, 1500, 1500, 40.0), (7, 8192, 8192, 7.5)])
def test_gpu_equals_oracle(gf, seed, n1, n2, th):
    case = sim3_case(seed, n1, n2, th=th)
    m = gf.ORBmatcher()
    for th_dist in (100, 40):
        nf, m12, vn1, vn2 = m.SearchBySim3(*case, th_dist=th_dist, bounds=BOUNDS)
        onf, om12, ovn1, ovn2 = O.search_by_sim3(case[0], case[1], case[2], case[3], BOUNDS, *case[4:], th_dist)
        assert np.array_equal(vn1, ovn1) and np.array_equal(vn2, ovn2)
        assert nf == onf and np.array_equal(m12, om12)


def test_empty_sides_and_inactive_queries(gf):
    k1, d1, k2, d2, q12, qd12, q21, qd21 = sim3_case(9, 200, 220)
    m = gf.ORBmatcher()
    nf, m12, vn1, vn2 = m.SearchBySim3(k1, d1, k2[:0], d2[:0], q12, qd12, q21[:0], qd21[:0], bounds=BOUNDS)
    assert nf == 0 and (m12 == -1).all() and (vn1 == -1).all() and len(vn2) == 0
    nf, m12, vn1, vn2 = m.SearchBySim3(k1[:0], d1[:0], k2, d2, q12[:0], qd12[:0], q21, qd21, bounds=BOUNDS)
    assert nf == 0 and len(m12) == 0 and (vn2 == -1).all()
    q12["flags"] = 0
    nf, m12, vn1, vn2 = m.SearchBySim3(k1, d1, k2, d2, q12, qd12, q21, qd21, bounds=BOUNDS)
    assert nf == 0 and (vn1 == -1).all() and (vn2 >= 0).any()
    assert np.array_equal(vn2, O.search_by_sim3(k1, d1, k2, d2, BOUNDS, q12, qd12, q21, qd21)[3])


def test_on_extracted_key_frames(gf):
    """two extracted frames of one synthetic scene as the key frames; each key's own descriptor as its map point's"""
    from synth import synth_stereo, PROJ_QUERY
    l, r = synth_stereo(21)
    ex = gf.ORBextractor()
    k1, d1 = ex(l); k2, d2 = ex(r)
    k1, d1, k2, d2 = np.ascontiguousarray(k1), np.ascontiguousarray(d1), np.ascontiguousarray(k2), np.ascontiguousarray(d2)
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)

    def q(k, shift):
        out = np.zeros(len(k), PROJ_QUERY)
        out["x"], out["y"] = k["x"] + np.float32(shift), k["y"]
        out["r"] = np.float32(30.0) * sf[k["octave"]]
        out["min_level"], out["max_level"] = k["octave"] - 1, k["octave"]
        out["flags"] = 1
        return out
    m = gf.ORBmatcher()
    got = m.SearchBySim3(k1, d1, k2, d2, q(k1, -25.0), d1, q(k2, 25.0), d2, bounds=BOUNDS)
    want = O.search_by_sim3(k1, d1, k2, d2, BOUNDS, q(k1, -25.0), d1, q(k2, 25.0), d2)
    assert got[0] == want[0] and got[0] > 100
    for a, b in zip(got[1:], want[1:]):
        assert np.array_equal(a, b)
