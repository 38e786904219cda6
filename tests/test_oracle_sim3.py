"""Oracle pin for ORBmatcher::SearchBySim3 (oracle/frame_grid.cpp: orc_search_by_sim3).

The reference's tests hold no vectors for it (parity unpinned by the reference), so the two search loops
(src/ORBmatcher.cc:1259-1337, :1339-1417) and the agreement test (:1420-1435) are restated here in pure Python,
literally, over the same already-projected queries; the C++ oracle has to reproduce it.  The second pin: the documented
equivalence with the projection entry (two searches with plain PQ_ACTIVE queries, nothing blocking, no ratio test)."""
import numpy as np
import pytest

from oracle import pyoracle as O
from synth import sim3_case
from test_oracle_grid import py_area, py_grid
from test_oracle_projsearch import BOUNDS, TH_HIGH, ham

INT_MAX = 2 ** 31 - 1


def py_one_way(keys, desc, q, qd, th_dist):
    grid = py_grid(keys, BOUNDS)
    vn = [-1] * len(q)
    for i in range(len(q)):
        if not q[i]["flags"] & 1:                       # no map point / already matched / bad / not in view (:1263-1292)
            continue
        lvl = int(q[i]["max_level"])                    # nPredictedLevel
        vind = py_area(keys, grid, q[i]["x"], q[i]["y"], np.float32(q[i]["r"]), -1, -1)
        if not vind:
            continue
        best, bi = INT_MAX, -1
        for idx in vind:
            if keys[idx]["octave"] < lvl - 1 or keys[idx]["octave"] > lvl:
                continue
            d = ham(qd[i], desc[idx])
            if d < best:
                best, bi = d, idx
        if best <= th_dist:
            vn[i] = bi
    return vn


def py_sim3(k1, d1, k2, d2, q12, qd12, q21, qd21, th_dist=TH_HIGH):
    vn1 = py_one_way(k2, d2, q12, qd12, th_dist)
    vn2 = py_one_way(k1, d1, q21, qd21, th_dist)
    m12 = [-1] * len(k1); found = 0
    for i1 in range(len(k1)):
        idx2 = vn1[i1]
        if idx2 >= 0 and vn2[idx2] == i1:
            m12[i1] = idx2; found += 1
    return found, m12, vn1, vn2


@pytest.mark.parametrize("seed,n1,n2,th", [(0, 300, 280, 7.5), (1, 500, 650, 7.5), (2, 64, 900, 10.0), (3, 700, 40, 7.5),
                                            (4, 1, 1, 7.5), (5, 400, 400, 3.0), (6, 350, 350, 30.0)])
def test_oracle_equals_literal_python(seed, n1, n2, th):
    case = sim3_case(seed, n1, n2, th=th)
    want = py_sim3(*case)
    nf, m12, vn1, vn2 = O.search_by_sim3(case[0], case[1], case[2], case[3], BOUNDS, *case[4:])
    assert list(vn1) == want[2] and list(vn2) == want[3]
    assert nf == want[0] and list(m12) == want[1]
    if min(n1, n2) >= 64:
        assert nf > 0.05 * min(n1, n2) and nf < sum(1 for v in want[2] if v >= 0)      # agreement really filters


def test_threshold_and_empty_sides():
    k1, d1, k2, d2, q12, qd12, q21, qd21 = sim3_case(9, 200, 220)
    for th in (0, 30, 255):
        assert O.search_by_sim3(k1, d1, k2, d2, BOUNDS, q12, qd12, q21, qd21, th)[0] == py_sim3(k1, d1, k2, d2, q12, qd12, q21, qd21, th)[0]
    nf, m12, vn1, vn2 = O.search_by_sim3(k1, d1, k2[:0], d2[:0], BOUNDS, q12, qd12, q21[:0], qd21[:0])
    assert nf == 0 and (m12 == -1).all() and (vn1 == -1).all() and len(vn2) == 0
    q12["flags"] = 0
    assert O.search_by_sim3(k1, d1, k2, d2, BOUNDS, q12, qd12, q21, qd21)[0] == 0


def test_equals_two_projection_searches():
    """include/gforb.h: each direction is the projection entry with PQ_ACTIVE queries, no gate, no ratio, no rotation
    check, levels nPredictedLevel - 1 .. nPredictedLevel -- query_key is vnMatch."""
    k1, d1, k2, d2, q12, qd12, q21, qd21 = sim3_case(11, 600, 560)
    nf, m12, vn1, vn2 = O.search_by_sim3(k1, d1, k2, d2, BOUNDS, q12, qd12, q21, qd21)
    _, _, _, qk12 = O.search_by_projection(k2, d2, None, None, BOUNDS, q12, qd12, TH_HIGH, 0.6, False, False)
    _, _, _, qk21 = O.search_by_projection(k1, d1, None, None, BOUNDS, q21, qd21, TH_HIGH, 0.6, False, False)
    assert np.array_equal(qk12, vn1) and np.array_equal(qk21, vn2)
