"""Oracle pin for ORBmatcher::SearchByBoW (oracle/frame_grid.cpp: orc_search_by_bow).

The oracle folds the reference's two forms into one loop over flattened DBoW2 feature vectors.  The reference's tests
hold no vectors for them (parity unpinned by the reference), so each form is restated here in pure Python, literally
and separately -- src/ORBmatcher.cc:270-404 (key frame against a frame, result indexed by the FRAME's features,
`bestDist1 <= TH_LOW`, rotation histogram of frame indices) and :635-768 (key frame against key frame, result indexed
by the first key frame's features, `bestDist1 < TH_LOW`, candidates need a good MapPoint) -- on std::map-like dicts,
and the folded oracle has to reproduce both."""
import bisect

import numpy as np

from oracle import pyoracle as O
from synth import bow_case, triangulation_case
from test_oracle_projsearch import F32, ham, rot_bin, three_maxima

TH_LOW = 50


def merge(fv1, fv2):
    """the while loop over two std::map iterators with lower_bound (:292-376, :666-743): yields matching node ids"""
    a, b = sorted(fv1), sorted(fv2)
    i = j = 0
    while i != len(a) and j != len(b):
        if a[i] == b[j]:
            yield a[i]
            i += 1; j += 1
        elif a[i] < b[j]:
            i = bisect.bisect_left(a, b[j], i)
        else:
            j = bisect.bisect_left(b, a[i], j)


def py_bow_frame(k1, d1, valid1, fv1, k2, d2, fv2, ratio, ori):
    """:270-404 -> vpMapPointMatches (per FRAME feature: key-frame feature index or -1), nmatches"""
    mp = [-1] * len(k2)
    hist = [[] for _ in range(30)]
    n = 0
    for node in merge(fv1, fv2):
        for i1 in fv1[node]:
            if not valid1[i1]:
                continue
            b1 = b2 = 256; bi = -1
            for i2 in fv2[node]:
                if mp[i2] >= 0:
                    continue
                d = ham(d1[i1], d2[i2])
                if d < b1:
                    b2, b1, bi = b1, d, i2
                elif d < b2:
                    b2 = d
            if b1 <= TH_LOW and F32(b1) < F32(ratio) * F32(b2):
                mp[bi] = i1
                if ori:
                    hist[rot_bin(k1[i1]["angle"], k2[bi]["angle"])].append(bi)
                n += 1
    if ori:
        ind = three_maxima([len(h) for h in hist])
        for i, h in enumerate(hist):
            if i not in ind:
                for j in h:
                    mp[j] = -1
                    n -= 1
    return n, mp


def py_bow_keyframes(k1, d1, valid1, fv1, k2, d2, valid2, fv2, ratio, ori):
    """:635-768 -> vpMatches12 (per first-key-frame feature: second-key-frame feature index or -1), nmatches"""
    m12 = [-1] * len(k1); matched2 = [False] * len(k2)
    hist = [[] for _ in range(30)]
    n = 0
    for node in merge(fv1, fv2):
        for i1 in fv1[node]:
            if not valid1[i1]:
                continue
            b1 = b2 = 256; bi = -1
            for i2 in fv2[node]:
                if matched2[i2] or not valid2[i2]:
                    continue
                d = ham(d1[i1], d2[i2])
                if d < b1:
                    b2, b1, bi = b1, d, i2
                elif d < b2:
                    b2 = d
            if b1 < TH_LOW and F32(b1) < F32(ratio) * F32(b2):
                m12[i1] = bi; matched2[bi] = True
                if ori:
                    hist[rot_bin(k1[i1]["angle"], k2[bi]["angle"])].append(i1)
                n += 1
    if ori:
        ind = three_maxima([len(h) for h in hist])
        for i, h in enumerate(hist):
            if i not in ind:
                for j in h:
                    m12[j] = -1
                    n -= 1
    return n, m12


CASES = [(0, 900, 1000, 300, 0.7, True), (1, 1200, 700, 150, 0.75, True), (2, 500, 500, 40, 0.9, True), (3, 600, 600, 300, 0.7, False),
         (4, 50, 0, 10, 0.7, True), (5, 0, 50, 10, 0.7, True), (6, 300, 300, 1, 0.8, True)]


def test_frame_form():
    for seed, n1, n2, nodes, ratio, ori in CASES:
        k1, d1, v1, fv1, k2, d2, v2, fv2 = bow_case(seed, n1, n2, nodes)
        nm, m12 = O.search_by_bow(k1, d1, v1, O.flatten_feature_vector(fv1), k2, d2, None, O.flatten_feature_vector(fv2), TH_LOW, ratio, ori)
        rn, mp = py_bow_frame(k1, d1, v1, fv1, k2, d2, fv2, ratio, ori)
        inv = [-1] * n1
        for i2, i1 in enumerate(mp):
            if i1 >= 0:
                inv[i1] = i2
        assert nm == rn and list(m12) == inv
        if n1 >= 500 and n2 >= 500 and nodes > 1:
            assert nm > 40


def test_key_frame_form():
    for seed, n1, n2, nodes, ratio, ori in CASES:
        k1, d1, v1, fv1, k2, d2, v2, fv2 = bow_case(seed + 20, n1, n2, nodes)
        nm, m12 = O.search_by_bow(k1, d1, v1, O.flatten_feature_vector(fv1), k2, d2, v2, O.flatten_feature_vector(fv2), TH_LOW - 1, ratio, ori)
        rn, rm12 = py_bow_keyframes(k1, d1, v1, fv1, k2, d2, v2, fv2, ratio, ori)
        assert nm == rn and list(m12) == rm12
        if n1 >= 500 and n2 >= 500 and nodes > 1:
            assert nm > 30


def py_triangulation(k1, d1, v1, st1, fv1, k2, d2, v2, st2, fv2, F12, ex, ey, sf, sig2, ori):
    """src/ORBmatcher.cc:770-935 + CheckDistEpipolarLine :251-268, float by float"""
    m12 = [-1] * len(k1)
    hist = [[] for _ in range(30)]
    n = 0
    for node in merge(fv1, fv2):
        for i1 in fv1[node]:
            if not v1[i1]:
                continue
            x1, y1 = F32(k1[i1]["x"]), F32(k1[i1]["y"])
            best, bi = TH_LOW, -1
            for i2 in fv2[node]:
                if not v2[i2]:
                    continue
                d = ham(d1[i1], d2[i2])
                if d > TH_LOW or d > best:
                    continue
                x2, y2, o2 = F32(k2[i2]["x"]), F32(k2[i2]["y"]), int(k2[i2]["octave"])
                if not st1[i1] and not st2[i2]:
                    dx, dy = F32(ex - x2), F32(ey - y2)
                    if F32(F32(dx * dx) + F32(dy * dy)) < F32(F32(100) * sf[o2]):
                        continue
                a = F32(F32(F32(x1 * F12[0, 0]) + F32(y1 * F12[1, 0])) + F12[2, 0])
                b = F32(F32(F32(x1 * F12[0, 1]) + F32(y1 * F12[1, 1])) + F12[2, 1])
                c = F32(F32(F32(x1 * F12[0, 2]) + F32(y1 * F12[1, 2])) + F12[2, 2])
                num = F32(F32(F32(a * x2) + F32(b * y2)) + c)
                den = F32(F32(a * a) + F32(b * b))
                if den == 0:
                    continue
                dsqr = F32(F32(num * num) / den)
                if float(dsqr) < 3.84 * float(sig2[o2]):
                    bi, best = i2, d
            if bi >= 0:
                m12[i1] = bi
                n += 1
                if ori:
                    hist[rot_bin(k1[i1]["angle"], k2[bi]["angle"])].append(i1)
    if ori:
        ind = three_maxima([len(h) for h in hist])
        for i, h in enumerate(hist):
            if i not in ind:
                for j in h:
                    m12[j] = -1
                    n -= 1
    return n, m12


def test_search_for_triangulation():
    for seed, n1, n2, nodes, ori in [(0, 900, 1000, 300, True), (1, 1200, 700, 100, True), (2, 600, 600, 300, False), (3, 40, 0, 10, True),
                                     (4, 400, 400, 1, True)]:
        c = triangulation_case(seed, n1, n2, nodes)
        k1, d1, v1, st1, fv1, k2, d2, v2, st2, fv2, F12, ex, ey, sf, sig2 = c
        nm, m12 = O.search_for_triangulation(k1, d1, v1, st1, O.flatten_feature_vector(fv1), k2, d2, v2, st2, O.flatten_feature_vector(fv2),
                                             F12, ex, ey, sf, sig2, TH_LOW, ori)
        rn, rm12 = py_triangulation(k1, d1, v1, st1, fv1, k2, d2, v2, st2, fv2, F12, ex, ey, sf, sig2, ori)
        assert nm == rn and list(m12) == rm12
        if n1 >= 600 and n2 >= 600 and nodes > 1:
            assert nm > 50
