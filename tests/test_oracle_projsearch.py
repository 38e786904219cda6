"""Oracle pin for ORBmatcher::SearchByProjection (oracle/frame_grid.cpp: orc_search_by_projection).

The oracle folds the reference's three per-frame forms into ONE loop over query records.  The reference's tests
hold no vectors for them (parity unpinned by the reference), so each form is restated here in pure Python, literally
and separately, from its own reference lines -- src/ORBmatcher.cc:155-241 (map points into a frame), :1440-1593
(last frame into the current frame) and :1595-1721 (key frame into the current frame) -- and the folded oracle has to
reproduce all three, with the flags the C ABI documents for each.
"""
import numpy as np

from oracle import pyoracle as O
from synth import projection_case
from test_oracle_grid import c_round, py_area, py_grid

F32 = np.float32
BOUNDS = (0.0, 752.0, 0.0, 480.0)
TH_HIGH = 100


def ham(a, b):
    return int(np.unpackbits(a ^ b).sum())


def three_maxima(sizes):
    m = [0, 0, 0]; ind = [-1, -1, -1]
    for i, s in enumerate(sizes):
        if s > m[0]:
            m = [s, m[0], m[1]]; ind = [i, ind[0], ind[1]]
        elif s > m[1]:
            m = [m[0], s, m[1]]; ind = [ind[0], i, ind[1]]
        elif s > m[2]:
            m[2] = s; ind[2] = i
    if m[1] < F32(0.1) * F32(m[0]):
        ind[1] = ind[2] = -1
    elif m[2] < F32(0.1) * F32(m[0]):
        ind[2] = -1
    return ind


def rot_bin(a_query, a_key):
    rot = F32(F32(a_query) - F32(a_key))
    if rot < 0:
        rot = F32(rot + F32(360))
    b = c_round(float(rot * (F32(1) / F32(30))))
    return 0 if b == 30 else b


class MP:                                    # the part of MapPoint the matcher looks at
    def __init__(self, q, nobs):
        self.q, self.nobs = q, nobs


def py_map_form(keys, desc, ur, occupied, q, qd, ratio):
    """src/ORBmatcher.cc:155-241.  mvpMapPoints[k] starts as a point with observations where occupied[k]."""
    grid = py_grid(keys, BOUNDS)
    mvp = [MP(-1, 1) if o else None for o in occupied]
    score = [0] * len(keys)
    nmatches = 0
    for iq in range(len(q)):
        if not q[iq]["flags"] & 1:                                   # mbTrackInView / isBad
            continue
        r = F32(q[iq]["r"])                                          # r * F.mvScaleFactors[nPredictedLevel]
        vind = py_area(keys, grid, q[iq]["x"], q[iq]["y"], r, int(q[iq]["min_level"]), int(q[iq]["max_level"]))
        if not vind:
            continue
        best = best2 = 256; lvl = lvl2 = -1; bi = -1
        for idx in vind:
            if mvp[idx] is not None and mvp[idx].nobs > 0:
                continue
            if ur[idx] > 0:
                er = abs(F32(q[iq]["xr"]) - F32(ur[idx]))
                if er > r:
                    continue
            d = ham(qd[iq], desc[idx])
            if d < best:
                best2, best, lvl2, lvl, bi = best, d, lvl, int(keys[idx]["octave"]), idx
            elif d < best2:
                lvl2, best2 = int(keys[idx]["octave"]), d
        if best <= TH_HIGH:
            if lvl == lvl2 and F32(best) > F32(ratio) * F32(best2):
                continue
            mvp[bi] = MP(iq, 1 if q[iq]["flags"] & 2 else 0)
            score[bi] = best
            nmatches += 1
    kq = np.array([m.q if m is not None else -1 for m in mvp], np.int32)
    return nmatches, kq, np.array(score, np.int32)


def py_frame_form(keys, desc, ur, occupied, q, qd, th_dist, any_assigned_blocks):
    """src/ORBmatcher.cc:1440-1593 (any_assigned_blocks False) and :1595-1721 (True: `if (mvpMapPoints[i2]) continue`,
    no mvuRight test).  Rotation histogram over the KEYS taken, three maxima, removal."""
    grid = py_grid(keys, BOUNDS)
    mvp = [MP(-1, 1) if o else None for o in occupied]
    cleared = set()
    nmatches = 0
    hist = [[] for _ in range(30)]
    for iq in range(len(q)):
        if not q[iq]["flags"] & 1:
            continue
        radius = F32(q[iq]["r"])
        vind = py_area(keys, grid, q[iq]["x"], q[iq]["y"], radius, int(q[iq]["min_level"]), int(q[iq]["max_level"]))
        if not vind:
            continue
        best, bi = 256, -1
        for i2 in vind:
            if any_assigned_blocks:
                if mvp[i2] is not None:
                    continue
            else:
                if mvp[i2] is not None and mvp[i2].nobs > 0:
                    continue
                if ur[i2] > 0:
                    er = abs(F32(q[iq]["xr"]) - F32(ur[i2]))
                    if er > radius:
                        continue
            d = ham(qd[iq], desc[i2])
            if d < best:
                best, bi = d, i2
        if best <= th_dist:
            mvp[bi] = MP(iq, 1 if q[iq]["flags"] & 2 else 0)
            nmatches += 1
            hist[rot_bin(q[iq]["angle"], keys[bi]["angle"])].append(bi)
    ind = three_maxima([len(h) for h in hist])
    for i, h in enumerate(hist):
        if i not in ind:
            for k in h:
                mvp[k] = None
                cleared.add(k)
                nmatches -= 1
    kq = np.array([(-2 if k in cleared else -1) if m is None else m.q for k, m in enumerate(mvp)], np.int32)
    return nmatches, kq


def py_sim3_form(keys, desc, occupied, q, qd, th_low):
    """src/ORBmatcher.cc:406-518 (loop closing: map points into a key frame through a Sim3).  KeyFrame::GetFeaturesInArea
    (src/KeyFrame.cc:877-916) has no level arguments; the level test sits in the candidate loop (:493-496)."""
    grid = py_grid(keys, BOUNDS)
    matched = [(-1 if not o else -3) for o in occupied]           # vpMatched[idx]: -1 = NULL
    nmatches = 0
    for iq in range(len(q)):
        if not q[iq]["flags"] & 1:
            continue
        radius = F32(q[iq]["r"])
        lvl = int(q[iq]["max_level"])                             # nPredictedLevel
        vind = py_area(keys, grid, q[iq]["x"], q[iq]["y"], radius, -1, -1)
        if not vind:
            continue
        best, bi = 256, -1
        for idx in vind:
            if matched[idx] != -1:
                continue
            kl = int(keys[idx]["octave"])
            if kl < lvl - 1 or kl > lvl:
                continue
            d = ham(qd[iq], desc[idx])
            if d < best:
                best, bi = d, idx
        if best <= th_low:
            matched[bi] = iq
            nmatches += 1
    return nmatches, np.array([m if m >= 0 else -1 for m in matched], np.int32)


def check_outputs(nm, kq, kd, qk, q, keys, desc, qd):
    """the convenience outputs agree with key_query: a query that still holds a key points at it"""
    for k in np.nonzero(kq >= 0)[0]:
        assert qk[kq[k]] == k and kd[k] == ham(qd[kq[k]], desc[k])
    assert nm >= int((kq >= 0).sum())                 # a key taken twice (first taker did not block) counts twice


def test_map_form():
    for seed, n, nq, th, ratio in [(0, 900, 700, 1.0, 0.8), (1, 500, 900, 3.0, 0.8), (2, 300, 300, 1.0, 0.6), (3, 60, 0, 1.0, 0.8),
                                   (4, 0, 40, 1.0, 0.8), (5, 1200, 400, 5.0, 0.9)]:
        keys, desc, ur, occ, q, qd = projection_case(seed, n, nq, form="map", th=th)
        nm, kq, kd, qk = O.search_by_projection(keys, desc, ur, occ, BOUNDS, q, qd, TH_HIGH, ratio, True, False)
        rn, rkq, rscore = py_map_form(keys, desc, ur, occ, q, qd, ratio)
        assert nm == rn and np.array_equal(kq, rkq)
        assert np.array_equal(kd[kq >= 0], rscore[kq >= 0])
        check_outputs(nm, kq, kd, qk, q, keys, desc, qd)
        if n >= 300 and nq >= 300:
            assert nm > nq // 4 and (q["flags"] & 2 == 0).any()


def test_last_frame_form():
    for seed, n, nq, th in [(10, 900, 800, 15.0), (11, 600, 600, 7.0), (12, 400, 900, 15.0), (13, 800, 500, 30.0), (14, 50, 0, 15.0)]:
        keys, desc, ur, occ, q, qd = projection_case(seed, n, nq, form="last", th=th, stereo=seed != 12)
        nm, kq, kd, qk = O.search_by_projection(keys, desc, ur, occ, BOUNDS, q, qd, TH_HIGH, 0.9, False, True)
        rn, rkq = py_frame_form(keys, desc, ur, occ, q, qd, TH_HIGH, False)
        assert nm == rn and np.array_equal(kq, rkq)
        if nq >= 500:
            assert nm > nq // 4 and (kq == -2).any()


def test_key_frame_form():
    for seed, n, nq, th, orb_dist in [(20, 900, 700, 10.0, 100), (21, 500, 800, 3.0, 64), (22, 700, 300, 10.0, 100)]:
        keys, desc, ur, occ, q, qd = projection_case(seed, n, nq, form="keyframe", th=th)
        nm, kq, kd, qk = O.search_by_projection(keys, desc, ur, occ, BOUNDS, q, qd, orb_dist, 0.9, False, True)
        rn, rkq = py_frame_form(keys, desc, ur, occ, q, qd, orb_dist, True)
        assert nm == rn and np.array_equal(kq, rkq)
        assert nm == int((kq >= 0).sum()) > nq // 8          # every taker blocks: no key is taken twice


def test_rotation_check_off_and_no_right_coordinates():
    keys, desc, ur, occ, q, qd = projection_case(30, 500, 500, form="last", th=15.0)
    nm, kq, _, qk = O.search_by_projection(keys, desc, None, None, BOUNDS, q, qd, TH_HIGH, 0.9, False, False)
    assert (kq != -2).all() and nm == int((qk >= 0).sum())
    nm2, kq2, _, _ = O.search_by_projection(keys, desc, np.full(len(keys), -1, F32), np.zeros(len(keys), np.uint8), BOUNDS, q, qd,
                                            TH_HIGH, 0.9, False, False)
    assert nm2 == nm and np.array_equal(kq, kq2)


def test_sim3_form():
    for seed, n, nq, th in [(40, 900, 700, 10.0), (41, 500, 900, 4.0), (42, 1200, 300, 10.0)]:
        keys, desc, ur, occ, q, qd = projection_case(seed, n, nq, form="sim3", th=th)
        nm, kq, kd, qk = O.search_by_projection(keys, desc, None, occ, BOUNDS, q, qd, 50, 0.9, False, False)
        rn, rkq = py_sim3_form(keys, desc, occ, q, qd, 50)
        assert nm == rn > nq // 10 and np.array_equal(kq, rkq)


def py_fuse_form(keys, desc, ur, q, qd, inv_sig2, th_low):
    """src/ORBmatcher.cc:937-1087, the candidate search: bestIdx per map point (-1: none within TH_LOW)"""
    grid = py_grid(keys, BOUNDS)
    out = []
    for iq in range(len(q)):
        if not q[iq]["flags"] & 1:
            out.append(-1)
            continue
        u, v, urq = F32(q[iq]["x"]), F32(q[iq]["y"]), F32(q[iq]["xr"])
        lvl = int(q[iq]["max_level"])
        best, bi = 256, -1
        for idx in py_area(keys, grid, u, v, F32(q[iq]["r"]), -1, -1):
            kl = int(keys[idx]["octave"])
            if kl < lvl - 1 or kl > lvl:
                continue
            ex, ey = F32(u - keys[idx]["x"]), F32(v - keys[idx]["y"])
            if ur[idx] >= 0:
                er = F32(urq - ur[idx])
                e2 = F32(F32(F32(ex * ex) + F32(ey * ey)) + F32(er * er))
                if float(F32(e2 * inv_sig2[kl])) > 7.8:
                    continue
            else:
                e2 = F32(F32(ex * ex) + F32(ey * ey))
                if float(F32(e2 * inv_sig2[kl])) > 5.99:
                    continue
            d = ham(qd[iq], desc[idx])
            if d < best:
                best, bi = d, idx
        out.append(bi if best <= th_low else -1)
    return np.array(out, np.int32)


def test_fuse_form():
    sf = (F32(1.2) ** np.arange(8)).astype(F32)
    inv = (F32(1) / (sf * sf)).astype(F32)
    for seed, n, nq, th in [(60, 900, 700, 3.0), (61, 500, 900, 3.0), (62, 1200, 300, 2.5)]:
        keys, desc, ur, occ, q, qd = projection_case(seed, n, nq, form="fuse", th=th)
        nm, kq, kd, qk = O.search_by_projection(keys, desc, ur, None, BOUNDS, q, qd, 50, 0.9, False, False, inv)
        want = py_fuse_form(keys, desc, ur, q, qd, inv, 50)
        assert np.array_equal(qk, want) and nm == int((want >= 0).sum()) > nq // 10
