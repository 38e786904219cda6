"""GPU parity for ORBmatcher::SearchByProjection (src/ORBmatcher.cc:155-241, 1440-1593, 1595-1721)
through the C ABI against the sequential CPU oracle (oracle/frame_grid.cpp: orc_search_by_projection, itself checked
against a literal per-form Python restatement in tests/test_oracle_projsearch.py).  Integer / index work and float
compares on identical float sequences: the bar is bit-exact, including which of two equal minima wins and every
take-over the reference's in-order loop produces."""
import importlib

import numpy as np
import pytest

from synth import projection_case

pytestmark = pytest.mark.gpu
BOUNDS = (0.0, 752.0, 0.0, 480.0)
F32 = np.float32

FORMS = {                      # form -> (th_dist, use_ratio, check_ori, nnratio)
    "map": (100, True, False, 0.8),
    "last": (100, False, True, 0.9),
    "keyframe": (100, False, True, 0.9),
    "sim3": (50, False, False, 0.9),              # :406-518, loop closing: TH_LOW, no ratio, no rotation check
}


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.fixture(scope="module")
def O():
    from oracle import pyoracle
    return pyoracle


def _check(gf, O, case, th_dist, use_ratio, ori, ratio, no_right=False, no_occ=False):
    keys, desc, ur, occ, q, qd = case
    ur = None if no_right else ur
    occ = None if no_occ else occ
    want = O.search_by_projection(keys, desc, ur, occ, BOUNDS, q, qd, th_dist, ratio, use_ratio, ori)
    m = gf.ORBmatcher(ratio, ori)
    got = m.SearchByProjection(keys, desc, q, qd, ur, occ, th_dist, use_ratio, ori, bounds=BOUNDS)
    assert got[0] == want[0]
    for g, w, name in zip(got[1:], want[1:], ("key_query", "key_dist", "query_key")):
        assert np.array_equal(g, w), name
    return want


@pytest.mark.parametrize("form", ["map", "last", "keyframe", "sim3"])
def test_forms_bit_exact(gf, O, form):
    th_dist, use_ratio, ori, ratio = FORMS[form]
    for seed, n, nq, th in [(0, 1000, 1000, 1.0), (1, 2000, 1500, 3.0), (2, 700, 2500, 15.0), (3, 1000, 300, 7.0),
                            (4, 8192, 8192, 5.0), (5, 400, 900, 30.0)]:
        th = th if form == "map" else max(th, 3.0)
        case = projection_case(seed, n, nq, form=form, th=th)
        nm, kq, kd, qk = _check(gf, O, case, th_dist, use_ratio, ori, ratio)
        assert nm > nq // 10
        if ori:
            assert (kq == -2).any()


def test_heavy_competition(gf, O):
    """Every query aims at one of 40 keys: nearly every step of the in-order replay meets a taken key."""
    for form in ("map", "last", "keyframe"):
        th_dist, use_ratio, ori, ratio = FORMS[form]
        keys, desc, ur, occ, q, qd = projection_case(40, 600, 1200, form=form, th=10.0, p_dup=0.0, p_free=0.3)
        rng = np.random.default_rng(5)
        tgt = rng.integers(0, 40, len(q))
        q["x"], q["y"] = keys["x"][tgt] + F32(0.5), keys["y"][tgt] - F32(0.25)
        q["r"] = F32(25.0)
        q["min_level"], q["max_level"] = 0, 7
        q["xr"] = np.where(ur[tgt] > 0, ur[tgt], 0).astype(F32)
        qd[:] = desc[tgt] ^ np.packbits(rng.random((len(q), 256)) < 0.04, axis=1)
        nm, kq, kd, qk = _check(gf, O, (keys, desc, ur, occ, q, qd), th_dist, use_ratio, ori, ratio)
        assert nm > 10


def test_edges(gf, O):
    c = projection_case(50, 500, 400, form="last", th=15.0)
    _check(gf, O, c, 100, False, False, 0.9)                                  # rotation check off
    _check(gf, O, c, 100, False, True, 0.9, no_right=True, no_occ=True)       # monocular frame, nothing held yet
    _check(gf, O, c, 30, True, True, 0.6)                                     # ratio AND rotation, tight threshold
    _check(gf, O, projection_case(51, 300, 0, form="map"), 100, True, False, 0.8)      # no queries
    _check(gf, O, projection_case(52, 0, 200, form="map"), 100, True, False, 0.8)      # empty frame
    keys, desc, ur, occ, q, qd = projection_case(53, 300, 300, form="map")
    q["flags"] &= ~1                                                                   # nothing in view
    nm, kq, _, qk = _check(gf, O, (keys, desc, ur, occ, q, qd), 100, True, False, 0.8)
    assert nm == 0 and (kq == -1).all() and (qk == -1).all()
    occ[:] = 1                                                                         # every key already held
    q["flags"] |= 1
    nm, kq, _, _ = _check(gf, O, (keys, desc, ur, occ, q, qd), 100, True, False, 0.8)
    assert nm == 0
    # complementary descriptors: distance 256 never beats the initial bestDist = 256 (:183, :1510)
    keys, desc, ur, occ, q, qd = projection_case(54, 200, 200, form="keyframe", th=10.0)
    qd[:100] = ~desc[np.arange(100) % 200]
    _check(gf, O, (keys, desc, ur, occ, q, qd), 255, False, True, 0.9)
    m = gf.ORBmatcher(0.8, False)
    with pytest.raises(gf.GforbError):                       # the reference would write mvpMapPoints[-1]
        m.SearchByProjection(keys, desc, q, qd, th_dist=256, bounds=BOUNDS)
    with pytest.raises(gf.GforbError):
        big = projection_case(55, 10, 8193, form="map")
        m.SearchByProjection(big[0], big[1], big[4], big[5], bounds=BOUNDS)


def test_on_extracted_frames(gf, O, fixtures):
    """Frames as the reference builds them: the right image's keypoints projected into the left frame (the stereo
    fixtures stand in for consecutive frames), last-frame form with th = 15 (src/Tracking.cc: TrackWithMotionModel)."""
    scale = (F32(1.2) ** np.arange(8)).astype(F32)
    for a, b in (("EuRoC_l", "EuRoC_r"), ("TUM_l", "TUM_r")):
        ex = gf.ORBextractor(2000)
        k1, d1 = ex(fixtures[a]); k2, d2 = ex(fixtures[b])
        h, w = fixtures[a].shape
        q = np.zeros(len(k2), gf.PROJ_QUERY_DTYPE)
        q["x"], q["y"], q["angle"] = k2["x"], k2["y"], k2["angle"]
        q["r"] = F32(15.0) * scale[k2["octave"]]
        q["min_level"], q["max_level"] = k2["octave"] - 1, k2["octave"] + 1
        q["xr"] = -1
        q["flags"] = gf.PQ_ACTIVE | gf.PQ_BLOCKS
        bounds = (0.0, float(w), 0.0, float(h))
        want = O.search_by_projection(k1, d1, None, None, bounds, q, d2, 100, 0.9, False, True)
        got = gf.ORBmatcher(0.9, True).SearchByProjection(k1, d1, q, d2, None, None, 100, False, True, bounds=bounds)
        assert got[0] == want[0] > 50
        for g, wv in zip(got[1:], want[1:]):
            assert np.array_equal(g, wv)


def test_batched_device_form_equals_single(gf, O):
    rng = np.random.default_rng(17)
    P, KS, QS = 7, 1300, 1500
    keys = np.zeros((P, KS), gf.KP_DTYPE); desc = np.zeros((P, KS, 32), np.uint8)
    ur = np.zeros((P, KS), F32); occ = np.zeros((P, KS), np.uint8)
    q = np.zeros((P, QS), gf.PROJ_QUERY_DTYPE); qd = np.zeros((P, QS, 32), np.uint8)
    n = np.zeros(P, np.int32); nq = np.zeros(P, np.int32)
    ref = []
    for p in range(P):
        a, b = (0, 0) if p == 3 else (int(rng.integers(200, KS + 1)), int(rng.integers(200, QS + 1)))
        if p == 5:
            a, b = KS, QS
        n[p], nq[p] = a, b
        c = projection_case(100 + p, a, b, form="last", th=15.0)
        keys[p, :a], desc[p, :a], ur[p, :a], occ[p, :a], q[p, :b], qd[p, :b] = c
        ref.append(O.search_by_projection(*c[:4], BOUNDS, c[4], c[5], 100, 0.9, False, True))
    D = gf.DeviceArray.from_host
    dk, dd, du, do, dn, dq, dqd, dnq = D(keys), D(desc), D(ur), D(occ), D(n), D(q), D(qd), D(nq)
    dkq = gf.DeviceArray((P, KS), np.int32); dkd = gf.DeviceArray((P, KS), np.int32)
    dqk = gf.DeviceArray((P, QS), np.int32); dnm = gf.DeviceArray((P,), np.int32)
    m = gf.ORBmatcher(0.9, True)
    m.SearchByProjectionBatchDevice(BOUNDS, P, dk, dd, du, do, dn, KS, dq, dqd, dnq, QS, 100, False, True, dkq, dkd, dqk, dnm)
    assert m.sync() > 0
    kq, kd, qk, nm = dkq.to_host(), dkd.to_host(), dqk.to_host(), dnm.to_host()
    for p in range(P):
        on, okq, okd, oqk = ref[p]
        assert nm[p] == on and np.array_equal(kq[p, :n[p]], okq) and np.array_equal(qk[p, :nq[p]], oqk), p
        assert np.array_equal(kd[p, :n[p]], okd), p


def test_device_resident_chain_from_the_batch_extractor(gf, O):
    """extract on the device -> search on the device: the frame side (counts, keypoints, descriptors) is read where
    gforb_batch left it (gforb_batch_device_results), every eye image as a frame, the other eye's keypoints as queries"""
    from synth import synth_stereo
    F, W, H = 3, 752, 480
    frames = [synth_stereo(200 + f, W, H) for f in range(F)]
    left = np.ascontiguousarray(np.stack([a for a, _ in frames])); right = np.ascontiguousarray(np.stack([b for _, b in frames]))
    sb = gf.StereoBatch(W, H, F, 1000, 1.2, 8, 20, 7)
    out = sb.alloc_outputs(F, pinned=False)
    sb.process(left, right, out, 47.9, 0.11)
    K = sb.K
    dev = sb.device_results()
    assert dev["K"] == K and all(dev[k] for k in ("counts", "keypoints", "descriptors", "u_right", "depth"))
    P = 2 * F
    n = out["counts"]
    scale = (F32(1.2) ** np.arange(8)).astype(F32)
    q = np.zeros((P, K), gf.PROJ_QUERY_DTYPE); qd = np.zeros((P, K, 32), np.uint8); nq = np.zeros(P, np.int32)
    for i in range(P):
        o = i ^ 1
        ko = out["keypoints"][o, :n[o]]
        nq[i] = n[o]
        for name in ("x", "y", "angle"):
            q[name][i, :n[o]] = ko[name]
        q["r"][i, :n[o]] = F32(15.0) * scale[ko["octave"]]
        q["min_level"][i, :n[o]] = ko["octave"] - 1; q["max_level"][i, :n[o]] = ko["octave"] + 1
        q["flags"][i, :n[o]] = gf.PQ_ACTIVE | gf.PQ_BLOCKS
        qd[i, :n[o]] = out["descriptors"][o, :n[o]]
    D = gf.DeviceArray.from_host
    dq, dqd, dnq = D(q), D(qd), D(nq)
    dkq = gf.DeviceArray((P, K), np.int32); dkd = gf.DeviceArray((P, K), np.int32)
    dqk = gf.DeviceArray((P, K), np.int32); dnm = gf.DeviceArray((P,), np.int32)
    m = gf.ORBmatcher(0.9, True)
    bounds = (0.0, float(W), 0.0, float(H))
    m.SearchByProjectionBatchDevice(bounds, P, dev["keypoints"], dev["descriptors"], None, None, dev["counts"], K, dq, dqd, dnq, K,
                                    100, False, True, dkq, dkd, dqk, dnm)
    assert m.sync() > 0
    kq, qk, nm = dkq.to_host(), dqk.to_host(), dnm.to_host()
    for i in range(P):
        want = O.search_by_projection(out["keypoints"][i, :n[i]], out["descriptors"][i, :n[i]], None, None, bounds,
                                      q[i, :nq[i]], qd[i, :nq[i]], 100, 0.9, False, True)
        assert nm[i] == want[0] > 20 and np.array_equal(kq[i, :n[i]], want[1]) and np.array_equal(qk[i, :nq[i]], want[3]), i


def test_fuse_candidate_search(gf, O):
    """ORBmatcher::Fuse (src/ORBmatcher.cc:937-1087): reprojection-error gate in the reference's float arithmetic, queries
    neither block nor see occupancy"""
    sf = (F32(1.2) ** np.arange(8)).astype(F32)
    inv = (F32(1) / (sf * sf)).astype(F32)
    for seed, n, nq, th in [(60, 1000, 1000, 3.0), (61, 2000, 1500, 3.0), (62, 700, 2500, 2.5), (63, 8192, 8192, 3.0)]:
        keys, desc, ur, occ, q, qd = projection_case(seed, n, nq, form="fuse", th=th)
        want = O.search_by_projection(keys, desc, ur, None, BOUNDS, q, qd, 50, 0.9, False, False, inv)
        m = gf.ORBmatcher(0.9, False)
        m.SetInvLevelSigma2(inv)
        got = m.SearchByProjection(keys, desc, q, qd, ur, None, 50, False, False, bounds=BOUNDS)
        assert got[0] == want[0] > nq // 10
        for g, w in zip(got[1:], want[1:]):
            assert np.array_equal(g, w)
