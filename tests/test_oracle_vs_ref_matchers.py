"""The pose-driven matcher forms: the restatement (oracle/frame_grid.cpp) against THE REFERENCE'S OWN CODE.

oracle/_ref/libgforb_ref.so carries, lifted verbatim by line range (oracle/ref_extract.py):
    ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono, ...)      src/ORBmatcher.cc:1440-1593
    ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)  :1595-1721
    ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)           :1214-1438
    ORBmatcher::SearchByBoW(pKF, F, ...) / (pKF1, pKF2, ...)                       :270-404, :635-768
    ORBmatcher::SearchForTriangulation + CheckDistEpipolarLine                     :770-935, :251-268
    MapPoint::PredictScale x2, Get{Min,Max}DistanceInvariance                       src/MapPoint.cc:530-574
    KeyFrame::GetFeaturesInArea, IsInImage                                          src/KeyFrame.cc:877-921
Those functions derive their windows themselves -- from poses, world points and per-point distance limits -- while the
restatement (like the product's C ABI) takes already-projected queries.  Each test builds a small 3D scene, lets the
reference's own text run on it, derives the same queries in numpy float32 (the arithmetic of the reference's preamble,
operation for operation; cv::Mat products accumulate in double as in oracle/ref_shim/cvshim.hpp) and demands that the
restatement assigns exactly the same map point to every key.
"""
import ctypes as C
import math

import numpy as np
import pytest

from oracle import pyoracle as O
from oracle import pyref
from synth import KEYPOINT, PROJ_QUERY, bow_case, triangulation_case

pytestmark = pytest.mark.skipif(not pyref.available("ref"), reason="oracle/_ref is not built and /root/reference is absent")

F32 = np.float32
BOUNDS = (0.0, 752.0, 0.0, 480.0)
CAM = np.array([458.654, 457.296, 367.215, 248.375, 47.90639384423901, 47.90639384423901 / 458.654], F32)   # fx fy cx cy mbf mb
vp, i32, f32 = C.c_void_p, C.c_int, C.c_float


@pytest.fixture(scope="module")
def L():
    R = pyref.load("ref")
    lib = R.lib()
    lib.ref_search_by_projection_last.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, vp, i32, f32, i32, i32, vp]
    lib.ref_search_by_projection_kf.argtypes = [vp, vp, vp, i32, vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, i32, f32, i32, i32, vp]
    lib.ref_search_by_sim3.argtypes = [vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, f32, vp, vp, vp, vp, vp, vp, vp, vp, i32, f32, vp]
    lib.ref_search_by_bow_kf_f.argtypes = [vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, f32, i32, vp]
    lib.ref_search_by_bow_kf_kf.argtypes = [vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, f32, i32, vp]
    lib.ref_search_for_triangulation.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32,
                                                 vp, vp, vp, i32, i32, vp, i32]
    return lib


def P(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def A(a, dt):
    return np.ascontiguousarray(a, dt)


# ---- the arithmetic of oracle/ref_shim/cvshim.hpp, in numpy -----------------------------------------------------------
def mm(a, b):
    """CV_32F product: each dot product accumulated in double in k order, stored as float"""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    if b.ndim == 1:
        b = b[:, None]
    out = np.zeros((a.shape[0], b.shape[1]))
    for k in range(a.shape[1]):
        out = out + a[:, k:k + 1] * b[k:k + 1, :]
    return out.astype(F32)


def norm2(v):
    s = 0.0
    for x in np.asarray(v, F32).ravel():
        s += float(x) * float(x)
    return math.sqrt(s)


def scale_factors():
    return O.Extractor().scale.astype(F32)          # mvScaleFactor, src/ORBextractor.cc:414-422


def rodrigues(rv):
    th = float(np.linalg.norm(rv)); k = np.asarray(rv, np.float64) / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)


def make_pose(rng, rot=0.03, trans=0.3):
    T = np.eye(4)
    T[:3, :3] = rodrigues(rng.normal(0, rot, 3) + 1e-4)
    T[:3, 3] = rng.normal(0, trans, 3)
    return T.astype(F32)


def project(T, X):
    """x3Dc = Rcw * x3Dw + tcw; u = fx * xc * invzc + cx (src/ORBmatcher.cc:1473-1483, all float)"""
    Rcw, tcw = T[:3, :3], T[:3, 3]
    c = (mm(Rcw, X)[:, 0] + tcw).astype(F32)
    invz = F32(1.0 / float(c[2]))
    u = F32(F32(F32(CAM[0] * c[0]) * invz) + CAM[2])
    v = F32(F32(F32(CAM[1] * c[1]) * invz) + CAM[3])
    return c, invz, u, v


def predict_scale(maxd, dist, nlevels, sf1):
    """MapPoint::PredictScale (src/MapPoint.cc:542-574); the scenes keep log(ratio) / logScale half a level away from an integer"""
    ratio = F32(F32(maxd) / F32(dist))
    val = math.log(float(ratio)) / math.log(float(sf1))
    assert abs(val - round(val)) > 0.2, "scene construction keeps PredictScale away from a rounding boundary"
    n = math.ceil(val)
    return 0 if n < 0 else (nlevels - 1 if n >= nlevels else n)


class Scene:
    """np map points seen from a camera at Tcw: world positions, descriptors, the level each is seen at, and a frame of keys
    that re-observes most of them (a few pixels off, noisy descriptor, octave off by one now and then) plus clutter."""

    def __init__(self, seed, npts=500, nclutter=250, w=752, h=480):
        rng = self.rng = np.random.default_rng(31000 + seed)
        self.sf = scale_factors()
        self.Tcw = make_pose(rng)
        Rcw, tcw = self.Tcw[:3, :3].astype(np.float64), self.Tcw[:3, 3].astype(np.float64)
        uv = np.stack([rng.random(npts) * (w + 60) - 30, rng.random(npts) * (h + 60) - 30], 1)
        z = rng.random(npts) * 18 + 2
        z[rng.random(npts) < 0.03] *= -1                                   # a few points behind the camera
        Xc = np.stack([(uv[:, 0] - CAM[2]) / CAM[0] * z, (uv[:, 1] - CAM[3]) / CAM[1] * z, z], 1)
        self.wpos = A(((Xc - tcw) @ Rcw), F32)                             # Rcw^T (Xc - tcw)
        self.pdesc = rng.integers(0, 256, (npts, 32), dtype=np.uint8)
        self.level = np.minimum(rng.geometric(0.35, npts) - 1, 7)
        self.pflag = (1 | np.where(rng.random(npts) < 0.04, 2, 0) | np.where(rng.random(npts) < 0.8, 4, 0)).astype(np.uint8)
        self.npts = npts
        # the observing frame
        seen = rng.random(npts) < 0.85
        n = int(seen.sum()) + nclutter
        k = np.zeros(n, KEYPOINT)
        d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        idx = np.nonzero(seen)[0]
        m = len(idx)
        k["x"][:m] = (uv[idx, 0] + rng.normal(0, 2.0, m)).astype(F32); k["y"][:m] = (uv[idx, 1] + rng.normal(0, 2.0, m)).astype(F32)
        k["octave"][:m] = np.clip(self.level[idx] + rng.choice([0, 0, 0, 1, -1], m), 0, 7)
        d[:m] = self.pdesc[idx] ^ np.packbits(rng.random((m, 256)) < 0.06, axis=1)
        k["x"][m:] = (rng.random(nclutter) * w).astype(F32); k["y"][m:] = (rng.random(nclutter) * h).astype(F32)
        k["octave"][m:] = rng.integers(0, 8, nclutter)
        k["angle"] = (rng.random(n) * 360).astype(F32)
        k["size"], k["class_id"] = 31, -1
        perm = rng.permutation(n)
        self.keys, self.desc = np.ascontiguousarray(k[perm]), np.ascontiguousarray(d[perm])
        self.key_of_point = np.full(npts, -1)
        inv = np.argsort(perm)
        self.key_of_point[idx] = inv[:m]
        self.n = n


def expected_points(key_query, query_point, cur_point):
    """what CurrentFrame.mvpMapPoints holds afterwards: the matched query's point, NULL where the rotation check removed a
    match, the pre-existing point where nothing happened"""
    out = np.where(key_query >= 0, query_point[np.maximum(key_query, 0)], np.where(key_query == -2, -1, cur_point))
    return out.astype(np.int32)


@pytest.mark.parametrize("seed,th,mono,ori", [(0, 15.0, 0, 1), (1, 7.0, 0, 1), (2, 15.0, 1, 1), (3, 30.0, 0, 0), (4, 15.0, 0, 1), (5, 4.0, 1, 0)])
def test_search_by_projection_last_frame(L, seed, th, mono, ori):
    S = Scene(seed)
    rng = S.rng
    sf = S.sf
    # the last frame: pose a little way off (forward / backward / sideways by seed), one key per surviving point
    Tlw = S.Tcw.copy()
    Tlw[:3, 3] += np.array([[0, 0, 0.5], [0, 0, -0.5], [0.2, 0, 0.02]][seed % 3], F32)
    held = np.nonzero(rng.random(S.npts) < 0.9)[0]
    nlast = len(held) + 40
    lkeys = np.zeros(nlast, KEYPOINT)
    last_point = np.full(nlast, -1, np.int32)
    last_point[:len(held)] = held
    lkeys["octave"][:len(held)] = S.level[held]
    lkeys["octave"][len(held):] = rng.integers(0, 8, 40)
    ck = S.key_of_point[held]
    lkeys["angle"][:len(held)] = np.where(ck >= 0, np.mod(S.keys["angle"][np.maximum(ck, 0)] + rng.normal(12, 4, len(held)), 360), rng.random(len(held)) * 360)
    lkeys["angle"][len(held):] = rng.random(40) * 360
    lkeys["x"], lkeys["y"] = (rng.random(nlast) * 752).astype(F32), (rng.random(nlast) * 480).astype(F32)
    lkeys["size"], lkeys["class_id"] = 31, -1
    last_outlier = (rng.random(nlast) < 0.05).astype(np.uint8)
    cu_right = np.where(rng.random(S.n) < 0.6, S.keys["x"] - (rng.random(S.n) * 40 + 1), -1).astype(F32)
    cur_point = np.where(rng.random(S.n) < 0.08, rng.integers(0, S.npts, S.n), -1).astype(np.int32)
    key_point = np.empty(S.n, np.int32)
    nm = L.ref_search_by_projection_last(P(S.keys), P(S.desc), P(cu_right), P(cur_point), S.n, P(lkeys), P(last_point), P(last_outlier), nlast,
                                         P(A(BOUNDS, F32)), P(CAM), P(sf), 8, P(A(S.Tcw, F32)), P(A(Tlw, F32)), P(S.wpos), P(S.pdesc),
                                         P(S.pflag), S.npts, th, mono, ori, P(key_point))
    # ---- the same preamble in numpy (src/ORBmatcher.cc:1452-1508)
    Rcw, tcw = S.Tcw[:3, :3], S.Tcw[:3, 3]
    twc = mm(-Rcw.T, tcw)[:, 0]
    tlc = (mm(Tlw[:3, :3], twc)[:, 0] + Tlw[:3, 3]).astype(F32)
    fwd = bool(tlc[2] > CAM[5]) and not mono
    bwd = bool(-tlc[2] > CAM[5]) and not mono
    q = np.zeros(nlast, PROJ_QUERY); qd = np.zeros((nlast, 32), np.uint8)
    for i in range(nlast):
        p = last_point[i]
        if p < 0 or last_outlier[i]:
            continue
        c, invz, u, v = project(S.Tcw, S.wpos[p])
        if invz < 0 or u < BOUNDS[0] or u > BOUNDS[1] or v < BOUNDS[2] or v > BOUNDS[3]:
            continue
        o = int(lkeys["octave"][i])
        lo, hi = (o, -1) if fwd else ((0, o) if bwd else (o - 1, o + 1))
        q[i] = (u, v, F32(F32(th) * sf[o]), F32(u - F32(CAM[4] * invz)), lkeys["angle"][i], lo, hi,
                O.PQ_ACTIVE | O.PQ_RIGHT_GATE | (O.PQ_BLOCKS if S.pflag[p] & 4 else 0))
        qd[i] = S.pdesc[p]
    occupied = ((cur_point >= 0) & ((S.pflag[np.maximum(cur_point, 0)] & 4) != 0)).astype(np.uint8)
    onm, kq, _, _ = O.search_by_projection(S.keys, S.desc, cu_right, occupied, BOUNDS, q, qd, 100, 0.9, False, bool(ori))
    assert (q["flags"] != 0).sum() > 200 and onm > 60
    assert onm == nm
    assert np.array_equal(expected_points(kq, last_point, cur_point), key_point)


@pytest.mark.parametrize("seed,th,orbdist,ori", [(0, 10.0, 100, 1), (1, 3.0, 64, 1), (2, 10.0, 100, 0), (3, 20.0, 100, 1)])
def test_search_by_projection_key_frame(L, seed, th, orbdist, ori):
    S = Scene(100 + seed)
    rng = S.rng
    sf = S.sf
    Rcw, tcw = S.Tcw[:3, :3], S.Tcw[:3, 3]
    Ow = mm(-Rcw.T, tcw)[:, 0]
    # distance limits that put each point on its level (PredictScale) -- or outside the invariance range now and then
    pdist = np.zeros((S.npts, 2), F32)
    for p in range(S.npts):
        dist = F32(norm2((S.wpos[p] - Ow).astype(F32)))
        pdist[p, 1] = F32(float(dist) * float(sf[1]) ** (S.level[p] - 0.5))
        pdist[p, 0] = F32(0.2 * float(dist) if rng.random() > 0.06 else 3.0 * float(dist))
    held = np.nonzero(rng.random(S.npts) < 0.9)[0]
    nkf = len(held) + 30
    kkeys = np.zeros(nkf, KEYPOINT)
    kf_point = np.full(nkf, -1, np.int32); kf_point[:len(held)] = held
    ck = S.key_of_point[held]
    kkeys["angle"][:len(held)] = np.where(ck >= 0, np.mod(S.keys["angle"][np.maximum(ck, 0)] + rng.normal(12, 4, len(held)), 360), rng.random(len(held)) * 360)
    kkeys["x"], kkeys["y"] = (rng.random(nkf) * 752).astype(F32), (rng.random(nkf) * 480).astype(F32)
    kkeys["size"], kkeys["class_id"] = 31, -1
    already = (rng.random(S.npts) < 0.07).astype(np.uint8)
    cur_point = np.where(rng.random(S.n) < 0.08, rng.integers(0, S.npts, S.n), -1).astype(np.int32)
    key_point = np.empty(S.n, np.int32)
    nm = L.ref_search_by_projection_kf(P(S.keys), P(S.desc), P(cur_point), S.n, P(kkeys), P(kf_point), nkf, P(already), P(A(BOUNDS, F32)),
                                       P(CAM), P(sf), 8, P(A(S.Tcw, F32)), P(S.wpos), P(S.pdesc), P(S.pflag), P(pdist), S.npts, th, orbdist,
                                       ori, P(key_point))
    q = np.zeros(nkf, PROJ_QUERY); qd = np.zeros((nkf, 32), np.uint8)
    for i in range(nkf):
        p = kf_point[i]
        if p < 0 or (S.pflag[p] & 2) or already[p]:
            continue
        c, invz, u, v = project(S.Tcw, S.wpos[p])
        if u < BOUNDS[0] or u > BOUNDS[1] or v < BOUNDS[2] or v > BOUNDS[3] or not np.isfinite(u):
            continue
        dist = F32(norm2((S.wpos[p] - Ow).astype(F32)))
        if dist < F32(F32(0.8) * pdist[p, 0]) or dist > F32(F32(1.2) * pdist[p, 1]):
            continue
        lvl = predict_scale(pdist[p, 1], dist, 8, sf[1])
        q[i] = (u, v, F32(F32(th) * sf[lvl]), 0, kkeys["angle"][i], lvl - 1, lvl + 1, O.PQ_ACTIVE | O.PQ_BLOCKS)
        qd[i] = S.pdesc[p]
    occupied = (cur_point >= 0).astype(np.uint8)
    onm, kq, _, _ = O.search_by_projection(S.keys, S.desc, None, occupied, BOUNDS, q, qd, orbdist, 0.9, False, bool(ori))
    assert (q["flags"] != 0).sum() > 150 and onm > 40
    assert onm == nm
    assert np.array_equal(expected_points(kq, kf_point, cur_point), key_point)


@pytest.mark.parametrize("seed,th", [(0, 7.5), (1, 7.5), (2, 15.0), (3, 3.0)])
def test_search_by_sim3(L, seed, th):
    S1, S2 = Scene(200 + seed, npts=450), Scene(300 + seed, npts=450)
    rng = S1.rng
    sf = S1.sf
    # KF2 lives in a world that is KF1's world seen through a Sim3: X_c1 = s12 R12 X_c2 + t12 links the camera frames
    s12 = F32(1.0 + rng.normal(0, 0.05))
    R12 = rodrigues(rng.normal(0, 0.02, 3) + 1e-4).astype(F32)
    t12 = rng.normal(0, 0.1, 3).astype(F32)
    np1, np2 = S1.npts, S2.npts
    wpos = np.concatenate([S1.wpos, S2.wpos]); pdesc = np.concatenate([S1.pdesc, S2.pdesc]); pflag = np.concatenate([S1.pflag, S2.pflag])
    kf1_point = np.full(S1.n, -1, np.int32); kf2_point = np.full(S2.n, -1, np.int32)
    for p in range(np1):
        if S1.key_of_point[p] >= 0 and rng.random() < 0.9:
            kf1_point[S1.key_of_point[p]] = p
    for p in range(np2):
        if S2.key_of_point[p] >= 0 and rng.random() < 0.9:
            kf2_point[S2.key_of_point[p]] = np1 + p
    # make the two key frames look at the same scene: KF2's keys / descriptors re-observe KF1's points through the Sim3
    T1, T2 = S1.Tcw, S2.Tcw
    sR21 = (np.float64(1.0 / float(s12)) * R12.T.astype(np.float64)).astype(F32)         # (1.0 / s12) * R12.t()
    t21 = mm(-sR21, t12)[:, 0]
    sR12 = (np.float64(float(s12)) * R12.astype(np.float64)).astype(F32)
    pdist = np.zeros((np1 + np2, 2), F32)
    q12 = np.zeros(S1.n, PROJ_QUERY); q21 = np.zeros(S2.n, PROJ_QUERY)
    qd12 = np.zeros((S1.n, 32), np.uint8); qd21 = np.zeros((S2.n, 32), np.uint8)
    matches_in = np.where((kf1_point >= 0) & (rng.random(S1.n) < 0.1), kf1_point, -1).astype(np.int32)     # already matched: skipped
    obs2 = np.full(np1 + np2, -1, np.int32)
    for i1 in np.nonzero(matches_in >= 0)[0][::2]:
        obs2[matches_in[i1]] = rng.integers(0, S2.n)                                      # ... and seen in KF2 at that key
    done2 = np.zeros(S2.n, bool)
    done2[obs2[matches_in[matches_in >= 0]][obs2[matches_in[matches_in >= 0]] >= 0]] = True

    def one_way(kf_point, Tw, sR, t, n_other_sf, q, qd, skip):
        for i in range(len(kf_point)):
            p = kf_point[i]
            if p < 0 or skip[i] or (pflag[p] & 2):
                continue
            c1 = (mm(Tw[:3, :3], wpos[p])[:, 0] + Tw[:3, 3]).astype(F32)
            c2 = (mm(sR, c1)[:, 0] + t).astype(F32)
            if c2[2] < 0.0:
                continue
            invz = F32(1.0 / float(c2[2]))
            x, y = F32(c2[0] * invz), F32(c2[1] * invz)
            u, v = F32(F32(CAM[0] * x) + CAM[2]), F32(F32(CAM[1] * y) + CAM[3])
            if not (u >= 0 and u < 752 and v >= 0 and v < 480):                           # KeyFrame::IsInImage, int bounds
                continue
            dist = F32(norm2(c2))
            lvl_want = int(rng.integers(0, 8))
            pdist[p, 1] = F32(float(dist) * float(sf[1]) ** (lvl_want - 0.5))
            pdist[p, 0] = F32(0.2 * float(dist) if rng.random() > 0.05 else 3.0 * float(dist))
            if dist < F32(F32(0.8) * pdist[p, 0]) or dist > F32(F32(1.2) * pdist[p, 1]):
                continue
            lvl = predict_scale(pdist[p, 1], dist, 8, sf[1])
            q[i] = (u, v, F32(F32(th) * n_other_sf[lvl]), 0, 0, lvl - 1, lvl, O.PQ_ACTIVE)
            qd[i] = pdesc[p]

    one_way(kf1_point, T1, sR21, t21, sf, q12, qd12, matches_in >= 0)
    one_way(kf2_point, T2, sR12, t12, sf, q21, qd21, done2)
    # plant agreement: KF2 keys near where KF1's points land, carrying their descriptors, and the other way round
    k2, d2, k1, d1 = S2.keys.copy(), S2.desc.copy(), S1.keys.copy(), S1.desc.copy()
    act = np.nonzero(q12["flags"] != 0)[0]
    tgt = rng.permutation(S2.n)[:len(act)]
    for i1, j in zip(act, tgt):
        k2["x"][j], k2["y"][j] = q12["x"][i1] + F32(rng.normal(0, 1.5)), q12["y"][i1] + F32(rng.normal(0, 1.5))
        k2["octave"][j] = max(int(q12["max_level"][i1]) - int(rng.integers(0, 2)), 0)
        d2[j] = qd12[i1] ^ np.packbits(rng.random(256) < 0.05)
        if q21["flags"][j]:
            k1["x"][i1], k1["y"][i1] = q21["x"][j] + F32(rng.normal(0, 1.5)), q21["y"][j] + F32(rng.normal(0, 1.5))
            k1["octave"][i1] = max(int(q21["max_level"][j]) - int(rng.integers(0, 2)), 0)
            d1[i1] = qd21[j] ^ np.packbits(rng.random(256) < 0.05)
    matches12 = np.empty(S1.n, np.int32)
    nf = L.ref_search_by_sim3(P(k1), P(d1), P(kf1_point), S1.n, P(k2), P(d2), P(kf2_point), S2.n, P(A(BOUNDS, F32)), P(CAM), P(sf), 8,
                              P(A(T1, F32)), P(A(T2, F32)), s12, P(A(R12, F32)), P(t12), P(matches_in), P(obs2), P(wpos), P(pdesc), P(pflag),
                              P(pdist), np1 + np2, th, P(matches12))
    onf, m12, vn1, vn2 = O.search_by_sim3(k1, d1, k2, d2, BOUNDS, q12, qd12, q21, qd21)
    assert (q12["flags"] != 0).sum() > 100 and (q21["flags"] != 0).sum() > 100 and onf > 20
    assert onf == nf
    want = np.where(m12 >= 0, kf2_point[np.maximum(m12, 0)], matches_in)                  # vpMatches12[i1] = vpMapPoints2[idx2]
    assert np.array_equal(want.astype(np.int32), matches12)


def _flat(fv):
    ids, st, ix = O.flatten_feature_vector(fv)
    return A(ids, np.int32), A(st, np.int32), A(ix, np.int32)


@pytest.mark.parametrize("seed,n1,n2,ori", [(0, 600, 650, 1), (1, 900, 400, 1), (2, 300, 300, 0), (3, 50, 1200, 1)])
def test_search_by_bow_both_forms(L, seed, n1, n2, ori):
    k1, d1, v1, fv1, k2, d2, v2, fv2 = bow_case(seed, n1, n2, nodes=120)
    rng = np.random.default_rng(seed)
    # map points: one per valid feature; bad ones are the features bow_case marks invalid half of the time
    pflag = []; kf1_point = np.full(n1, -1, np.int32); kf2_point = np.full(n2, -1, np.int32)
    for i in range(n1):
        if v1[i] or rng.random() < 0.5:
            kf1_point[i] = len(pflag); pflag.append(1 if v1[i] else 3)
    for j in range(n2):
        if v2[j] or rng.random() < 0.5:
            kf2_point[j] = len(pflag); pflag.append(1 if v2[j] else 3)
    pflag = A(pflag, np.uint8); npnt = len(pflag)
    pdesc = np.zeros((npnt, 32), np.uint8)
    a1, a2 = _flat(fv1), _flat(fv2)
    # (pKF, F): valid2 plays no part, th_dist = TH_LOW (<=)
    f_point = np.empty(n2, np.int32)
    nm = L.ref_search_by_bow_kf_f(P(k1), P(d1), P(kf1_point), n1, P(a1[0]), P(a1[1]), P(a1[2]), len(a1[0]), P(k2), P(d2), n2, P(a2[0]),
                                  P(a2[1]), P(a2[2]), len(a2[0]), P(A(BOUNDS, F32)), P(pdesc), P(pflag), npnt, 0.7, ori, P(f_point))
    onm, m12 = O.search_by_bow(k1, d1, v1, a1, k2, d2, None, a2, 50, 0.7, bool(ori))
    want = np.full(n2, -1, np.int32)
    want[m12[m12 >= 0]] = kf1_point[np.nonzero(m12 >= 0)[0]]
    assert onm == nm and onm > 0.15 * min(n1, n2) and np.array_equal(want, f_point)
    # (pKF1, pKF2): candidates need a good map point too, th_dist = TH_LOW - 1 (<)
    matches12 = np.empty(n1, np.int32)
    nm = L.ref_search_by_bow_kf_kf(P(k1), P(d1), P(kf1_point), n1, P(a1[0]), P(a1[1]), P(a1[2]), len(a1[0]), P(k2), P(d2), P(kf2_point), n2,
                                   P(a2[0]), P(a2[1]), P(a2[2]), len(a2[0]), P(A(BOUNDS, F32)), P(pdesc), P(pflag), npnt, 0.7, ori, P(matches12))
    onm, m12 = O.search_by_bow(k1, d1, v1, a1, k2, d2, v2, a2, 49, 0.7, bool(ori))
    want = np.where(m12 >= 0, kf2_point[np.maximum(m12, 0)], -1).astype(np.int32)
    assert onm == nm and onm > 0.1 * min(n1, n2) and np.array_equal(want, matches12)


@pytest.mark.parametrize("seed,n1,n2,only_stereo", [(0, 700, 700, 0), (1, 400, 900, 0), (2, 600, 500, 1)])
def test_search_for_triangulation(L, seed, n1, n2, only_stereo):
    k1, d1, v1, s1, fv1, k2, d2, v2, s2, fv2, F12, ex, ey, sf, sig2 = triangulation_case(seed, n1, n2, nodes=150)
    sf = scale_factors()
    # poses whose epipole (the projection of KF1's centre in KF2, :776-783) is a known float pair
    T1 = np.eye(4, dtype=F32)
    T2 = make_pose(np.random.default_rng(seed), rot=0.01, trans=0.4)
    T2[2, 3] = F32(abs(T2[2, 3]) + 0.5)
    Ow1 = mm(-T1[:3, :3].T, T1[:3, 3])[:, 0]
    c2 = (mm(T2[:3, :3], Ow1)[:, 0] + T2[:3, 3]).astype(F32)
    invz = F32(F32(1.0) / c2[2])
    ex = F32(F32(F32(CAM[0] * c2[0]) * invz) + CAM[2]); ey = F32(F32(F32(CAM[1] * c2[1]) * invz) + CAM[3])
    ur1 = np.where(s1 != 0, np.maximum(k1["x"] - 5, 0), -1).astype(F32); ur2 = np.where(s2 != 0, np.maximum(k2["x"] - 5, 0), -1).astype(F32)
    has1 = (1 - v1).astype(np.uint8); has2 = (1 - v2).astype(np.uint8)                  # valid = the key holds no map point yet
    a1, a2 = _flat(fv1), _flat(fv2)
    cap = max(n1, 1)
    pairs = np.empty((cap, 2), np.int32)
    npairs = L.ref_search_for_triangulation(P(k1), P(d1), P(has1), P(ur1), n1, P(a1[0]), P(a1[1]), P(a1[2]), len(a1[0]), P(k2), P(d2), P(has2),
                                            P(ur2), n2, P(a2[0]), P(a2[1]), P(a2[2]), len(a2[0]), P(A(BOUNDS, F32)), P(CAM), P(sf), 8,
                                            P(A(T1, F32)), P(A(T2, F32)), P(A(F12, F32)), only_stereo, 1, P(pairs), cap)
    valid1 = v1 & (s1 if only_stereo else 1)                  # bOnlyStereo drops the monocular keys of BOTH key frames (:822-824, :842-844)
    valid2 = v2 & (s2 if only_stereo else 1)
    onm, m12 = O.search_for_triangulation(k1, d1, valid1.astype(np.uint8), s1, a1, k2, d2, valid2.astype(np.uint8), s2, a2, F12, ex, ey, sf,
                                          (sf * sf).astype(F32), 50, True)
    idx = np.nonzero(m12 >= 0)[0]
    want = np.stack([idx, m12[idx]], 1).astype(np.int32)
    assert onm == npairs and npairs > 0.05 * min(n1, n2)
    assert np.array_equal(want, pairs[:npairs])


@pytest.mark.parametrize("seed,th", [(0, 10), (1, 10), (2, 4), (3, 25)])
def test_search_by_projection_loop_closing(L, seed, th):
    """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:406-518): the Sim3 is decomposed, the
    candidate points pass the depth, image, invariance-range and 60-degree viewing tests, then take keys in order."""
    L.ref_search_by_projection_loop.argtypes = [vp, vp, i32, vp, vp, vp, i32, vp, vp, i32, vp, vp, vp, vp, vp, vp, i32, i32, vp]
    S = Scene(400 + seed)
    rng = S.rng
    sf = S.sf
    scale = 1.0 + 0.2 * rng.random()
    Scw = S.Tcw.astype(np.float64).copy()
    Scw[:3, :] *= scale                                                   # [s R | s t]: the same projections as Tcw
    Scw = Scw.astype(F32)
    sR = Scw[:3, :3]
    scw = F32(math.sqrt(float(sum(float(x) * float(x) for x in sR[0]))))
    Rcw = (sR.astype(np.float64) / float(scw)).astype(F32)
    tcw = (Scw[:3, 3].astype(np.float64) / float(scw)).astype(F32)
    Ow = mm(-Rcw.T, tcw)[:, 0]
    pdist = np.zeros((S.npts, 2), F32)
    pnormal = np.zeros((S.npts, 3), F32)
    for p in range(S.npts):
        PO = (S.wpos[p] - Ow).astype(F32)
        dist = F32(norm2(PO))
        pdist[p, 1] = F32(float(dist) * float(sf[1]) ** (S.level[p] - 0.5))
        pdist[p, 0] = F32(0.2 * float(dist) if rng.random() > 0.06 else 3.0 * float(dist))
        n = PO.astype(np.float64) / max(float(dist), 1e-9)
        if rng.random() < 0.15:                                           # seen from the side: fails the viewing-angle test
            n = np.cross(n, [0.3, 1.0, 0.2]); n /= np.linalg.norm(n)
        pnormal[p] = n.astype(F32)
    order = rng.permutation(S.npts)[: int(0.9 * S.npts)].astype(np.int32)
    matched_in = np.where(rng.random(S.n) < 0.1, rng.integers(0, S.npts, S.n), -1).astype(np.int32)
    matched = np.empty(S.n, np.int32)
    nm = L.ref_search_by_projection_loop(P(S.keys), P(S.desc), S.n, P(A(BOUNDS, F32)), P(CAM), P(sf), 8, P(A(Scw, F32)), P(order), len(order),
                                         P(matched_in), P(S.wpos), P(S.pdesc), P(S.pflag | 1), P(pdist), P(pnormal), S.npts, th, P(matched))
    found = set(matched_in[matched_in >= 0].tolist())
    q = np.zeros(len(order), PROJ_QUERY); qd = np.zeros((len(order), 32), np.uint8)
    for j, p in enumerate(order):
        if (S.pflag[p] & 2) or int(p) in found:
            continue
        c = (mm(Rcw, S.wpos[p])[:, 0] + tcw).astype(F32)
        if c[2] < 0.0:
            continue
        invz = F32(F32(1) / c[2])
        u = F32(F32(CAM[0] * F32(c[0] * invz)) + CAM[2]); v = F32(F32(CAM[1] * F32(c[1] * invz)) + CAM[3])
        if not (u >= 0 and u < 752 and v >= 0 and v < 480):
            continue
        PO = (S.wpos[p] - Ow).astype(F32)
        dist = F32(norm2(PO))
        if dist < F32(F32(0.8) * pdist[p, 0]) or dist > F32(F32(1.2) * pdist[p, 1]):
            continue
        dot = 0.0
        for a, b in zip(PO, pnormal[p]):
            dot += float(a) * float(b)
        if dot < 0.5 * float(dist):
            continue
        lvl = predict_scale(pdist[p, 1], dist, 8, sf[1])
        q[j] = (u, v, F32(F32(th) * sf[lvl]), 0, 0, lvl - 1, lvl, O.PQ_ACTIVE | O.PQ_BLOCKS)
        qd[j] = S.pdesc[p]
    occupied = (matched_in >= 0).astype(np.uint8)
    onm, kq, _, _ = O.search_by_projection(S.keys, S.desc, None, occupied, BOUNDS, q, qd, 50, 0.75, False, False)
    assert (q["flags"] != 0).sum() > 120 and onm > 40
    assert onm == nm
    assert np.array_equal(expected_points(kq, order, matched_in), matched)


@pytest.mark.parametrize("seed,th", [(0, 3.0), (1, 3.0), (2, 6.0), (3, 1.5)])
def test_fuse_candidate_search(L, seed, th):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:937-1087): which key each candidate point lands on (the
    reprojection-error gate in float, 7.8 / 5.99 compared in double).  The reference's map surgery behind it (AddObservation /
    AddMapPoint / Replace) is recorded by the stand-ins and has to name exactly the keys the restatement reports."""
    L.ref_fuse.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, i32, vp, vp, vp, vp, vp, vp, i32, f32, vp, i32, vp]
    S = Scene(500 + seed)
    rng = S.rng
    sf = S.sf
    Rcw, tcw = S.Tcw[:3, :3], S.Tcw[:3, 3]
    Ow = mm(-Rcw.T, tcw)[:, 0]
    pdist = np.zeros((S.npts, 2), F32); pnormal = np.zeros((S.npts, 3), F32)
    for p in range(S.npts):
        PO = (S.wpos[p] - Ow).astype(F32)
        dist = F32(norm2(PO))
        pdist[p, 1] = F32(float(dist) * float(sf[1]) ** (S.level[p] - 0.5))
        pdist[p, 0] = F32(0.2 * float(dist) if rng.random() > 0.06 else 3.0 * float(dist))
        n = PO.astype(np.float64) / max(float(dist), 1e-9)
        if rng.random() < 0.1:
            n = np.cross(n, [0.3, 1.0, 0.2]); n /= np.linalg.norm(n)
        pnormal[p] = n.astype(F32)
    # keys sit ~2 px from the projections (Scene): tighten a share of them so the error gate passes and fails both ways
    ur = np.full(S.n, -1, F32)
    for p in range(S.npts):
        k = S.key_of_point[p]
        if k < 0:
            continue
        c, invz, u, v = project(S.Tcw, S.wpos[p])
        if not np.isfinite(u) or c[2] <= 0:
            continue
        if rng.random() < 0.6:
            S.keys["x"][k] = u + F32(rng.normal(0, 0.7)); S.keys["y"][k] = v + F32(rng.normal(0, 0.7))
        if rng.random() < 0.5:
            ur[k] = F32(u - F32(CAM[4] * invz)) + F32(rng.normal(0, 1.0))
    kf_point = np.where(rng.random(S.n) < 0.3, rng.integers(0, S.npts, S.n), -1).astype(np.int32)
    order = rng.permutation(S.npts)[: int(0.9 * S.npts)].astype(np.int32)
    order[rng.random(len(order)) < 0.03] = -1
    in_kf = (rng.random(S.npts) < 0.08).astype(np.uint8)
    events = np.zeros((len(order) + 8, 3), np.int32); nev = C.c_int(0)
    nf = L.ref_fuse(P(S.keys), P(S.desc), P(ur), P(kf_point), S.n, P(A(BOUNDS, F32)), P(CAM), P(sf), 8, P(A(S.Tcw, F32)), P(order), len(order),
                    P(in_kf), P(S.wpos), P(S.pdesc), P(S.pflag), P(pdist), P(pnormal), S.npts, th, P(events), len(events), C.byref(nev))
    q = np.zeros(len(order), PROJ_QUERY); qd = np.zeros((len(order), 32), np.uint8)
    for j, p in enumerate(order):
        if p < 0 or (S.pflag[p] & 2) or in_kf[p]:
            continue
        c = (mm(Rcw, S.wpos[p])[:, 0] + tcw).astype(F32)
        if c[2] < 0.0:
            continue
        invz = F32(F32(1) / c[2])
        u = F32(F32(CAM[0] * F32(c[0] * invz)) + CAM[2]); v = F32(F32(CAM[1] * F32(c[1] * invz)) + CAM[3])
        if not (u >= 0 and u < 752 and v >= 0 and v < 480):
            continue
        xr = F32(u - F32(CAM[4] * invz))
        PO = (S.wpos[p] - Ow).astype(F32)
        dist = F32(norm2(PO))
        if dist < F32(F32(0.8) * pdist[p, 0]) or dist > F32(F32(1.2) * pdist[p, 1]):
            continue
        dot = 0.0
        for a, b in zip(PO, pnormal[p]):
            dot += float(a) * float(b)
        if dot < 0.5 * float(dist):
            continue
        lvl = predict_scale(pdist[p, 1], dist, 8, sf[1])
        q[j] = (u, v, F32(F32(th) * sf[lvl]), xr, 0, lvl - 1, lvl, O.PQ_ACTIVE | O.PQ_REPROJ_GATE)
        qd[j] = S.pdesc[p]
    inv_sig2 = (F32(1.0) / (sf * sf).astype(F32)).astype(F32)
    onm, _, _, qk = O.search_by_projection(S.keys, S.desc, ur, None, BOUNDS, q, qd, 50, 0.6, False, False, inv_sig2)
    assert (q["flags"] != 0).sum() > 120 and onm > 40
    assert onm == nf
    # replay the surgery the reference would have done, from the restatement's keys
    held = kf_point.copy()
    want = []
    for j, p in enumerate(order):
        k = qk[j]
        if k < 0:
            continue
        other = held[k]
        if other >= 0:
            if not (S.pflag[other] & 2):
                obs_other, obs_p = (1 if S.pflag[other] & 4 else 0), (1 if S.pflag[p] & 4 else 0)
                want.append((1, p, other) if obs_other > obs_p else (1, other, p))
        else:
            want.append((0, p, k)); held[k] = p
    assert nev.value == len(want)
    assert [tuple(e) for e in events[:nev.value].tolist()] == [tuple(int(x) for x in w) for w in want]


def test_compute_distinctive_descriptors(L):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:397-462), lifted verbatim: the descriptor with the least
    median distance to the others, first on ties -- sets of 1 .. 60 observations, near-duplicates, bad key frames."""
    L.ref_distinctive_descriptors.argtypes = [vp, vp, vp, i32, vp, vp]
    rng = np.random.default_rng(77)
    sizes = np.concatenate([[1, 2, 2, 3, 60], rng.integers(1, 40, 200), [0]])
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    desc = np.zeros((int(start[-1]), 32), np.uint8)
    for p, n in enumerate(sizes):
        if n:
            centre = rng.integers(0, 256, 32, dtype=np.uint8)
            desc[start[p]:start[p + 1]] = centre ^ np.packbits(rng.random((n, 256)) < rng.choice([0.02, 0.1, 0.4]), axis=1)
            if n > 3 and rng.random() < 0.3:
                desc[start[p] + 1] = desc[start[p]]                       # exact duplicates: ties
    bad = (rng.random(len(desc)) < 0.1).astype(np.uint8)
    best = np.zeros((len(sizes), 32), np.uint8); has = np.zeros(len(sizes), np.uint8)
    L.ref_distinctive_descriptors(P(desc), P(start), P(bad), len(sizes), P(best), P(has))
    # the product / restatement entry takes the descriptors of the good key frames only (the caller's filter, :418-424)
    keep = bad == 0
    kstart = np.concatenate([[0], np.cumsum([int(keep[start[p]:start[p + 1]].sum()) for p in range(len(sizes))])]).astype(np.int32)
    kdesc = np.ascontiguousarray(desc[keep])
    bi, bm = O.distinctive_descriptors(kdesc, kstart)
    for p in range(len(sizes)):
        n = kstart[p + 1] - kstart[p]
        assert bool(has[p]) == (n > 0), p
        if n:
            assert np.array_equal(best[p], kdesc[kstart[p] + bi[p]]), p
    assert has.sum() > 190
