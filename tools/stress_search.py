#!/usr/bin/env python3
"""Randomised GPU-vs-oracle stress of the sequential-semantics searches (SearchByProjection all forms, SearchByBoW both
forms, SearchForTriangulation, SearchForInitialization): many seeds, sizes, competition levels.  Exits non-zero on the
first mismatch.  usage: python tools/stress_search.py [cases]"""
import importlib, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from synth import bow_case, projection_case, triangulation_case
gf = importlib.import_module("gf-orb-slam2_b200")
from oracle import pyoracle as O

CASES = int(sys.argv[1]) if len(sys.argv) > 1 else 200
B = (0.0, 752.0, 0.0, 480.0)
rng = np.random.default_rng(2024)
sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
inv = (np.float32(1) / (sf * sf)).astype(np.float32)
FORMS = {"map": (100, True, False), "last": (100, False, True), "keyframe": (100, False, True), "sim3": (50, False, False),
         "fuse": (50, False, False)}
bad = 0
for c in range(CASES):
    form = list(FORMS)[c % 5]
    th_dist, use_ratio, ori = FORMS[form]
    n = int(rng.choice([50, 300, 1000, 2500, 5000, 8192])); nq = int(rng.choice([40, 500, 1000, 3000, 4097, 8192]))
    th = float(rng.choice([1.0, 3.0, 7.0, 15.0, 40.0]))
    case = projection_case(10_000 + c, n, nq, form=form, th=th, p_dup=float(rng.choice([0.0, 0.2, 0.6, 0.95])),
                           p_free=float(rng.choice([0.0, 0.15, 0.5])), p_occupied=float(rng.choice([0.0, 0.1, 0.5])),
                           flip=float(rng.choice([0.02, 0.07, 0.15])))
    keys, desc, ur, occ, q, qd = case
    if form == "fuse":
        occ = None
    ratio = float(rng.choice([0.6, 0.8, 0.9]))
    want = O.search_by_projection(keys, desc, ur, occ, B, q, qd, th_dist, ratio, use_ratio, ori, inv)
    m = gf.ORBmatcher(ratio, ori); m.SetInvLevelSigma2(inv)
    got = m.SearchByProjection(keys, desc, q, qd, ur, occ, th_dist, use_ratio, ori, bounds=B)
    ok = got[0] == want[0] and all(np.array_equal(g, w) for g, w in zip(got[1:], want[1:]))
    if not ok:
        print("MISMATCH projection", form, c, n, nq, th); bad += 1
for c in range(CASES // 2):
    n1 = int(rng.choice([100, 1000, 3000, 8000])); n2 = int(rng.choice([100, 1000, 3000, 8000]))
    nodes = int(rng.choice([1, 5, 50, 500, 3000]))
    k1, d1, v1, fv1, k2, d2, v2, fv2 = bow_case(20_000 + c, n1, n2, nodes, dup=float(rng.choice([0.0, 0.25, 0.8])))
    kf = bool(c & 1); ori = bool(c & 2); ratio = float(rng.choice([0.6, 0.75, 0.9]))
    th = 49 if kf else 50
    want = O.search_by_bow(k1, d1, v1, O.flatten_feature_vector(fv1), k2, d2, v2 if kf else None, O.flatten_feature_vector(fv2), th, ratio, ori)
    got = gf.ORBmatcher(ratio, ori).SearchByBoW(k1, d1, v1, fv1, k2, d2, v2 if kf else None, fv2, th)
    if not (got[0] == want[0] and np.array_equal(got[1], want[1])):
        print("MISMATCH bow", c, n1, n2, nodes); bad += 1
    t = triangulation_case(30_000 + c, n1, n2, nodes)
    want = O.search_for_triangulation(t[0], t[1], t[2], t[3], O.flatten_feature_vector(t[4]), t[5], t[6], t[7], t[8],
                                      O.flatten_feature_vector(t[9]), t[10], t[11], t[12], t[13], t[14], 50, ori)
    got = gf.ORBmatcher(0.6, ori).SearchForTriangulation(*t[:10], t[10], t[11], t[12], t[13], t[14], 50)
    if not (got[0] == want[0] and np.array_equal(got[1], want[1])):
        print("MISMATCH triangulation", c, n1, n2, nodes); bad += 1
from test_oracle_grid import _pair
ninit = 0
for c in range(CASES // 2):
    n1 = int(rng.choice([60, 400, 1000, 2500, 5000, 8192])); n2 = int(rng.choice([60, 400, 1000, 2500, 5000, 8192]))
    r2 = np.random.default_rng(40_000 + c)
    k1, d1, k2, d2 = _pair(r2, n1, n2, flip=float(rng.choice([0.03, 0.06, 0.12])), p0=float(rng.choice([0.3, 0.6, 1.0])))
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
    win = int(rng.choice([10, 30, 100, 400])); ratio = float(rng.choice([0.6, 0.9, 0.95])); ori = bool(c & 1)
    wn, wm12, wprev = O.search_for_initialization(k1, d1, k2, d2, B, prev, win, ratio, ori)
    gp = prev.copy()
    gn, gm12 = gf.ORBmatcher(ratio, ori).SearchForInitialization(k1, d1, k2, d2, gp, win, bounds=B)
    ninit += 1
    if not (gn == wn and np.array_equal(gm12, wm12) and np.array_equal(gp, wprev)):
        print("MISMATCH init", c, n1, n2, win, ratio); bad += 1
print("init cases:", ninit)
print("stress:", CASES, "projection +", CASES // 2, "bow +", CASES // 2, "triangulation cases,", bad, "mismatches")
sys.exit(1 if bad else 0)
