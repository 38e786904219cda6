"""Oracle pin for the Frame feature grid and ORBmatcher::SearchForInitialization (oracle/frame_grid.cpp).

The reference's tests hold no vectors for these functions, so the C++ oracle is checked here against an
independent pure-Python restatement written from the same reference lines (src/Frame.cc:461-476,593-656;
src/ORBmatcher.cc:520-633,1723-1764) on small seeded cases.
"""
import math

import numpy as np

from oracle import pyoracle as O

F32 = np.float32
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
               ("octave", "<i4"), ("class_id", "<i4")])


def make_keys(rng, n, w=752, h=480, spread=0.0, p_level0=0.5):
    k = np.zeros(n, KP)
    k["x"] = (rng.random(n) * (w + 2 * spread) - spread).astype(F32)
    k["y"] = (rng.random(n) * (h + 2 * spread) - spread).astype(F32)
    k["angle"] = (rng.random(n) * 360).astype(F32)
    k["octave"] = np.where(rng.random(n) < p_level0, 0, rng.integers(1, 8, n))
    k["size"], k["class_id"] = 31, -1
    return k


def c_round(v):                      # C round(): half away from zero
    return int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)


def py_grid(keys, bounds):
    minx, maxx, miny, maxy = (F32(b) for b in bounds)
    winv, hinv = F32(64) / (maxx - minx), F32(48) / (maxy - miny)
    cells = {}
    for i, k in enumerate(keys):
        px, py = c_round(float((k["x"] - minx) * winv)), c_round(float((k["y"] - miny) * hinv))
        if 0 <= px < 64 and 0 <= py < 48:
            cells.setdefault((px, py), []).append(i)
    return cells, minx, miny, winv, hinv


def py_area(keys, grid, x, y, r, lo, hi):
    cells, minx, miny, winv, hinv = grid
    x, y, r = F32(x), F32(y), F32(r)
    x0 = max(0, math.floor(float((x - minx - r) * winv)))
    x1 = min(63, math.ceil(float((x - minx + r) * winv)))
    y0 = max(0, math.floor(float((y - miny - r) * hinv)))
    y1 = min(47, math.ceil(float((y - miny + r) * hinv)))
    if x0 >= 64 or x1 < 0 or y0 >= 48 or y1 < 0:
        return []
    check = lo > 0 or hi >= 0
    out = []
    for ix in range(x0, x1 + 1):
        for iy in range(y0, y1 + 1):
            for i in cells.get((ix, iy), []):
                o = keys[i]["octave"]
                if check and (o < lo or (hi >= 0 and o > hi)):
                    continue
                if abs(keys[i]["x"] - x) < r and abs(keys[i]["y"] - y) < r:
                    out.append(i)
    return out


def py_search(k1, d1, k2, d2, bounds, prev, win, ratio, ori):
    grid = py_grid(k2, bounds)
    m12 = [-1] * len(k1); m21 = [-1] * len(k2); md = [2 ** 31 - 1] * len(k2)
    hist = [[] for _ in range(30)]
    prev = prev.copy()
    for i1 in range(len(k1)):
        if k1[i1]["octave"] > 0:
            continue
        best = best2 = 2 ** 31 - 1; bi = -1
        for i2 in py_area(k2, grid, prev[i1, 0], prev[i1, 1], win, 0, 0):
            d = int(np.unpackbits(d1[i1] ^ d2[i2]).sum())
            if md[i2] <= d:
                continue
            if d < best:
                best2, best, bi = best, d, i2
            elif d < best2:
                best2 = d
        if best <= 50 and F32(best) < F32(best2) * F32(ratio):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
            m12[i1], m21[bi], md[bi] = bi, i1, best
            if ori:
                rot = F32(k1[i1]["angle"] - k2[bi]["angle"])
                if rot < 0:
                    rot = F32(rot + F32(360))
                b = c_round(float(rot * (F32(1) / F32(30))))
                hist[0 if b == 30 else b].append(i1)
    if ori:
        sizes = [len(h) for h in hist]
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
        for i, h in enumerate(hist):
            if i not in ind:
                for j in h:
                    m12[j] = -1
    for i1, j in enumerate(m12):
        if j >= 0:
            prev[i1] = (k2[j]["x"], k2[j]["y"])
    return sum(j >= 0 for j in m12), np.array(m12, np.int32), prev


BOUNDS = (0.0, 752.0, 0.0, 480.0)


def test_grid_assignment_and_area_queries():
    rng = np.random.default_rng(0)
    keys = make_keys(rng, 700, spread=30.0)            # some keys fall outside the image bounds (undistorted case)
    start, idx = O.assign_features_to_grid(keys, BOUNDS)
    cells = py_grid(keys, BOUNDS)[0]
    assert start[-1] == len(idx) == sum(len(v) for v in cells.values()) < len(keys)
    for ix in range(64):
        for iy in range(48):
            c = ix * 48 + iy
            assert list(idx[start[c]:start[c + 1]]) == cells.get((ix, iy), [])
    grid = py_grid(keys, BOUNDS)
    for q in range(200):
        x, y = rng.random() * 900 - 70, rng.random() * 600 - 60
        r = float(rng.choice([5, 15.5, 40, 100, 2000]))
        lo, hi = [(-1, -1), (0, 0), (2, -1), (1, 3), (0, 7)][q % 5]
        assert list(O.features_in_area(keys, BOUNDS, x, y, r, lo, hi)) == py_area(keys, grid, x, y, r, lo, hi)


def _pair(rng, n1, n2, flip=0.06, p0=0.6):
    """Frame 2 = frame 1 moved by a few pixels with noisy descriptors, plus clutter with near-duplicate
    descriptors so that several level-0 queries compete for the same target (the vMatchedDistance rule)."""
    k1 = make_keys(rng, n1, p_level0=p0)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    m = min(n1, n2)
    k2 = make_keys(rng, n2, p_level0=p0)
    k2[:m] = k1[:m]
    k2["x"][:m] += rng.normal(0, 6, m).astype(F32); k2["y"][:m] += rng.normal(0, 6, m).astype(F32)
    k2["angle"][:m] = np.mod(k1["angle"][:m] + rng.normal(8, 3, m), 360).astype(F32)
    wild = rng.random(m) < 0.15
    k2["angle"][:m][wild] = (rng.random(int(wild.sum())) * 360).astype(F32)
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    d2[:m] = d1[:m] ^ np.packbits(rng.random((m, 256)) < flip, axis=1)
    dup = rng.integers(0, m, n1 // 5)                  # queries that copy another query's descriptor and place
    src = rng.integers(0, m, n1 // 5)
    d1[dup] = d1[src] ^ np.packbits(rng.random((len(dup), 256)) < 0.02, axis=1)
    k1["x"][dup] = k1["x"][src] + 3; k1["y"][dup] = k1["y"][src] - 2; k1["octave"][dup] = 0
    perm = rng.permutation(n2)
    return k1, d1, k2[perm], d2[perm]


def test_search_for_initialization_matches_python_restatement():
    rng = np.random.default_rng(1)
    for n1, n2, win, ratio, ori in [(300, 320, 100, 0.9, True), (250, 200, 30, 0.6, True), (200, 260, 100, 0.9, False),
                                    (40, 0, 100, 0.9, True), (0, 50, 100, 0.9, True), (180, 180, 400, 0.95, True)]:
        k1, d1, k2, d2 = _pair(rng, n1, n2) if n1 and n2 else (make_keys(rng, n1), rng.integers(0, 256, (n1, 32), dtype=np.uint8),
                                                               make_keys(rng, n2), rng.integers(0, 256, (n2, 32), dtype=np.uint8))
        prev = np.stack([k1["x"], k1["y"]], axis=1).astype(F32)
        n, m12, p = O.search_for_initialization(k1, d1, k2, d2, BOUNDS, prev, win, ratio, ori)
        rn, rm12, rp = py_search(k1, d1, k2, d2, BOUNDS, prev, win, ratio, ori)
        assert n == rn and np.array_equal(m12, rm12) and np.array_equal(p, rp)
        if n1 >= 180 and n2 >= 180:
            assert n > 20
