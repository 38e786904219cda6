"""GPU: the header-only C++ mirror (include/gforb/ORBextractor.hpp) driven like Frame::Frame (two threads, then
ComputeStereoMatches) must give byte-identical results to the ctypes path; plus the TMA helper probe."""
import importlib
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

from conftest import B, BF

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _fnv(a):
    h = 1469598103934665603
    for b in np.ascontiguousarray(a).view(np.uint8).ravel().tolist():
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def test_cpp_shim_matches_ctypes(fixtures):
    exe = ROOT / "tests/cpp/_build/shim_test"
    assert exe.exists(), "run __graft_entry__.build() first"
    out = subprocess.run([str(exe), str(ROOT / "tests/data/EuRoC_l.pgm"), str(ROOT / "tests/data/EuRoC_r.pgm")],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    m = re.search(r"nL=(\d+) nR=(\d+) nmatched=(\d+) levels=(\d+) dist=(\d+)", out.stdout)
    hs = dict(re.findall(r"(\w+)=([0-9a-f]{16})", out.stdout))
    gf = importlib.import_module("gf-orb-slam2_b200")
    exL, exR = gf.ORBextractor(), gf.ORBextractor()
    kL, dL = exL(fixtures["EuRoC_l"]); kR, dR = exR(fixtures["EuRoC_r"])
    s = gf.ComputeStereoMatches(exL, exR, kL, dL, kR, dR, np.float32(BF), np.float32(np.float32(BF) / np.float32(435.2046959714599)))
    assert (int(m[1]), int(m[2]), int(m[3]), int(m[4])) == (len(kL), len(kR), s["nmatched"], 8)
    assert int(m[5]) == gf.DescriptorDistance(dL[0], dR[0])
    assert int(hs["kL"], 16) == _fnv(kL) and int(hs["dL"], 16) == _fnv(dL)
    assert int(hs["kR"], 16) == _fnv(kR) and int(hs["dR"], 16) == _fnv(dR)
    assert int(hs["u"], 16) == _fnv(s["mvuRight"]) and int(hs["z"], 16) == _fnv(s["mvDepth"])
    assert "empty ok=1" in out.stdout
    # SearchForInitialization + feature grid through the C++ mirror == the ctypes mirror
    mi = re.search(r"init nmatches=(\d+) .* area=(\d+) ", out.stdout)
    prev = np.stack([kL["x"], kL["y"]], axis=1).astype(np.float32)
    n, m12 = gf.ORBmatcher(0.9, True).SearchForInitialization(kL, dL, kR, dR, prev, 100, width=752, height=480)
    area = gf.FrameGrid(kR, width=752, height=480).GetFeaturesInArea([(376.0, 240.0, 60.0, 0, 2)])[0]
    assert (int(mi[1]), int(mi[2])) == (n, len(area)) and n > 0 and len(area) > 0
    assert int(hs["m12"], 16) == _fnv(m12) and int(hs["prev"], 16) == _fnv(prev)
    assert int(hs["area_hash"], 16) == _fnv(area.astype(np.uint64))
    # last-frame SearchByProjection through the C++ mirror == the ctypes mirror
    mp = re.search(r"proj nmatches=(\d+) ", out.stdout)
    q = np.zeros(len(kR), gf.PROJ_QUERY_DTYPE)
    sc = np.ones(len(kR), np.float32)
    for lvl in range(1, 8):
        sc[kR["octave"] >= lvl] *= np.float32(1.2)
    q["x"], q["y"], q["angle"] = kR["x"], kR["y"], kR["angle"]
    q["r"] = np.float32(15.0) * sc
    q["xr"] = kR["x"] - np.float32(20.0)
    q["min_level"], q["max_level"] = kR["octave"] - 1, kR["octave"] + 1
    q["flags"] = gf.PQ_ACTIVE | gf.PQ_BLOCKS | gf.PQ_RIGHT_GATE
    nm, kq, kd, qk = gf.ORBmatcher(0.9, True).SearchByProjection(kL, dL, q, dR, s["mvuRight"], None, 100, False, True,
                                                                 width=752, height=480)
    assert int(mp[1]) == nm > 0
    assert int(hs["kq"], 16) == _fnv(kq) and int(hs["kd"], 16) == _fnv(kd) and int(hs["qk"], 16) == _fnv(qk)
    # SearchByBoW through the C++ mirror == the ctypes mirror
    mbw = re.search(r"bow nmatches=(\d+) ", out.stdout)

    def fv(d):
        o = {}
        for i, row in enumerate(d):
            o.setdefault(int(row[0] & 0xF0), []).append(i)
        return o
    nb, mb = gf.ORBmatcher(0.7, True).SearchByBoW(kL, dL, None, fv(dL), kR, dR, None, fv(dR))
    assert int(mbw[1]) == nb > 0 and int(hs["mb"], 16) == _fnv(mb)
    # SearchForTriangulation through the C++ mirror == the ctypes mirror
    mt = re.search(r"tri nmatches=(\d+) ", out.stdout)
    sf = np.ones(8, np.float32)
    for l in range(1, 8):
        sf[l] = sf[l - 1] * np.float32(1.2)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    nt, m12t = gf.ORBmatcher(0.7, True).SearchForTriangulation(kL, dL, None, None, fv(dL), kR, dR, None, None, fv(dR), F12, 400.0, 240.0,
                                                               sf, (sf * sf).astype(np.float32))
    idx = np.nonzero(m12t >= 0)[0]
    pairs = np.stack([idx, m12t[idx]], axis=1).astype(np.uint64)
    assert int(mt[1]) == nt > 0 and int(hs["pairs"], 16) == _fnv(pairs)
    # SearchBySim3 through the C++ mirror == the ctypes mirror
    ms = re.search(r"sim3 nfound=(\d+) ", out.stdout)

    def sq(k, shift):
        o = np.zeros(len(k), gf.PROJ_QUERY_DTYPE)
        scl = np.ones(len(k), np.float32)
        for lvl in range(1, 8):
            scl[k["octave"] >= lvl] *= np.float32(1.2)
        o["x"], o["y"] = k["x"] + np.float32(shift), k["y"]
        o["r"] = np.float32(30.0) * scl
        o["min_level"], o["max_level"] = k["octave"] - 1, k["octave"]
        o["flags"] = gf.PQ_ACTIVE
        return o
    nf, m12s, _, _ = gf.ORBmatcher(0.9, True).SearchBySim3(kL, dL, kR, dR, sq(kL, -25.0), dL, sq(kR, 25.0), dR, width=752, height=480)
    assert int(ms[1]) == nf > 0 and int(hs["ms"], 16) == _fnv(m12s)
    # MultiIndexHashing through the C++ mirror == the ctypes mirror
    mm = re.search(r"mih bits=(\d+) inserted=(\d+) ", out.stdout)
    g = gf.MultiIndexHashing(32)
    okh, _ = g.insert(dL, np.arange(len(kL), dtype=np.int32) + 1000)
    cand, cnt = g.query(dR, 8, 20)
    flat = np.concatenate([np.concatenate([[c], row[:c]]) for row, c in zip(cand, cnt)]).astype(np.int32)
    assert (int(mm[1]), int(mm[2])) == (8, len(kL))
    assert int(hs["okh"], 16) == _fnv(okh) and int(hs["cands"], 16) == _fnv(flat)


def test_tma_probe():
    exe = ROOT / "tests/cpp/_build/tma_probe"
    assert exe.exists(), "run __graft_entry__.build() first"
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "TMA PROBE OK" in out.stdout, out.stdout + out.stderr
