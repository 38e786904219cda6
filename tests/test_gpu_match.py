"""GPU parity of the brute-force Hamming matcher (BASELINE.json configs[3]) against the oracle's sequential
restatement of the reference loop (src/ORBmatcher.cc:562-576).  Integer work: bit-exact."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.fixture(scope="module")
def O():
    from oracle import pyoracle
    return pyoracle


def make_sets(n, m, seed, flip=0.1):
    """A random; B = A[perm] (cycled/truncated to m) with each bit flipped w.p. `flip`."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    perm = rng.permutation(max(n, 1))
    src = a[perm[np.arange(m) % max(n, 1)]] if n else rng.integers(0, 256, (m, 32), dtype=np.uint8)
    noise = np.packbits(rng.random((m, 256)) < flip, axis=1)
    return a, src ^ noise


def check(gf, O, a, b, ratio, th):
    m = gf.ORBmatcher(ratio)
    got = m.BruteForce(a, b, th)
    ref = O.match_bruteforce(a, b, ratio, th, nthreads=8)
    for k in ("best_idx", "best_dist", "second_dist", "match"):
        assert np.array_equal(got[k], ref[k]), k
    return got


@pytest.mark.parametrize("n,m", [(1000, 1000), (1, 1), (1, 5000), (129, 255), (128, 256), (2048, 511), (4096, 4096), (300, 70000)])
@pytest.mark.parametrize("ratio,th", [(0.6, 50), (0.9, 100)])
def test_bruteforce_vs_oracle(gf, O, n, m, ratio, th):
    a, b = make_sets(n, m, 1000 * n + m)
    got = check(gf, O, a, b, ratio, th)
    if n >= 1000 and m >= n:
        assert (got["match"] >= 0).mean() > 0.9       # the planted neighbours (~26 bits) are found


def test_bruteforce_ties_and_duplicates(gf, O):
    """equal minima: the first index wins and the second-best equals the best (multiset semantics of the loop)"""
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    b = np.concatenate([a[::-1], a[::-1], a[:7]])        # every query has >= 2 exact copies in b
    got = check(gf, O, a, b, 0.6, 50)
    assert (got["best_dist"] == 0).all() and (got["second_dist"] == 0).all() and (got["match"] == -1).all()
    assert np.array_equal(got["best_idx"][7:], 63 - np.arange(7, 64))
    # few distinct values -> massive distance ties across tile and segment boundaries
    a2 = np.repeat(rng.integers(0, 2, (200, 1), dtype=np.uint8) * 255, 32, axis=1)
    b2 = np.repeat(rng.integers(0, 2, (3000, 1), dtype=np.uint8) * 255, 32, axis=1)
    check(gf, O, a2, b2, 0.9, 100)


def test_bruteforce_empty_and_small(gf, O):
    m = gf.ORBmatcher(0.6)
    a, b = make_sets(10, 1, 3)
    r = m.BruteForce(a, b, 256)
    assert (r["best_idx"] == 0).all() and (r["second_dist"] == np.iinfo(np.int32).max).all()
    assert (r["match"] == 0).all()                       # (float)INT_MAX * ratio dominates, as in the reference
    check(gf, O, a, b, 0.6, 256)
    r = m.BruteForce(a, np.zeros((0, 32), np.uint8), 50)
    assert (r["best_idx"] == -1).all() and (r["best_dist"] == np.iinfo(np.int32).max).all() and (r["match"] == -1).all()
    r = m.BruteForce(np.zeros((0, 32), np.uint8), b, 50)
    assert len(r["best_idx"]) == 0


def test_bruteforce_device_form_and_on_extracted_descriptors(gf, O, fixtures):
    exL, exR = gf.ORBextractor(), gf.ORBextractor()
    _, dL = exL(fixtures["EuRoC_l"]); _, dR = exR(fixtures["EuRoC_r"])
    ref = O.match_bruteforce(dL, dR, 0.9, 100)
    m = gf.ORBmatcher(0.9)
    A, B = gf.DeviceArray.from_host(dL), gf.DeviceArray.from_host(dR)
    outs = [gf.DeviceArray((len(dL),), np.int32) for _ in range(4)]
    m.BruteForceDevice(A, len(dL), B, len(dR), 100, *outs)
    assert m.sync() > 0
    for k, o in zip(("best_idx", "best_dist", "second_dist", "match"), outs):
        assert np.array_equal(o.to_host(), ref[k]), k
    # distances agree with the host DescriptorDistance
    i = int(np.argmax(ref["match"] >= 0))
    assert gf.DescriptorDistance(dL[i], dR[ref["best_idx"][i]]) == ref["best_dist"][i]


def test_distinctive_descriptors_vs_oracle(gf, O):
    """MapPoint::ComputeDistinctiveDescriptors over many map points: bit-exact incl. median ties"""
    rng = np.random.default_rng(17)
    sizes = [1, 2, 3, 4, 5, 0, 7, 32, 33, 64, 100, 2, 0, 257, 1024] + [int(v) for v in rng.integers(1, 60, 300)]
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    desc = np.zeros((start[-1], 32), np.uint8)
    for p, n in enumerate(sizes):                                   # a cluster per point: base + a few flipped bits, some duplicates
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        d = base ^ np.packbits(rng.random((n, 256)) < rng.choice([0.02, 0.1, 0.3]), axis=1)
        if n > 3:
            d[n // 2] = d[0]
        desc[start[p]:start[p + 1]] = d
    m = gf.ORBmatcher()
    bi, bm = gf.ComputeDistinctiveDescriptors(m, desc, start)
    obi, obm = O.distinctive_descriptors(desc, start)
    assert np.array_equal(bi, obi) and np.array_equal(bm, obm)
    assert bi[5] == -1 and bm[5] == np.iinfo(np.int32).max
    with pytest.raises(gf.GforbError):
        gf.ComputeDistinctiveDescriptors(m, np.zeros((1025, 32), np.uint8), np.array([0, 1025], np.int32))


@pytest.mark.parametrize("n,m", [(2048, 2048), (4100, 3000), (2500, 10007), (8192, 8192), (2048, 70000)])
def test_tensor_core_path_vs_oracle(gf, O, n, m):
    """Sets of >= 2,048 descriptors a side take the tcgen05 path (k_match_tc.cu: bits expanded to +-1 int8, dot product on the
    tensor cores, best / second-best off TMEM): the same four outputs, bit for bit, incl. ragged last tiles and row blocks."""
    a, b = make_sets(n, m, 7 * n + m)
    for ratio, th in ((0.6, 50), (0.9, 100)):
        check(gf, O, a, b, ratio, th)


def test_tensor_core_path_ties(gf, O):
    """few distinct descriptors: equal minima everywhere, across tiles, segments and accumulator stages -- first index wins"""
    rng = np.random.default_rng(11)
    vals = rng.integers(0, 256, (5, 32), dtype=np.uint8)
    a = vals[rng.integers(0, 5, 3000)]
    b = vals[rng.integers(0, 5, 5000)]
    got = check(gf, O, a, b, 0.6, 50)
    assert (got["best_dist"] == 0).all() and (got["second_dist"] == 0).all()
