"""gforb_match_bruteforce_static: the allocation-free form of the brute-force Hamming matcher (its scratch is a fixed
device array inside the library; csrc/k_match_static.cu).  No counterpart in the reference.
Collected last on purpose.

Checked through the C ABI against the existing entry point gforb_match_bruteforce_device on the same device inputs:
integer outputs, so the tolerance is the one every matcher test here uses -- bit-exact -- on two calls in a row (the
second with other sizes, so stale scratch of the first would show), plus the stated limits."""
import importlib

import numpy as np
import pytest

ERR_ARG, ERR_CAPACITY = -1, -2          # include/gforb.h


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


def _sets(rng, na, nb):
    a = rng.integers(0, 256, (na, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (nb, 32), dtype=np.uint8)
    k = min(na, nb) // 2
    if k:                                    # near-duplicates of train rows: real matches, ties on distance
        idx = rng.integers(0, nb, k)
        a[:k] = b[idx]
        flips = rng.integers(0, 32, k)
        a[np.arange(k), flips] ^= np.uint8(1) << rng.integers(0, 8, k).astype(np.uint8)
    return a, b


def _run_both(gf, m, a, b, th):
    D = gf.DeviceArray
    da, db = D.from_host(a), D.from_host(b)
    n = len(a)
    names = ("best_idx", "best_dist", "second_dist", "match")
    ref = {k: D((max(n, 1),), np.int32) for k in names}
    got = {k: D.from_host(np.full(max(n, 1), -7, np.int32)) for k in names}
    m.BruteForceDevice(da, n, db, len(b), th, *(ref[k] for k in names))
    m.sync()
    gf.ORBmatcher.BruteForceStatic(da, n, db, len(b), 0.6, th, *(got[k] for k in names))
    # NULL stream = the legacy default stream: the blocking cudaMemcpy of to_host() below is ordered behind the kernels
    return {k: ref[k].to_host()[:n] for k in names}, {k: got[k].to_host()[:n] for k in names}


@pytest.mark.gpu
def test_static_matcher_equals_device_matcher_twice(gf, monkeypatch):
    monkeypatch.setenv("GFORB_NO_TC", "1")      # the static form runs the POPC kernels; both paths give identical bits anyway
    m = gf.ORBmatcher(0.6)
    rng = np.random.default_rng(77)
    # calls in a row, different shapes; the last at the stated maximum n_a
    for na, nb, th in ((3000, 5000, 50), (777, 12345, 100), (gf.MATCH_STATIC_MAX_QUERIES, 1500, 50)):
        a, b = _sets(rng, na, nb)
        ref, got = _run_both(gf, m, a, b, th)
        for k in ref:
            assert np.array_equal(ref[k], got[k]), (na, nb, k)
        assert (got["match"] >= 0).sum() > 100


@pytest.mark.gpu
def test_static_matcher_limits_and_edges(gf):
    D = gf.DeviceArray
    m = gf.ORBmatcher(0.6)
    rng = np.random.default_rng(78)
    a, b = _sets(rng, 64, 1)
    ref, got = _run_both(gf, m, a, b, 50)                             # one train row: second_dist = INT32_MAX
    for k in ref:
        assert np.array_equal(ref[k], got[k]), k
    da, db = D.from_host(a), D.from_host(b)
    out = D((gf.MATCH_STATIC_MAX_QUERIES + 1,), np.int32)
    import ctypes as C
    L = gf.lib
    rc = L.gforb_match_bruteforce_static(C.c_void_p(da.ptr), gf.MATCH_STATIC_MAX_QUERIES + 1, C.c_void_p(db.ptr), 1, 0.6, 50,
                                         C.c_void_p(out.ptr), None, None, None, None)
    assert rc == ERR_CAPACITY and b"65,536" in L.gforb_last_error_string()
    rc = L.gforb_match_bruteforce_static(C.c_void_p(da.ptr), 64, C.c_void_p(db.ptr), (1 << 22) + 1, 0.6, 50,
                                         C.c_void_p(out.ptr), None, None, None, None)
    assert rc == ERR_CAPACITY
    assert L.gforb_match_bruteforce_static(None, 0, None, 0, 0.6, 50, None, None, None, None, None) == 0      # empty: no-op
    assert L.gforb_match_bruteforce_static(None, 5, None, 0, 0.6, 50, None, None, None, None, None) == ERR_ARG
