"""Multi-index hashing of 256-bit descriptors (reference include/Hashing.h:171-216, src/Hashing.cc):
the restatement (oracle/mih.cpp) against THE REFERENCE'S OWN src/Hashing.cc, compiled unmodified into
oracle/_ref/libgforb_ref_mih.so -- inserts in several batches, re-inserts, both query forms, the `dynamics` offset,
8-, 16- and odd-width substrings -- and the substring keys against a bit-level numpy statement of the layout."""
import numpy as np
import pytest

from oracle import pymih


def descriptors(rng, n, clusters=40, flip=0.04):
    """n descriptors around `clusters` centres: buckets hold several entries, and near-duplicates share most substrings"""
    centres = rng.integers(0, 256, (clusters, 32), dtype=np.uint8)
    d = centres[rng.integers(0, clusters, n)] ^ np.packbits(rng.random((n, 256)) < flip, axis=1)
    return np.ascontiguousarray(d)


@pytest.mark.parametrize("T", [32, 16, 24, 20, 13])
def test_keys_are_the_documented_substrings(T):
    rng = np.random.default_rng(T)
    d = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    m = pymih.MultiIndexHashing(T)
    bits = 256 // T
    assert m.getSubstringLen() == bits
    keys = m.hash(d)
    # table i = bits [i * bits, (i + 1) * bits) of the descriptor read as little-endian 32-bit words, LSB first
    allbits = np.unpackbits(d, axis=1, bitorder="little")
    for i in range(T):
        w = allbits[:, i * bits:(i + 1) * bits].astype(np.uint64)
        want = (w << np.arange(bits, dtype=np.uint64)).sum(1)
        assert np.array_equal(keys[:, i].astype(np.uint64), want), (T, i)


@pytest.fixture(scope="module")
def have_ref():
    if not pymih.available("ref"):
        pytest.skip("oracle/_ref/libgforb_ref_mih.so was not built")


@pytest.mark.parametrize("T", [32, 16, 24, 13])
def test_restatement_equals_reference(have_ref, T):
    rng = np.random.default_rng(100 + T)
    a, b = pymih.MultiIndexHashing(T, "oracle"), pymih.MultiIndexHashing(T, "ref")
    n = 3000
    d = descriptors(rng, n)
    ids = rng.permutation(100000)[:n].astype(np.int32)
    hashed = {}
    pos = 0
    for batch in (1, 700, 64, 1500, 735):                       # several inserts: chains grow across batches
        sl = slice(pos, pos + batch); pos += batch
        ra, rb = a.insert(d[sl], ids[sl]), b.insert(d[sl], ids[sl])
        assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1])
        assert ra[0].all() and (ra[1] == (2 ** T - 1 if T < 32 else 0xFFFFFFFF)).all()
        for i, h in zip(ids[sl], ra[1]):
            hashed[int(i)] = h
    # re-insert: already hashed everywhere -> nothing happens; with some flags cleared (MapPoint copied / replaced,
    # src/MapPoint.cc:163) and a changed descriptor -> lands in other buckets, or meets itself as the bucket's latest
    again = rng.integers(0, n, 400)
    hm = np.array([hashed[int(ids[i])] for i in again], np.uint32)
    ra, rb = a.insert(d[again], ids[again], hm), b.insert(d[again], ids[again], hm)
    assert np.array_equal(ra[0], rb[0]) and not ra[0].any()
    hm2 = hm & rng.integers(0, 2 ** 32, len(again), dtype=np.uint64).astype(np.uint32)
    d2 = d[again] ^ np.packbits(rng.random((len(again), 256)) < 0.01, axis=1)
    ra, rb = a.insert(d2, ids[again], hm2), b.insert(d2, ids[again], hm2)
    assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1]) and ra[0].any()
    same = np.full(5, ids[7], np.int32)                           # the same point five times in a row, flags clear:
    ra, rb = a.insert(np.repeat(d[7:8], 5, 0), same), b.insert(np.repeat(d[7:8], 5, 0), same)   # latest->mnId == mnId
    assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1])

    q = np.concatenate([d[rng.integers(0, n, 300)] ^ np.packbits(rng.random((300, 256)) < 0.03, axis=1),
                        rng.integers(0, 256, (50, 32), dtype=np.uint8)])
    for dyn_steps in (0, 3, 25):
        for _ in range(dyn_steps):
            assert a.updateDynamics(0.01) == b.updateDynamics(0.01)
        dyn = a.updateDynamics(0.0)                               # |e| < MIN_TIME_DIFF: returns the current value
        for nactive, nret in ((8, 20), (0, 20), (1, 1), (min(T, 12), 5), (3, 0)):
            nret = max(nret, -dyn)            # nretrieves + dynamics < 0 is undefined in the reference (Bucket::back, it + n)
            ca, na = a.query(q, nactive, nret); cb, nb = b.query(q, nactive, nret)
            assert np.array_equal(na, nb) and np.array_equal(ca, cb), (T, dyn_steps, nactive, nret)
            if (nactive, nret) == (8, 20) and dyn_steps == 0:
                assert na.max() > 20 and (ca[:, 0] >= 0).sum() > 250           # near-duplicates find their bucket mates
        for nf in (np.full(T, 20), np.full(T, 60), rng.integers(-1, 40, T), np.zeros(T, np.int32)):
            nf = np.where(nf > 0, np.maximum(nf, -dyn), nf)      # same: a positive nfeatures[k] with nfeatures[k] + dynamics < 0
            ra, rb = a.query2(q, nf), b.query2(q, nf)
            for x, y in zip(ra, rb):
                assert np.array_equal(x, y), (T, dyn_steps, nf)
    for e in (-0.01, -0.5, 0.0005, -0.01, -0.01):
        assert a.updateDynamics(e) == b.updateDynamics(e)
    for _ in range(30):
        assert a.updateDynamics(-1.0) == b.updateDynamics(-1.0)         # saturates at MAX_BUCKET_SIZE - NUM_CANDIDATE_FEATURES
    ra, rb = a.query2(q, np.full(T, 45)), b.query2(q, np.full(T, 45))
    assert ra[3].max() > 0                                               # the additional-candidate branch is exercised
    for x, y in zip(ra, rb):
        assert np.array_equal(x, y)


def test_unsupported_layouts_are_rejected():
    for T in (0, 33, 8, 12):          # mvbHashed holds 32 flags; 2^(256 / T) buckets per table must be allocatable
        with pytest.raises(ValueError):
            pymih.MultiIndexHashing(T)
