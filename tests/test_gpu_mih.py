"""Multi-index hashing on the device (gforb_mih_*, k_mih.cu) == the restatement (oracle/mih.cpp) == THE REFERENCE'S OWN
src/Hashing.cc (oracle/_ref/libgforb_ref_mih.so, when shipped): substring keys, insertion over several batches incl.
re-inserts and one > 8,192-entry batch, both query forms, the dynamics offset; device-resident chaining from the
extractor's descriptors."""
import importlib

import numpy as np
import pytest

from oracle import pymih
from test_oracle_mih import descriptors

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


def _others(T):
    out = [pymih.MultiIndexHashing(T, "oracle")]
    if pymih.available("ref"):
        out.append(pymih.MultiIndexHashing(T, "ref"))
    return out


@pytest.mark.parametrize("T", [32, 16, 24, 20])
def test_keys(gf, T):
    rng = np.random.default_rng(T)
    d = rng.integers(0, 256, (1000, 32), dtype=np.uint8)
    g = gf.MultiIndexHashing(T)
    assert g.getSubstringLen() == 256 // T
    assert np.array_equal(g.hash(d), pymih.MultiIndexHashing(T, "oracle").hash(d))
    assert g.hash(d[:0]).shape == (0, T)


@pytest.mark.parametrize("T", [32, 16, 24])
def test_insert_and_queries(gf, T):
    rng = np.random.default_rng(200 + T)
    g, others = gf.MultiIndexHashing(T), _others(T)
    n = 15000
    d = descriptors(rng, n, clusters=120)
    ids = rng.permutation(1000000)[:n].astype(np.int32)
    hashed = np.zeros(n, np.uint32)
    pos = 0
    for batch in (1, 700, 64, 9500, 3000, 1735):                  # 9,500 > one launch's 8,192: chunks keep the order
        sl = slice(pos, pos + batch); pos += batch
        ok, hm = g.insert(d[sl], ids[sl])
        for o in others:
            ook, ohm = o.insert(d[sl], ids[sl])
            assert np.array_equal(ok, ook) and np.array_equal(hm, ohm), (o.variant, batch)
        hashed[sl] = hm
    again = rng.integers(0, n, 900)
    hm2 = hashed[again] & rng.integers(0, 2 ** 32, len(again), dtype=np.uint64).astype(np.uint32)
    d2 = d[again] ^ np.packbits(rng.random((len(again), 256)) < 0.01, axis=1)
    ok, hm = g.insert(d2, ids[again], hm2)
    same = np.full(5, ids[7], np.int32)
    ok5, hm5 = g.insert(np.repeat(d[7:8], 5, 0), same)
    for o in others:
        ook, ohm = o.insert(d2, ids[again], hm2)
        assert np.array_equal(ok, ook) and np.array_equal(hm, ohm) and ok.any(), o.variant
        ook, ohm = o.insert(np.repeat(d[7:8], 5, 0), same)
        assert np.array_equal(ok5, ook) and np.array_equal(hm5, ohm), o.variant

    q = np.concatenate([d[rng.integers(0, n, 2000)] ^ np.packbits(rng.random((2000, 256)) < 0.03, axis=1),
                        rng.integers(0, 256, (100, 32), dtype=np.uint8)])
    for dyn_steps in (0, 3, 25):
        for _ in range(dyn_steps):
            dyn = g.updateDynamics(0.01)
            for o in others:
                assert o.updateDynamics(0.01) == dyn
        dyn = g.updateDynamics(0.0)
        for nactive, nret in ((8, 20), (0, 20), (1, 1), (min(T, 12), 5), (3, 0)):
            nret = max(nret, -dyn)                                  # nretrieves + dynamics < 0: undefined in the reference
            cg, ng = g.query(q, nactive, nret)
            for o in others:
                co, no = o.query(q, nactive, nret)
                assert np.array_equal(ng, no) and np.array_equal(cg, co), (o.variant, T, dyn_steps, nactive, nret)
        for nf in (np.full(T, 20), np.full(T, 60), rng.integers(-1, 40, T), np.zeros(T, np.int32)):
            nf = np.where(nf > 0, np.maximum(nf, -dyn), nf)
            rg = g.query_split(q, nf)
            for o in others:
                for x, y in zip(rg, o.query2(q, nf)):
                    assert np.array_equal(x, y), (o.variant, T, dyn_steps)
    for _ in range(40):
        dyn = g.updateDynamics(-1.0)
        for o in others:
            assert o.updateDynamics(-1.0) == dyn
    assert dyn == 0
    rg = g.query_split(q, np.full(T, 45))
    assert rg[3].max() > 0
    for o in others:
        for x, y in zip(rg, o.query2(q, np.full(T, 45))):
            assert np.array_equal(x, y), o.variant
    with pytest.raises(gf.GforbError):                              # a row too short for the worst case is refused, not truncated
        cand = np.zeros((len(q), 3), np.int32); cnt = np.zeros(len(q), np.int32)
        gf._check(gf.lib.gforb_mih_query(g._h, gf._p(q), len(q), 8, 20, gf._p(cand), 3, gf._p(cnt)))


def test_unsupported_layouts(gf):
    for T in (0, 8, 12, 15, 33):
        with pytest.raises(gf.GforbError):
            gf.MultiIndexHashing(T)


def test_device_chain_from_extracted_descriptors(gf):
    """descriptors of a batch run never leave the device: gforb_batch_device_results -> insert_device -> query_device"""
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))
    import synthgen
    from conftest import BF, B
    n, w, h = 2, 752, 480
    hl, hr = synthgen.host_frames(3, n, w, h)
    sb = gf.StereoBatch(w, h, n)
    out = sb.alloc_outputs(n, pinned=False)
    sb.process(hl, hr, out, BF, B)
    res = sb.device_results()
    K = res["K"]
    nL = int(out["counts"][0]); nQ = int(out["counts"][2])
    g = gf.MultiIndexHashing(32)
    ent = gf.DeviceArray.from_host(np.arange(nL, dtype=np.int32))
    hashed = gf.DeviceArray.from_host(np.zeros(nL, np.uint32)); ok = gf.DeviceArray((nL,), np.uint8)
    g.insert_device(res["descriptors"], ent, nL, hashed, ok)                       # image 0 = left eye of frame 0
    stride = 8 * 20
    cand = gf.DeviceArray.from_host(np.full((nQ, stride), -1, np.int32)); cnt = gf.DeviceArray((nQ,), np.int32)
    g.query_device(res["descriptors"] + 2 * K * 32, nQ, 8, 20, cand, stride, cnt)   # image 2 = left eye of frame 1
    g.sync()
    o = pymih.MultiIndexHashing(32, "oracle")
    ook, _ = o.insert(out["descriptors"][0, :nL], np.arange(nL, dtype=np.int32))
    assert np.array_equal(ok.to_host(), ook)
    co, no = o.query(out["descriptors"][2, :nQ], 8, 20)
    assert np.array_equal(cnt.to_host(), no) and np.array_equal(cand.to_host(), co) and no.max() > 0
