#!/usr/bin/env python3
"""One ORBmatcher::SearchByBoW call (key frame against a frame, src/ORBmatcher.cc:270-404) on the device, host to host,
next to the sequential CPU oracle on the same inputs.  Prints one JSON line.
usage: python tools/bench_bow_search.py [features] [nodes]"""
import importlib, json, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from synth import bow_case
gf = importlib.import_module("gf-orb-slam2_b200")
from oracle import pyoracle as O

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
NODES = int(sys.argv[2]) if len(sys.argv) > 2 else 700
k1, d1, v1, fv1, k2, d2, v2, fv2 = bow_case(1, N, N, NODES)
f1, f2 = gf.flatten_feature_vector(fv1), gf.flatten_feature_vector(fv2)
m = gf.ORBmatcher(0.7, True)
D = gf.DeviceArray.from_host
dev = [D(a) for a in (k1, d1, v1, f1[0], f1[1], f1[2], k2, d2, f2[0], f2[1], f2[2])]
dm = gf.DeviceArray((N,), np.int32); dn = gf.DeviceArray((1,), np.int32)
L = gf._L
dev_ms, host_s = [], []
for it in range(23):
    gf._check(L.gforb_search_by_bow_device(m._h, dev[0].ptr, dev[1].ptr, dev[2].ptr, N, dev[3].ptr, dev[4].ptr, dev[5].ptr, len(f1[0]),
                                           dev[6].ptr, dev[7].ptr, None, N, dev[8].ptr, dev[9].ptr, dev[10].ptr, len(f2[0]),
                                           50, 1, dm.ptr, dn.ptr, None))
    ms = m.sync()
    t0 = time.perf_counter()
    nm, m12 = m.SearchByBoW(k1, d1, v1, f1, k2, d2, None, f2, 50)
    t1 = time.perf_counter()
    if it >= 3:
        dev_ms.append(ms); host_s.append(t1 - t0)
t0 = time.perf_counter()
for _ in range(20):
    on, om12 = O.search_by_bow(k1, d1, v1, f1, k2, d2, None, f2, 50, 0.7, True)
cpu_s = (time.perf_counter() - t0) / 20
ok = on == nm and np.array_equal(om12, m12) and np.array_equal(dm.to_host(), m12) and int(dn.to_host()[0]) == nm
print(json.dumps({"what": "ORBmatcher::SearchByBoW (key frame against a frame), one call", "features_per_frame": N, "nodes": NODES,
                  "nmatches": int(nm), "us_device": float(np.median(dev_ms)) * 1e3, "us_host_to_host": float(np.median(host_s)) * 1e6,
                  "cpu_oracle_us_1thread": cpu_s * 1e6, "parity_vs_oracle": bool(ok), "launches": 2}))
