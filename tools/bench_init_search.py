#!/usr/bin/env python3
"""Throughput of the batched device SearchForInitialization (grid build + search, one CTA per frame pair) on keypoints
extracted from the bench's synthetic stereo frames (left eye = frame 1, right eye = frame 2, window 100), next to the
sequential CPU oracle on the same pairs.  Prints one JSON line.  usage: python tools/bench_init_search.py [pairs]"""
import importlib, json, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import bench
gf = importlib.import_module("gf-orb-slam2_b200")
from oracle import pyoracle as O

P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
left, right = bench.make_frames(P)
sb = gf.StereoBatch(bench.W, bench.H, P, bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH)
out = sb.alloc_outputs(P)
sb.process(left, right, out, bench.BF, bench.B)
K = sb.K
k1 = np.ascontiguousarray(out["keypoints"][0::2]); k2 = np.ascontiguousarray(out["keypoints"][1::2])
d1 = np.ascontiguousarray(out["descriptors"][0::2]); d2 = np.ascontiguousarray(out["descriptors"][1::2])
n1 = np.ascontiguousarray(out["counts"][0::2]); n2 = np.ascontiguousarray(out["counts"][1::2])
prev = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], axis=2).astype(np.float32))
bounds = (0.0, float(bench.W), 0.0, float(bench.H))
D = gf.DeviceArray.from_host
dk1, dd1, dn1, dk2, dd2, dn2 = D(k1), D(d1), D(n1), D(k2), D(d2), D(n2)
dm12 = gf.DeviceArray((P, K), np.int32); dnm = gf.DeviceArray((P,), np.int32)
m = gf.ORBmatcher(0.9, True)
times = []
for it in range(13):
    dprev = D(prev)                       # vbPrevMatched is in/out: fresh copy per run
    m.SearchForInitializationBatchDevice(bounds, P, dk1, dd1, dn1, K, dk2, dd2, dn2, K, dprev, 100, dm12, dnm)
    ms = m.sync()
    if it >= 3:
        times.append(ms)
nm = dnm.to_host(); m12 = dm12.to_host()
# CPU oracle, single thread, bounded sample
S = min(P, 64)
t0 = time.perf_counter()
ok = True
for p in range(S):
    on, om12, _ = O.search_for_initialization(k1[p, :n1[p]], d1[p, :n1[p]], k2[p, :n2[p]], d2[p, :n2[p]], bounds,
                                              prev[p, :n1[p]], 100, 0.9, True)
    ok &= on == nm[p] and np.array_equal(om12, m12[p, :n1[p]])
cpu_s = (time.perf_counter() - t0) / S
ms = float(np.median(times))
lvl0 = float(np.mean([(k1[p, :n1[p]]["octave"] == 0).sum() for p in range(P)]))
print(json.dumps({"what": "ORBmatcher::SearchForInitialization, batched device form", "pairs": P, "keys_per_frame": float(n1.mean()),
                  "level0_queries_per_pair": lvl0, "window": 100, "ms_per_batch": ms, "pairs_per_s": P / ms * 1e3,
                  "us_per_pair_amortised": ms * 1e3 / P, "mean_matches": float(nm.mean()), "parity_vs_oracle_sample": bool(ok),
                  "cpu_oracle_us_per_pair_1thread": cpu_s * 1e6, "launches": 3}))
