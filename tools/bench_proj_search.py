#!/usr/bin/env python3
"""Throughput and single-frame latency of the device SearchByProjection (grid build + window scan + in-order
resolution) on keypoints extracted from the bench's synthetic stereo frames: the right eye's keypoints play the last
frame's map points projected into the left eye's frame (last-frame form, th = 15, rotation check;
src/ORBmatcher.cc:1440-1593), next to the sequential CPU oracle on the same frames.  Prints one JSON line.
usage: python tools/bench_proj_search.py [frames]"""
import importlib, json, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import bench
gf = importlib.import_module("gf-orb-slam2_b200")
from oracle import pyoracle as O

P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
TH = 15.0
left, right = bench.make_frames(P)
sb = gf.StereoBatch(bench.W, bench.H, P, bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH)
out = sb.alloc_outputs(P)
sb.process(left, right, out, bench.BF, bench.B)
K = sb.K
keys = np.ascontiguousarray(out["keypoints"][0::2]); k2 = np.ascontiguousarray(out["keypoints"][1::2])
desc = np.ascontiguousarray(out["descriptors"][0::2]); qdesc = np.ascontiguousarray(out["descriptors"][1::2])
n = np.ascontiguousarray(out["counts"][0::2]); nq = np.ascontiguousarray(out["counts"][1::2])
ur = np.ascontiguousarray(out["u_right"], np.float32).reshape(P, K)
scale = (np.float32(bench.SCALE) ** np.arange(bench.NLEVELS)).astype(np.float32)
def make_queries(kq):
    q = np.zeros((P, K), gf.PROJ_QUERY_DTYPE)
    q["x"], q["y"], q["angle"] = kq["x"], kq["y"], kq["angle"]
    o = np.clip(kq["octave"], 0, bench.NLEVELS - 1)
    q["r"] = np.float32(TH) * scale[o]
    q["min_level"], q["max_level"] = o - 1, o + 1
    q["xr"] = kq["x"] - 20.0
    q["flags"] = gf.PQ_ACTIVE | gf.PQ_BLOCKS | gf.PQ_RIGHT_GATE
    return q


if len(sys.argv) > 2 and sys.argv[2] == "tracking":
    # the tracking case: the frame's own keypoints as the last frame's points after a small motion (3 px sigma, 5 % of
    # the descriptor bits flipped), so nearly every query finds its key -- TrackWithMotionModel's regime
    rng = np.random.default_rng(7)
    k2 = keys.copy()
    k2["x"] += rng.normal(0, 3, k2.shape).astype(np.float32); k2["y"] += rng.normal(0, 3, k2.shape).astype(np.float32)
    qdesc = desc ^ np.packbits(rng.random(desc.shape[:2] + (256,)) < 0.05, axis=2)
    nq = n.copy()
    scenario = "tracking: own keypoints moved by N(0, 3 px), 5 % descriptor bits flipped"
else:
    scenario = "stereo pair as consecutive frames: right-eye keypoints into the left-eye frame"
q = make_queries(k2)
bounds = (0.0, float(bench.W), 0.0, float(bench.H))
D = gf.DeviceArray.from_host
dk, dd, du, dn, dq, dqd, dnq = D(keys), D(desc), D(ur), D(n), D(q), D(qdesc), D(nq)
dkq = gf.DeviceArray((P, K), np.int32); dkd = gf.DeviceArray((P, K), np.int32)
dqk = gf.DeviceArray((P, K), np.int32); dnm = gf.DeviceArray((P,), np.int32)
m = gf.ORBmatcher(0.9, True)


def run(frames):
    t = []
    for it in range(13):
        m.SearchByProjectionBatchDevice(bounds, frames, dk, dd, du, None, dn, K, dq, dqd, dnq, K, 100, False, True, dkq, dkd, dqk, dnm)
        ms = m.sync()
        if it >= 3:
            t.append(ms)
    return float(np.median(t))


ms_one = run(1)
ms = run(P)
kq, qk, nm = dkq.to_host(), dqk.to_host(), dnm.to_host()
S = min(P, 64)
ok = True
t0 = time.perf_counter()
for p in range(S):
    on, okq, _, oqk = O.search_by_projection(keys[p, :n[p]], desc[p, :n[p]], ur[p, :n[p]], None, bounds, q[p, :nq[p]], qdesc[p, :nq[p]],
                                             100, 0.9, False, True)
    ok &= on == nm[p] and np.array_equal(okq, kq[p, :n[p]]) and np.array_equal(oqk, qk[p, :nq[p]])
cpu_s = (time.perf_counter() - t0) / S
# host-pointer form, one frame: what a tracker that keeps its frames on the host pays per call
t = []
for it in range(13):
    t0 = time.perf_counter()
    m.SearchByProjection(keys[0, :n[0]], desc[0, :n[0]], q[0, :nq[0]], qdesc[0, :nq[0]], ur[0, :n[0]], None, 100, False, True, bounds=bounds)
    if it >= 3:
        t.append(time.perf_counter() - t0)
print(json.dumps({"what": "ORBmatcher::SearchByProjection (last-frame form, th 15, rotation check), device", "scenario": scenario,
                  "frames": P, "keys_per_frame": float(n.mean()), "queries_per_frame": float(nq.mean()),
                  "ms_per_batch": ms, "frames_per_s": P / ms * 1e3, "us_per_frame_amortised": ms * 1e3 / P,
                  "us_one_frame_device": ms_one * 1e3, "us_one_frame_host_to_host": float(np.median(t)) * 1e6,
                  "mean_matches": float(nm.mean()), "parity_vs_oracle_sample": bool(ok),
                  "cpu_oracle_us_per_frame_1thread": cpu_s * 1e6, "launches": 3}))
