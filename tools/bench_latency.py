#!/usr/bin/env python3
"""Single-frame latency of the drop-in path (what Tracking sees, one stereo frame at a time, host buffers in / out):
  (a) reference-shaped: two ORBextractor handles called from two threads (src/Frame.cc:268-271) + ComputeStereoMatches;
  (b) the fused one-frame batch entry gforb_stereo_frames_batch(n = 1);
  (c) the CPU oracle, reference-shaped (2 threads, one per eye, serial match) on the same frames.
Prints one JSON line.  usage: python tools/bench_latency.py [frames]"""
import importlib, json, sys, threading, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import bench
gf = importlib.import_module("gf-orb-slam2_b200")
from oracle import pyoracle as O

N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
left, right = bench.make_frames(N)
exL, exR = gf.ORBextractor(bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH), gf.ORBextractor(bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH)


def frame_threads(i):
    res = [None, None]
    def run(k, ex, im): res[k] = ex(im)
    ts = [threading.Thread(target=run, args=(0, exL, left[i])), threading.Thread(target=run, args=(1, exR, right[i]))]
    for t in ts: t.start()
    for t in ts: t.join()
    (kL, dL), (kR, dR) = res
    return gf.ComputeStereoMatches(exL, exR, kL, dL, kR, dR, bench.BF, bench.B)


def frame_serial(i):
    kL, dL = exL(left[i]); kR, dR = exR(right[i])
    return gf.ComputeStereoMatches(exL, exR, kL, dL, kR, dR, bench.BF, bench.B)


sb = gf.StereoBatch(bench.W, bench.H, 1, bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH)
out = sb.alloc_outputs(1)
pl = gf.PinnedArray((1, bench.H, bench.W), np.uint8); pr = gf.PinnedArray((1, bench.H, bench.W), np.uint8)


def frame_batch1(i):
    pl.array[0] = left[i]; pr.array[0] = right[i]
    sb.process(pl.array, pr.array, out, bench.BF, bench.B)


def timeit(fn):
    for i in range(5): fn(i % N)
    ts = []
    for i in range(N):
        t0 = time.perf_counter(); fn(i); ts.append((time.perf_counter() - t0) * 1e3)
    return {"median_ms": round(float(np.median(ts)), 4), "p95_ms": round(float(np.percentile(ts, 95)), 4)}


# (d) a tracking thread's per-frame front-end: the fused one-frame entry, then TrackWithMotionModel's
# SearchByProjection(CurrentFrame, LastFrame, th = 15) with the previous frame's left keypoints as the projected points
matcher = gf.ORBmatcher(0.9, True)
scale = (np.float32(bench.SCALE) ** np.arange(bench.NLEVELS)).astype(np.float32)
last = {}


def frame_track(i):
    frame_batch1(i)
    n = int(out["counts"][0])
    keys, desc, ur = out["keypoints"][0, :n], out["descriptors"][0, :n], out["u_right"][0, :n]
    if last:
        matcher.SearchByProjection(keys, desc, last["q"], last["qd"], ur, None, 100, False, True, width=bench.W, height=bench.H)
    q = np.zeros(n, gf.PROJ_QUERY_DTYPE)                      # the caller's projection step (kept outside the library)
    o = np.clip(keys["octave"], 0, bench.NLEVELS - 1)
    q["x"], q["y"], q["angle"] = keys["x"], keys["y"], keys["angle"]
    q["r"] = np.float32(15.0) * scale[o]
    q["min_level"], q["max_level"] = o - 1, o + 1
    q["xr"] = ur
    q["flags"] = gf.PQ_ACTIVE | gf.PQ_BLOCKS | gf.PQ_RIGHT_GATE
    last["q"], last["qd"] = q, desc.copy()


res = {"what": "one 752x480 stereo frame, 1000 kp, host in / host out", "frames": N,
       "two_handles_two_threads": timeit(frame_threads), "two_handles_serial": timeit(frame_serial),
       "batch_of_one": timeit(frame_batch1),
       "batch_of_one_plus_search_by_projection": timeit(frame_track)}
imgs = np.empty((16,) + left.shape[1:], np.uint8); imgs[0::2] = left[:8]; imgs[1::2] = right[:8]      # frame f = images 2f, 2f+1
secs = O.bench_stereo(imgs, 32, 2, shape=1)
res["cpu_oracle_reference_shaped_ms"] = round((secs[0] if isinstance(secs, tuple) else secs) / 32 * 1e3, 3)
print(json.dumps(res))
