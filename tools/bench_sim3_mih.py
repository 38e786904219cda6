#!/usr/bin/env python3
"""Side bench of the two rows widened in round 2, next to the CPU restatement and (where built) the reference's own code:
  * ORBmatcher::SearchBySim3 (gforb_search_by_sim3): one key-frame pair, host to host and device time;
  * HASHING::MultiIndexHashing (gforb_mih_*): batched insertion and the Tracking-shaped query (8 active tables, 20
    entries each, src/Tracking.cc:3015) over descriptor batches.
Prints one JSON line.  usage: python tools/bench_sim3_mih.py > profiles/sim3_mih_r2.json"""
import importlib, json, os, sys, time
_OUT = os.dup(1); os.dup2(2, 1)      # the reference's MultiIndexHashing constructor prints on stdout: keep it to the one JSON line
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from synth import sim3_case
from test_oracle_mih import descriptors
gf = importlib.import_module("gf-orb-slam2_b200")
from oracle import pyoracle as O, pymih

BOUNDS = (0.0, 752.0, 0.0, 480.0)
out = {}


def best(f, reps=20):
    ts = []
    for _ in range(reps):
        t = time.perf_counter(); r = f(); ts.append(time.perf_counter() - t)
    return min(ts), r


# ---- SearchBySim3
m = gf.ORBmatcher()
for n in (1000, 2000, 8000):
    case = sim3_case(1, n, n)
    m.SearchBySim3(*case, bounds=BOUNDS)
    t_gpu, g = best(lambda: m.SearchBySim3(*case, bounds=BOUNDS))
    dev_ms = m.sync()
    t_cpu, o = best(lambda: O.search_by_sim3(case[0], case[1], case[2], case[3], BOUNDS, *case[4:]), 5)
    assert g[0] == o[0] and np.array_equal(g[1], o[1])
    out[f"sim3_{n}x{n}"] = {"gpu_host_to_host_ms": round(t_gpu * 1e3, 4), "gpu_device_ms": round(dev_ms, 4),
                            "cpu_oracle_ms": round(t_cpu * 1e3, 3), "nfound": int(g[0]), "launches": 7}

# ---- multi-index hashing
rng = np.random.default_rng(0)
for T in (32, 16):
    n, mq = 200000, 100000
    d = descriptors(rng, n, clusters=4000)
    ids = np.arange(n, dtype=np.int32)
    q = d[rng.integers(0, n, mq)] ^ np.packbits(rng.random((mq, 256)) < 0.03, axis=1)
    g = gf.MultiIndexHashing(T)
    t_ins, _ = best(lambda: g.insert(d[:1000], ids[:1000] + 10 ** 7), 1)          # warm-up: allocations, opt-in
    g = gf.MultiIndexHashing(T)
    t0 = time.perf_counter(); ok, _ = g.insert(d, ids); t_ins = time.perf_counter() - t0
    dd = gf.DeviceArray.from_host(q); dc = gf.DeviceArray.from_host(np.full((mq, 160), -1, np.int32)); dn = gf.DeviceArray((mq,), np.int32)
    g.query_device(dd, mq, 8, 20, dc, 160, dn); g.sync()
    t0 = time.perf_counter()
    for _ in range(10):
        g.query_device(dd, mq, 8, 20, dc, 160, dn)
    g.sync(); t_q = (time.perf_counter() - t0) / 10
    t_qh, (cg, ng) = best(lambda: g.query(q[:1000], 8, 20), 10)
    rec = {"insert_entries": n, "gpu_insert_host_to_host_ms": round(t_ins * 1e3, 2), "gpu_insert_M_entries_per_s": round(n / t_ins / 1e6, 2),
           "query_descriptors": mq, "gpu_query_device_ms": round(t_q * 1e3, 3), "gpu_query_M_desc_per_s": round(mq / t_q / 1e6, 1),
           "gpu_query_1000_host_to_host_ms": round(t_qh * 1e3, 3), "mean_candidates_per_query": float(dn.to_host().mean())}
    for variant in ("oracle", "ref"):
        if not pymih.available(variant):
            continue
        o = pymih.MultiIndexHashing(T, variant)
        t0 = time.perf_counter(); ook, _ = o.insert(d, ids); t_oi = time.perf_counter() - t0
        t0 = time.perf_counter(); co, no = o.query(q, 8, 20); t_oq = time.perf_counter() - t0
        assert np.array_equal(ok, ook) and np.array_equal(dn.to_host(), no) and np.array_equal(dc.to_host(), co)
        rec[f"cpu_{variant}_insert_ms"] = round(t_oi * 1e3, 1); rec[f"cpu_{variant}_query_ms"] = round(t_oq * 1e3, 1)
    out[f"mih_T{T}"] = rec
os.write(_OUT, (json.dumps(out) + "\n").encode())
