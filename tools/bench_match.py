"""BASELINE.json configs[3]: brute-force 256-bit Hamming sweep N x N, N = 1k .. 64k, best / second-best + ratio test.
Device-resident descriptors; device time from CUDA events around the two kernels (gforb_matcher_sync).  Reports
G pairs/s and the fraction of the POPC-issue roofline (8 POPC per pair, 16 POPC lanes/clk/SM measured by
tools/int_peaks.cu).  Usage: python tools/bench_match.py [--max 65536] [--cpu] > profiles/match_sweep_rN.json"""
import argparse
import os
import importlib
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max", type=int, default=65536)
    ap.add_argument("--cpu", action="store_true", help="also time the oracle loop (all host threads) on N <= 8192")
    ap.add_argument("--sms", type=int, default=148); ap.add_argument("--mhz", type=float, default=1965.0)
    a = ap.parse_args()
    gf = importlib.import_module("gf-orb-slam2_b200")
    from test_gpu_match import make_sets
    peak_pairs = a.sms * a.mhz * 1e6 * 16 / 8        # POPC-bound pairs/s
    # tensor roofline of the tcgen05 path (sets of >= 2,048 a side): 2 * 256 int8 ops per pair against the dense int8 rate,
    # taken as twice the MEASURED dense bf16 rate (MEASURED_PEAKS.json: cuBLAS 8192^3; int8 runs at 2x bf16 on this part)
    peaks = ROOT / "MEASURED_PEAKS.json"
    bf16 = json.loads(peaks.read_text())["bf16_tflops"] if peaks.exists() else 1600.0
    int8_tops = 2.0 * bf16
    rows = []
    n = 1024
    while n <= a.max:
        A, B = make_sets(n, n, n)
        m = gf.ORBmatcher(0.6)
        dA, dB = gf.DeviceArray.from_host(A), gf.DeviceArray.from_host(B)
        outs = [gf.DeviceArray((n,), np.int32) for _ in range(4)]
        for _ in range(3):
            m.BruteForceDevice(dA, n, dB, n, 50, *outs)
        m.sync()
        ms = []
        for _ in range(5):
            m.BruteForceDevice(dA, n, dB, n, 50, *outs)
            ms.append(m.sync())
        t = float(np.median(ms))
        t0 = time.perf_counter(); r = m.BruteForce(A, B, 50); e2e = time.perf_counter() - t0
        row = dict(n=n, m=n, ms=round(t, 4), gpairs_s=round(n * n / t / 1e6, 1), popc_roofline_frac=round(n * n / (t * 1e-3) / peak_pairs, 3),
                   e2e_ms=round(e2e * 1e3, 3), matched_frac=round(float((r["match"] >= 0).mean()), 4))
        if n >= 2048 and not os.environ.get("GFORB_NO_TC"):
            tops = 2.0 * 256 * n * n / (t * 1e-3) / 1e12
            row.update(path="tcgen05 kind::i8", int8_tops=round(tops, 1), tensor_roofline_frac=round(tops / int8_tops, 3))
        else:
            row.update(path="popc")
        if a.cpu and n <= 8192:
            from oracle import pyoracle as O
            th = os.cpu_count() or 1
            t0 = time.perf_counter(); ref = O.match_bruteforce(A, B, 0.6, 50, nthreads=th); c = time.perf_counter() - t0
            row.update(cpu_ms=round(c * 1e3, 2), cpu_threads=th, cpu_gpairs_s=round(n * n / c / 1e9, 3),
                       parity=bool(all(np.array_equal(r[k], ref[k]) for k in ref)))
        rows.append(row)
        print(json.dumps(row), file=sys.stderr)
        n *= 2
    print(json.dumps(dict(workload="configs[3]: brute-force Hamming N x N, ratio 0.6, TH_LOW 50", popc_peak_gpairs_s=round(peak_pairs / 1e9, 1),
                          int8_peak_tops=round(int8_tops, 1), int8_peak_source="2 x MEASURED_PEAKS.json bf16_tflops (dense)", rows=rows)))


if __name__ == "__main__":
    main()
