#!/usr/bin/env python3
"""End-to-end (host buffers in/out) throughput vs pipeline chunk size, plus the raw pinned H2D/D2H rates of the box.
usage: python tools/e2e_sweep.py [frames]"""
import importlib, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch
import bench
gf = importlib.import_module("gf-orb-slam2_b200")
F = int(sys.argv[1]) if len(sys.argv) > 1 else 256
left, right = bench.make_frames(F)
sb = gf.StereoBatch(bench.W, bench.H, F, bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH)
out = sb.alloc_outputs(F)
pl = gf.PinnedArray((F, bench.H, bench.W), np.uint8); pr = gf.PinnedArray((F, bench.H, bench.W), np.uint8)
pl.array[:] = left; pr.array[:] = right
# raw copy rates
n = 2 * F * bench.W * bench.H
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, (a, b_) in (("H2D", (d, h)), ("D2H", (h, d))):
    a.copy_(b_, non_blocking=True); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): a.copy_(b_, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(f"{name}: {5 * n / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
for chunk in (0, 16, 32, 64):
    if chunk > F: break
    sb.set_chunk(chunk)
    for _ in range(2): sb.process(pl.array, pr.array, out, bench.BF, bench.B)
    t0 = time.perf_counter()
    for _ in range(8): sb.process(pl.array, pr.array, out, bench.BF, bench.B)
    dt = (time.perf_counter() - t0) / 8
    print(f"chunk {chunk if chunk else 'auto':>4}: {dt * 1e3:7.3f} ms/step  {F / dt:9.0f} stereo frames/s")
