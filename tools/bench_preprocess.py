#!/usr/bin/env python3
"""Device time of the fused pre-processing kernel (remap rectification of 512 raw 752x480 gray images straight into the
batch input buffer) with its HBM fraction.  Prints one JSON line.  usage: python tools/bench_preprocess.py"""
import importlib, json, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from synth import rectify_maps
gf = importlib.import_module("gf-orb-slam2_b200")
W, H, N = 752, 480, 512
rng = np.random.default_rng(0)
res = {}
peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (ROOT / "MEASURED_PEAKS.json").exists() else 6650.0
for name, fmt, ch, dt in (("gray8+remap", 0, 0, np.uint8), ("gray16+remap", 1, 0, np.uint16), ("rgb8+remap", 2, 3, np.uint8)):
    shape = (N, H, W, ch) if ch else (N, H, W)
    src = gf.DeviceArray.from_host(rng.integers(0, 256, shape).astype(dt))
    dst = gf.DeviceArray((N, H, W), np.uint8)
    r = gf.Rectifier(*rectify_maps(W, H))
    bpp = gf._PIX_BPP[fmt]
    ts = []
    for it in range(13):
        r.batch_device(src, W, H, W * bpp, W * H * bpp, fmt, N, dst, W, W * H)
        ms = r.sync()
        if it >= 3:
            ts.append(ms)
    ms = float(np.median(ts))
    alg = N * W * H * (bpp + 1) + 8 * W * H
    res[name] = {"ms": round(ms, 4), "images_per_s": round(N / ms * 1e3), "alg_bytes": alg, "gbs": round(alg / ms / 1e6, 1),
                 "frac_hbm": round(alg / ms / 1e6 / peak, 4)}
print(json.dumps({"what": "gforb_preprocess_batch_device, 512 images 752x480", "hbm_peak_gbs": peak, "cases": res}))
