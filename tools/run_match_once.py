#!/usr/bin/env python3
"""one brute-force match call on device-resident random descriptors (for ncu captures): run_match_once.py N [reps]"""
import importlib, sys
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
gf = importlib.import_module("gf-orb-slam2_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rng = np.random.default_rng(0)
a = rng.integers(0, 256, (n, 32), dtype=np.uint8)
b = a[rng.permutation(n)] ^ np.packbits(rng.random((n, 256)) < 0.1, axis=1)
m = gf.ORBmatcher(0.6)
da, db = gf.DeviceArray.from_host(a), gf.DeviceArray.from_host(b)
o = [gf.DeviceArray((n,), np.int32) for _ in range(4)]
for _ in range(reps):
    m.BruteForceDevice(da, n, db, n, 50, *o)
    print(n, m.sync(), "ms")
