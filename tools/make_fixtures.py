#!/usr/bin/env python3
"""Decode the reference's four PNG fixtures (test/*.png) to 8-bit binary PGM under tests/data/.

Runs only in the build container (needs /root/reference and cv2).  Decoding uses
cv2.imread(..., IMREAD_GRAYSCALE), i.e. exactly what the reference's own test does
(test/test_Stereo.cpp:120-135); the 16-bit TUM-VI images come out as their high byte.
"""
import sys
from pathlib import Path
import cv2

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
out = Path(__file__).resolve().parent.parent / "tests/data"
out.mkdir(parents=True, exist_ok=True)
for name in ("EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r"):
    im = cv2.imread(str(ref / "test" / f"{name}.png"), cv2.IMREAD_GRAYSCALE)
    assert im is not None and im.dtype.name == "uint8" and im.ndim == 2
    h, w = im.shape
    (out / f"{name}.pgm").write_bytes(b"P5\n%d %d\n255\n" % (w, h) + im.tobytes())
    print(name, w, h, int(im.mean()))
