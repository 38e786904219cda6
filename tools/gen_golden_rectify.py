#!/usr/bin/env python3
"""Golden vectors for the pre-processing path (tests/golden/rectify.npz), made with the genuine OpenCV in this image
(cv2 4.13) on the reference's own fixtures -- run in the build container (needs /root/reference and cv2):
  * cv2.remap(INTER_LINEAR) of test/EuRoC_l.png / EuRoC_r.png with the deterministic maps of tests/synth.py:rectify_maps
    (the reference's camera YAMLs are not in its tree, test/test_Stereo.cpp:87-95, so no real calibration exists);
  * a 256x256 16-bit crop of test/TUM_l.png with cv2.imread(IMREAD_GRAYSCALE)'s 8-bit rendering of the same crop,
    and its remap;
  * cv2.cvtColor(remap(colour)) of a synthetic colour image derived from EuRoC_l.
Also records a checksum of the maps so that a machine computing different map bits is detected."""
import sys, zlib
from pathlib import Path
import numpy as np
import cv2
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
from synth import rectify_maps
ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference") / "test"
out = {}
l = cv2.imread(str(ref / "EuRoC_l.png"), cv2.IMREAD_GRAYSCALE); r = cv2.imread(str(ref / "EuRoC_r.png"), cv2.IMREAD_GRAYSCALE)
h, w = l.shape
m1l, m2l = rectify_maps(w, h)
m1r, m2r = rectify_maps(w, h, k1=-0.26, k2=0.05, tilt=-0.009, shift=(-4.5, 1.75), scale=1.04)
out["maps_crc"] = np.array([zlib.crc32(m.tobytes()) for m in (m1l, m2l, m1r, m2r)], np.uint32)
out["EuRoC_l_rect"] = cv2.remap(l, m1l, m2l, cv2.INTER_LINEAR)
out["EuRoC_r_rect"] = cv2.remap(r, m1r, m2r, cv2.INTER_LINEAR)
t16 = cv2.imread(str(ref / "TUM_l.png"), cv2.IMREAD_UNCHANGED); t8 = cv2.imread(str(ref / "TUM_l.png"), cv2.IMREAD_GRAYSCALE)
assert t16.dtype == np.uint16 and t8.dtype == np.uint8
crop = np.ascontiguousarray(t16[128:384, 128:384])
out["TUM_l_crop16"] = crop
out["TUM_l_crop8"] = np.ascontiguousarray(t8[128:384, 128:384])                      # what imread(GRAYSCALE) hands the reference
print("imread(GRAYSCALE) == high byte:", bool(np.array_equal(t8, (t16 >> 8).astype(np.uint8))))
mc = rectify_maps(256, 256, shift=(1.5, 0.25))
out["TUM_l_crop_rect"] = cv2.remap(out["TUM_l_crop8"], mc[0], mc[1], cv2.INTER_LINEAR)
col = np.stack([l, np.roll(l, 7, axis=1), 255 - np.roll(l, 3, axis=0)], axis=-1)      # synthetic RGB from the fixture
out["colour_rect_gray"] = cv2.cvtColor(cv2.remap(col, m1l, m2l, cv2.INTER_LINEAR), cv2.COLOR_RGB2GRAY)
np.savez_compressed(ROOT / "tests/golden/rectify.npz", **out)
print({k: (v.shape, str(v.dtype)) for k, v in out.items()})
