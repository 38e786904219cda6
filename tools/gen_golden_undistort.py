#!/usr/bin/env python3
"""Golden vectors for keypoint undistortion (tests/golden/undistort.npz), made with the genuine cv2.undistortPoints of
this image on seeded points (integer pixel positions, sub-pixel positions, points far outside the image).
Cameras: EuRoC cam0 literature intrinsics (the reference's YAMLs are absent from its tree), a 5-coefficient variant, and a
stereo form with R / P as cv::stereoRectify would hand them to Frame (src/Frame.cc:688).  Run in the build container."""
import sys
from pathlib import Path
import numpy as np
import cv2
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from test_undistort import CAMERAS, points
out = {}
for name, (K, D, R, P) in CAMERAS.items():
    pts = points(name)
    out[name + "_in"] = pts
    out[name + "_out"] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K.astype(np.float32), D.astype(np.float32), R=R,
                                             P=K.astype(np.float32) if P is None else P).reshape(-1, 2)
    b = np.array([[0, 0], [752, 0], [0, 480], [752, 480], [0, 240], [376, 0], [376, 480], [752, 240]], np.float32)
    out[name + "_border"] = cv2.undistortPoints(b.reshape(-1, 1, 2), K.astype(np.float32), D.astype(np.float32), R=R,
                                                P=K.astype(np.float32) if P is None else P).reshape(-1, 2)
np.savez_compressed(ROOT / "tests/golden/undistort.npz", **out)
print({k: v.shape for k, v in out.items()})
