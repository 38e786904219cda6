"""CPU-side checks of the product library: it loads, exports every symbol include/gforb.h declares, and the
host-only entries (tables, Hamming, argument validation) behave.  No compute calls -- there is no GPU here."""
import ctypes as C
import importlib
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
gf = importlib.import_module("gf-orb-slam2_b200")


def test_library_exports_every_declared_symbol():
    hdr = (ROOT / "include/gforb.h").read_text()
    declared = set(re.findall(r"GFORB_API[^;(]*?\b(gforb_\w+)\s*\(", hdr))
    assert len(declared) >= 29
    lib = C.CDLL(gf.LIBRARY_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"libgforb.so does not export {name}"
    assert set(gf._SIGS) == declared


def test_version_and_error_string():
    assert gf.lib.gforb_version() == 1
    assert gf.lib.gforb_last_error_string() is not None


def test_keypoint_layout_is_cv_keypoint():
    assert gf.KP_DTYPE.itemsize == 28
    assert [gf.KP_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == \
        [0, 4, 8, 12, 16, 20, 24]


def test_constructor_tables_match_oracle():
    from oracle import pyoracle as O
    for args in [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (800, 1.2, 8, 20, 7), (500, 1.5, 5, 15, 5),
                 (1000, 2.0, 4, 20, 7), (300, 1.1, 12, 20, 7)]:
        ex = gf.ORBextractor(*args)
        ox = O.Extractor(*args)
        assert np.array_equal(ex.GetScaleFactors(), ox.scale)
        assert np.array_equal(ex.GetInverseScaleFactors(), ox.inv_scale)
        assert np.array_equal(ex.GetScaleSigmaSquares(), ox.sigma2)
        assert np.array_equal(ex.GetInverseScaleSigmaSquares(), ox.inv_sigma2)
        assert np.array_equal(ex.mnFeaturesPerLevel, ox.quota)
        assert np.array_equal(ex.umax, ox.umax)
        assert ex.GetLevels() == args[2] and ex.GetInitThres() == args[3] and ex.GetMinThres() == args[4]


def test_bad_parameters_are_rejected():
    for args in [(1000, 1.0, 8, 20, 7), (1000, 1.2, 0, 20, 7), (1000, 1.2, 13, 20, 7), (-1, 1.2, 8, 20, 7)]:
        with pytest.raises(gf.GforbError) as e:
            gf.ORBextractor(*args)
        assert e.value.code == gf.ERR_ARG


def test_descriptor_distance_host():
    from oracle import pyoracle as O
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    for x, y in zip(a, b):
        assert gf.DescriptorDistance(x, y) == O.hamming256(x, y) == int(np.unpackbits(x ^ y).sum())


def test_empty_image_is_silent_noop():
    ex = gf.ORBextractor()
    k, d = ex(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    n = C.c_int(7)
    assert gf.lib.gforb_extract(ex._h, None, 0, 0, 0, None, None, 0, C.byref(n)) == 0 and n.value == 0


def test_max_keypoints_and_too_small_image():
    ex = gf.ORBextractor()
    assert gf.lib.gforb_extractor_max_keypoints(ex._h, 752, 480) >= 1000 + 8 * 4
    assert gf.lib.gforb_extractor_max_keypoints(ex._h, 120, 100) == gf.ERR_ARG   # level 7 would be 33x28


def test_stereo_before_extract_is_state_error():
    exL, exR = gf.ORBextractor(), gf.ORBextractor()
    k = np.zeros(4, gf.KP_DTYPE); d = np.zeros((4, 32), np.uint8)
    with pytest.raises(gf.GforbError) as e:
        gf.ComputeStereoMatches(exL, exR, k, d, k, d, 40.0, 0.1)
    assert e.value.code == gf.ERR_STATE


def test_nvtx_build_flag_compiles(tmp_path):
    """`make NVTX=1` wraps every stage in a named NVTX range (the reference's USE_NVTX / PUSH_RANGE,
    include/Util_cuda.hpp:39-67).  Compile-only check of the two translation units that carry the ranges."""
    import subprocess
    csrc = ROOT / "gf-orb-slam2_b200" / "csrc"
    for tu in ("engine.cu",):
        obj = tmp_path / (tu + ".o")
        subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-std=c++17",
                               "--fmad=false", "-Xcompiler", "-fPIC", "-DGFORB_NVTX", "-c", str(csrc / tu), "-o", str(obj)])
        syms = subprocess.run(["nm", "-C", str(obj)], capture_output=True, text=True).stdout
        assert "nvtxRangePushEx" in syms or "nvtxGlobals" in syms, "NVTX ranges were compiled away"
