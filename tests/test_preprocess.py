"""Pre-processing in front of the extractor: 16-bit -> 8-bit, cv::remap rectification, colour -> gray.
CPU part: the oracle reproduces the cv2-made golden vectors (tools/gen_golden_rectify.py).  GPU part (-m gpu): the CUDA
kernel through the C ABI is bit-identical to the oracle and to the golden vectors, single and batched; a rectified pair
then runs through extraction + stereo matching on the device without touching the host."""
import importlib
import zlib
from pathlib import Path

import numpy as np
import pytest

from conftest import B, BF
from synth import rectify_maps, synth_stereo
from oracle import pyoracle as O

GOLD = Path(__file__).resolve().parent / "golden" / "rectify.npz"
R_ARGS = dict(k1=-0.26, k2=0.05, tilt=-0.009, shift=(-4.5, 1.75), scale=1.04)


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def _colour(l):
    return np.ascontiguousarray(np.stack([l, np.roll(l, 7, axis=1), 255 - np.roll(l, 3, axis=0)], axis=-1))


def _cases(fixtures, gold):
    l, r = fixtures["EuRoC_l"], fixtures["EuRoC_r"]
    h, w = l.shape
    ml, mr = rectify_maps(w, h), rectify_maps(w, h, **R_ARGS)
    crc = [zlib.crc32(m.tobytes()) for m in (*ml, *mr)]
    assert crc == gold["maps_crc"].tolist(), "rectify_maps computed different bits than the golden run"
    mc = rectify_maps(256, 256, shift=(1.5, 0.25))
    return [("EuRoC_l", l, O.PIX_GRAY8, ml, gold["EuRoC_l_rect"]), ("EuRoC_r", r, O.PIX_GRAY8, mr, gold["EuRoC_r_rect"]),
            ("TUM16", gold["TUM_l_crop16"], O.PIX_GRAY16, (None, None), gold["TUM_l_crop8"]),
            ("TUM16 rect", gold["TUM_l_crop16"], O.PIX_GRAY16, mc, gold["TUM_l_crop_rect"]),
            ("colour", _colour(l), O.PIX_RGB8, ml, gold["colour_rect_gray"])]


def test_oracle_reproduces_cv2_golden(fixtures, gold):
    for name, src, fmt, (mx, my), want in _cases(fixtures, gold):
        assert np.array_equal(O.preprocess(src, fmt, mx, my), want), name


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.mark.gpu
def test_gpu_preprocess_matches_golden_and_oracle(gf, fixtures, gold):
    for name, src, fmt, (mx, my), want in _cases(fixtures, gold):
        got = gf.Rectifier(mx, my)(src, fmt)
        assert np.array_equal(got, want), name
    rng = np.random.default_rng(9)
    for (h, w, dh, dw) in [(97, 131, 97, 131), (120, 168, 75, 101), (64, 64, 200, 333), (33, 47, 33, 47)]:
        mx = (rng.random((dh, dw), dtype=np.float32) * (w + 30) - 15).astype(np.float32)
        my = (rng.random((dh, dw), dtype=np.float32) * (h + 30) - 15).astype(np.float32)
        mx[0, : min(dw, 64)] = (np.arange(min(dw, 64), dtype=np.float32) / 32 - 1.25)
        rect, plain = gf.Rectifier(mx, my), gf.Rectifier()
        for fmt, ch, dt in ((0, 0, np.uint8), (1, 0, np.uint16), (2, 3, np.uint8), (3, 3, np.uint8), (4, 4, np.uint8), (5, 4, np.uint8)):
            shape = (h, w, ch) if ch else (h, w)
            im = rng.integers(0, np.iinfo(dt).max + 1, shape).astype(dt)
            assert np.array_equal(rect(im, fmt), O.preprocess(im, fmt, mx, my)), (fmt, h, w)
            assert np.array_equal(plain(im, fmt), O.preprocess(im, fmt)), (fmt, h, w)
        strided = np.ascontiguousarray(rng.integers(0, 256, (h, w + 13), dtype=np.uint8))[:, :w]      # pitch != width
        assert np.array_equal(rect(strided, 0), O.preprocess(np.ascontiguousarray(strided), 0, mx, my))
    r = gf.Rectifier()
    with pytest.raises(gf.GforbError):
        gf._check(gf.lib.gforb_preprocess(r._h, None, 10, 10, 10, 0, None, 10))


@pytest.mark.gpu
def test_gpu_rectify_into_batch_then_extract_and_match(gf):
    """Raw (unrectified) stereo frames are uploaded once; the rectifier writes straight into the batch's device input
    buffer; extraction + stereo matching follow on the device.  Result == oracle on the oracle-rectified images."""
    n, w, h = 6, 752, 480
    ml, mr = rectify_maps(w, h), rectify_maps(w, h, **R_ARGS)
    raw = [synth_stereo(300 + i, w, h, nshapes=300) for i in range(n)]
    sb = gf.StereoBatch(w, h, n)
    ptr, stride, pitch = sb.device_input()
    rl, rr = gf.Rectifier(*ml), gf.Rectifier(*mr)
    dl = gf.DeviceArray.from_host(np.stack([p[0] for p in raw])); dr = gf.DeviceArray.from_host(np.stack([p[1] for p in raw]))
    # batch layout: image 2f = left, 2f+1 = right  -> destination image stride 2 * stride
    rl.batch_device(dl, w, h, w, w * h, gf.PIX_GRAY8, n, ptr, pitch, 2 * stride)
    rr.batch_device(dr, w, h, w, w * h, gf.PIX_GRAY8, n, ptr + stride, pitch, 2 * stride)
    assert rl.sync() > 0 and rr.sync() > 0
    sb.run(n, BF, B)
    out = sb.alloc_outputs(n, pinned=False)
    sb.download(n, out); sb.sync()
    for f in (0, n - 1):
        il, ir = O.preprocess(raw[f][0], 0, *ml), O.preprocess(raw[f][1], 0, *mr)
        oxL, oxR = O.Extractor(), O.Extractor()
        kL, dL, _ = oxL.extract(il); kR, dR, _ = oxR.extract(ir)
        ref = O.stereo_match(oxL, oxR, kL, dL, kR, dR, BF, B)
        nL, nR = out["counts"][2 * f], out["counts"][2 * f + 1]
        assert nL == len(kL) and nR == len(kR)
        for fld in ("x", "y", "size", "angle", "response", "octave"):
            assert np.array_equal(out["keypoints"][2 * f, :nL][fld], kL[fld]), fld
        assert np.array_equal(out["descriptors"][2 * f, :nL], dL) and np.array_equal(out["descriptors"][2 * f + 1, :nR], dR)
        assert np.array_equal(out["u_right"][f, :nL], ref["uRight"]) and np.array_equal(out["depth"][f, :nL], ref["depth"])
