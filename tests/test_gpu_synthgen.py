"""Bench input generator (tools/synthgen.cu) and the caller-owned device input of the batch engine:
the device kernels and the host loop write identical bytes for a seed; gforb_batch_run from an external device buffer
(gforb_batch_set_device_input) == the same frames uploaded from the host; the Python mirror rejects arrays the C entry
would read out of bounds."""
import importlib
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
from conftest import BF, B


def test_synthgen_host_is_seeded_and_distinct():
    import synthgen
    l, r = synthgen.host_frames(5, 3, 320, 200, threads=2)
    l2, r2 = synthgen.host_frames(6, 2, 320, 200, threads=1)
    assert np.array_equal(l[1:], l2) and np.array_equal(r[1:], r2)         # frame = f(seed) alone
    assert (l[0] != l[1]).mean() > 0.5 and (l[0] != r[0]).mean() > 0.3
    assert 60 < l.mean() < 160 and l.std() > 30


@pytest.mark.gpu
@pytest.mark.parametrize("w,h", [(752, 480), (1280, 720), (333, 257)])
def test_synthgen_device_equals_host(w, h):
    import synthgen
    gf = importlib.import_module("gf-orb-slam2_b200")
    n = 5
    hl, hr = synthgen.host_frames(40, n, w, h)
    dev = gf.DeviceArray((2 * n, h, w), np.uint8)
    synthgen.device_batch_input(40, n, w, h, dev.ptr)
    got = dev.to_host()                                    # gforb_device_copy synchronises the device
    assert np.array_equal(got[0::2], hl) and np.array_equal(got[1::2], hr)


@pytest.mark.gpu
def test_batch_runs_from_external_device_input():
    import synthgen
    gf = importlib.import_module("gf-orb-slam2_b200")
    n, w, h = 6, 752, 480
    hl, hr = synthgen.host_frames(7, n, w, h)
    sb = gf.StereoBatch(w, h, n)
    a, b_ = sb.alloc_outputs(n, pinned=False), sb.alloc_outputs(n, pinned=False)
    sb.process(hl, hr, a, BF, B)
    dev = gf.DeviceArray((2 * n, h, w), np.uint8)
    synthgen.device_batch_input(7, n, w, h, dev.ptr)
    dev.to_host()
    sb.upload(hl[::-1].copy(), hr[::-1].copy())            # the batch's own buffer now holds other frames
    sb.set_device_input(dev.ptr)
    with pytest.raises(gf.GforbError):
        sb.upload(hl, hr)                                  # host-input entries refuse while an external input is set
    sb.run(n, BF, B)
    sb.download(n, b_)
    sb.sync()
    for k in a:
        assert a[k].tobytes() == b_[k].tobytes(), k
    sb.set_device_input(None)
    sb.process(hl, hr, b_, BF, B)
    assert a["depth"].tobytes() == b_["depth"].tobytes()


@pytest.mark.gpu
def test_python_mirror_validates_buffers():
    gf = importlib.import_module("gf-orb-slam2_b200")
    sb = gf.StereoBatch(256, 192, 2)
    out = sb.alloc_outputs(2, pinned=False)
    good = np.zeros((2, 192, 256), np.uint8)
    for bad in (good[:, :, ::2], good.astype(np.float32), np.zeros((2, 256, 192), np.uint8), np.zeros((3, 192, 256), np.uint8),
                good.transpose(0, 2, 1)):
        with pytest.raises(ValueError):
            sb.process(bad, bad, out, BF, B)
    small = dict(out, depth=np.zeros((1, sb.K), np.float32))
    with pytest.raises(ValueError):
        sb.process(good, good, small, BF, B)
    sb.process(good, good, out, BF, B)
