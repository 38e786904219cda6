"""GPU parity against THE REFERENCE'S OWN CODE (oracle/_ref/libgforb_ref.so: src/ORBextractor.cc compiled unmodified,
Frame::ComputeStereoMatches[_Undistorted] lifted verbatim; see tests/test_oracle_vs_ref.py for how it is built and why
it runs under an address-monotonic operator new).  The prebuilt library travels to the GPU box; nothing here reads
/root/reference.  Bit-exact on every field: keypoint records (28 bytes each, in order), descriptors, padded pyramid
levels, FAST candidates, mvuRight, mvDepth.
"""
import importlib

import numpy as np
import pytest

from conftest import B, BF
from synth import synth_stereo

pytestmark = pytest.mark.gpu
NAMES = ("EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r")


@pytest.fixture(scope="module")
def gf():
    return importlib.import_module("gf-orb-slam2_b200")


@pytest.fixture(scope="module")
def R():
    from oracle import pyref
    if not pyref.available("ref"):
        pytest.skip("oracle/_ref/libgforb_ref.so was not shipped")
    m = pyref.load("ref")
    m.monotonic_new(1)
    return m


@pytest.mark.parametrize("name", NAMES)
def test_extract_fixture_vs_reference(gf, R, fixtures, name):
    im = fixtures[name]
    ex, rx = gf.ORBextractor(), R.Extractor()
    k, d = ex(im)
    rk, rd, _ = rx.extract(im)
    for l in range(8):
        assert np.array_equal(ex.pyramid_level(l, with_border=True), rx.level(l, padded=True)), f"pyramid level {l}"
        assert np.array_equal(ex.debug_candidates(l), rx.last_candidates(l)), f"FAST candidates level {l}"
    assert np.ascontiguousarray(k).tobytes() == rk.tobytes()
    assert np.array_equal(d, rd)


@pytest.mark.parametrize("pair", ["EuRoC", "TUM"])
@pytest.mark.parametrize("mode", [0, 1])
def test_stereo_fixture_vs_reference(gf, R, fixtures, pair, mode):
    left, right = fixtures[pair + "_l"], fixtures[pair + "_r"]
    exL, exR = gf.ORBextractor(), gf.ORBextractor()
    kL, dL = exL(left); kR, dR = exR(right)
    rxL, rxR = R.Extractor(), R.Extractor()
    rkL, rdL, _ = rxL.extract(left); rkR, rdR, _ = rxR.extract(right)
    got = gf.ComputeStereoMatches(exL, exR, kL, dL, kR, dR, BF, B, mode=mode)
    want = R.stereo_match(rxL, rxR, rkL, rdL, rkR, rdR, BF, B, mode=mode)
    assert want["ret"] >= 0
    assert np.asarray(got["mvuRight"]).tobytes() == want["uRight"].tobytes()
    assert np.asarray(got["mvDepth"]).tobytes() == want["depth"].tobytes()
    assert got["nmatched"] == want["ret"]


def test_batch_vs_reference(gf, R):
    """The batched entry (the one bench.py times) on 6 synthetic stereo frames: every eye, every output."""
    n, w, h = 6, 752, 480
    frames = [synth_stereo(900 + i, w, h) for i in range(n)]
    left = np.stack([f[0] for f in frames]); right = np.stack([f[1] for f in frames])
    sb = gf.StereoBatch(w, h, n)
    out = sb.alloc_outputs(n)
    sb.process(left, right, out, BF, B)
    for i in range(n):
        rxL, rxR = R.Extractor(), R.Extractor()
        rkL, rdL, _ = rxL.extract(left[i]); rkR, rdR, _ = rxR.extract(right[i])
        rs = R.stereo_match(rxL, rxR, rkL, rdL, rkR, rdR, BF, B)
        nL, nR = int(out["counts"][2 * i]), int(out["counts"][2 * i + 1])
        assert (nL, nR) == (len(rkL), len(rkR))
        assert out["keypoints"][2 * i, :nL].tobytes() == rkL.tobytes()
        assert out["keypoints"][2 * i + 1, :nR].tobytes() == rkR.tobytes()
        assert np.array_equal(out["descriptors"][2 * i, :nL], rdL) and np.array_equal(out["descriptors"][2 * i + 1, :nR], rdR)
        assert out["u_right"][i, :nL].tobytes() == rs["uRight"].tobytes()
        assert out["depth"][i, :nL].tobytes() == rs["depth"].tobytes()


def test_engines_of_different_footprints_coexist(gf, R, fixtures):
    """The reference's monocular Tracking keeps mpIniORBextractor (2 * nFeatures) and mpORBextractorLeft (nFeatures)
    alive together and alternates between them (src/Tracking.cc:247-254); a second image size on top.  The kernels'
    shared-memory opt-in is a per-kernel attribute: it must only ever grow."""
    a = gf.ORBextractor(2000, 1.2, 8, 20, 7)
    ka, da = a(fixtures["EuRoC_l"])
    b = gf.ORBextractor(1000, 1.2, 8, 20, 7)
    kb, db = b(fixtures["EuRoC_l"])
    c = gf.ORBextractor(500, 1.2, 8, 20, 7)
    kc, dc = c(fixtures["TUM_l"])
    ka2, da2 = a(fixtures["EuRoC_l"])                      # the big one again, after smaller engines configured
    kb2, db2 = b(fixtures["TUM_r"])                        # and another image size through a live engine
    assert np.ascontiguousarray(ka).tobytes() == np.ascontiguousarray(ka2).tobytes() and np.array_equal(da, da2)
    for ex_params, img, k, d in ((2000, "EuRoC_l", ka2, da2), (1000, "TUM_r", kb2, db2), (500, "TUM_l", kc, dc)):
        rk, rd, _ = R.Extractor(ex_params, 1.2, 8, 20, 7).extract(fixtures[img])
        assert np.ascontiguousarray(k).tobytes() == rk.tobytes() and np.array_equal(d, rd)


def test_blurred_levels_bit_exact(gf, fixtures):
    """a9 (src/ORBextractor.cc:1153-1155): the GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) copy of every level, pixel for
    pixel, against the cv2-pinned primitive the reference build links (oracle cvprim) -- directly, not only through the
    descriptors sampled from it."""
    from oracle import pyoracle as O
    for name, args in (("EuRoC_l", (1000, 1.2, 8, 20, 7)), ("TUM_r", (1500, 1.3, 6, 20, 7))):
        im = fixtures[name]
        ex = gf.ORBextractor(*args); ox = O.Extractor(*args)
        ex(im); ox.compute_pyramid(im)
        for l in range(args[2]):
            assert np.array_equal(ex.debug_blurred_level(l), O.gaussian_blur(ox.level(l))), (name, l)


# EuRoC.yaml literature values (parameters only): cam0 / cam1 intrinsics, distortion, rectifying rotations, projection
_EUROC = dict(
    KL=[[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]], DL=[-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05],
    RL=[[0.999966347530033, -0.001422739138722922, 0.008079580483432283],
        [0.001365741834644127, 0.9999741760894847, 0.007055629199258132],
        [-0.008089410156878961, -0.007044357138835809, 0.9999424675829176]],
    KR=[[457.587, 0, 379.999], [0, 456.134, 255.238], [0, 0, 1]], DR=[-0.28368365, 0.07451284, -0.00010473, -3.555907e-05],
    RR=[[0.9999633526194376, -0.003625811871560086, 0.007755443660172947],
        [0.003680398547259526, 0.9999684752771629, -0.007035845251224894],
        [-0.007729688520722713, 0.007064130529506649, 0.999945173484644]],
    P=[[435.2046959714599, 0, 367.4517211914062, 0], [0, 435.2046959714599, 252.2008514404297, 0], [0, 0, 1, 0]])


def test_undistorted_stereo_on_the_euroc_pair(gf, R, fixtures):
    """a14 on fixture-derived keys (the default build's chain, src/Frame.cc:61-147): extract both EuRoC eyes, undistort both
    key sets on the device (UndistortKeyPointsStereo with the EuRoC.yaml coefficients), then
    ComputeStereoMatches_Undistorted on the undistorted keys -- against the restatement AND the reference's own lifted
    member (oracle/_ref), bit for bit."""
    from oracle import pyoracle as O
    E = {k: np.asarray(v, np.float64) for k, v in _EUROC.items()}
    left, right = fixtures["EuRoC_l"], fixtures["EuRoC_r"]
    exL, exR = gf.ORBextractor(), gf.ORBextractor()
    kL, dL = exL(left); kR, dR = exR(right)
    m = gf.ORBmatcher()
    unL = gf.UndistortKeyPoints(m, gf.Camera.make(E["KL"], E["DL"], E["RL"], E["P"]), kL)
    unR = gf.UndistortKeyPoints(m, gf.Camera.make(E["KR"], E["DR"], E["RR"], E["P"]), kR)
    for un, k, K, D, Rm in ((unL, kL, E["KL"], E["DL"], E["RL"]), (unR, kR, E["KR"], E["DR"], E["RR"])):
        xy = O.undistort_points(np.stack([k["x"], k["y"]], 1), K, D, Rm, E["P"])
        assert np.array_equal(np.stack([un["x"], un["y"]], 1), xy)
        assert np.abs(un["x"] - k["x"]).max() > 1.0                      # the coefficients really move the keys
    rxL, rxR = R.Extractor(), R.Extractor()
    rxL.extract(left); rxR.extract(right)
    oxL, oxR = O.Extractor(), O.Extractor()
    oxL.extract(left); oxR.extract(right)
    for min_x in (0.0, 25.0):
        got = gf.ComputeStereoMatches(exL, exR, unL, dL, unR, dR, BF, B, mode=1, min_x=min_x)
        for name, mod, a, b_ in (("restatement", O, oxL, oxR), ("reference", R, rxL, rxR)):
            want = mod.stereo_match(a, b_, np.ascontiguousarray(unL), dL, np.ascontiguousarray(unR), dR, BF, B, mode=1, min_x=min_x)
            assert want["ret"] > 50, name
            assert np.asarray(got["mvuRight"]).tobytes() == want["uRight"].tobytes(), name
            assert np.asarray(got["mvDepth"]).tobytes() == want["depth"].tobytes(), name
            assert got["nmatched"] == want["ret"], name
