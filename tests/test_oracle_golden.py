"""Oracle vs golden vectors (tests/golden/, produced by tools/gen_golden_cv2.py: the reference's control
flow over genuine cv2 primitives).  Bit-exact on every field: this is what pins the oracle."""
import numpy as np
import pytest

from conftest import B, BF
from oracle import pyoracle as O

NAMES = ("EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r")


def test_tables():
    ex = O.Extractor(1000, 1.2, 8, 20, 7)
    assert list(ex.quota) == [217, 181, 151, 126, 105, 87, 73, 60]
    assert list(ex.umax) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert ex.scale[0] == 1 and abs(ex.scale[7] - 1.2 ** 7) < 1e-5
    ex2 = O.Extractor(2000, 1.2, 8, 20, 7)
    assert list(ex2.quota) == [434, 362, 302, 251, 209, 175, 145, 122]


@pytest.mark.parametrize("name", NAMES)
def test_extract_matches_golden(fixtures, golden, name):
    ex = O.Extractor()
    k, d, lc = ex.extract(fixtures[name])
    g = golden[name]
    assert len(k) == len(g["kps"])
    for f in g["kps"].dtype.names:
        assert np.array_equal(k[f], g["kps"][f]), f
    assert np.array_equal(d, g["desc"])
    assert [len(ex.last_candidates(l)) for l in range(8)] == list(g["cand_counts"])
    for l in (0, 3, 7):
        assert np.array_equal(ex.last_candidates(l), g[f"cand{l}"])
    for l in (1, 4, 7):
        assert np.array_equal(ex.level(l), g[f"pyr{l}"])


def test_euroc_level_sizes_and_candidates(fixtures):
    ex = O.Extractor()
    ex.extract(fixtures["EuRoC_l"])
    sizes = [ex.level(l).shape[::-1] for l in range(8)]
    assert sizes == [(752, 480), (627, 400), (522, 333), (435, 278), (363, 231), (302, 193), (252, 161), (210, 134)]
    # survey-time probe of the reference algorithm on this image
    assert [len(ex.last_candidates(l)) for l in range(8)] == [2688, 1409, 1018, 715, 524, 383, 309, 218]


@pytest.mark.parametrize("pair", ["EuRoC", "TUM"])
def test_stereo_matches_golden(fixtures, golden, pair):
    exL, exR = O.Extractor(), O.Extractor()
    kL, dL, _ = exL.extract(fixtures[pair + "_l"])
    kR, dR, _ = exR.extract(fixtures[pair + "_r"])
    r = O.stereo_match(exL, exR, kL, dL, kR, dR, BF, B)
    g = golden["stereo_" + pair]
    assert np.array_equal(r["best_dist"], g["best_dist"])
    assert np.array_equal(r["best_idx"], g["best_idx"])
    assert np.array_equal(r["uRight"], g["uRight"])
    assert np.array_equal(r["depth"], g["depth"])
    assert r["ret"] == int((g["depth"] > 0).sum())


def test_pad_frame_is_reflect101(fixtures):
    ex = O.Extractor()
    ex.compute_pyramid(fixtures["EuRoC_l"])
    for l in (0, 5):
        assert np.array_equal(ex.level(l, padded=True), O.border_reflect101(ex.level(l), 19))


def test_hamming_popcount():
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    for x, y in zip(a, b):
        assert O.hamming256(x, y) == int(np.unpackbits(x ^ y).sum())
    assert O.hamming256(a[0], a[0]) == 0 and O.hamming256(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


def test_octree_edge_cases():
    # empty, single point, all points identical column, N smaller than initial nodes
    assert len(O.distribute_octree(np.zeros((0, 3), np.int32), 16, 736, 16, 464, 100)) == 0
    assert list(O.distribute_octree(np.array([[5, 5, 30]], np.int32), 16, 736, 16, 464, 100)) == [0]
    pts = np.array([[10, y, 20 + (y % 7)] for y in range(3, 400, 3)], np.int32)
    keep = O.distribute_octree(pts, 16, 736, 16, 464, 50)
    assert 50 <= len(keep) <= 53 and len(set(keep.tolist())) == len(keep)
    keep1 = O.distribute_octree(pts, 16, 736, 16, 464, 1)
    assert len(keep1) >= 1


def test_extract_empty_and_flat():
    ex = O.Extractor()
    k, d, _ = ex.extract(np.full((480, 752), 128, np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
