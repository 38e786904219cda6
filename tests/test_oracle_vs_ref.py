"""The restatement (oracle/) held against THE REFERENCE'S OWN CODE (oracle/_ref), bit for bit.

oracle/_ref/libgforb_ref.so is /root/reference/src/ORBextractor.cc compiled whole and unmodified, plus the member
functions oracle/ref_extract.py lifts verbatim by line range out of src/Frame.cc, include/Frame.h and
src/ORBmatcher.cc, over a cv::-compatible shim whose image primitives are pinned against cv2 (tests/test_oracle_cv2.py).
Both libraries export the entry points of oracle/gforb_oracle.h; every test below feeds both the same inputs and
demands identical bits.  This is what turns "the oracle reads faithful" into "the oracle IS the reference":
    ORBextractor ctor tables            src/ORBextractor.cc:409-469
    ComputePyramid                      :1176-1201      (padded levels, frame included)
    cell loop + FAST + retry            :767-831        (vToDistributeKeys, in emission order)
    DistributeOctTree / DivideNode      :481-763        (fixtures, synthetic frames, random and adversarial inputs)
    operator()                          :1112-1174      (keys in order, all 7 fields, descriptors)
    Frame::ComputeStereoMatches         src/Frame.cc:889-1078
    ..._Undistorted + PrepareStereoCandidates           src/Frame.cc:1167-1316, include/Frame.h:230-263
    AssignFeaturesToGrid / GetFeaturesInArea / PosInGrid  src/Frame.cc:461-476,593-658
    ORBmatcher::SearchForInitialization + ComputeThreeMaxima + DescriptorDistance  src/ORBmatcher.cc:520-633,1723-1784
    ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>, th)                  src/ORBmatcher.cc:155-249

One thing had to be settled first: the reference breaks quadtree size ties by HEAP ADDRESS (it sorts
pair<int, ExtractorNode*>, :591,684), so under glibc malloc it returns different keypoints for the same image from
one call to the next (test_reference_is_allocator_dependent).  Inside the entry points the reference build therefore
runs under an address-monotonic operator new (oracle/ref_shim/mono_new.cpp), which makes the tie rule "the node
created later is split first" -- the rule the restatement and the CUDA path froze.

On the GPU box /root/reference does not exist; the prebuilt libgforb_ref.so travels with the snapshot.  The
`-march=native` build of the reference (its own CMake flags, FMA contraction allowed) is only compared where it was
built, i.e. where /root/reference exists.
"""
import numpy as np
import pytest

from oracle import pyoracle as O
from oracle import pyref
import synth
from conftest import B, BF

pytestmark = pytest.mark.skipif(not pyref.available("ref"), reason="oracle/_ref is not built and /root/reference is absent")

F32 = np.float32
BOUNDS = (0.0, 752.0, 0.0, 480.0)


@pytest.fixture(scope="module")
def R():
    m = pyref.load("ref")
    m.monotonic_new(1)
    return m


@pytest.fixture(scope="module")
def RN():
    if not O.REFERENCE_ROOT.exists():
        pytest.skip("the -march=native build of the reference is only valid on the machine that built it")
    m = pyref.load("ref_native")
    m.monotonic_new(1)
    return m


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def assert_same_extraction(mo, mr, img, params=(1000, 1.2, 8, 20, 7), tag=""):
    eo, er = mo.Extractor(*params), mr.Extractor(*params)
    ko, do, lo = eo.extract(img)
    kr, dr, lr = er.extract(img)
    nlev = params[2]
    for l in range(nlev):
        assert np.array_equal(eo.level(l, True), er.level(l, True)), f"{tag}: padded pyramid level {l}"
        assert np.array_equal(eo.last_candidates(l), er.last_candidates(l)), f"{tag}: FAST candidates level {l}"
    assert np.array_equal(lo, lr), f"{tag}: per-level counts {lo} vs {lr}"
    assert bits(ko) == bits(kr), f"{tag}: keypoints (28-byte records, in order)"
    assert np.array_equal(do, dr), f"{tag}: descriptors"
    return eo, er, ko, do


# ------------------------------------------------------------------------------------------------ extractor
PARAM_SETS = [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (500, 1.2, 8, 20, 7), (1500, 1.1, 12, 12, 5),
              (300, 1.5, 4, 30, 10), (1200, 1.3, 6, 25, 7), (4000, 1.2, 8, 20, 7)]


@pytest.mark.parametrize("params", PARAM_SETS)
def test_constructor_tables(R, params):
    eo, er = O.Extractor(*params), R.Extractor(*params)
    for f in ("scale", "inv_scale", "sigma2", "inv_sigma2", "quota", "umax", "pattern"):
        assert bits(getattr(eo, f)) == bits(getattr(er, f)), f


@pytest.mark.parametrize("name", ["EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r"])
def test_fixtures(R, fixtures, name):
    assert_same_extraction(O, R, fixtures[name], tag=name)


@pytest.mark.parametrize("name", ["EuRoC_l", "TUM_l"])
@pytest.mark.parametrize("params", PARAM_SETS[1:])
def test_fixtures_parameter_sets(R, fixtures, name, params):
    assert_same_extraction(O, R, fixtures[name], params, tag=f"{name} {params}")


@pytest.mark.parametrize("name", ["EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r"])
def test_fixtures_native_flags(R, RN, fixtures, name):
    """The reference built with its own flags (-O3 -march=native, contraction allowed) returns the same bits as the
    contraction-off build: no float expression on the path changes under FMA on these images."""
    assert_same_extraction(R, RN, fixtures[name], tag=name + " native")


SIZES = [(752, 480), (640, 480), (512, 512), (1280, 720), (376, 240), (333, 217), (160, 120)]


def fit_levels(params, w, h):
    """Fewer levels where the smallest one would fall under 64 px: below ~33 px a level has no cell grid and the
    reference divides by zero (nIni = round(width / 0), src/ORBextractor.cc:547) and dies in vector::resize."""
    nf, sf, nl, a, b = params
    while nl > 1 and min(w, h) / sf ** (nl - 1) < 64:
        nl -= 1
    return (nf, sf, nl, a, b)


def _random_case(c, rng):
    """the distribution tools/stress_extract.py walks on the GPU (sizes with every w % 4, 1-8 levels, scale 1.1-2.0, 10-4,000
    features, FAST thresholds down to 0), so that what the CUDA path is stressed against is itself pinned by the reference"""
    nlevels = int(rng.integers(1, 9))
    scale = float(rng.choice([1.2, 1.2, 1.2, 1.1, 1.3, 1.5, 2.0]))
    while 48 * scale ** (nlevels - 1) > 600:
        nlevels -= 1
    smallest = 48 * scale ** (nlevels - 1)
    w = int(rng.integers(max(int(smallest) + 8, 96), 1000)); h = int(rng.integers(max(int(smallest) + 8, 96), 760))
    nfeat = int(rng.choice([10, 50, 300, 1000, 2000, 4000]))
    ini = int(rng.choice([20, 20, 12, 40, 3, 0, 120])); mn = int(rng.choice([7, 7, 5, ini, 0, 1]))
    if min(ini, mn) <= 3:
        w, h = min(w, int(smallest) + 8 + 200), min(h, int(smallest) + 8 + 140)
    return w, h, (nfeat, scale, nlevels, ini, min(mn, ini)), int(rng.integers(60, 400))


@pytest.mark.parametrize("block", range(6))
def test_random_sizes_and_parameters(R, block):
    rng = np.random.default_rng(9000 + block)
    for c in range(6):
        w, h, params, nshapes = _random_case(c, rng)
        if w < h:            # portrait levels make nIni = round(width / height) = 0 in the reference: a division by zero and an
            w, h = h, w      # out-of-range vector index (src/ORBextractor.cc:543-569) -- it crashes; frozen here as "no keypoints"
        left, right = synth.synth_stereo(5000 + 10 * block + c, w, h, nshapes=nshapes)
        assert_same_extraction(O, R, left, params, tag=f"block {block} case {c} {w}x{h} {params}")
        a = stereo_inputs(O, left, right, params); b = stereo_inputs(R, left, right, params)
        so = O.stereo_match(*a, BF, B); sr = R.stereo_match(*b, BF, B)
        assert so["ret"] == sr["ret"] and bits(so["uRight"]) == bits(sr["uRight"]) and bits(so["depth"]) == bits(sr["depth"])


@pytest.mark.parametrize("seed", range(56))
def test_synthetic_frames(R, seed):
    w, h = SIZES[seed % len(SIZES)]
    params = fit_levels(PARAM_SETS[(seed // len(SIZES)) % len(PARAM_SETS)], w, h)
    img = synth.synth_stereo(seed, w, h, nshapes=max(40, 400 * w * h // (752 * 480)))[seed & 1]
    assert_same_extraction(O, R, img, params, tag=f"seed {seed} {w}x{h} {params}")


def test_degenerate_images(R):
    rng = np.random.default_rng(5)
    flat = np.full((480, 752), 127, np.uint8)
    noise = rng.integers(0, 256, (240, 376), dtype=np.uint8)                 # every cell saturated with corners
    faint = (120 + rng.integers(0, 9, (480, 752))).astype(np.uint8)          # only the minThFAST retry finds anything
    checker = (((np.add.outer(np.arange(480) // 8, np.arange(752) // 8)) & 1) * 255).astype(np.uint8)  # equal scores
    for tag, img in (("flat", flat), ("noise", noise), ("faint", faint), ("checker", checker)):
        assert_same_extraction(O, R, img, tag=tag)
    ko, do, _ = R.Extractor().extract(flat)
    assert len(ko) == 0 and do.shape == (0, 32)


def test_row_step_and_reuse(R, fixtures):
    """A non-contiguous source (cv::Mat step > cols) and one extractor reused across sizes (src/ORBextractor.cc:1176-1201
    re-allocates every level every call)."""
    big = np.zeros((480, 800), np.uint8)
    big[:, :752] = fixtures["EuRoC_l"]
    view = big[:, :752]
    eo, er = O.Extractor(), R.Extractor()
    for img in (fixtures["TUM_l"], view, fixtures["EuRoC_r"][:300, :400].copy()):
        a = eo.extract(np.ascontiguousarray(img))
        b = er.extract(np.ascontiguousarray(img))
        assert bits(a[0]) == bits(b[0]) and np.array_equal(a[1], b[1])


# ------------------------------------------------------------------------------------------------ quadtree alone
def octree_case(seed):
    rng = np.random.default_rng(4000 + seed)
    kind = seed % 6
    W, H = (720, 448) if seed % 3 else (int(rng.integers(40, 1300)), int(rng.integers(40, 700)))
    H = min(H, int(W / 0.6))        # nIni = round(W / H) must be >= 1: the reference indexes vpIniNodes[0] (:547-570)
    n = int(rng.integers(1, 6000))
    if kind == 0:                                  # uniform
        x, y = rng.integers(0, W, n), rng.integers(0, H, n)
    elif kind == 1:                                # clustered: deep splits in a few places
        c = rng.integers(0, [W, H], (8, 2))
        p = c[rng.integers(0, 8, n)] + rng.normal(0, 9, (n, 2))
        x, y = np.clip(p[:, 0], 0, W - 1).astype(int), np.clip(p[:, 1], 0, H - 1).astype(int)
    elif kind == 2:                                # a lattice: very many nodes of exactly equal size -> tie order decides
        step = int(rng.integers(3, 12))
        gx, gy = np.meshgrid(np.arange(1, W - 1, step), np.arange(1, H - 1, step))
        x, y = gx.ravel(), gy.ravel()
        n = len(x)
    elif kind == 3:                                # duplicates of few positions (nodes that can never be separated)
        k = int(rng.integers(1, 40))
        px, py = rng.integers(0, W, k), rng.integers(0, H, k)
        i = rng.integers(0, k, n)
        x, y = px[i], py[i]
    elif kind == 4:                                # on the split lines and the box edges
        x = rng.choice([0, W // 2 - 1, W // 2, W // 2 + 1, W // 4, W - 2, W - 1], n)   # x == W would index vpIniNodes[nIni]
        y = rng.integers(0, H, n)
    else:                                          # few points
        n = int(rng.integers(1, 12))
        x, y = rng.integers(0, W, n), rng.integers(0, H, n)
    order = rng.permutation(n)
    resp = rng.integers(7, 40 if kind in (2, 3) else 256, n)      # equal responses too: "first one wins" (:744-760)
    xyr = np.stack([x[order], y[order], resp], 1).astype(np.int32)
    N = int(rng.choice([1, 5, 60, 217, 434, 1000, 3000]))
    return xyr, 16, 16 + W, 16, 16 + H, N


@pytest.mark.parametrize("seed", range(240))
def test_distribute_octree_random(R, seed):
    xyr, minX, maxX, minY, maxY, N = octree_case(seed)
    a = O.distribute_octree(xyr, minX, maxX, minY, maxY, N)
    b = R.distribute_octree(xyr, minX, maxX, minY, maxY, N)
    assert np.array_equal(a, b), f"seed {seed}: kept candidates / their order differ ({len(a)} vs {len(b)})"


def test_reference_is_allocator_dependent(fixtures):
    """Documents why the monotonic allocator is needed: with plain malloc the reference's result depends on the heap's
    history.  Nothing is asserted about HOW it differs (that is the allocator's business) -- only that everything
    upstream of the quadtree is unaffected and that every kept key is one of the FAST candidates."""
    M = pyref.load("ref")
    M.monotonic_new(0)
    try:
        img = fixtures["TUM_l"]
        ex = M.Extractor()
        eo = O.Extractor()
        eo.extract(img)
        outs = set()
        for _ in range(6):
            k, d, lc = ex.extract(img)
            outs.add(bits(k))
            for l in range(8):
                assert np.array_equal(ex.level(l, True), eo.level(l, True))
                cand = ex.last_candidates(l)
                assert np.array_equal(cand, eo.last_candidates(l))
                kl = k[k["octave"] == l]
                s = eo.scale[l]
                have = {(int(c[0]) + 16, int(c[1]) + 16) for c in cand}
                got = {(int(round(float(x) / float(s))), int(round(float(y) / float(s)))) for x, y in zip(kl["x"], kl["y"])}
                assert got <= have
        print(f"\n[reference under glibc malloc] {len(outs)} distinct keypoint sets in 6 calls on one image")
    finally:
        M.monotonic_new(1)


# ------------------------------------------------------------------------------------------------ stereo
def stereo_inputs(mod, left, right, params=(1000, 1.2, 8, 20, 7)):
    exL, exR = mod.Extractor(*params), mod.Extractor(*params)
    kL, dL, _ = exL.extract(left)
    kR, dR, _ = exR.extract(right)
    return exL, exR, kL, dL, kR, dR


def assert_same_stereo(R, left, right, bf=BF, b=B, mode=0, min_x=0.0, tag=""):
    a = stereo_inputs(O, left, right)
    c = stereo_inputs(R, left, right)
    so = O.stereo_match(*a, bf, b, mode=mode, min_x=min_x)
    sr = R.stereo_match(*c, bf, b, mode=mode, min_x=min_x)
    assert sr["ret"] >= 0, f"{tag}: reference wrapper refused the input ({sr['ret']})"
    assert so["ret"] == sr["ret"], f"{tag}: match count {so['ret']} vs {sr['ret']}"
    assert bits(so["uRight"]) == bits(sr["uRight"]), f"{tag}: mvuRight"
    assert bits(so["depth"]) == bits(sr["depth"]), f"{tag}: mvDepth"
    return so


@pytest.mark.parametrize("pair", ["EuRoC", "TUM"])
def test_stereo_classic_fixtures(R, fixtures, pair):
    so = assert_same_stereo(R, fixtures[pair + "_l"], fixtures[pair + "_r"], tag=pair)
    assert (so["depth"] > 0).sum() > 0


@pytest.mark.parametrize("seed", range(12))
def test_stereo_classic_synthetic(R, seed):
    w, h = [(752, 480), (640, 480), (512, 512)][seed % 3]
    left, right = synth.synth_stereo(100 + seed, w, h, dmax=[60.0, 30.0, 90.0][seed % 3])
    bf = [BF, 35.0, 80.0][seed % 3]
    so = assert_same_stereo(R, left, right, bf=bf, b=bf / 435.2, tag=f"seed {seed}")
    assert (so["depth"] > 0).sum() > 50            # SAD + parabola really ran


def test_stereo_classic_identical_and_unrelated(R, fixtures):
    img = fixtures["EuRoC_l"]
    assert_same_stereo(R, img, img, tag="identical eyes (disparity 0 -> the 0.01 clamp)")
    assert_same_stereo(R, fixtures["EuRoC_l"], fixtures["TUM_l"][:480, :].repeat(2, 1)[:, :752].copy(), tag="unrelated eyes")


@pytest.mark.parametrize("seed", range(8))
def test_stereo_undistorted(R, seed):
    left, right = synth.synth_stereo(200 + seed)
    assert_same_stereo(R, left, right, mode=1, min_x=[0.0, 25.5, -10.0, 300.0][seed % 4], tag=f"seed {seed}")


@pytest.mark.parametrize("pair", ["EuRoC", "TUM"])
def test_stereo_undistorted_fixtures(R, fixtures, pair):
    assert_same_stereo(R, fixtures[pair + "_l"], fixtures[pair + "_r"], mode=1, tag=pair)


def test_stereo_undistorted_shifted_keys(R):
    """Undistortion moves keys off the pixel grid and outside the image: the row clamp of PrepareStereoCandidates and
    the vL range test (src/Frame.cc:1211, include/Frame.h:253-254) only act then."""
    left, right = synth.synth_stereo(321)
    a = stereo_inputs(O, left, right)
    c = stereo_inputs(R, left, right)
    rng = np.random.default_rng(1)
    kL, kR = a[2].copy(), a[4].copy()
    for k in (kL, kR):
        k["x"] += rng.normal(0, 3, len(k)).astype(F32)
        k["y"] = (k["y"] * F32(1.12) - F32(30) + rng.normal(0, 0.7, len(k)).astype(F32)).astype(F32)
    so = O.stereo_match(a[0], a[1], kL, a[3], kR, a[5], BF, B, mode=1, min_x=-3.0)
    sr = R.stereo_match(c[0], c[1], kL, c[3], kR, c[5], BF, B, mode=1, min_x=-3.0)
    assert so["ret"] == sr["ret"] and bits(so["uRight"]) == bits(sr["uRight"]) and bits(so["depth"]) == bits(sr["depth"])
    assert ((kL["y"] < 0) | (kL["y"] > 479)).any()


@pytest.mark.parametrize("seed", range(4))
def test_stereo_undistorted_map_point_windows(R, seed):
    """Keys that carry a map point search a +-DISPARITY_THRES window around its predicted disparity
    (src/Frame.cc:1220-1231, run here through the reference's own WorldToCameraPoint with an identity pose).  The
    restatement takes the window as two arrays; they are formed below with the reference's float expressions."""
    left, right = synth.synth_stereo(400 + seed)
    a = stereo_inputs(O, left, right)
    c = stereo_inputs(R, left, right)
    nL = len(a[2])
    rng = np.random.default_rng(seed)
    z = np.full(nL, np.nan, F32)
    pick = rng.random(nL) < 0.6
    z[pick] = (rng.random(int(pick.sum())) * 12 + 0.3).astype(F32)
    z[rng.random(nL) < 0.05] = F32(-1.0)                        # behind the camera: WorldToCameraPoint says no
    bf, b = F32(BF), F32(B)
    min_d = np.zeros(nL, F32)
    max_d = np.full(nL, bf / b, F32)
    ok = ~np.isnan(z) & (z > 0)
    disp = (bf / z[ok]).astype(F32)
    min_d[ok] = np.maximum(disp - F32(50.0), F32(0))
    max_d[ok] = np.minimum(disp + F32(50.0), bf / b)
    so = O.stereo_match(*a, BF, B, mode=1, min_d=min_d, max_d=max_d)
    L = R.lib()
    u = np.empty(nL, F32); d = np.empty(nL, F32)
    prm = R.StereoParams(BF, B, 1, 100, 50, 0.0)
    import ctypes as C
    ret = L.ref_stereo_match_undistorted_mp(c[0].h, c[1].h, C.byref(prm), R._p(c[2]), nL, R._p(c[3]), R._p(c[4]), len(c[4]),
                                            R._p(c[5]), R._p(z), R._p(u), R._p(d))
    assert ret == so["ret"] and bits(u) == bits(so["uRight"]) and bits(d) == bits(so["depth"])


def test_hamming(R):
    rng = np.random.default_rng(0)
    for _ in range(2000):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert O.hamming256(a, b) == R.hamming256(a, b) == int(np.unpackbits(a ^ b).sum())


# ------------------------------------------------------------------------------------------------ grid + searches
def random_keys(rng, n, spread=30.0, p0=0.5):
    k = np.zeros(n, O.KP_DTYPE)
    k["x"] = (rng.random(n) * (752 + 2 * spread) - spread).astype(F32)
    k["y"] = (rng.random(n) * (480 + 2 * spread) - spread).astype(F32)
    k["angle"] = (rng.random(n) * 360).astype(F32)
    k["octave"] = np.where(rng.random(n) < p0, 0, rng.integers(1, 8, n))
    k["size"], k["class_id"] = 31, -1
    return k


@pytest.mark.parametrize("seed", range(10))
def test_grid_and_area(R, seed):
    rng = np.random.default_rng(seed)
    bounds = BOUNDS if seed % 2 == 0 else (-12.5, 770.25, -9.0, 497.5)       # undistorted image bounds
    keys = random_keys(rng, int(rng.integers(0, 3000)))
    so, io = O.assign_features_to_grid(keys, bounds)
    sr, ir = R.assign_features_to_grid(keys, bounds)
    assert np.array_equal(so, sr) and np.array_equal(io, ir)
    for _ in range(200):
        x, y = rng.random() * 900 - 70, rng.random() * 600 - 60
        r = float(rng.choice([1.0, 7.5, 15.0, 40.0, 100.0, 1000.0]))
        lo, hi = [(-1, -1), (0, 0), (2, 3), (1, -1), (0, 7), (3, 2)][int(rng.integers(0, 6))]
        assert np.array_equal(O.features_in_area(keys, bounds, x, y, r, lo, hi), R.features_in_area(keys, bounds, x, y, r, lo, hi))


def init_case(rng, n1, n2, flip=0.06):
    k1, k2 = random_keys(rng, n1, spread=0.0, p0=0.7), random_keys(rng, n2, spread=0.0, p0=0.7)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    m = min(n1, n2)
    if m:
        k2["x"][:m] = k1["x"][:m] + rng.normal(0, 12, m).astype(F32)
        k2["y"][:m] = k1["y"][:m] + rng.normal(0, 12, m).astype(F32)
        k2["octave"][:m] = k1["octave"][:m]
        k2["angle"][:m] = np.mod(k1["angle"][:m] + rng.normal(15, 5, m), 360).astype(F32)
        wild = rng.random(m) < 0.1
        k2["angle"][:m][wild] = (rng.random(int(wild.sum())) * 360).astype(F32)
        d2[:m] = d1[:m] ^ np.packbits(rng.random((m, 256)) < flip, axis=1)
        dup = rng.integers(0, m, m // 5)                       # several queries after one candidate (vMatchedDistance rule)
        d1[dup] = d1[rng.integers(0, m, len(dup))]
        p = rng.permutation(n2)
        k2, d2 = k2[p], d2[p]
    return k1, d1, k2, d2


@pytest.mark.parametrize("seed", range(24))
def test_search_for_initialization(R, seed):
    rng = np.random.default_rng(700 + seed)
    n1, n2 = [(1000, 1000), (2000, 1800), (300, 900), (0, 50), (40, 0), (1, 1)][seed % 6]
    k1, d1, k2, d2 = init_case(rng, n1, n2)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(F32)
    win = [100, 30, 10][seed % 3]
    ratio = [0.9, 0.6, 0.75][(seed // 3) % 3]
    ori = seed % 4 != 3
    no, mo, po = O.search_for_initialization(k1, d1, k2, d2, BOUNDS, prev.copy(), win, ratio, ori)
    nr, mr, pr = R.search_for_initialization(k1, d1, k2, d2, BOUNDS, prev.copy(), win, ratio, ori)
    assert no == nr and np.array_equal(mo, mr) and bits(po) == bits(pr)
    if n1 >= 300 and n2 >= 300:
        assert no > 20


@pytest.mark.parametrize("seed", range(24))
def test_search_by_projection_map_points(R, seed):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>, th): the restatement takes already-derived windows
    (radius, level range, flags); the reference derives them from the map point's tracking fields.  The derivation is
    done below with the reference's float expressions (r = RadiusByViewingCos(viewCos) [* th]; r * scale[level])."""
    import ctypes as C
    n, nq = [(1000, 900), (2000, 2500), (150, 60), (0, 20), (30, 0)][seed % 5]
    th = [1.0, 3.0, 5.0][seed % 3]
    ratio = [0.8, 0.9][seed % 2]
    keys, desc, u_right, occupied, q, qdesc = synth.projection_case(seed, n, nq, form="map", th=th)
    rng = np.random.default_rng(seed)
    scale = (F32(1.2) ** np.arange(8)).astype(F32)
    lvl = np.asarray(q["max_level"], np.int32)
    view_cos = np.where(rng.random(nq) < 0.5, F32(0.9995), F32(0.93)).astype(F32)
    r = np.where(view_cos > 0.998, F32(2.5), F32(4.0)).astype(F32)
    if th != 1.0:
        r = (r * F32(th)).astype(F32)
    q = q.copy()
    q["r"] = (r * scale[lvl]).astype(F32) if nq else q["r"]
    in_view = ((q["flags"] & 1) != 0).astype(np.uint8)
    bad = np.zeros(nq, np.uint8)
    flip_bad = rng.random(nq) < 0.03
    bad[flip_bad] = 1
    q["flags"] = np.where(flip_bad, q["flags"] & ~1, q["flags"])
    observations = np.where((q["flags"] & 2) != 0, rng.integers(1, 9, nq), 0).astype(np.int32)
    key_obs = np.where(occupied != 0, 1, -1).astype(np.int32)
    nm_o, kq_o, kd_o, qk_o = O.search_by_projection(keys, desc, u_right, occupied, BOUNDS, q, qdesc, th_dist=100,
                                                    nnratio=ratio, use_ratio=True, check_ori=False)
    kp = np.full(max(n, 1), -9, np.int32); kd = np.zeros(max(n, 1), np.int32)
    b = np.asarray(BOUNDS, F32)
    args = [np.ascontiguousarray(x) for x in (keys, desc, u_right, key_obs, b, scale, in_view, bad, lvl, view_cos,
                                              q["x"].astype(F32), q["y"].astype(F32), q["xr"].astype(F32), observations, qdesc)]
    p = [R._p(x) for x in args]
    nm_r = R.lib().ref_search_by_projection_mp(p[0], p[1], p[2], p[3], n, p[4], p[5], 8, p[6], p[7], p[8], p[9], p[10],
                                               p[11], p[12], p[13], p[14], nq, th, ratio, R._p(kp), R._p(kd))
    assert nm_o == nm_r
    assert np.array_equal(kq_o[:n], kp[:n]) and np.array_equal(kd_o[:n], kd[:n])
    if n >= 1000:
        assert nm_o > 100
