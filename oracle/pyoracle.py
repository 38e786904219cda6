"""ORACLE -- TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_build/libgforb_oracle.so (the CPU restatement of the reference's ORB
front-end).  Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; the product package never imports it.
"""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


class StereoParams(C.Structure):
    _fields_ = [("bf", C.c_float), ("b", C.c_float), ("mode", C.c_int),
                ("th_high", C.c_int), ("th_low", C.c_int), ("min_x", C.c_float)]


_VARIANT = "oracle"     # "oracle" = the restatement; "ref" / "ref_native" = the reference's own code (oracle/pyref.py)
REFERENCE_ROOT = Path("/root/reference")


def build(force=False):
    if _VARIANT == "oracle":
        so = _HERE / "_build" / "libgforb_oracle.so"
        if force or not so.exists():
            subprocess.check_call(["make", "-C", str(_HERE)] + (["-B"] if force else []),
                                  stdout=subprocess.DEVNULL)
        return so
    # oracle/_ref: compiled from the reference sources where they lie; only the prebuilt files exist on the GPU box
    so = _HERE / "_ref" / f"libgforb_{_VARIANT}.so"
    if (force or not so.exists()) and REFERENCE_ROOT.exists():
        subprocess.check_call(["make", "-C", str(_HERE), "ref"] + (["-B"] if force else []),
                              stdout=subprocess.DEVNULL)
    if not so.exists():
        raise FileNotFoundError(f"{so} is missing and {REFERENCE_ROOT} is not here to build it from")
    return so


class _Partial:
    """ctypes library view that tolerates entry points a variant does not export (the reference build has no
    counterpart for the restatement-only helpers); touching a missing one raises AttributeError on use."""

    def __init__(self, cdll):
        object.__setattr__(self, "_cdll", cdll)

    def __getattr__(self, name):
        try:
            return getattr(self._cdll, name)
        except AttributeError:
            if _VARIANT == "oracle":
                raise
            return _Missing(name)


class _Missing:
    def __init__(self, name):
        self.name = name
        self.argtypes = self.restype = None

    def __call__(self, *a):
        raise AttributeError(f"{self.name} is not exported by the '{_VARIANT}' library")


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    so = build()
    try:
        L = C.CDLL(str(so))
    except OSError:
        so = build(force=True)
        L = C.CDLL(str(so))
    L = _Partial(L)
    u8p, ip, fp = C.c_void_p, C.c_void_p, C.c_void_p
    sz = C.c_size_t
    L.orc_cv_round_f.argtypes = [C.c_float]; L.orc_cv_round_f.restype = C.c_int
    L.orc_cv_round_d.argtypes = [C.c_double]; L.orc_cv_round_d.restype = C.c_int
    L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]; L.orc_fast_atan2.restype = C.c_float
    L.orc_border_reflect101.argtypes = [u8p, C.c_int, C.c_int, sz, u8p, sz, C.c_int]
    L.orc_border_reflect101.restype = None
    L.orc_resize_linear_8u.argtypes = [u8p, C.c_int, C.c_int, sz, u8p, C.c_int, C.c_int, sz]
    L.orc_resize_linear_8u.restype = None
    L.orc_gaussian_blur_7x7_s2.argtypes = [u8p, C.c_int, C.c_int, sz, u8p, sz]
    L.orc_gaussian_blur_7x7_s2.restype = None
    L.orc_fast9_16.argtypes = [u8p, C.c_int, C.c_int, sz, C.c_int, C.c_int, ip, C.c_int]
    L.orc_fast9_16.restype = C.c_int
    L.orc_hamming256.argtypes = [u8p, u8p]; L.orc_hamming256.restype = C.c_int
    L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.orc_extractor_create.restype = C.c_void_p
    L.orc_extractor_destroy.argtypes = [C.c_void_p]; L.orc_extractor_destroy.restype = None
    L.orc_extractor_tables.argtypes = [C.c_void_p, fp, fp, fp, fp, ip, ip]
    L.orc_extractor_tables.restype = None
    L.orc_extractor_pattern.argtypes = [C.c_void_p, C.c_void_p]; L.orc_extractor_pattern.restype = None
    L.orc_compute_pyramid.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, sz]
    L.orc_compute_pyramid.restype = C.c_int
    L.orc_pyramid_level.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                    C.POINTER(C.c_int), C.POINTER(sz)]
    L.orc_pyramid_level.restype = C.c_int
    L.orc_detect_level.argtypes = [C.c_void_p, C.c_int, ip, C.c_int]; L.orc_detect_level.restype = C.c_int
    L.orc_distribute_octree.argtypes = [ip, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, ip, C.c_int]
    L.orc_distribute_octree.restype = C.c_int
    L.orc_extract.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, sz, C.c_void_p, u8p, C.c_int, ip]
    L.orc_extract.restype = C.c_int
    L.orc_last_candidates.argtypes = [C.c_void_p, C.c_int, ip, C.c_int]
    L.orc_last_candidates.restype = C.c_int
    L.orc_stereo_match.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(StereoParams),
                                   C.c_void_p, C.c_int, u8p, C.c_void_p, C.c_int, u8p,
                                   fp, fp, fp, fp, ip, ip, ip]
    L.orc_stereo_match.restype = C.c_int
    L.orc_bench_stereo.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_int, C.c_float, C.c_int, C.c_int, C.c_int,
                                   C.c_float, C.c_float, C.c_int, C.POINTER(C.c_longlong)]
    L.orc_bench_stereo.restype = C.c_double
    L.orc_match_bruteforce.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_float, C.c_int, ip, ip, ip, ip, C.c_int]
    L.orc_match_bruteforce.restype = None
    L.orc_assign_features_to_grid.argtypes = [C.c_void_p, C.c_int, fp, ip, ip]
    L.orc_assign_features_to_grid.restype = C.c_int
    L.orc_features_in_area.argtypes = [C.c_void_p, C.c_int, fp, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
                                       ip, C.c_int]
    L.orc_features_in_area.restype = C.c_int
    L.orc_search_for_initialization.argtypes = [C.c_void_p, u8p, C.c_int, C.c_void_p, u8p, C.c_int, fp, fp, C.c_int,
                                                C.c_float, C.c_int, ip]
    L.orc_search_for_initialization.restype = C.c_int
    L.orc_search_by_projection.argtypes = [C.c_void_p, u8p, fp, u8p, C.c_int, fp, C.c_void_p, u8p, C.c_int, C.c_int,
                                           C.c_float, C.c_int, C.c_int, fp, ip, ip, ip]
    L.orc_search_by_projection.restype = C.c_int
    if hasattr(L, "orc_search_by_sim3"):
        L.orc_search_by_sim3.argtypes = [C.c_void_p, u8p, C.c_int, C.c_void_p, u8p, C.c_int, fp, C.c_void_p, u8p, C.c_void_p, u8p,
                                         C.c_int, ip, ip, ip]
        L.orc_search_by_sim3.restype = C.c_int
    L.orc_search_by_bow.argtypes = [C.c_void_p, u8p, u8p, C.c_int, ip, ip, ip, C.c_int, C.c_void_p, u8p, u8p, C.c_int, ip, ip, ip,
                                    C.c_int, C.c_int, C.c_float, C.c_int, ip]
    L.orc_search_by_bow.restype = C.c_int
    L.orc_search_for_triangulation.argtypes = [C.c_void_p, u8p, u8p, u8p, C.c_int, ip, ip, ip, C.c_int, C.c_void_p, u8p, u8p, u8p,
                                               C.c_int, ip, ip, ip, C.c_int, fp, C.c_float, C.c_float, fp, fp, C.c_int, C.c_int, ip]
    L.orc_search_for_triangulation.restype = C.c_int
    L.orc_preprocess.argtypes = [C.c_void_p, C.c_int, C.c_int, sz, C.c_int, fp, fp, C.c_int, C.c_int, u8p, sz]
    L.orc_preprocess.restype = C.c_int
    L.orc_undistort_matrix.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]; L.orc_undistort_matrix.restype = None
    L.orc_undistort_points.argtypes = [fp, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, fp]
    L.orc_undistort_points.restype = None
    L.orc_distinctive_descriptors.argtypes = [u8p, ip, C.c_int, ip, ip]; L.orc_distinctive_descriptors.restype = None
    if _VARIANT != "oracle":
        L.ref_stereo_match_undistorted_mp.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(StereoParams), C.c_void_p, C.c_int,
                                                      u8p, C.c_void_p, C.c_int, u8p, fp, fp, fp]
        L.ref_stereo_match_undistorted_mp.restype = C.c_int
        L.ref_search_by_projection_mp.argtypes = [C.c_void_p, u8p, fp, ip, C.c_int, fp, fp, C.c_int, u8p, u8p, ip, fp, fp,
                                                  fp, fp, ip, u8p, C.c_int, C.c_float, C.c_float, ip, ip]
        L.ref_search_by_projection_mp.restype = C.c_int
    _LIB = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.ndim == 2
    return a


# ------------------------------------------------------------------ primitives
def cv_round(v):
    return lib().orc_cv_round_d(float(v))


def cv_round_f(v):
    return lib().orc_cv_round_f(float(np.float32(v)))


def fast_atan2(y, x):
    return np.float32(lib().orc_fast_atan2(float(np.float32(y)), float(np.float32(x))))


def border_reflect101(img, border):
    img = _img(img)
    h, w = img.shape
    out = np.empty((h + 2 * border, w + 2 * border), np.uint8)
    lib().orc_border_reflect101(_p(img), w, h, img.strides[0], _p(out), out.strides[0], border)
    return out


def resize_linear(img, dw, dh):
    img = _img(img)
    h, w = img.shape
    out = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_8u(_p(img), w, h, img.strides[0], _p(out), dw, dh, out.strides[0])
    return out


def gaussian_blur(img):
    img = _img(img)
    h, w = img.shape
    out = np.empty((h, w), np.uint8)
    lib().orc_gaussian_blur_7x7_s2(_p(img), w, h, img.strides[0], _p(out), out.strides[0])
    return out


def fast9(img, threshold, nonmax=True):
    img = _img(img)
    h, w = img.shape
    cap = max(16, w * h)
    buf = np.empty((cap, 3), np.int32)
    n = lib().orc_fast9_16(_p(img), w, h, img.strides[0], threshold, int(nonmax), _p(buf), cap)
    return buf[:n].copy()


def hamming256(a, b):
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return lib().orc_hamming256(_p(a), _p(b))


def distribute_octree(xyr, minX, maxX, minY, maxY, N):
    xyr = np.ascontiguousarray(xyr, np.int32).reshape(-1, 3)
    keep = np.empty(max(len(xyr), 1), np.int32)
    n = lib().orc_distribute_octree(_p(xyr), len(xyr), minX, maxX, minY, maxY, N, _p(keep), len(keep))
    return keep[:n].copy()


# ------------------------------------------------------------------ extractor
class Extractor:
    """Mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:76-168 in the reference)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = lib().orc_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        if not self.h:
            raise ValueError("bad extractor parameters")
        n = nlevels
        self.scale = np.empty(n, np.float32); self.inv_scale = np.empty(n, np.float32)
        self.sigma2 = np.empty(n, np.float32); self.inv_sigma2 = np.empty(n, np.float32)
        self.quota = np.empty(n, np.int32); self.umax = np.empty(16, np.int32)
        lib().orc_extractor_tables(self.h, _p(self.scale), _p(self.inv_scale), _p(self.sigma2),
                                   _p(self.inv_sigma2), _p(self.quota), _p(self.umax))
        self.pattern = np.empty((512, 2), np.int8)
        lib().orc_extractor_pattern(self.h, _p(self.pattern))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_extractor_destroy(self.h)
            self.h = None

    def compute_pyramid(self, img):
        img = _img(img)
        h, w = img.shape
        rc = lib().orc_compute_pyramid(self.h, _p(img), w, h, img.strides[0])
        assert rc == 0
        return [self.level(l) for l in range(self.nlevels)]

    def level(self, l, padded=False):
        ptr = C.c_void_p(); w = C.c_int(); h = C.c_int(); pitch = C.c_size_t()
        rc = lib().orc_pyramid_level(self.h, l, C.byref(ptr), C.byref(w), C.byref(h), C.byref(pitch))
        assert rc == 0
        W, H, P = w.value, h.value, pitch.value
        base = ptr.value - 19 * P - 19
        buf = np.ctypeslib.as_array((C.c_uint8 * (P * (H + 38))).from_address(base)).reshape(H + 38, P)
        return buf.copy() if padded else buf[19:19 + H, 19:19 + W].copy()

    def detect_level(self, l, cap=200000):
        buf = np.empty((cap, 3), np.int32)
        n = lib().orc_detect_level(self.h, l, _p(buf), cap)
        assert n <= cap
        return buf[:n].copy()

    def last_candidates(self, l, cap=200000):
        buf = np.empty((cap, 3), np.int32)
        n = lib().orc_last_candidates(self.h, l, _p(buf), cap)
        assert 0 <= n <= cap
        return buf[:n].copy()

    def extract(self, img, cap=None):
        """operator()(image, mask, keypoints, descriptors): returns (kps[KP_DTYPE], desc[n,32], level_counts)."""
        img = _img(img)
        h, w = img.shape
        cap = cap or (self.nfeatures * 2 + 64)
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        lc = np.zeros(self.nlevels, np.int32)
        n = lib().orc_extract(self.h, _p(img), w, h, img.strides[0], _p(kps), _p(desc), cap, _p(lc))
        assert n <= cap, (n, cap)
        return kps[:n].copy(), desc[:n].copy(), lc


def stereo_match(exL, exR, kL, dL, kR, dR, bf, b, mode=0, th_high=100, th_low=50, min_x=0.0,
                 min_d=None, max_d=None):
    """Frame::ComputeStereoMatches (mode 0) / _Undistorted (mode 1). Returns dict."""
    kL = np.ascontiguousarray(kL); kR = np.ascontiguousarray(kR)
    dL = np.ascontiguousarray(dL, np.uint8); dR = np.ascontiguousarray(dR, np.uint8)
    nL, nR = len(kL), len(kR)
    u = np.empty(max(nL, 1), np.float32); z = np.empty(max(nL, 1), np.float32)
    bd = np.empty(max(nL, 1), np.int32); bi = np.empty(max(nL, 1), np.int32); sad = np.empty(max(nL, 1), np.int32)
    prm = StereoParams(bf, b, mode, th_high, th_low, min_x)
    if min_d is not None:
        min_d = np.ascontiguousarray(min_d, np.float32); max_d = np.ascontiguousarray(max_d, np.float32)
    r = lib().orc_stereo_match(exL.h, exR.h, C.byref(prm), _p(kL), nL, _p(dL), _p(kR), nR, _p(dR),
                               _p(min_d), _p(max_d), _p(u), _p(z), _p(bd), _p(bi), _p(sad))
    return dict(ret=r, uRight=u[:nL], depth=z[:nL], best_dist=bd[:nL], best_idx=bi[:nL], sad=sad[:nL])


def bench_stereo(images, nframes, nthreads, shape=0, nfeatures=1000, scale_factor=1.2, nlevels=8,
                 ini_th=20, min_th=7, bf=47.90639384423901, b=47.90639384423901 / 435.2046959714599,
                 do_match=True):
    images = np.ascontiguousarray(images, np.uint8)
    n, h, w = images.shape
    chk = C.c_longlong(0)
    secs = lib().orc_bench_stereo(_p(images), n, w, h, nframes, nthreads, shape, nfeatures, scale_factor,
                                  nlevels, ini_th, min_th, bf, b, int(do_match), C.byref(chk))
    return secs, chk.value


def match_bruteforce(desc_a, desc_b, ratio=0.6, threshold=50, nthreads=1):
    """The reference's best / second-best loop + acceptance tests (src/ORBmatcher.cc:562-576) over all of desc_b."""
    a = np.ascontiguousarray(desc_a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(desc_b, np.uint8).reshape(-1, 32)
    n = max(len(a), 1)
    bi = np.empty(n, np.int32); bd = np.empty(n, np.int32); sd = np.empty(n, np.int32); m = np.empty(n, np.int32)
    lib().orc_match_bruteforce(_p(a), len(a), _p(b), len(b), ratio, threshold, _p(bi), _p(bd), _p(sd), _p(m), nthreads)
    k = len(a)
    return dict(best_idx=bi[:k], best_dist=bd[:k], second_dist=sd[:k], match=m[:k])


GRID_COLS, GRID_ROWS = 64, 48


def _bounds(bounds):
    return np.ascontiguousarray(bounds, dtype=np.float32)


def assign_features_to_grid(keys, bounds):
    """Frame::AssignFeaturesToGrid (src/Frame.cc:461-476) as CSR over mGrid[ix][iy] (ix major)."""
    keys = np.ascontiguousarray(keys); b = _bounds(bounds)
    start = np.zeros(GRID_COLS * GRID_ROWS + 1, np.int32); idx = np.zeros(max(len(keys), 1), np.int32)
    n = lib().orc_assign_features_to_grid(_p(keys), len(keys), _p(b), _p(start), _p(idx))
    return start, idx[:n]


def features_in_area(keys, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (src/Frame.cc:593-643)."""
    keys = np.ascontiguousarray(keys); b = _bounds(bounds)
    out = np.zeros(max(len(keys), 1), np.int32)
    n = lib().orc_features_in_area(_p(keys), len(keys), _p(b), x, y, r, min_level, max_level, _p(out), len(out))
    return out[:n]


def search_for_initialization(k1, d1, k2, d2, bounds, prev_xy, window_size=100, nnratio=0.9, check_ori=True):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:520-633).  Returns (nmatches, matches12, prev_xy')."""
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    d1 = np.ascontiguousarray(d1, dtype=np.uint8); d2 = np.ascontiguousarray(d2, dtype=np.uint8)
    prev = np.array(prev_xy, dtype=np.float32, order="C", copy=True).reshape(len(k1), 2)
    m12 = np.full(max(len(k1), 1), -1, np.int32)
    n = lib().orc_search_for_initialization(_p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(_bounds(bounds)),
                                            _p(prev), int(window_size), nnratio, int(bool(check_ori)), _p(m12))
    return n, m12[:len(k1)], prev


PIX_GRAY8, PIX_GRAY16, PIX_RGB8, PIX_BGR8, PIX_RGBA8, PIX_BGRA8 = range(6)


def preprocess(src, fmt, mapx=None, mapy=None):
    """[16-bit -> 8-bit] -> cv::remap(INTER_LINEAR, constant 0 border) -> cvtColor(*2GRAY); see oracle/preprocess.cpp."""
    src = np.ascontiguousarray(src)
    sh, sw = src.shape[:2]
    if mapx is not None:
        mapx = np.ascontiguousarray(mapx, np.float32); mapy = np.ascontiguousarray(mapy, np.float32)
        dh, dw = mapx.shape
    else:
        dh, dw = sh, sw
    dst = np.empty((dh, dw), np.uint8)
    rc = lib().orc_preprocess(_p(src), sw, sh, src.strides[0], fmt, _p(mapx), _p(mapy), dw, dh, _p(dst), dw)
    assert rc == 0
    return dst


def undistort_points(xy, K, dist, R=None, P=None):
    """cv::undistortPoints(xy, xy, K, dist, R, P) on n x 2 float32 points (src/Frame.cc:688,710,745); P = None -> K
    is NOT implied: pass P=K for Frame::UndistortKeyPoints (identity new camera matrix otherwise, as OpenCV does)."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    K9 = np.ascontiguousarray(np.asarray(K, np.float64).reshape(3, 3))
    d8 = np.zeros(8, np.float64); dd = np.asarray(dist, np.float64).ravel(); d8[:len(dd)] = dd
    R9 = np.eye(3) if R is None else np.ascontiguousarray(np.asarray(R, np.float64).reshape(3, 3))
    P9 = np.eye(3) if P is None else np.ascontiguousarray(np.asarray(P, np.float64)[:3, :3])
    RR = np.zeros(9, np.float64)
    lib().orc_undistort_matrix(_p(np.ascontiguousarray(P9)), _p(np.ascontiguousarray(R9)), _p(RR))
    out = np.empty_like(xy)
    lib().orc_undistort_points(_p(xy), len(xy), _p(K9), _p(d8), _p(RR), _p(out))
    return out


def distinctive_descriptors(desc, set_start):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:397-458) -> (best_idx, best_median) per set."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); ss = np.ascontiguousarray(set_start, np.int32)
    n = len(ss) - 1
    bi = np.empty(max(n, 1), np.int32); bm = np.empty(max(n, 1), np.int32)
    lib().orc_distinctive_descriptors(_p(desc), _p(ss), n, _p(bi), _p(bm))
    return bi[:n], bm[:n]


PROJ_QUERY = np.dtype([("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("xr", "<f4"), ("angle", "<f4"),
                       ("min_level", "<i4"), ("max_level", "<i4"), ("flags", "<i4")])
PQ_ACTIVE, PQ_BLOCKS, PQ_RIGHT_GATE, PQ_REPROJ_GATE = 1, 2, 4, 8


def search_by_projection(keys, desc, u_right, occupied, bounds, queries, qdesc, th_dist=100, nnratio=0.8, use_ratio=True,
                         check_ori=False, inv_level_sigma2=None):
    """ORBmatcher::SearchByProjection over already-projected queries (src/ORBmatcher.cc:155-241, 1440-1593, 1595-1721;
    oracle/frame_grid.cpp).  Returns (nmatches, key_query, key_dist, query_key)."""
    keys = np.ascontiguousarray(keys); desc = np.ascontiguousarray(desc, np.uint8)
    queries = np.ascontiguousarray(queries, PROJ_QUERY); qdesc = np.ascontiguousarray(qdesc, np.uint8)
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    n, nq = len(keys), len(queries)
    kq = np.empty(max(n, 1), np.int32); kd = np.empty(max(n, 1), np.int32); qk = np.empty(max(nq, 1), np.int32)
    inv = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    nm = lib().orc_search_by_projection(_p(keys), _p(desc), _p(ur), _p(oc), n, _p(_bounds(bounds)), _p(queries), _p(qdesc), nq,
                                        int(th_dist), nnratio, int(bool(use_ratio)), int(bool(check_ori)), _p(inv), _p(kq), _p(kd),
                                        _p(qk))
    return nm, kq[:n], kd[:n], qk[:nq]


def search_by_sim3(k1, d1, k2, d2, bounds, q12, qd12, q21, qd21, th_dist=100):
    """ORBmatcher::SearchBySim3 over already-projected queries (src/ORBmatcher.cc:1214-1438; oracle/frame_grid.cpp).
    Returns (nFound, match12, vnMatch1, vnMatch2)."""
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    d1 = np.ascontiguousarray(d1, np.uint8); d2 = np.ascontiguousarray(d2, np.uint8)
    q12 = np.ascontiguousarray(q12, PROJ_QUERY); q21 = np.ascontiguousarray(q21, PROJ_QUERY)
    qd12 = np.ascontiguousarray(qd12, np.uint8); qd21 = np.ascontiguousarray(qd21, np.uint8)
    n1, n2 = len(k1), len(k2)
    assert len(q12) == n1 and len(q21) == n2
    m1 = np.empty(max(n1, 1), np.int32); m2 = np.empty(max(n2, 1), np.int32); m12 = np.empty(max(n1, 1), np.int32)
    nf = lib().orc_search_by_sim3(_p(k1), _p(d1), n1, _p(k2), _p(d2), n2, _p(_bounds(bounds)), _p(q12), _p(qd12), _p(q21), _p(qd21),
                                  int(th_dist), _p(m1), _p(m2), _p(m12))
    return nf, m12[:n1], m1[:n1], m2[:n2]


def flatten_feature_vector(fv):
    """DBoW2::FeatureVector as {node id: [feature indices]} -> (node_ids ascending, node_start CSR, node_idx)"""
    ids = np.array(sorted(fv), np.int32)
    start = np.zeros(len(ids) + 1, np.int32)
    if len(ids):
        start[1:] = np.cumsum([len(fv[i]) for i in ids])
    idx = np.array([j for i in ids for j in fv[i]], np.int32)
    return ids, start, idx


def search_by_bow(k1, d1, valid1, fv1, k2, d2, valid2, fv2, th_dist=50, nnratio=0.7, check_ori=True):
    """ORBmatcher::SearchByBoW (src/ORBmatcher.cc:270-404 with valid2=None, th_dist=TH_LOW; :635-768 with TH_LOW - 1).
    fvX = flatten_feature_vector(...) triples.  Returns (nmatches, matches12)."""
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    d1 = np.ascontiguousarray(d1, np.uint8); d2 = np.ascontiguousarray(d2, np.uint8)
    v1 = None if valid1 is None else np.ascontiguousarray(valid1, np.uint8)
    v2 = None if valid2 is None else np.ascontiguousarray(valid2, np.uint8)
    i1, s1, x1 = (np.ascontiguousarray(a, np.int32) for a in fv1); i2, s2, x2 = (np.ascontiguousarray(a, np.int32) for a in fv2)
    m12 = np.empty(max(len(k1), 1), np.int32)
    nm = lib().orc_search_by_bow(_p(k1), _p(d1), _p(v1), len(k1), _p(i1), _p(s1), _p(x1), len(i1), _p(k2), _p(d2), _p(v2), len(k2),
                                 _p(i2), _p(s2), _p(x2), len(i2), int(th_dist), nnratio, int(bool(check_ori)), _p(m12))
    return nm, m12[:len(k1)]


def search_for_triangulation(k1, d1, valid1, stereo1, fv1, k2, d2, valid2, stereo2, fv2, F12, ex, ey, scale_factors, level_sigma2,
                             th_low=50, check_ori=True):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:770-935).  Returns (nmatches, matches12)."""
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    d1 = np.ascontiguousarray(d1, np.uint8); d2 = np.ascontiguousarray(d2, np.uint8)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    v1, s1, v2, s2 = u8(valid1), u8(stereo1), u8(valid2), u8(stereo2)
    i1, t1, x1 = (np.ascontiguousarray(a, np.int32) for a in fv1); i2, t2, x2 = (np.ascontiguousarray(a, np.int32) for a in fv2)
    F = np.ascontiguousarray(F12, np.float32).reshape(9)
    sf = np.ascontiguousarray(scale_factors, np.float32); sg = np.ascontiguousarray(level_sigma2, np.float32)
    m12 = np.empty(max(len(k1), 1), np.int32)
    nm = lib().orc_search_for_triangulation(_p(k1), _p(d1), _p(v1), _p(s1), len(k1), _p(i1), _p(t1), _p(x1), len(i1), _p(k2), _p(d2),
                                            _p(v2), _p(s2), len(k2), _p(i2), _p(t2), _p(x2), len(i2), _p(F), float(ex), float(ey),
                                            _p(sf), _p(sg), int(th_low), int(bool(check_ori)), _p(m12))
    return nm, m12[:len(k1)]


def monotonic_new(on):
    """Reference builds only (oracle/pyref.py): strictly increasing operator-new addresses inside the entry points,
    which turns the reference's address-ordered quadtree ties (src/ORBextractor.cc:591,684) into a function of the
    input; see oracle/ref_shim/mono_new.cpp."""
    L = lib()
    L.ref_monotonic_new.argtypes = [C.c_int]
    L.ref_monotonic_new.restype = None
    L.ref_monotonic_new(int(bool(on)))
