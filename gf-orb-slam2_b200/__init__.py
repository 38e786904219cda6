"""gforb -- B200-native ORB stereo front-end (host-side Python mirror of the reference interface).

Thin ctypes binding of ``libgforb.so`` (C ABI in ``include/gforb.h``; hand-written sm_100a kernels in
``csrc/``).  Class and method names follow the reference (qiaozhijian/gf-orb-slam2):

* :class:`ORBextractor`  -- ``ORB_SLAM2::ORBextractor`` (include/ORBextractor.h:76-168)
* :func:`DescriptorDistance` -- ``ORBmatcher::DescriptorDistance`` (src/ORBmatcher.cc:1768-1784)
* :func:`ComputeStereoMatches` -- ``Frame::ComputeStereoMatches`` / ``_Undistorted`` (src/Frame.cc:889-1078, 1167-1316)
* :class:`ORBmatcher` -- ``ORB_SLAM2::ORBmatcher(nnratio, checkOri)`` (include/ORBmatcher.h:46): brute-force
  best / second-best search with the ratio and threshold tests of src/ORBmatcher.cc:562-576
* :class:`StereoBatch` -- the batched ``Frame::Frame(imLeft, imRight, ...)`` front-end (src/Frame.cc:248-326)

There is no CPU fallback: a missing library raises at import, a missing CUDA device raises on first use.
The package directory name contains '-', so import it with
``importlib.import_module("gf-orb-slam2_b200")``.
"""
import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_SO = Path(os.environ["GFORB_LIBRARY"]) if os.environ.get("GFORB_LIBRARY") else _HERE / "libgforb.so"   # override: A/B builds of kernel variants (tools/build_variant.sh)
if not _SO.exists():
    raise ImportError(f"{_SO} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(nvcc, sm_100a).  gforb has no CPU fallback.")
_L = C.CDLL(str(_SO))

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
TH_HIGH, TH_LOW = 100, 50
STEREO_CLASSIC, STEREO_UNDISTORTED = 0, 1
OK, ERR_ARG, ERR_CAPACITY, ERR_CUDA, ERR_STATE, ERR_OVERFLOW = 0, -1, -2, -3, -4, -5


class GforbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gforb error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32)]


class Limits(C.Structure):
    _fields_ = [("max_candidates_per_level", C.c_int32)]


class StereoParams(C.Structure):
    _fields_ = [("bf", C.c_float), ("b", C.c_float), ("mode", C.c_int32), ("th_high", C.c_int32),
                ("th_low", C.c_int32), ("min_x", C.c_float)]


_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t
_SIGS = {
    "gforb_last_error_string": (C.c_char_p, []),
    "gforb_version": (_i, []),
    "gforb_extractor_create": (_i, [C.POINTER(Params), C.POINTER(Limits), C.POINTER(_vp)]),
    "gforb_extractor_destroy": (None, [_vp]),
    "gforb_extractor_tables": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gforb_extractor_params": (_i, [_vp, C.POINTER(Params)]),
    "gforb_extract": (_i, [_vp, _vp, _i, _i, _sz, _vp, _vp, _i, C.POINTER(_i)]),
    "gforb_extractor_max_keypoints": (_i, [_vp, _i, _i]),
    "gforb_compute_pyramid": (_i, [_vp, _vp, _i, _i, _sz]),
    "gforb_pyramid_level": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_i), C.POINTER(_i), C.POINTER(_sz)]),
    "gforb_download_pyramid": (_i, [_vp, _i, _vp, _sz, _i]),
    "gforb_debug_candidates": (_i, [_vp, _i, _vp, _i, C.POINTER(_i)]),
    "gforb_debug_blurred_level": (_i, [_vp, _i, _vp, _sz]),
    "gforb_debug_level_counts": (_i, [_vp, _vp]),
    "gforb_hamming256": (_i, [_vp, _vp]),
    "gforb_stereo_match": (_i, [_vp, _vp, C.POINTER(StereoParams), _vp, _i, _vp, _vp, _i, _vp, _vp, _vp,
                                _vp, _vp, C.POINTER(_i)]),
    "gforb_debug_stereo": (_i, [_vp, _i, _vp, _vp, _vp]),
    "gforb_batch_create": (_i, [C.POINTER(Params), C.POINTER(Limits), _i, _i, _i, C.POINTER(_vp)]),
    "gforb_batch_destroy": (None, [_vp]),
    "gforb_batch_set_stream": (_i, [_vp, _vp]),
    "gforb_batch_capacity": (_i, [_vp, C.POINTER(_i), C.POINTER(_i)]),
    "gforb_batch_device_input": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz), C.POINTER(_sz)]),
    "gforb_batch_set_device_input": (_i, [_vp, _vp]),
    "gforb_batch_device_results": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                                        C.POINTER(_i)]),
    "gforb_batch_upload": (_i, [_vp, _vp, _vp, _i]),
    "gforb_batch_run": (_i, [_vp, _i, C.POINTER(StereoParams)]),
    "gforb_batch_download": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "gforb_batch_sync": (_i, [_vp]),
    "gforb_stereo_frames_batch": (_i, [_vp, _vp, _vp, _i, C.POINTER(StereoParams), _vp, _vp, _vp, _vp, _vp]),
    "gforb_stereo_frames_batch_async": (_i, [_vp, _vp, _vp, _i, C.POINTER(StereoParams), _vp, _vp, _vp, _vp, _vp]),
    "gforb_batch_launches_per_run": (_i, [_vp, _i]),
    "gforb_batch_last_launches": (_i, [_vp]),
    "gforb_batch_set_chunk": (_i, [_vp, _i]),
    "gforb_batch_reset_graphs": (_i, [_vp]),
    "gforb_batch_debug_copy_only": (_i, [_vp, _i]),
    "gforb_batch_stage_timing": (_i, [_vp, _i]),
    "gforb_batch_stage_times": (_i, [_vp, _vp]),
    "gforb_matcher_create": (_i, [C.c_float, C.POINTER(_vp)]),
    "gforb_matcher_destroy": (None, [_vp]),
    "gforb_match_bruteforce": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "gforb_match_bruteforce_device": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "gforb_match_bruteforce_static": (_i, [_vp, _i, _vp, _i, C.c_float, _i, _vp, _vp, _vp, _vp, _vp]),
    "gforb_matcher_sync": (_i, [_vp, C.POINTER(C.c_float)]),
    "gforb_assign_features_to_grid": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "gforb_features_in_area": (_i, [_vp, _vp, _vp, _i, _vp, _i, _vp, _vp, _i]),
    "gforb_search_for_initialization": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _vp, C.POINTER(_i)]),
    "gforb_search_for_initialization_batch_device": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i,
                                                          _vp, _vp, _vp]),
    "gforb_search_by_projection": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, C.POINTER(_i)]),
    "gforb_search_by_projection_batch_device": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i,
                                                     _vp, _vp, _vp, _vp, _vp]),
    "gforb_search_by_sim3": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, C.POINTER(_i)]),
    "gforb_search_by_sim3_device": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "gforb_search_by_bow": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp,
                                 C.POINTER(_i)]),
    "gforb_search_by_bow_device": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp,
                                        _vp, _vp]),
    "gforb_search_for_triangulation": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i,
                                            _vp, _i, _i, _vp, C.POINTER(_i)]),
    "gforb_search_for_triangulation_device": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp,
                                                   _i, _vp, _i, _i, _vp, _vp, _vp]),
    "gforb_matcher_set_inv_level_sigma2": (_i, [_vp, _vp, _i]),
    "gforb_rectifier_create": (_i, [_vp, _vp, _i, _i, C.POINTER(_vp)]),
    "gforb_rectifier_destroy": (None, [_vp]),
    "gforb_preprocess": (_i, [_vp, _vp, _i, _i, _sz, _i, _vp, _sz]),
    "gforb_preprocess_batch_device": (_i, [_vp, _vp, _i, _i, _sz, _sz, _i, _i, _vp, _sz, _sz, _vp]),
    "gforb_rectifier_sync": (_i, [_vp, C.POINTER(C.c_float)]),
    "gforb_undistort_keypoints": (_i, [_vp, _vp, _vp, _i, _vp]),
    "gforb_undistort_keypoints_batch_device": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp]),
    "gforb_compute_image_bounds": (_i, [_vp, _vp, _i, _i, _vp]),
    "gforb_distinctive_descriptors": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "gforb_mih_create": (_i, [_i, _i, C.POINTER(_vp)]),
    "gforb_mih_destroy": (None, [_vp]),
    "gforb_mih_substring_len": (_i, [_vp]),
    "gforb_mih_update_dynamics": (_i, [_vp, C.c_double]),
    "gforb_mih_dynamics": (_i, [_vp]),
    "gforb_mih_hash": (_i, [_vp, _vp, _i, _vp]),
    "gforb_mih_insert": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "gforb_mih_insert_device": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "gforb_mih_query": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _vp]),
    "gforb_mih_query_device": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _vp, _vp]),
    "gforb_mih_query_split": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i]),
    "gforb_mih_sync": (_i, [_vp]),
    "gforb_device_alloc": (_i, [C.POINTER(_vp), _sz]),
    "gforb_device_free": (None, [_vp]),
    "gforb_device_copy": (_i, [_vp, _vp, _sz, _i]),
    "gforb_host_alloc": (_i, [C.POINTER(_vp), _sz]),
    "gforb_host_free": (None, [_vp]),
}
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(_L, _name)       # AttributeError here = the library does not export a declared symbol
    _fn.restype, _fn.argtypes = _res, _args

lib = _L
LIBRARY_PATH = str(_SO)


def _check(rc):
    if rc != 0:
        raise GforbError(rc, _L.gforb_last_error_string().decode())


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _stereo_params(bf, b, mode, th_high, th_low, min_x):
    return StereoParams(bf, b, mode, th_high, th_low, min_x)


def DescriptorDistance(a, b):
    """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1768-1784)."""
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    assert a.size == 32 and b.size == 32
    return _L.gforb_hamming256(_p(a), _p(b))


class DeviceArray:
    """cudaMalloc'ed buffer with numpy upload/download (device-resident inputs for ORBmatcher / benches)."""

    def __init__(self, shape, dtype):
        self.shape, self.dtype = tuple(shape), np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._ptr = _vp()
        _check(_L.gforb_device_alloc(C.byref(self._ptr), self.nbytes))

    @property
    def ptr(self):
        return self._ptr.value

    @classmethod
    def from_host(cls, a):
        a = np.ascontiguousarray(a)
        d = cls(a.shape, a.dtype)
        if a.nbytes:
            _check(_L.gforb_device_copy(d._ptr, _p(a), a.nbytes, 0))
        return d

    def to_host(self):
        out = np.empty(self.shape, self.dtype)
        if self.nbytes:
            _check(_L.gforb_device_copy(_p(out), self._ptr, self.nbytes, 1))
        return out

    def __del__(self):
        p = getattr(self, "_ptr", None)
        if p and _L is not None:
            _L.gforb_device_free(p)
            self._ptr = None


MATCH_STATIC_MAX_QUERIES = 65536      # GFORB_MATCH_STATIC_MAX_QUERIES (include/gforb.h)


class ORBmatcher:
    """ORB_SLAM2::ORBmatcher(float nnratio = 0.6, bool checkOri = true) (include/ORBmatcher.h:46).  Carries
    mfNNratio; BruteForce() is the matcher's nearest / second-nearest loop (src/ORBmatcher.cc:562-576) over all
    train descriptors, on the device."""
    TH_HIGH, TH_LOW = TH_HIGH, TH_LOW
    DescriptorDistance = staticmethod(DescriptorDistance)

    def __init__(self, nnratio=0.6, checkOri=True):
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)
        self._h = _vp()
        _check(_L.gforb_matcher_create(self.mfNNratio, C.byref(self._h)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _L is not None:
            _L.gforb_matcher_destroy(h)
            self._h = None

    def BruteForce(self, desc_a, desc_b, th=TH_LOW):
        """host arrays [nA,32], [nB,32] u8 -> dict(best_idx, best_dist, second_dist, match) (int32[nA])"""
        a = np.ascontiguousarray(desc_a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(desc_b, np.uint8).reshape(-1, 32)
        n = len(a)
        out = {k: np.empty(n, np.int32) for k in ("best_idx", "best_dist", "second_dist", "match")}
        _check(_L.gforb_match_bruteforce(self._h, _p(a) if n else None, n, _p(b) if len(b) else None, len(b), th,
                                         _p(out["best_idx"]), _p(out["best_dist"]), _p(out["second_dist"]),
                                         _p(out["match"])))
        return out

    def BruteForceDevice(self, dev_a, n_a, dev_b, n_b, th, dev_best_idx, dev_best_dist, dev_second_dist, dev_match,
                         stream=None):
        """device pointers (ints or DeviceArray); asynchronous on `stream` (a cudaStream_t value) or the matcher's own"""
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        _check(_L.gforb_match_bruteforce_device(self._h, ptr(dev_a), n_a, ptr(dev_b), n_b, th, ptr(dev_best_idx),
                                                ptr(dev_best_dist), ptr(dev_second_dist), ptr(dev_match),
                                                _vp(stream) if stream else None))

    @staticmethod
    def BruteForceStatic(dev_a, n_a, dev_b, n_b, nnratio, th, dev_best_idx, dev_best_dist, dev_second_dist, dev_match,
                         stream=None):
        """gforb_match_bruteforce_static: the same matcher without a handle and without any allocation -- its scratch is a
        fixed 4 MiB device array inside the library (n_a <= MATCH_STATIC_MAX_QUERIES; overlapping calls race).  No
        counterpart in the reference.  Device pointers; asynchronous on `stream` (None = the default stream)."""
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        _check(_L.gforb_match_bruteforce_static(ptr(dev_a), n_a, ptr(dev_b), n_b, nnratio, th, ptr(dev_best_idx),
                                                ptr(dev_best_dist), ptr(dev_second_dist), ptr(dev_match),
                                                _vp(stream) if stream else None))

    def sync(self):
        """wait for the last device-form call; returns its device time in ms"""
        ms = C.c_float(0)
        _check(_L.gforb_matcher_sync(self._h, C.byref(ms)))
        return ms.value

    def SearchForInitialization(self, keys1, desc1, keys2, desc2, vbPrevMatched, windowSize=10, bounds=None,
                                width=None, height=None):
        """ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
        (src/ORBmatcher.cc:520-633).  keysX/descX = FX.mvKeysUn / FX.mDescriptors; vbPrevMatched [n1,2] float32 is
        updated in place like the reference's vector.  Returns (nmatches, vnMatches12)."""
        b = _bounds(bounds, width, height)
        k1 = np.ascontiguousarray(keys1, KP_DTYPE); k2 = np.ascontiguousarray(keys2, KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        assert len(d1) == len(k1) and len(d2) == len(k2)
        assert vbPrevMatched.dtype == np.float32 and vbPrevMatched.shape == (len(k1), 2) and vbPrevMatched.flags.c_contiguous
        m12 = np.full(len(k1), -1, np.int32)
        n = _i(0)
        _check(_L.gforb_search_for_initialization(self._h, _p(b), _p(k1) if len(k1) else None, _p(d1) if len(k1) else None,
                                                  len(k1), _p(k2) if len(k2) else None, _p(d2) if len(k2) else None, len(k2),
                                                  _p(vbPrevMatched) if len(k1) else None, int(windowSize),
                                                  int(self.mbCheckOrientation), _p(m12) if len(k1) else None, C.byref(n)))
        return n.value, m12

    def SearchForInitializationBatchDevice(self, bounds, n_pairs, dev_k1, dev_d1, dev_n1, stride1, dev_k2, dev_d2, dev_n2,
                                           stride2, dev_prev, windowSize, dev_m12, dev_nm, stream=None):
        """device pointers (ints or DeviceArray), one CTA per frame pair; asynchronous"""
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        b = _bounds(bounds, None, None)
        _check(_L.gforb_search_for_initialization_batch_device(
            self._h, _p(b), n_pairs, ptr(dev_k1), ptr(dev_d1), ptr(dev_n1), stride1, ptr(dev_k2), ptr(dev_d2), ptr(dev_n2),
            stride2, ptr(dev_prev), int(windowSize), int(self.mbCheckOrientation), ptr(dev_m12), ptr(dev_nm),
            _vp(stream) if stream else None))

    def SearchByProjection(self, keys, desc, queries, query_desc, u_right=None, occupied=None, th_dist=TH_HIGH,
                           use_ratio=True, check_ori=None, bounds=None, width=None, height=None):
        """ORBmatcher::SearchByProjection, the three per-frame forms (src/ORBmatcher.cc:155-241, 1440-1593, 1595-1721)
        over already-projected queries (PROJ_QUERY_DTYPE records in the reference's iteration order; include/gforb.h
        says how each caller fills them).  keys/desc/u_right = F.mvKeysUn / mDescriptors / mvuRight; occupied[k] != 0
        where mvpMapPoints[k] already blocks.  check_ori None = the matcher's mbCheckOrientation (frame forms); the
        map form (use_ratio=True) has no rotation check in the reference: pass check_ori=False there.
        Returns (nmatches, key_query, key_dist, query_key)."""
        b = _bounds(bounds, width, height)
        k = np.ascontiguousarray(keys, KP_DTYPE); d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        q = np.ascontiguousarray(queries, PROJ_QUERY_DTYPE); qd = np.ascontiguousarray(query_desc, np.uint8).reshape(-1, 32)
        assert len(d) == len(k) and len(qd) == len(q)
        ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        assert (ur is None or len(ur) == len(k)) and (oc is None or len(oc) == len(k))
        n, nq = len(k), len(q)
        kq = np.empty(n, np.int32); kd = np.empty(n, np.int32); qk = np.empty(nq, np.int32)
        nm = _i(0)
        ori = self.mbCheckOrientation if check_ori is None else bool(check_ori)
        _check(_L.gforb_search_by_projection(self._h, _p(b), _p(k) if n else None, _p(d) if n else None,
                                             _p(ur) if (n and ur is not None) else None, _p(oc) if (n and oc is not None) else None,
                                             n, _p(q) if nq else None, _p(qd) if nq else None, nq, int(th_dist), int(bool(use_ratio)),
                                             int(ori), _p(kq) if n else None, _p(kd) if n else None, _p(qk) if nq else None,
                                             C.byref(nm)))
        return nm.value, kq, kd, qk

    def SearchBySim3(self, keys1, desc1, keys2, desc2, queries12, qdesc12, queries21, qdesc21, th_dist=TH_HIGH, bounds=None,
                     width=None, height=None):
        """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1214-1438) over already-projected queries: queries12[i1] = map
        point i1 of KF1 projected into KF2 through the Sim3 (flags = PQ_ACTIVE or 0, levels nPredictedLevel - 1 ..
        nPredictedLevel, r = th * mvScaleFactors[level]), qdesc12 = the map points' descriptors; queries21 / qdesc21 the
        other way round.  Returns (nFound, match12, vnMatch1, vnMatch2): match12[i1] = agreed key of KF2 or -1."""
        b = _bounds(bounds, width, height)
        k1 = np.ascontiguousarray(keys1, KP_DTYPE); d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        k2 = np.ascontiguousarray(keys2, KP_DTYPE); d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        q12 = np.ascontiguousarray(queries12, PROJ_QUERY_DTYPE); qd12 = np.ascontiguousarray(qdesc12, np.uint8).reshape(-1, 32)
        q21 = np.ascontiguousarray(queries21, PROJ_QUERY_DTYPE); qd21 = np.ascontiguousarray(qdesc21, np.uint8).reshape(-1, 32)
        n1, n2 = len(k1), len(k2)
        assert len(d1) == len(q12) == len(qd12) == n1 and len(d2) == len(q21) == len(qd21) == n2
        m1 = np.full(n1, -1, np.int32); m2 = np.full(n2, -1, np.int32); m12 = np.full(n1, -1, np.int32)
        nf = _i(0)
        P = lambda a, n: _p(a) if n else None
        _check(_L.gforb_search_by_sim3(self._h, _p(b), P(k1, n1), P(d1, n1), n1, P(k2, n2), P(d2, n2), n2, P(q12, n1), P(qd12, n1),
                                       P(q21, n2), P(qd21, n2), int(th_dist), P(m1, n1), P(m2, n2), P(m12, n1), C.byref(nf)))
        return nf.value, m12, m1, m2

    def SetInvLevelSigma2(self, inv_level_sigma2):
        """pKF->mvInvLevelSigma2 for the reprojection-error gate of ORBmatcher::Fuse (PQ_REPROJ_GATE queries)"""
        t = np.ascontiguousarray(inv_level_sigma2, np.float32)
        _check(_L.gforb_matcher_set_inv_level_sigma2(self._h, _p(t), len(t)))

    def SearchByProjectionBatchDevice(self, bounds, n_frames, dev_keys, dev_desc, dev_u_right, dev_occupied, dev_n, key_stride,
                                      dev_queries, dev_qdesc, dev_nq, query_stride, th_dist, use_ratio, check_ori,
                                      dev_key_query, dev_key_dist, dev_query_key, dev_nmatches, stream=None):
        """device pointers (ints or DeviceArray); frame f at + f * stride; asynchronous (3 launches)"""
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        b = _bounds(bounds, None, None)
        _check(_L.gforb_search_by_projection_batch_device(
            self._h, _p(b), n_frames, ptr(dev_keys), ptr(dev_desc), ptr(dev_u_right), ptr(dev_occupied), ptr(dev_n), key_stride,
            ptr(dev_queries), ptr(dev_qdesc), ptr(dev_nq), query_stride, int(th_dist), int(bool(use_ratio)), int(bool(check_ori)),
            ptr(dev_key_query), ptr(dev_key_dist), ptr(dev_query_key), ptr(dev_nmatches), _vp(stream) if stream else None))

    def SearchByBoW(self, keys1, desc1, valid1, featvec1, keys2, desc2, valid2, featvec2, th_dist=TH_LOW):
        """ORBmatcher::SearchByBoW (src/ORBmatcher.cc:270-404: valid2=None, th_dist=TH_LOW; :635-768: th_dist=TH_LOW - 1).
        featvecX = DBoW2::FeatureVector as {node id: [feature indices]} or an already flattened
        (node_ids, node_start, node_idx) triple.  Returns (nmatches, matches12)."""
        k1 = np.ascontiguousarray(keys1, KP_DTYPE); k2 = np.ascontiguousarray(keys2, KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        assert len(d1) == len(k1) and len(d2) == len(k2)
        v1 = None if valid1 is None else np.ascontiguousarray(valid1, np.uint8)
        v2 = None if valid2 is None else np.ascontiguousarray(valid2, np.uint8)
        i1, s1, x1 = flatten_feature_vector(featvec1); i2, s2, x2 = flatten_feature_vector(featvec2)
        n1, n2 = len(k1), len(k2)
        m12 = np.full(n1, -1, np.int32)
        nm = _i(0)
        opt = lambda a: _p(a) if (a is not None and len(a)) else None
        _check(_L.gforb_search_by_bow(self._h, opt(k1), opt(d1), opt(v1), n1, opt(i1), opt(s1) if len(i1) else None, opt(x1), len(i1),
                                      opt(k2), opt(d2), opt(v2), n2, opt(i2), opt(s2) if len(i2) else None, opt(x2), len(i2),
                                      int(th_dist), int(self.mbCheckOrientation), opt(m12), C.byref(nm)))
        return nm.value, m12

    def SearchForTriangulation(self, keys1, desc1, valid1, stereo1, featvec1, keys2, desc2, valid2, stereo2, featvec2, F12, ex, ey,
                               scale_factors, level_sigma2, th_low=TH_LOW):
        """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:770-935).  validX: feature has no MapPoint (and is
        stereo if bOnlyStereo); stereoX: mvuRight >= 0 (None: monocular).  Returns (nmatches, vMatches12)."""
        k1 = np.ascontiguousarray(keys1, KP_DTYPE); k2 = np.ascontiguousarray(keys2, KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
        v1, s1, v2, s2 = u8(valid1), u8(stereo1), u8(valid2), u8(stereo2)
        i1, t1, x1 = flatten_feature_vector(featvec1); i2, t2, x2 = flatten_feature_vector(featvec2)
        ep = Epipolar()
        ep.F12[:] = [float(v) for v in np.asarray(F12, np.float32).reshape(9)]
        ep.ex, ep.ey = float(ex), float(ey)
        sf = np.asarray(scale_factors, np.float32); sg = np.asarray(level_sigma2, np.float32)
        ep.n_levels = len(sf)
        ep.scale_factors[:len(sf)] = [float(v) for v in sf]; ep.level_sigma2[:len(sg)] = [float(v) for v in sg]
        n1, n2 = len(k1), len(k2)
        m12 = np.full(n1, -1, np.int32)
        nm = _i(0)
        opt = lambda a: _p(a) if (a is not None and len(a)) else None
        _check(_L.gforb_search_for_triangulation(self._h, opt(k1), opt(d1), opt(v1), opt(s1), n1, opt(i1), opt(t1) if len(i1) else None,
                                                 opt(x1), len(i1), opt(k2), opt(d2), opt(v2), opt(s2), n2, opt(i2),
                                                 opt(t2) if len(i2) else None, opt(x2), len(i2), C.byref(ep), int(th_low),
                                                 int(self.mbCheckOrientation), opt(m12), C.byref(nm)))
        return nm.value, m12


class Epipolar(C.Structure):
    """gforb_epipolar"""
    _fields_ = [("F12", C.c_float * 9), ("ex", C.c_float), ("ey", C.c_float), ("n_levels", C.c_int32),
                ("scale_factors", C.c_float * 32), ("level_sigma2", C.c_float * 32)]


def flatten_feature_vector(fv):
    """DBoW2::FeatureVector ({node id: [feature indices]}) -> (node_ids ascending, node_start CSR, node_idx) int32 arrays"""
    if isinstance(fv, tuple):
        return tuple(np.ascontiguousarray(a, np.int32) for a in fv)
    ids = np.array(sorted(fv), np.int32)
    start = np.zeros(len(ids) + 1, np.int32)
    if len(ids):
        start[1:] = np.cumsum([len(fv[i]) for i in ids])
    idx = np.array([j for i in ids for j in fv[i]], np.int32)
    return ids, start, idx


PROJ_QUERY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("xr", "<f4"), ("angle", "<f4"),
                             ("min_level", "<i4"), ("max_level", "<i4"), ("flags", "<i4")])
PQ_ACTIVE, PQ_BLOCKS, PQ_RIGHT_GATE, PQ_REPROJ_GATE = 1, 2, 4, 8


FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48      # include/Frame.h:92-93


def _bounds(bounds, width, height):
    """{mnMinX, mnMaxX, mnMinY, mnMaxY}; without distortion the image rectangle (src/Frame.cc:760-887)."""
    if bounds is None:
        bounds = (0.0, float(width), 0.0, float(height))
    return np.ascontiguousarray(bounds, np.float32)


class FrameGrid:
    """The feature grid of ORB_SLAM2::Frame: AssignFeaturesToGrid (src/Frame.cc:461-476) at construction,
    GetFeaturesInArea (src/Frame.cc:593-643) as a batch of queries.  mGrid[ix][iy] == cell(ix, iy)."""

    def __init__(self, keys, bounds=None, width=None, height=None, matcher=None):
        self._m = matcher or ORBmatcher()
        self._b = _bounds(bounds, width, height)
        self._keys = np.ascontiguousarray(keys, KP_DTYPE)
        n = len(self._keys)
        self.cell_start = np.zeros(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1, np.int32)
        idx = np.zeros(max(n, 1), np.int32)
        _check(_L.gforb_assign_features_to_grid(self._m._h, _p(self._b), _p(self._keys) if n else None, n,
                                                _p(self.cell_start), _p(idx)))
        self.cell_idx = idx[: self.cell_start[-1]]

    def cell(self, ix, iy):
        c = ix * FRAME_GRID_ROWS + iy
        return self.cell_idx[self.cell_start[c]: self.cell_start[c + 1]]

    def GetFeaturesInArea(self, queries, cap=None):
        """queries: iterable of (x, y, r[, minLevel, maxLevel]) -> list of int32 index arrays (reference order)"""
        q = np.zeros(len(queries), AREA_QUERY_DTYPE)
        for i, t in enumerate(queries):
            q[i] = tuple(t) + (-1, -1) if len(t) == 3 else tuple(t)
        n, nq = len(self._keys), len(q)
        cap = max(int(cap if cap is not None else nq * max(n, 1)), 1)
        start = np.zeros(nq + 1, np.int32); out = np.zeros(cap, np.int32)
        _check(_L.gforb_features_in_area(self._m._h, _p(self._b), _p(self._keys) if n else None, n, _p(q) if nq else None, nq,
                                         _p(start), _p(out), cap))
        return [out[start[i]: start[i + 1]] for i in range(nq)]


AREA_QUERY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4")])


class ORBextractor:
    """ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 max_candidates_per_level=0):
        self._h = _vp()
        prm = Params(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        lim = Limits(max_candidates_per_level)
        _check(_L.gforb_extractor_create(C.byref(prm), C.byref(lim), C.byref(self._h)))
        self.nfeatures, self.scaleFactor, self.nlevels = nfeatures, scaleFactor, nlevels
        self.iniThFAST, self.minThFAST = iniThFAST, minThFAST
        n = nlevels
        self._scale = np.empty(n, np.float32); self._inv = np.empty(n, np.float32)
        self._sig = np.empty(n, np.float32); self._isig = np.empty(n, np.float32)
        self.mnFeaturesPerLevel = np.empty(n, np.int32); self.umax = np.empty(16, np.int32)
        _check(_L.gforb_extractor_tables(self._h, _p(self._scale), _p(self._inv), _p(self._sig), _p(self._isig),
                                         _p(self.mnFeaturesPerLevel), _p(self.umax)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _L is not None:
            _L.gforb_extractor_destroy(h)
            self._h = None

    # getters, include/ORBextractor.h:93-119
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return self.scaleFactor
    def GetInitThres(self): return self.iniThFAST
    def GetMinThres(self): return self.minThFAST
    def GetScaleFactors(self): return self._scale.copy()
    def GetInverseScaleFactors(self): return self._inv.copy()
    def GetScaleSigmaSquares(self): return self._sig.copy()
    def GetInverseScaleSigmaSquares(self): return self._isig.copy()

    def __call__(self, image, mask=None):
        """operator()(image, mask, keypoints, descriptors) -> (keypoints[KP_DTYPE], descriptors[n,32] u8).
        The mask is ignored, as in the reference."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise GforbError(ERR_ARG, "image must be 2-D uint8 (CV_8UC1)")
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = _L.gforb_extractor_max_keypoints(self._h, w, h)
        if cap < 0:
            _check(cap)
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        n = _i(0)
        _check(_L.gforb_extract(self._h, _p(image), w, h, image.strides[0], _p(kps), _p(desc), cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def ComputePyramid(self, image):
        if image.dtype != np.uint8 or image.ndim != 2:
            raise GforbError(ERR_ARG, "image must be 2-D uint8 (CV_8UC1)")
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        _check(_L.gforb_compute_pyramid(self._h, _p(image), w, h, image.strides[0]))

    def level_geometry(self, level):
        ptr, w, h, pitch = _vp(), _i(), _i(), _sz()
        _check(_L.gforb_pyramid_level(self._h, level, C.byref(ptr), C.byref(w), C.byref(h), C.byref(pitch)))
        return ptr.value, w.value, h.value, pitch.value

    def pyramid_level(self, level, with_border=False):
        """Host copy of mvImagePyramid[level] (with its 19-px REFLECT_101 frame if asked)."""
        _, w, h, _ = self.level_geometry(level)
        if with_border:
            w, h = w + 38, h + 38
        out = np.empty((h, w), np.uint8)
        _check(_L.gforb_download_pyramid(self._h, level, _p(out), out.strides[0], int(with_border)))
        return out

    def debug_blurred_level(self, level):
        """Host copy of the blurred level the last call's descriptors were sampled from (src/ORBextractor.cc:1153-1155)."""
        _, w, h, _ = self.level_geometry(level)
        out = np.empty((h, w), np.uint8)
        _check(_L.gforb_debug_blurred_level(self._h, level, _p(out), out.strides[0]))
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    def debug_candidates(self, level, cap=400000):
        buf = np.empty((cap, 3), np.int32); n = _i(0)
        _check(_L.gforb_debug_candidates(self._h, level, _p(buf), cap, C.byref(n)))
        return buf[:n.value].copy()

    def debug_level_counts(self):
        c = np.empty(self.nlevels, np.int32)
        _check(_L.gforb_debug_level_counts(self._h, _p(c)))
        return c


def ComputeStereoMatches(exL, exR, kpsL, descL, kpsR, descR, bf, b, mode=STEREO_CLASSIC, th_high=TH_HIGH,
                         th_low=TH_LOW, min_x=0.0, min_d=None, max_d=None, debug=False):
    """Frame::ComputeStereoMatches (mode CLASSIC) / ComputeStereoMatches_Undistorted (mode UNDISTORTED).

    Returns dict(mvuRight, mvDepth, nmatched[, best_dist, best_idx, sad])."""
    kpsL = np.ascontiguousarray(kpsL, KP_DTYPE); kpsR = np.ascontiguousarray(kpsR, KP_DTYPE)
    descL = np.ascontiguousarray(descL, np.uint8); descR = np.ascontiguousarray(descR, np.uint8)
    nL, nR = len(kpsL), len(kpsR)
    u = np.full(max(nL, 1), -1, np.float32); z = np.full(max(nL, 1), -1, np.float32)
    if min_d is not None:
        min_d = np.ascontiguousarray(min_d, np.float32); max_d = np.ascontiguousarray(max_d, np.float32)
    prm = _stereo_params(bf, b, mode, th_high, th_low, min_x)
    nm = _i(0)
    _check(_L.gforb_stereo_match(exL._h, exR._h, C.byref(prm), _p(kpsL), nL, _p(descL), _p(kpsR), nR, _p(descR),
                                 _p(min_d), _p(max_d), _p(u), _p(z), C.byref(nm)))
    out = dict(mvuRight=u[:nL], mvDepth=z[:nL], nmatched=nm.value)
    if debug and nL:
        bd = np.empty(nL, np.int32); bi = np.empty(nL, np.int32); sad = np.empty(nL, np.int32)
        _check(_L.gforb_debug_stereo(exL._h, nL, _p(bd), _p(bi), _p(sad)))
        out.update(best_dist=bd, best_idx=bi, sad=sad)
    return out


class PinnedArray:
    """numpy view over cudaHostAlloc'ed memory (for async H2D/D2H in StereoBatch)."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.nbytes = int(np.prod(shape)) * self.dtype.itemsize
        self._ptr = _vp()
        _check(_L.gforb_host_alloc(C.byref(self._ptr), max(self.nbytes, 16)))
        buf = (C.c_uint8 * max(self.nbytes, 16)).from_address(self._ptr.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(shape))).reshape(shape)

    def __del__(self):
        p = getattr(self, "_ptr", None)
        if p and _L is not None:        # _L is None while the interpreter shuts down: the OS reclaims the pages
            self.array = None
            _L.gforb_host_free(p)
            self._ptr = None


class StereoBatch:
    """Batched stereo front-end: for each frame, ExtractORB(left), ExtractORB(right), ComputeStereoMatches
    (src/Frame.cc:248-326), all on the device."""

    def __init__(self, width, height, max_frames, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20,
                 minThFAST=7, max_candidates_per_level=0):
        self._h = _vp()
        prm = Params(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        lim = Limits(max_candidates_per_level)
        _check(_L.gforb_batch_create(C.byref(prm), C.byref(lim), width, height, max_frames, C.byref(self._h)))
        self.width, self.height, self.max_frames = width, height, max_frames
        mf, K = _i(), _i()
        _check(_L.gforb_batch_capacity(self._h, C.byref(mf), C.byref(K)))
        self.K = K.value

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _L is not None:
            _L.gforb_batch_destroy(h)
            self._h = None

    def set_stream(self, cuda_stream_ptr):
        _check(_L.gforb_batch_set_stream(self._h, _vp(cuda_stream_ptr) if cuda_stream_ptr else None))

    def device_input(self):
        ptr, stride, pitch = _vp(), _sz(), _sz()
        _check(_L.gforb_batch_device_input(self._h, C.byref(ptr), C.byref(stride), C.byref(pitch)))
        return ptr.value, stride.value, pitch.value

    def set_device_input(self, dev_ptr):
        """run from caller-owned device images (int address, the batch's own layout); None / 0 = the batch's buffer"""
        _check(_L.gforb_batch_set_device_input(self._h, dev_ptr or None))

    def device_results(self):
        """device pointers of the results of the last run, in gforb_batch_download's layout:
        dict(counts, keypoints, descriptors, u_right, depth -> int addresses, K -> key stride)"""
        c, k, d, u, z, K = _vp(), _vp(), _vp(), _vp(), _vp(), _i(0)
        _check(_L.gforb_batch_device_results(self._h, C.byref(c), C.byref(k), C.byref(d), C.byref(u), C.byref(z), C.byref(K)))
        return dict(counts=c.value, keypoints=k.value, descriptors=d.value, u_right=u.value, depth=z.value, K=K.value)

    def _check_images(self, left, right):
        """The C entries read n*h*w tight bytes from each pointer: anything else (a slice, a transposed view, another
        dtype or size, more frames than the batch holds) must be an error here, not a silent out-of-bounds read."""
        for name, a in (("left", left), ("right", right)):
            if not isinstance(a, np.ndarray) or a.dtype != np.uint8 or a.ndim != 3 or not a.flags.c_contiguous:
                raise ValueError(f"{name}: expected a C-contiguous uint8 array [n, {self.height}, {self.width}]")
        n = left.shape[0]
        if left.shape != (n, self.height, self.width) or right.shape != left.shape:
            raise ValueError(f"left {left.shape} / right {right.shape}: expected [n, {self.height}, {self.width}] for both")
        if not 1 <= n <= self.max_frames:
            raise ValueError(f"{n} frames: the batch holds 1..{self.max_frames}")
        return n

    def _check_outputs(self, out, n):
        """out[...] are written with the batch's key stride K for 2n images / n frames: smaller buffers would overrun."""
        K = self.K
        need = dict(counts=((2 * n,), np.int32), keypoints=((2 * n, K), KP_DTYPE), descriptors=((2 * n, K, 32), np.uint8),
                    u_right=((n, K), np.float32), depth=((n, K), np.float32))
        for k, a in out.items():
            if k not in need:
                raise ValueError(f"out[{k!r}]: unknown output")
            if a is None:
                continue
            shape, dt = need[k]
            if (not isinstance(a, np.ndarray) or a.dtype != np.dtype(dt) or not a.flags.c_contiguous or a.ndim != len(shape)
                    or a.shape[0] < shape[0] or a.shape[1:] != shape[1:]):
                raise ValueError(f"out[{k!r}]: expected a C-contiguous {np.dtype(dt)} array of at least {shape}, got "
                                 f"{getattr(a, 'shape', None)} {getattr(a, 'dtype', None)}")

    def upload(self, left, right):
        n = self._check_images(left, right)
        _check(_L.gforb_batch_upload(self._h, _p(left), _p(right), n))

    def run(self, n_frames, bf=None, b=None, mode=STEREO_CLASSIC, th_high=TH_HIGH, th_low=TH_LOW, min_x=0.0):
        prm = None if bf is None else C.byref(_stereo_params(bf, b, mode, th_high, th_low, min_x))
        _check(_L.gforb_batch_run(self._h, n_frames, prm))

    def launches_per_run(self, with_stereo=True):
        return _L.gforb_batch_launches_per_run(self._h, int(with_stereo))

    def last_launches(self):
        """kernels enqueued by the last run()/process() call"""
        return _L.gforb_batch_last_launches(self._h)

    def set_chunk(self, frames):
        """process() pipelines H2D / kernels / D2H in chunks of `frames` stereo frames (0 = automatic)"""
        _check(_L.gforb_batch_set_chunk(self._h, int(frames)))

    def debug_copy_only(self, on):
        """measurement switch: process()/submit() move their bytes as usual but launch no kernel (copy-only ceiling)"""
        _check(_L.gforb_batch_debug_copy_only(self._h, int(bool(on))))

    def reset_graphs(self):
        """drop the CUDA graphs cached for single-chunk process()/submit() calls (keyed on host pointer values)"""
        _check(_L.gforb_batch_reset_graphs(self._h))

    STAGES = ("pyramid", "fast", "quadtree", "blur", "describe", "stereo")

    def stage_timing(self, enable=True):
        _check(_L.gforb_batch_stage_timing(self._h, int(enable)))

    def stage_times(self):
        ms = np.zeros(6, np.float32)
        _check(_L.gforb_batch_stage_times(self._h, _p(ms)))
        return dict(zip(self.STAGES, ms.tolist()))

    def alloc_outputs(self, n_frames, pinned=True):
        K = self.K
        shapes = dict(counts=((2 * n_frames,), np.int32), keypoints=((2 * n_frames, K), KP_DTYPE),
                      descriptors=((2 * n_frames, K, 32), np.uint8), u_right=((n_frames, K), np.float32),
                      depth=((n_frames, K), np.float32))
        if pinned:
            pins = {k: PinnedArray(s, d) for k, (s, d) in shapes.items()}
            self.__dict__.setdefault("_pins", []).append(pins)      # pinned buffers live as long as the batch
            return {k: p.array for k, p in pins.items()}
        return {k: np.empty(s, d) for k, (s, d) in shapes.items()}

    def download(self, n_frames, out):
        self._check_outputs(out, n_frames)
        _check(_L.gforb_batch_download(self._h, n_frames, _p(out.get("counts")), _p(out.get("keypoints")),
                                       _p(out.get("descriptors")), _p(out.get("u_right")), _p(out.get("depth"))))

    def sync(self):
        _check(_L.gforb_batch_sync(self._h))

    def process(self, left, right, out, bf, b, mode=STEREO_CLASSIC):
        """End-to-end: host images in, host results out (upload + run + download + sync)."""
        n = self._check_images(left, right)
        self._check_outputs(out, n)
        prm = _stereo_params(bf, b, mode, TH_HIGH, TH_LOW, 0.0)
        _check(_L.gforb_stereo_frames_batch(self._h, _p(left), _p(right), n, C.byref(prm), _p(out.get("counts")),
                                            _p(out.get("keypoints")), _p(out.get("descriptors")),
                                            _p(out.get("u_right")), _p(out.get("depth"))))

    def submit(self, left, right, out, bf, b, mode=STEREO_CLASSIC):
        """process() without the final wait (gforb_stereo_frames_batch_async): `left`, `right` and `out` must stay
        alive and untouched until sync() returns.  Alternate two StereoBatch objects to overlap consecutive batches."""
        n = self._check_images(left, right)
        self._check_outputs(out, n)
        prm = _stereo_params(bf, b, mode, TH_HIGH, TH_LOW, 0.0)
        self._inflight = (left, right, out)
        _check(_L.gforb_stereo_frames_batch_async(self._h, _p(left), _p(right), n, C.byref(prm), _p(out.get("counts")),
                                                  _p(out.get("keypoints")), _p(out.get("descriptors")),
                                                  _p(out.get("u_right")), _p(out.get("depth"))))


PIX_GRAY8, PIX_GRAY16, PIX_RGB8, PIX_BGR8, PIX_RGBA8, PIX_BGRA8 = range(6)
_PIX_BPP = (1, 2, 3, 3, 4, 4)


class Rectifier:
    """The pre-processing the reference runs in front of the extractor, fused: [16-bit -> high byte] -> cv::remap(M1, M2,
    INTER_LINEAR) (ros_stereo.cc:290-295, test_Stereo.cpp:197-210) -> cvtColor(*2GRAY) (src/Tracking.cc:366-382).
    M1/M2: float32 [h, w] maps of one camera (None: no remap)."""

    def __init__(self, M1=None, M2=None):
        self._h = _vp()
        self.shape = None
        if M1 is not None:
            M1 = np.ascontiguousarray(M1, np.float32); M2 = np.ascontiguousarray(M2, np.float32)
            assert M1.shape == M2.shape and M1.ndim == 2
            self.shape = M1.shape
            _check(_L.gforb_rectifier_create(_p(M1), _p(M2), M1.shape[1], M1.shape[0], C.byref(self._h)))
        else:
            _check(_L.gforb_rectifier_create(None, None, 0, 0, C.byref(self._h)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _L is not None:
            _L.gforb_rectifier_destroy(h)
            self._h = None

    def __call__(self, image, pixel_format=PIX_GRAY8):
        """host image [h, w] (gray8 / gray16) or [h, w, 3|4] u8 -> rectified 8-bit gray image"""
        assert image.strides[-1] == image.itemsize and (image.ndim == 2 or image.strides[1] == image.shape[2])
        sh, sw = image.shape[:2]
        assert image.itemsize * (image.shape[2] if image.ndim == 3 else 1) == _PIX_BPP[pixel_format]
        dh, dw = self.shape or (sh, sw)
        out = np.empty((dh, dw), np.uint8)
        _check(_L.gforb_preprocess(self._h, _p(image), sw, sh, image.strides[0], pixel_format, _p(out), dw))
        return out

    def batch_device(self, dev_src, src_w, src_h, src_pitch, src_stride, pixel_format, n, dev_dst, dst_pitch, dst_stride,
                     stream=None):
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        _check(_L.gforb_preprocess_batch_device(self._h, ptr(dev_src), src_w, src_h, src_pitch, src_stride, pixel_format, n,
                                                ptr(dev_dst), dst_pitch, dst_stride, _vp(stream) if stream else None))

    def sync(self):
        ms = C.c_float(0)
        _check(_L.gforb_rectifier_sync(self._h, C.byref(ms)))
        return ms.value


class Camera(C.Structure):
    """gforb_camera: mK, mDistCoef (k1 k2 p1 p2 k3 k4 k5 k6), mR, left 3x3 of mP -- doubles, row-major."""
    _fields_ = [("K", C.c_double * 9), ("dist", C.c_double * 8), ("R", C.c_double * 9), ("P", C.c_double * 9)]

    @classmethod
    def make(cls, K, dist, R=None, P=None):
        K = np.asarray(K, np.float64).reshape(3, 3)
        d = np.zeros(8); dd = np.asarray(dist, np.float64).ravel(); d[:len(dd)] = dd
        R = np.eye(3) if R is None else np.asarray(R, np.float64).reshape(3, 3)
        P = K if P is None else np.asarray(P, np.float64)[:3, :3]          # Frame::UndistortKeyPoints passes P = mK
        c = cls()
        c.K[:] = K.ravel().tolist(); c.dist[:] = d.tolist(); c.R[:] = R.ravel().tolist(); c.P[:] = np.ascontiguousarray(P).ravel().tolist()
        return c


def UndistortKeyPoints(matcher, camera, keys):
    """Frame::UndistortKeyPoints / UndistortKeyPointsStereo (src/Frame.cc:670-758): mvKeys -> mvKeysUn."""
    keys = np.ascontiguousarray(keys, KP_DTYPE)
    out = np.empty_like(keys)
    _check(_L.gforb_undistort_keypoints(matcher._h, C.byref(camera), _p(keys) if len(keys) else None, len(keys),
                                        _p(out) if len(keys) else None))
    return out


def ComputeImageBounds(matcher, camera, cols, rows):
    """Frame::ComputeImageBounds[Stereo] (src/Frame.cc:760-887) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    b = np.zeros(4, np.float32)
    _check(_L.gforb_compute_image_bounds(matcher._h, C.byref(camera), cols, rows, _p(b)))
    return tuple(float(v) for v in b)


def ComputeDistinctiveDescriptors(matcher, desc, set_start):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:397-458) for many map points at once: set p = rows
    set_start[p]:set_start[p+1] of desc.  Returns (best_idx within each set, best_median)."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); ss = np.ascontiguousarray(set_start, np.int32)
    n = len(ss) - 1
    bi = np.empty(max(n, 1), np.int32); bm = np.empty(max(n, 1), np.int32)
    _check(_L.gforb_distinctive_descriptors(matcher._h, _p(desc) if len(desc) else None, _p(ss), n, _p(bi), _p(bm)))
    return bi[:n], bm[:n]


class MultiIndexHashing:
    """HASHING::MultiIndexHashing(feature_length = 256, table_num) (include/Hashing.h:171-216), the descriptor side, on the
    device: substring keys, batched insertion, the two candidate queries.  Entries are int32 handles (MapPoint::mnId)."""

    def __init__(self, table_num=32):
        self._h = _vp()
        _check(_L.gforb_mih_create(256, table_num, C.byref(self._h)))
        self.table_num = table_num

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _L is not None:
            _L.gforb_mih_destroy(h)
            self._h = None

    def getSubstringLen(self):
        return _L.gforb_mih_substring_len(self._h)

    def updateDynamics(self, e):
        return _L.gforb_mih_update_dynamics(self._h, float(e))

    def hash(self, desc):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        keys = np.empty((len(d), self.table_num), np.uint32)
        _check(_L.gforb_mih_hash(self._h, _p(d) if len(d) else None, len(d), _p(keys) if len(d) else None))
        return keys

    def insert(self, desc, entry, hashed=None):
        """returns (insert_success[n], hashed[n] afterwards)"""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); e = np.ascontiguousarray(entry, np.int32)
        assert len(d) == len(e)
        hm = np.zeros(len(d), np.uint32) if hashed is None else np.ascontiguousarray(hashed, np.uint32).copy()
        ok = np.zeros(len(d), np.uint8)
        if len(d):
            _check(_L.gforb_mih_insert(self._h, _p(d), _p(e), len(d), _p(hm), _p(ok)))
        return ok, hm

    def insert_device(self, dev_desc, dev_entry, n, dev_hashed, dev_success, stream=None):
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        _check(_L.gforb_mih_insert_device(self._h, ptr(dev_desc), ptr(dev_entry), n, ptr(dev_hashed), ptr(dev_success),
                                          _vp(stream) if stream else None))

    def query(self, qdesc, nactive, nretrieve):
        """(cand[m, stride] padded with -1, count[m]): MultiIndexHashing::query(pDesc, candidates, nactive, nretrieve)"""
        d = np.ascontiguousarray(qdesc, np.uint8).reshape(-1, 32)
        stride = max(1, (nactive or self.table_num) * max(nretrieve, 0))
        cand = np.full((len(d), stride), -1, np.int32); cnt = np.zeros(len(d), np.int32)
        if len(d):
            _check(_L.gforb_mih_query(self._h, _p(d), len(d), nactive, nretrieve, _p(cand), stride, _p(cnt)))
        return cand, cnt

    def query_device(self, dev_qdesc, m, nactive, nretrieve, dev_cand, stride, dev_count, stream=None):
        ptr = lambda x: None if x is None else _vp(x.ptr if isinstance(x, DeviceArray) else x)
        _check(_L.gforb_mih_query_device(self._h, ptr(dev_qdesc), m, nactive, nretrieve, ptr(dev_cand), stride, ptr(dev_count),
                                         _vp(stream) if stream else None))

    def query_split(self, qdesc, nfeatures):
        """(cand, count, addi, addi_count): MultiIndexHashing::query(pDesc, candidates, addi_candidates, nfeatures)"""
        d = np.ascontiguousarray(qdesc, np.uint8).reshape(-1, 32)
        nf = np.ascontiguousarray(nfeatures, np.int32)
        assert len(nf) == self.table_num
        stride = max(1, int(np.maximum(nf, 0).sum()))
        cand = np.full((len(d), stride), -1, np.int32); cnt = np.zeros(len(d), np.int32)
        addi = np.full((len(d), stride), -1, np.int32); acnt = np.zeros(len(d), np.int32)
        if len(d):
            _check(_L.gforb_mih_query_split(self._h, _p(d), len(d), _p(nf), _p(cand), _p(cnt), _p(addi), _p(acnt), stride))
        return cand, cnt, addi, acnt

    def sync(self):
        _check(_L.gforb_mih_sync(self._h))
