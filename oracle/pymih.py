"""ORACLE -- TEST INFRASTRUCTURE ONLY.

ctypes binding of the multi-index hashing entry points (oracle/mih.cpp), over either library that exports them:
  "oracle"  oracle/_build/libgforb_oracle.so      the restatement
  "ref"     oracle/_ref/libgforb_ref_mih.so       the reference's own src/Hashing.cc, compiled unmodified (oracle/ref_mih.cpp)
"""
import ctypes as C
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIBS = {}


def available(variant):
    return (_HERE / "_ref" / "libgforb_ref_mih.so").exists() if variant == "ref" else True


def _lib(variant):
    if variant not in _LIBS:
        if variant == "oracle":
            from oracle import pyoracle
            pyoracle.lib()                                   # builds it when missing
            so = _HERE / "_build" / "libgforb_oracle.so"
        else:
            so = _HERE / "_ref" / "libgforb_ref_mih.so"
        L = C.CDLL(str(so))
        vp, i32 = C.c_void_p, C.c_int
        L.orc_mih_create.restype = vp; L.orc_mih_create.argtypes = [i32, i32]
        L.orc_mih_destroy.restype = None; L.orc_mih_destroy.argtypes = [vp]
        L.orc_mih_substring_len.restype = i32; L.orc_mih_substring_len.argtypes = [vp]
        if variant == "oracle":
            L.orc_mih_hash.restype = None; L.orc_mih_hash.argtypes = [vp, vp, i32, vp]
        L.orc_mih_insert.restype = None; L.orc_mih_insert.argtypes = [vp, vp, vp, i32, vp, vp]
        L.orc_mih_query.restype = i32; L.orc_mih_query.argtypes = [vp, vp, i32, i32, i32, vp, i32, vp]
        L.orc_mih_query2.restype = i32; L.orc_mih_query2.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, i32]
        L.orc_mih_update_dynamics.restype = i32; L.orc_mih_update_dynamics.argtypes = [vp, C.c_double]
        _LIBS[variant] = L
    return _LIBS[variant]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class MultiIndexHashing:
    """HASHING::MultiIndexHashing(feature_length = 256, table_num) over integer entry ids (MapPoint::mnId)."""

    def __init__(self, table_num=32, variant="oracle"):
        self.L, self.table_num, self.variant = _lib(variant), table_num, variant
        self.h = self.L.orc_mih_create(256, table_num)
        if not self.h:
            raise ValueError(f"unsupported table_num {table_num}")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_mih_destroy(self.h); self.h = None

    def getSubstringLen(self):
        return self.L.orc_mih_substring_len(self.h)

    def hash(self, desc):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        keys = np.empty((len(d), self.table_num), np.uint32)
        self.L.orc_mih_hash(self.h, _p(d), len(d), _p(keys))
        return keys

    def insert(self, desc, entry, hashed=None):
        """returns (insert_success[n], hashed[n] after) -- hashed: mvbHashed bit masks carried by the caller (None: all clear)"""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); e = np.ascontiguousarray(entry, np.int32)
        hm = np.zeros(len(d), np.uint32) if hashed is None else np.ascontiguousarray(hashed, np.uint32).copy()
        ok = np.zeros(len(d), np.uint8)
        self.L.orc_mih_insert(self.h, _p(d), _p(e), len(d), _p(hm), _p(ok))
        return ok, hm

    def query(self, qdesc, nactive, nretrieve):
        d = np.ascontiguousarray(qdesc, np.uint8).reshape(-1, 32)
        stride = max(1, (nactive or self.table_num) * max(nretrieve, 0))
        cand = np.full((len(d), stride), -1, np.int32); cnt = np.zeros(len(d), np.int32)
        worst = self.L.orc_mih_query(self.h, _p(d), len(d), nactive, nretrieve, _p(cand), stride, _p(cnt))
        assert worst <= stride
        return cand, cnt

    def query2(self, qdesc, nfeatures):
        d = np.ascontiguousarray(qdesc, np.uint8).reshape(-1, 32)
        nf = np.ascontiguousarray(nfeatures, np.int32)
        assert len(nf) == self.table_num
        stride = max(1, int(np.maximum(nf, 0).sum()))
        cand = np.full((len(d), stride), -1, np.int32); cnt = np.zeros(len(d), np.int32)
        addi = np.full((len(d), stride), -1, np.int32); acnt = np.zeros(len(d), np.int32)
        worst = self.L.orc_mih_query2(self.h, _p(d), len(d), _p(nf), _p(cand), _p(cnt), _p(addi), _p(acnt), stride)
        assert worst <= stride
        return cand, cnt, addi, acnt

    def updateDynamics(self, e):
        return self.L.orc_mih_update_dynamics(self.h, float(e))
